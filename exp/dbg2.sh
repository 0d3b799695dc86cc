#!/bin/sh
# usage: exp/dbg2.sh variant "bench args" [ENV=val ...]
cd "$(dirname "$0")/.."
v=$1; a=$2; shift; shift
echo "== $v $a $*" >> gpurun_out/dbg.txt
env "$@" RT_DEBUG_BOUNCES=1 RT_B200_LIB=$PWD/exp/librt_$v.so python bench.py --steps 2 --warmup 3 --no-cpu-baseline $a 2>&1 >/dev/null | grep bounce | tail -20 | head -9 >> gpurun_out/dbg.txt
