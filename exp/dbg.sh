#!/bin/sh
# usage: exp/dbg.sh variant [ENV=val ...]: per-bounce timings of the last C3 frame
cd "$(dirname "$0")/.."
v=$1; shift
echo "== $v $*" >> gpurun_out/dbg.txt
env "$@" RT_DEBUG_BOUNCES=1 RT_B200_LIB=$PWD/exp/librt_$v.so python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | grep bounce | tail -20 | head -9 >> gpurun_out/dbg.txt
