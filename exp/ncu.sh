#!/bin/sh
# usage: exp/ncu.sh variant tag "bench args" [ENV=val...] -> gpurun_out/ncu_<tag>.ncu-rep (k_trace launches of the first frame)
cd "$(dirname "$0")/.."
v=$1; tag=$2; a=$3; shift; shift; shift
env "$@" RT_B200_LIB=$PWD/exp/librt_$v.so ncu --set full --import-source on --clock-control none -k regex:k_trace -c 9 -f -o gpurun_out/ncu_$tag \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline $a > gpurun_out/ncu_$tag.log 2>&1
tail -3 gpurun_out/ncu_$tag.log
