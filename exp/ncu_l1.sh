#!/bin/sh
# --set full of the first 14 traversal launches (bounces 0-3 of both sub-frames) of one C3 frame -> gpurun_out/$1.ncu-rep
cd "$(dirname "$0")/.."
ncu --set full --import-source on --clock-control none -k 'regex:k_trace|k_shadow' -s 180 -c 14 -f -o gpurun_out/$1 \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/$1.log 2>&1
tail -2 gpurun_out/$1.log | cut -c1-200
