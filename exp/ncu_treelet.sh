#!/bin/sh
cd "$(dirname "$0")/.."
ncu --set full --import-source on --clock-control none -k 'regex:k_treelet_round' -s 2 -c 3 -f -o gpurun_out/treelet \
  python bench.py --workload c5a --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/treelet.log 2>&1
tail -2 gpurun_out/treelet.log | cut -c1-200
