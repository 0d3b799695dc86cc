#!/bin/sh
cd "$(dirname "$0")/.."
ncu --set full --import-source on --clock-control none -k 'regex:k_trace|k_shadow' -c 8 -f -o gpurun_out/c5a_trav \
  python bench.py --workload c5a --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/c5a_trav.log 2>&1
tail -2 gpurun_out/c5a_trav.log | cut -c1-200
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/fin_c3.json 2> gpurun_out/fin_c3.err; echo "c3 rc=$?"
