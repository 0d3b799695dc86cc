#!/bin/sh
# re-entry check of HEAD on one B200: gpu tests, plain bench, launch list, --set full of the traversal kernels
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/re_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/re_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/re_c3.json 2> gpurun_out/re_c3.err; echo "c3 rc=$?"; cut -c1-400 gpurun_out/re_c3.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/re_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/re_launches.log 2>&1; echo "launches rc=$?"
timeout 900 ncu --set full --import-source on --clock-control none -k 'regex:k_trace|k_shadow' -c 18 -f -o gpurun_out/re_trav \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/re_trav.log 2>&1; echo "full rc=$?"
