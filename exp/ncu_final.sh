#!/bin/sh
# final captures of the round: launch list of one bench command, then --set full of the traversal kernels of one frame
# (two sub-frames x 9 bounces x (k_trace + k_shadow) = 36 launches)
cd "$(dirname "$0")/.."
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/final_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/final_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/final_launches.log 2>&1
ncu --set full --import-source on --clock-control none -k 'regex:k_trace|k_shadow' -c 36 -f -o gpurun_out/final_trav \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/final_trav.log 2>&1
tail -2 gpurun_out/final_trav.log | cut -c1-200
