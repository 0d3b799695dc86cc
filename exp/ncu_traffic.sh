#!/bin/sh
# round 2: launch list of one bench command, then --set full of the traversal kernels of ONE C3 frame
# (2 sub-frames x 9 bounces x (k_trace + k_shadow) = 36 launches) -> gpurun_out/r02_trav.ncu-rep
cd "$(dirname "$0")/.."
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02_plain.json 2> gpurun_out/r02_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_launches_c3.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02_launches.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r02_plain1.json 2>> gpurun_out/r02_plain.err &&
ncu --set full --clock-control none -k 'regex:k_trace|k_shadow' -s 180 -c 36 -f -o gpurun_out/r02_trav \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r02_trav.log 2>&1
tail -2 gpurun_out/r02_trav.log | cut -c1-200
# the report itself is just over gpurun's 64 MiB limit for gpurun_out/: bring its raw page back as csv instead
ncu -i gpurun_out/r02_trav.ncu-rep --page raw --csv > gpurun_out/r02_trav_raw.csv 2>/dev/null && rm -f gpurun_out/r02_trav.ncu-rep
