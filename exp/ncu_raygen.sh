#!/bin/sh
# --set full of two k_raygen and two k_resolve launches of a C3 frame; raw page back as csv
cd "$(dirname "$0")/.."
ncu --set full --import-source on --clock-control none -k 'regex:k_raygen|k_resolve' -s 20 -c 5 -f -o gpurun_out/r02_raygen \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r02_raygen.log 2>&1
ncu -i gpurun_out/r02_raygen.ncu-rep --page raw --csv > gpurun_out/r02_raygen_raw.csv 2>/dev/null
ls -la gpurun_out/r02_raygen.ncu-rep
