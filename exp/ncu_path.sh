#!/bin/sh
# --set full capture of the fused path kernel (both sub-frames of one C3 frame, after the warm-up frames)
cd "$(dirname "$0")/.."
export RT_FUSED=1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_path_plain.log 2>&1 &&
ncu --set full --import-source on --clock-control none -k 'regex:k_path' -s 6 -c 1 -f -o gpurun_out/r02_path \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_path.log 2>&1
tail -2 gpurun_out/ncu_path.log | cut -c1-200
