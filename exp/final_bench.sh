#!/bin/sh
# round-end measurements on one B200: every single-GPU workload + the reference arm
cd "$(dirname "$0")/.."
python bench.py --steps 10 --warmup 3 > gpurun_out/fin_c3.json 2> gpurun_out/fin_c3.err; echo "c3 rc=$?"
python bench.py --steps 10 --warmup 3 --accel 3 --no-cpu-baseline > gpurun_out/fin_c3_wide8.json 2>/dev/null; echo "c3 wide8 rc=$?"
python bench.py --steps 10 --warmup 3 --workload c2 --no-cpu-baseline > gpurun_out/fin_c2.json 2>/dev/null; echo "c2 rc=$?"
python bench.py --steps 5 --warmup 3 --workload c5a --no-cpu-baseline > gpurun_out/fin_c5a.json 2>/dev/null; echo "c5a rc=$?"
python bench.py --steps 5 --warmup 3 --workload c5a --accel 3 --no-cpu-baseline > gpurun_out/fin_c5a_wide8.json 2>/dev/null; echo "c5a wide8 rc=$?"
python bench.py --steps 5 --warmup 3 --workload c5b --no-cpu-baseline > gpurun_out/fin_c5b.json 2>/dev/null; echo "c5b rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/fin_ref.json 2>/dev/null; echo "ref rc=$?"
