#!/bin/sh
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/hf_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/hf_tests.log
for m in host-frame copy; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e $m > gpurun_out/hf_n1_$m.json 2> gpurun_out/hf_n1_$m.err; echo "n1 $m rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/hf_n1_$m.json')); print('$m', d['frame_ms'], d['e2e'], d.get('host_frame_equals_device_frame'))" || tail -5 gpurun_out/hf_n1_$m.err
done
