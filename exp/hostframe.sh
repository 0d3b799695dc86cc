#!/bin/sh
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/hf_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/hf_tests.log
for er in 1 0; do for m in host-frame copy; do
  RT_EARLY_RESOLVE=$er timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e $m > gpurun_out/hf_n1_${m}_$er.json 2> gpurun_out/hf_n1_${m}_$er.err; echo "n1 $m early=$er rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/hf_n1_${m}_$er.json')); print('$m early=$er', d['frame_ms'], d['e2e']['ms_per_step'], d['e2e']['value'], d.get('host_frame_equals_device_frame'), d['stage_ms'])" || tail -5 gpurun_out/hf_n1_${m}_$er.err
done; done
