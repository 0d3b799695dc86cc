#!/bin/sh
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/bt_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/bt_tests.log
for w in c3 c5a; do
timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bt_$w.json 2> gpurun_out/bt_$w.err; echo "$w rc=$?"
python -c "import json; d=json.load(open('gpurun_out/bt_$w.json')); print('$w build_ms', d['build_ms'], 'frame', d['frame_ms'], 'nodes', d['roofline']['nodes_visited'])"
done
