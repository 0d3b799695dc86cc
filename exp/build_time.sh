#!/bin/sh
# build times: C3 (bench default) and C5a with the phase breakdown
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/bt_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/bt_tests.log
RT_DEBUG_BUILD=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bt_c3.json 2> gpurun_out/bt_c3.err; echo "c3 rc=$?"
grep "build " gpurun_out/bt_c3.err | head -40
python -c "import json; d=json.load(open('gpurun_out/bt_c3.json')); print('c3 build_ms', d['build_ms'], 'frame', d['frame_ms'], 'nodes', d['roofline']['nodes_visited'])"
RT_DEBUG_BUILD=1 timeout 600 python bench.py --workload c5a --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bt_c5a.json 2> gpurun_out/bt_c5a.err; echo "c5a rc=$?"
grep "build " gpurun_out/bt_c5a.err | tail -24
python -c "import json; d=json.load(open('gpurun_out/bt_c5a.json')); print('c5a build_ms', d['build_ms'], 'frame', d['frame_ms'], 'nodes', d['roofline']['nodes_visited'])"
