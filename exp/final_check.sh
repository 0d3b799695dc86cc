#!/bin/sh
# what the driver runs at round end, on one B200: gpu tests, smoke(), default bench, reference arm
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/fc_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/fc_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/fc_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/fc_smoke.log
timeout 600 python bench.py > gpurun_out/fc_bench.json 2> gpurun_out/fc_bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/fc_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/fc_ref.json 2>/dev/null; echo "ref rc=$?"; cut -c1-200 gpurun_out/fc_ref.json
