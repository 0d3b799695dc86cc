#!/bin/sh
# usage: exp/build_ab3.sh v1 v2 ... -> C5a (binary layout) build phases per variant library
cd "$(dirname "$0")/.."
for v in "$@"; do
  echo "== $v"
  RT_B200_LIB=$PWD/exp/librt_$v.so RT_DEBUG_BUILD=1 python bench.py --workload c5a --accel 1 --steps 2 --warmup 1 --no-cpu-baseline 2> gpurun_out/build_$v.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('  build_ms', d['build_ms'], 'frame', d['frame_ms'], 'nodes', r['nodes_visited'])"
  grep "build treelet" gpurun_out/build_$v.err | awk '{printf "%s ", $(NF-1)} END {print ""}'
done
