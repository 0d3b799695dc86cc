#!/bin/sh
# treelet refinement: one thread per treelet (RT_TREELET_COOP=0) against eight lanes per treelet; trees must be identical (same nodes visited)
cd "$(dirname "$0")/.."
for w in c3 c5a; do for coop in 0 1; do
  echo "== $w RT_TREELET_COOP=$coop"
  RT_TREELET_COOP=$coop RT_DEBUG_BUILD=1 python bench.py --workload $w --accel 1 --steps 2 --warmup 1 --no-cpu-baseline 2> gpurun_out/build_$w$coop.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('  build_ms', d['build_ms'], 'frame', d['frame_ms'], 'nodes', r['nodes_visited'], 'tris', r['tris_tested'])"
  grep "build " gpurun_out/build_$w$coop.err | awk '{a[$2" "$3]+=$(NF-1)} END {for (k in a) printf "   %s %.3f ms\n", k, a[k]}' | sort | head -20
done; done
