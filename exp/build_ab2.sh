#!/bin/sh
# treelet refinement rooted only at nodes with >= RT_TREELET_MIN_LEAVES triangles below them: build time against nodes visited
cd "$(dirname "$0")/.."
for w in c3 c5a; do for ml in 0 3 4 5 7; do
  echo "== $w RT_TREELET_MIN_LEAVES=$ml"
  RT_TREELET_MIN_LEAVES=$ml python bench.py --workload $w --accel 1 --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('  build_ms', d['build_ms'], 'frame', d['frame_ms'], 'nodes', r['nodes_visited'], 'tris', r['tris_tested'])"
done; done
