#!/bin/sh
# usage: [EXTRA="--accel 4"] exp/run_wl.sh workload v1 v2 ... -> frame ms per variant library exp/librt_<v>.so
cd "$(dirname "$0")/.."
w=$1; shift
for v in "$@"; do
  RT_B200_LIB=$PWD/exp/librt_$v.so python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline $EXTRA 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; print('$w $v $EXTRA', 'frame', d['frame_ms'], 'e2e', d['e2e']['ms_per_step'], d['stage_ms'], 'nodes %.1fM tris %.1fM' % (r['nodes_visited']/1e6, r['tris_tested']/1e6))"
done
