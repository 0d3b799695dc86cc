#!/bin/sh
# usage: exp/run_wl.sh workload v1 v2 ... -> frame ms per variant library
cd "$(dirname "$0")/.."
w=$1; shift
for v in "$@"; do
  RT_B200_LIB=$PWD/exp/librt_$v.so python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$w $v', 'frame', d['frame_ms'], 'e2e', d['e2e']['ms_per_step'], d['stage_ms'])"
done
