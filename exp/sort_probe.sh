#!/bin/sh
# ray reordering between bounces, per-bounce closest-hit launch times (one sub-frame, separate shading kernels: what the sort path needs)
cd "$(dirname "$0")/.."
for cfg in "RT_SORT_FROM=1000" "RT_SORT_FROM=3 RT_SORT_TO=8" "RT_SORT_FROM=3 RT_SORT_TO=8 RT_SORT_BITS=24" "RT_SORT_FROM=1 RT_SORT_TO=8" "RT_SORT_FROM=3 RT_SORT_TO=4"; do
  echo "== $cfg"
  env $cfg RT_FUSE_SHADE=0 RT_SUBFRAMES=1 RT_DEBUG_BOUNCES=1 python bench.py --steps 1 --warmup 2 --no-cpu-baseline 2>&1 >/dev/null | grep "^sub" | tail -9 | awk '{split($5,a,"\\.\\."); printf "b%s %.3f  ", $4, a[2]-a[1]} END {print ""}'
  env $cfg RT_FUSE_SHADE=0 RT_SUBFRAMES=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('  frame', d['frame_ms'], d['stage_ms'])"
done
