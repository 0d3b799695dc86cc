#!/bin/sh
# per-bounce launch intervals of one C3 frame for a variant library and bench arguments: exp/bounces.sh <variant> [bench args]
cd "$(dirname "$0")/.."
v=$1; shift
RT_DEBUG_BOUNCES=1 RT_B200_LIB=$PWD/exp/librt_$v.so python bench.py --steps 1 --warmup 2 --no-cpu-baseline "$@" 2>&1 >/dev/null | grep "^sub" | tail -18 | awk '{printf "%s %s %s %s  trace %8.3f ms  rays %s\n", $1,$2,$3,$4, ($6+0 > 0 ? 0 : 0) , $0}' | sed 's/trace *0.000 ms  rays //' | cut -c1-150
