#!/bin/sh
# usage: exp/sweep_shard.sh "ENV=.." ...   C3 full frame and a 1/8 shard of C4 per setting (exp/shard_probe.py, PROBE_SHORT)
cd "$(dirname "$0")/.."
for cfg in "$@"; do
  echo "== $cfg"
  env $cfg PROBE_SHORT=1 timeout 300 python exp/shard_probe.py 2>&1 | grep -E "C3 full|C4 shard" 
done
