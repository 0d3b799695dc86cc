#!/bin/sh
# usage: exp/run_shard.sh v1 v2 ... -> C3 full / C4 shard 0/8 / C4 full per variant in gpurun_out/shard.txt
cd "$(dirname "$0")/.."
for v in "$@"; do
  echo "== $v" >> gpurun_out/shard.txt
  PROBE_SHORT=1 RT_B200_LIB=$PWD/exp/librt_$v.so timeout 200 python exp/shard_probe.py >> gpurun_out/shard.txt 2>&1
done
cat gpurun_out/shard.txt
