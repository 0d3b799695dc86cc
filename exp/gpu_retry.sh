#!/bin/sh
# usage: exp/gpu_retry.sh <timeout_s> <log> <command...>   retries while the pod answers busy (rc 3)
T=$1; LOG=$2; shift 2
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  /usr/local/graft/bin/gpurun --timeout $T -- "$@" > $LOG 2>&1
  rc=$?
  if grep -q "status=transient\|retry in a few minutes" $LOG; then sleep 120; continue; fi
  exit $rc
done
exit 3
