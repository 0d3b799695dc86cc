#!/bin/sh
# usage: exp/ab.sh "workloads" v1 v2 ...  -> frame ms of every variant library exp/librt_<v>.so on every workload
cd "$(dirname "$0")/.."
W=$1; shift
for w in $W; do sh exp/run_wl.sh $w "$@"; done
