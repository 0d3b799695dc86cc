#!/bin/sh
cd "$(dirname "$0")/.."
for g in 0 4 7 12 24; do
  echo "== RT_TREELET_MIN_LEAVES=$g"
  RT_TREELET_MIN_LEAVES=$g sh exp/sweep_accel.sh c3 1
  RT_TREELET_MIN_LEAVES=$g sh exp/sweep_accel.sh c5a 0
done
