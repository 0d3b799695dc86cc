#!/bin/sh
cd "$(dirname "$0")/.."
for lib in "" "$PWD/exp/librt_w4m24.so"; do
  echo "== lib ${lib:-default (64 registers)}"
  for w in c3 c5a c5b; do RT_B200_LIB=$lib sh exp/sweep_accel.sh $w 4; done
done
