#!/bin/bash
# second shared-memory allocation pass (BT_TC2_SMEM2): residual ring 2 -> 4 sub-tiles where the roundings left room
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "128 512 1 1 0 128 28" "256 1024 1 1 0 128 14" "512 2048 1 1 0 128 7"; do
  for R in 0 1; do
    echo "=== $CFG res=$R"
    run BT_RES=$R BT_TC2_SMEM2=0
    run BT_RES=$R BT_TC2_SMEM2=1 BT_TC2_LOG=1
  done
done
for CFG in "256 256 3 1 1 128 14" "1024 256 1 1 0 128 14" "512 256 1 1 0 128 28"; do
  echo "=== $CFG"
  run BT_TC2_SMEM2=0
  run BT_TC2_SMEM2=1
done
