#!/bin/bash
# experiment: streaming vs stationary / CL on the wide short-K expansions
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1; }
for CFG in "128 512 1 1 0 128 28" "256 1024 1 1 0 128 14" "512 2048 1 1 0 128 7"; do
  echo "=== $CFG (+res)"
  run BT_RES=1
  run BT_RES=1 BT_TC2_STATIONARY=1
  run BT_RES=1 BT_TC2_CL=1
  run BT_RES=1 BT_TC2_STATIONARY=1 BT_TC2_MT=4
  run BT_RES=1 BT_TC2_RES4=0
done
for CFG in "256 512 1 2 0 128 56" "512 1024 1 2 0 128 28" "1024 2048 1 2 0 128 14" "64 256 1 1 0 128 56"; do
  echo "=== $CFG (no res)"
  run BT_RES=0
  run BT_RES=0 BT_TC2_STATIONARY=1
  run BT_RES=0 BT_TC2_CL=1
done
