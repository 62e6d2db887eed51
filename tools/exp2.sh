#!/bin/bash
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "64 256 1 1 0 128 56" "128 512 1 1 0 128 28" "256 1024 1 1 0 128 14" "512 2048 1 1 0 128 7"; do
  for R in 0 1; do
  echo "=== $CFG res=$R"
  run BT_RES=$R BT_TC2_EPI8=0
  run BT_RES=$R BT_TC2_EPI8=1
  run BT_RES=$R BT_TC2_EPI8=1 BT_TC2_STATIONARY=1
  done
done
