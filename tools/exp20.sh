#!/bin/bash
# residual input added by the tensor core (BT_TC2_RESMMA) against the epilogue path
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1 BT_RES=1
run() { echo -n "$* :: "; env "$@" timeout 120 python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "64 256 1 1 0 128 56" "128 512 1 1 0 128 28" "256 1024 1 1 0 128 14" "512 2048 1 1 0 128 7"; do
  echo "=== $CFG +res"
  run BT_TC2_RESMMA=0
  run BT_TC2_RESMMA=1 BT_TC2_LOG=1
done
