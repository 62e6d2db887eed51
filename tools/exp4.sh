#!/bin/bash
# pre-sampled weights (BT_TC2_WPRE) against the fused sampling, streamed layers of ResNet-50 (32 samples x 128 images)
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "128 128 3 1 1 128 28" "256 256 3 1 1 128 14" "512 512 3 1 1 128 7" "1024 256 1 1 0 128 14" "2048 512 1 1 0 128 7" "1024 512 1 1 0 128 14" "256 512 1 2 0 128 56"; do
  echo "=== $CFG"
  run BT_TC2_WPRE=0
  run BT_TC2_WPRE=1
  run BT_TC2_WPRE=1 BT_TC2_MT=2
  run BT_TC2_WPRE=1 BT_TC2_MT=4
done
for CFG in "128 512 1 1 0 128 28" "256 1024 1 1 0 128 14" "512 2048 1 1 0 128 7"; do
  for R in 0 1; do
  echo "=== $CFG res=$R"
  run BT_RES=$R BT_TC2_WPRE=0
  run BT_RES=$R BT_TC2_WPRE=1
  run BT_RES=$R BT_TC2_WPRE=1 BT_TC2_MT=4
  done
done
echo "=== 64->256 @56: streaming order instead of the stationary panel (A read once)"
CFG="64 256 1 1 0 128 56"
for R in 0 1; do
  run BT_RES=$R
  run BT_RES=$R BT_TC2_STATIONARY=0
  run BT_RES=$R BT_TC2_STATIONARY=0 BT_TC2_WPRE_MIN=1
done
