#!/bin/bash
# eight epilogue warps on pre-sampled layers (BT_TC2_EPI8_WPRE_NKB: longest reduction, in k-blocks, that gets them)
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "128 512 1 1 0 128 28" "256 1024 1 1 0 128 14" "512 2048 1 1 0 128 7" "256 512 1 2 0 128 56" "512 1024 1 2 0 128 28" "1024 2048 1 2 0 128 14"; do
  for R in 0 1; do
  echo "=== $CFG res=$R"
  run BT_RES=$R BT_TC2_EPI8_WPRE_NKB=0
  run BT_RES=$R BT_TC2_EPI8_WPRE_NKB=64
  run BT_RES=$R BT_TC2_EPI8_WPRE_NKB=64 BT_TC2_MT=2
  run BT_RES=$R BT_TC2_EPI8_WPRE_NKB=64 BT_TC2_MT=4
  done
done
for CFG in "512 256 1 1 0 128 28" "1024 256 1 1 0 128 14" "2048 512 1 1 0 128 7" "256 256 3 1 1 128 14" "128 128 3 1 1 128 28"; do
  echo "=== $CFG"
  run BT_TC2_EPI8_WPRE_NKB=0
  run BT_TC2_EPI8_WPRE_NKB=64
done
