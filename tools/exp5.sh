#!/bin/bash
# round-robin item assignment (BT_TC2_RR) on the streamed layers with several n-blocks; pre-sampled W at 7x7 with it
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "1024 256 1 1 0 128 14" "512 256 1 1 0 128 28" "1024 512 1 1 0 128 14" "2048 512 1 1 0 128 7" "1024 2048 1 2 0 128 14" "512 1024 1 2 0 128 28" "256 256 3 1 1 128 14" "512 512 3 1 1 128 7" "512 512 3 2 1 128 14"; do
  echo "=== $CFG"
  run BT_TC2_RR=0
  run BT_TC2_RR=1
  run BT_TC2_RR=1 BT_TC2_WPRE=0
  run BT_TC2_RR=1 BT_TC2_WPRE_MIN=1
  run BT_TC2_RR=1 BT_TC2_WPRE_MIN=1 BT_TC2_MT=4
done
for CFG in "512 2048 1 1 0 128 7" "256 1024 1 1 0 128 14"; do
  for R in 0 1; do
  echo "=== $CFG res=$R"
  run BT_RES=$R BT_TC2_RR=0
  run BT_RES=$R BT_TC2_RR=1
  run BT_RES=$R BT_TC2_RR=1 BT_TC2_WPRE_MIN=1
  done
done
echo "=== cycle accounting 1024->256 @14"
python tc_prof.py 1024 256 1 1 0 128 14 32 2>&1 | tail -8
BT_TC2_RR=0 python tc_prof.py 1024 256 1 1 0 128 14 32 2>&1 | tail -8
echo "=== cycle accounting 256->256 3x3 @14"
python tc_prof.py 256 256 3 1 1 128 14 32 2>&1 | tail -8
