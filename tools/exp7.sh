#!/bin/bash
# cycle accounting of the epilogue-heavy expansions with pre-sampled W
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1 BT_TC2_LOG=1
for CFG in "256 1024 1 1 0 128 14" "128 512 1 1 0 128 28" "512 2048 1 1 0 128 7"; do
  for R in 0 1; do
    echo "=== $CFG res=$R"
    BT_RES=$R python tc_prof.py $CFG 32 2>&1 | grep -v "^$" | tail -9
  done
done
echo "=== 64 64 3x3 @56"
python tc_prof.py 64 64 3 1 1 128 56 32 2>&1 | tail -9
echo "=== 64 256 @56 res=1"
BT_RES=1 python tc_prof.py 64 256 1 1 0 128 56 32 2>&1 | tail -9
