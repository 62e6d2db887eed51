#!/bin/bash
# 256-column tiles (BT_TC2_BN256) on the long reductions into >= 256 channels
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "256 256 3 1 1 128 14" "512 512 3 1 1 128 7" "512 512 3 2 1 128 14" "1024 256 1 1 0 128 14" "1024 512 1 1 0 128 14" "2048 512 1 1 0 128 7" "512 256 1 1 0 128 28" "1024 2048 1 2 0 128 14" "512 1024 1 2 0 128 28"; do
  echo "=== $CFG"
  run BT_TC2_BN256=0
  run BT_TC2_BN256=1 BT_TC2_LOG=1
done
CFG="512 2048 1 1 0 128 7"
for R in 0 1; do
echo "=== $CFG res=$R"
run BT_RES=$R BT_TC2_BN256=0
run BT_RES=$R BT_TC2_BN256=1
done
cd ..
for v in 0 1; do echo -n "bench BT_TC2_BN256=$v :: "; BT_TC2_BN256=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-140; done
