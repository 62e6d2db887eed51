#!/bin/bash
# BN = 256 on shorter reductions (BT_TC2_BN256_NKB = fewest k-blocks that get 256-column tiles)
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "256 1024 1 1 0 128 14" "256 512 1 2 0 128 56"; do
  for R in 0 1; do
    echo "=== $CFG res=$R"
    run BT_RES=$R BT_TC2_BN256_NKB=8
    run BT_RES=$R BT_TC2_BN256_NKB=4
  done
done
cd ..
for v in 8 4 2; do echo -n "bench BT_TC2_BN256_NKB=$v :: "; BT_TC2_BN256_NKB=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-140; done
python tools/layer_table.py 32 > gpurun_out/layer_table_mc32.txt 2>&1; tail -32 gpurun_out/layer_table_mc32.txt
