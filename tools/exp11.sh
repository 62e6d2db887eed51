#!/bin/bash
# per-GPU work of the 8-GPU run on one GPU (4 of the 32 samples): MT chosen by item spread (BT_TC2_MTQ)
cd "$(dirname "$0")/.."
for v in 0 1; do echo -n "samples=4 BT_TC2_MTQ=$v :: "; BT_TC2_MTQ=$v python bench.py --samples 4 --mc-batch 4 --steps 20 --warmup 5 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-160; done
for v in 0 1; do echo -n "samples=8 BT_TC2_MTQ=$v :: "; BT_TC2_MTQ=$v python bench.py --samples 8 --mc-batch 8 --steps 20 --warmup 5 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-160; done
echo -n "samples=4 BT_TC2_WPRE=0 :: "; BT_TC2_WPRE=0 python bench.py --samples 4 --mc-batch 4 --steps 20 --warmup 5 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-160
echo -n "samples=4 BT_TC2_RR=0 :: "; BT_TC2_RR=0 python bench.py --samples 4 --mc-batch 4 --steps 20 --warmup 5 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-160
python tools/layer_table.py 4 > gpurun_out/layer_table_mc4.txt 2>&1; tail -32 gpurun_out/layer_table_mc4.txt
