#!/bin/bash
cd "$(dirname "$0")/.."
for v in 1 2; do echo -n "BT_TC2_WPRE_PDL=$v :: "; BT_TC2_WPRE_PDL=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-140; done
echo -n "BT_PDL=0 :: "; BT_PDL=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-140
