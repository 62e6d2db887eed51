#!/bin/bash
# weight-gradient kernel: pipeline depth (2 CTAs per SM at 3 stages) and waves of CTAs
cd "$(dirname "$0")/.."
for E in "X=0" "BT_WG_STAGES=3" "BT_WG_STAGES=3 BT_WG_WAVES=4" "BT_WG_WAVES=1" "BT_WG_WAVES=3" "BT_WG_STAGES=2 BT_WG_WAVES=4"; do echo -n "train $E :: "; env $E python bench.py --workload train --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'ms', round(d['ms_per_step'],2))"; done
