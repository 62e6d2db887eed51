#!/bin/bash
# weight-gradient kernel: split count rounded down to whole waves (BT_WG_ROUND_UP=1 restores the old rounding)
cd "$(dirname "$0")/.."
for v in 1 0; do echo -n "train BT_WG_ROUND_UP=$v :: "; BT_WG_ROUND_UP=$v python bench.py --workload train --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'ms', round(d['ms_per_step'],2), 'loss', d['loss'])"; done
