#!/bin/bash
# pre-sampled weights on the training path (BT_TC2_WPRE_TRAIN): Flipout forward + backward data
cd "$(dirname "$0")/.."
for v in 0 1; do echo -n "train BT_TC2_WPRE_TRAIN=$v :: "; BT_TC2_WPRE_TRAIN=$v python bench.py --workload train --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'ms', round(d['ms_per_step'],2), 'loss', d['loss'])"; done
python tools/train_step.py 2>&1 | tail -6
