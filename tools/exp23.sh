#!/bin/bash
# eight epilogue warps on the stem (stacked sample pairs): BT_TC2_EPI8_SPN
cd "$(dirname "$0")"
export BT_AFF=1
run() { echo -n "$* :: "; env "$@" timeout 120 python one_layer.py $CFG $S 5 2>&1 | tail -1 | sed 's/.*\] //'; }
CFG="3 64 7 2 3 128 224"
for S in 32 4; do echo "=== stem, $S samples"; run BT_TC2_EPI8_SPN=0; run BT_TC2_EPI8_SPN=1 BT_TC2_LOG=1; done
cd ..
for v in 0 1; do echo -n "bench BT_TC2_EPI8_SPN=$v :: "; BT_TC2_EPI8_SPN=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'e2e', round(d['e2e']['value']))"; done
