#!/bin/bash
# the stem (x shared by the samples): sample-fastest order with pre-sampled W (default) against one stationary panel per sample pair
cd "$(dirname "$0")"
export BT_AFF=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
CFG="3 64 7 2 3 128 224"
run BT_TC2_SINNER=0
run BT_TC2_LOG=1
CFG="3 64 7 2 3 128 224"
echo "4 samples:"; for E in BT_TC2_SINNER=0 X=1; do echo -n "$E :: "; env $E python one_layer.py $CFG 4 5 2>&1 | tail -1 | sed 's/.*\] //'; done
cd ..
for E in "BT_TC2_SINNER=0" "X=0"; do echo -n "bench $E :: "; env $E python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'e2e', round(d['e2e']['value']))"; done
