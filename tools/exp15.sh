#!/bin/bash
# which switch costs the end-to-end arm (host batch copied H2D on a side stream during the step)
cd "$(dirname "$0")/.."
pr() { python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'e2e', round(d['e2e']['value']))"; }
for E in "X=0" "BT_TC2_PAIR=0" "BT_TC2_WPRE_PDL=1" "BT_PDL=0" "BT_TC2_BN256=0" "BT_TC2_SMEM2=0"; do
  echo -n "$E :: "; env $E python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | pr
done
