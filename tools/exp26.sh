#!/bin/bash
# L2 prefetch cursor of the residual loader (BT_TC2_RES_PF = items ahead)
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1 BT_RES=1
run() { echo -n "$* :: "; env "$@" timeout 120 python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "64 256 1 1 0 128 56" "128 512 1 1 0 128 28" "256 1024 1 1 0 128 14" "512 2048 1 1 0 128 7"; do
  echo "=== $CFG +res"
  run BT_TC2_RES_PF=0
  run BT_TC2_RES_PF=2
  run BT_TC2_RES_PF=4
  run BT_TC2_RES_PF=8
done
cd ..
for v in 0 2 4; do echo -n "bench BT_TC2_RES_PF=$v :: "; BT_TC2_RES_PF=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'e2e', round(d['e2e']['value']))"; done
