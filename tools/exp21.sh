#!/bin/bash
# second A-loading thread (BT_TC2_LOADERS=2: the idle residual-loader warp loads ring 1)
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" timeout 120 python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "64 64 3 1 1 128 56" "128 128 3 1 1 128 28" "256 256 3 1 1 128 14" "1024 256 1 1 0 128 14" "256 64 1 1 0 128 56" "512 128 1 1 0 128 28" "64 256 1 1 0 128 56"; do
  echo "=== $CFG"
  run BT_TC2_LOADERS=1
  run BT_TC2_LOADERS=2
done
cd ..
for v in 1 2; do echo -n "bench BT_TC2_LOADERS=$v :: "; BT_TC2_LOADERS=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'e2e', round(d['e2e']['value']))"; done
