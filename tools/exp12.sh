#!/bin/bash
# single-CTA kernel launched as clusters of two (BT_TC2_PAIR): the CTAs that share A tiles under the round-robin order on one TPC
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "256 256 3 1 1 128 14" "512 512 3 1 1 128 7" "1024 256 1 1 0 128 14" "1024 512 1 1 0 128 14" "2048 512 1 1 0 128 7" "512 256 1 1 0 128 28"; do
  echo "=== $CFG"
  run BT_TC2_PAIR=0
  run BT_TC2_PAIR=1
done
CFG="256 1024 1 1 0 128 14"
echo "=== $CFG res=1"
run BT_RES=1 BT_TC2_PAIR=0
run BT_RES=1 BT_TC2_PAIR=1
cd ..
for v in 0 1; do echo -n "bench BT_TC2_PAIR=$v :: "; BT_TC2_PAIR=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras 2>/dev/null | cut -c1-140; done
