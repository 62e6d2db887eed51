#!/bin/bash
# what the epilogue-heavy expansions cost without each of their streams (BT_TC2_PROF=1 build, BT_TC2_DEBUG bits:
# 1 no output stores, 4 no MMAs, 8 no A loads, 64 no residual loads, 128 no W loads)
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1 BT_TC2_PROF=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
for CFG in "256 1024 1 1 0 128 14" "128 512 1 1 0 128 28"; do
  for R in 0 1; do
    echo "=== $CFG res=$R"
    for D in 0 1 4 8 64 128 65 73 201 205; do run BT_RES=$R BT_TC2_DEBUG=$D; done
  done
done
CFG="256 256 3 1 1 128 14"
echo "=== $CFG"
for D in 0 1 4 8 128 136 137; do run BT_TC2_DEBUG=$D; done
CFG="1024 256 1 1 0 128 14"
echo "=== $CFG"
for D in 0 1 4 8 128 136 137; do run BT_TC2_DEBUG=$D; done
