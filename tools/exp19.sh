#!/bin/bash
cd "$(dirname "$0")"
export BT_AFF=1 BT_XS=1
run() { echo -n "$* :: "; env "$@" python one_layer.py $CFG 32 5 2>&1 | tail -1 | sed 's/.*\] //'; }
CFG="64 256 1 1 0 128 56"; run BT_RES=0; run BT_RES=1
CFG="64 64 1 1 0 128 56"; run X=1
CFG="128 512 1 1 0 128 28"; run BT_RES=1
CFG="256 1024 1 1 0 128 14"; run BT_RES=1
CFG="256 256 3 1 1 128 14"; run X=1
