#!/bin/bash
cd "$(dirname "$0")"
export BT_AFF=1 BT_TC2_LOG=1
echo "=== stem 32 samples"; python tc_prof.py 3 64 7 2 3 128 224 32 2>&1 | tail -9
echo "=== 64->256 @56 no residual (XS)"; BT_XS=1 python tc_prof.py 64 256 1 1 0 128 56 32 2>&1 | tail -9
echo "=== 64->64 1x1 @56 (XS)"; BT_XS=1 python tc_prof.py 64 64 1 1 0 128 56 32 2>&1 | tail -9
