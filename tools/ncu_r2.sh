#!/bin/bash
# ncu --set full capture of one launch of each kernel family (gpurun -- tools/ncu_r2.sh); reports -> gpurun_out/r2_*.ncu-rep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cap() {  # name kernel-regex skip
  python tools/ncu_targets.py $1 > gpurun_out/ncu_plain_$1_$4.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o gpurun_out/r2_$1_$4 python tools/ncu_targets.py $1 > gpurun_out/ncu_$1_$4.log 2>&1
  echo "$1 $2 rc=$?"
}
cap kl kl_normal_multi_kernel 2 kl
cap stem tc2_fwd_kernel 2 fwd
cap stem tc2_patch 2 patch
cap maxpool maxpool2d_nhwc 2 pool
cap flipout tc2_fwd_kernel 2 flip
cap tf32 tc2_fwd_kernel 2 tf32
cap hbm1x1 tc2_fwd_kernel 2 fwd
cap conv3x3 tc2_fwd_kernel 2 fwd
cap pack tc2_pack_kernel 2 pack
cap wgrad tc_wgrad_kernel 2 wgrad
cap lstm lstm_step_kernel 4 lstm
ls -la gpurun_out/*.ncu-rep
