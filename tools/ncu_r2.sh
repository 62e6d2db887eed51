#!/bin/bash
# ncu --set full capture of one launch of each kernel family (gpurun -- tools/ncu_r2.sh); reports -> gpurun_out/r2_*.ncu-rep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cap() {  # name kernel-regex skip
  python tools/ncu_targets.py $1 > gpurun_out/ncu_plain_$1_$4.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o gpurun_out/r2_$1_$4 python tools/ncu_targets.py $1 > gpurun_out/ncu_$1_$4.log 2>&1
  rc=$?
  # the reports are ~20 MB each and gpurun_out/ travels back only below 64 MiB: summarise on the box, drop the report
  if [ -f gpurun_out/r2_$1_$4.ncu-rep ]; then
    ncu -i gpurun_out/r2_$1_$4.ncu-rep --page raw --csv > gpurun_out/r2_$1_$4_raw.csv 2>/dev/null
    python tools/ncu_extract.py gpurun_out/r2_$1_$4.ncu-rep gpurun_out/r2_$1_$4 "ncu --set full, one launch of $2 in tools/ncu_targets.py $1 (launch skip $3)" > /dev/null 2>&1
    rm -f gpurun_out/r2_$1_$4.ncu-rep
  fi
  echo "$1 $2 rc=$rc"
}
cap kl kl_normal_multi_kernel 2 kl
cap stem tc2_fwd_kernel 2 fwd
cap stem tc2_patch 2 patch
cap maxpool maxpool2d 2 pool
cap flipout tc2_fwd_kernel 2 flip
cap tf32 tc2_fwd_kernel 2 tf32
cap hbm1x1 tc2_fwd_kernel 2 fwd
cap conv3x3 tc2_fwd_kernel 2 fwd
cap conv3x3 tc2_presample_kernel 2 presample
cap conv3x3_128 tc2_fwd_kernel 2 fwd
cap exp_res tc2_fwd_kernel 2 fwd
cap pack tc2_pack_kernel 2 pack
cap wgrad tc_wgrad_kernel 2 wgrad
cap lstm lstm_step_kernel 4 lstm
du -sh gpurun_out
