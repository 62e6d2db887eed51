#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_gpu.py tests/test_tc_backward_gpu.py -x -q -m gpu > gpurun_out/t_res.log 2>&1; echo "tc tests rc=$?"; tail -4 gpurun_out/t_res.log
bash tools/$1.sh > gpurun_out/$1.txt 2>&1; cat gpurun_out/$1.txt
