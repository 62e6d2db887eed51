mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_backward_gpu.py -x -q 2>&1 | tail -25 > gpurun_out/bwd_test.log
cat gpurun_out/bwd_test.log
