mkdir -p gpurun_out
BT_TC2_LOG=1 timeout 300 python -m pytest tests/test_tc_backward_gpu.py -x -q 2>&1 | tail -40 > gpurun_out/bwd_test.log
cat gpurun_out/bwd_test.log
