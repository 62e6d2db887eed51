mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_configs_gpu.py -x -q 2>&1 | tail -30 ) > gpurun_out/t.log; cat gpurun_out/t.log
