mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_transpose_gpu.py -q 2>&1 | tail -60 ) > gpurun_out/t.log; cat gpurun_out/t.log
