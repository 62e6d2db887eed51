mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 > gpurun_out/gpu_tests.log
cat gpurun_out/gpu_tests.log
timeout 300 python tools/flipout_layers.py 128 5 > gpurun_out/flipout_layers.txt 2>&1
cat gpurun_out/flipout_layers.txt
