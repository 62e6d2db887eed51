mkdir -p gpurun_out
(timeout 300 python tools/linear_sweep.py 4096 1; timeout 300 python tools/linear_sweep.py 4096 8) > gpurun_out/linear_sweep.txt 2>&1
tail -14 gpurun_out/linear_sweep.txt
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
