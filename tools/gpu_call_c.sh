#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
bash tools/${1:-exp7}.sh > gpurun_out/${1:-exp7}.txt 2>&1; cat gpurun_out/${1:-exp7}.txt
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras > gpurun_out/bench_new.json 2> gpurun_out/bench_new.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_new.json
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
