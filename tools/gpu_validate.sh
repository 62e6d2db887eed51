#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras > gpurun_out/bench_new.json 2> gpurun_out/bench_new.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_new.json
python tools/ncu_targets.py maxpool > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:maxpool -c 6 python tools/ncu_targets.py maxpool 2>&1 | grep -E "maxpool|duration|dram__bytes" | tail -9
