#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/bench_train_1gpu.json 2> gpurun_out/bench_train.err; echo "train rc=$?"; cut -c1-300 gpurun_out/bench_train_1gpu.json
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-eager --no-extras --no-graph > gpurun_out/bench_nograph.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'tc2_|tc_w|maxpool2d_nhwc|mc_accumulate|mc_finalize|kl_normal|add_relu|simt_' -c 400 --csv --log-file gpurun_out/r2_launches_raw.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-eager --no-extras --no-graph > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
