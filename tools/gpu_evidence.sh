#!/bin/bash
# final evidence pass: launch list of one step, per-kernel ncu captures, the full default bench line and the reference arm
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-eager --no-extras --no-graph > gpurun_out/bench_nograph.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'tc2_|tc_w|maxpool2d|mc_accumulate|mc_finalize|kl_normal|add_relu|simt_' -c 480 --csv --log-file gpurun_out/r2_launches_raw.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-eager --no-extras --no-graph > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
bash tools/ncu_r2.sh
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "full bench rc=$?"; cut -c1-300 gpurun_out/bench_full.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"; cut -c1-400 gpurun_out/bench_reference.json
