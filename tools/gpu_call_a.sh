#!/bin/bash
# round-2 GPU call: the allocation-free KL test alone, the whole GPU suite, a short bench, the launch list, the per-kernel captures
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
if [ "$1" = tests ]; then
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
fi
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_short.json
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-eager --no-extras --no-graph > gpurun_out/bench_nograph.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_raw.csv \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-eager --no-extras --no-graph > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
bash tools/ncu_r2.sh
