#!/bin/bash
# the whole GPU suite, an experiment script (optional), a short bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -4 gpurun_out/t_gpu.log
if [ -n "$1" ]; then bash tools/$1.sh > gpurun_out/$1.txt 2>&1; cat gpurun_out/$1.txt; fi
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-eager --no-extras > gpurun_out/bench_new.json 2> gpurun_out/bench_new.err; echo "bench rc=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_new.json')); print(round(d['value']), 'e2e', round(d['e2e']['value']), d['clocks'])"
