#!/bin/bash
# end-of-round check: smoke(), the GPU suite, the default bench line, the reference arm
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu suite rc=$?"; tail -3 gpurun_out/t_gpu.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "full bench rc=$?"; python -c "import json; d=json.load(open('gpurun_out/bench_full.json')); print(round(d['value']), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'mixed', round(d['roofline']['mixed_bound_frac'],3), d['clocks'], 'cpu', round(d['cpu_baseline']['value'],2), 'eager', round(d['eager_b200']['bf16_channels_last']['value']))"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"; cut -c1-160 gpurun_out/bench_reference.json
python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/bench_train_1gpu.json 2> gpurun_out/bench_train.err; echo "train rc=$?"; cut -c1-200 gpurun_out/bench_train_1gpu.json
