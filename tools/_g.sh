mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
BT_FOLD=1 timeout 200 python tools/layer_table.py 8 > gpurun_out/table.txt 2>&1; cat gpurun_out/table.txt | cut -c1-100
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -1 gpurun_out/bench.json | cut -c1-300
