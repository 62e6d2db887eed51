timeout 120 python -m pytest tests/test_tc_gpu.py -x -q 2>&1 | tail -3
timeout 120 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 120 python tools/layer_table.py 8 2>&1 | tail -27
timeout 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | cut -c1-200
