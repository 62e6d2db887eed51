mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_tc_gpu.py tests/test_fuse_gpu.py -x -q 2>&1 | tail -15
export BT_AFF=1 BT_XS=1
for shape in "64 256 1 1 0 128 56 8 10" "256 256 3 1 1 128 14 8 10"; do timeout 100 python tools/one_layer.py $shape; done
BT_FOLD=1 timeout 200 python tools/layer_table.py 8 > gpurun_out/table.txt 2>&1; cat gpurun_out/table.txt | cut -c1-100
timeout 200 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 | cut -c1-200
