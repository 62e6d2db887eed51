mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_tc_gpu.py tests/test_fuse_gpu.py -x -q 2>&1 | tail -3
export BT_AFF=1 BT_XS=1
for shape in "64 256 1 1 0 128 56 8 10" "256 256 3 1 1 128 14 8 10" "512 512 3 1 1 128 7 8 10"; do timeout 100 python tools/one_layer.py $shape; done
timeout 200 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 | cut -c1-200
