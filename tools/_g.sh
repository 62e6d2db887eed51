mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_tc_gpu.py tests/test_rng_mc_gpu.py tests/test_fuse_gpu.py -x -q 2>&1 | tail -8 ) > gpurun_out/t.log; cat gpurun_out/t.log
export BT_AFF=1 BT_TC2_LOG=1
for si in 0 1; do echo "SINNER=$si"; BT_TC2_SINNER=$si timeout 120 python tools/one_layer.py 3 64 7 2 3 128 224 32 3 2>&1 | tail -2; done
unset BT_TC2_LOG
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/bench2.json 2> gpurun_out/bench2.err; cut -c1-200 gpurun_out/bench2.json; tail -3 gpurun_out/bench2.err
