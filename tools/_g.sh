mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 ) > gpurun_out/t.log; cat gpurun_out/t.log
export BT_AFF=1 BT_TC2_LOG=1
for sp in 0 1; do echo "SPN=$sp"; BT_TC2_SPN=$sp timeout 120 python tools/one_layer.py 3 64 7 2 3 128 224 32 3 2>&1 | tail -2; done
unset BT_TC2_LOG BT_AFF
for sp in 0 1; do BT_TC2_SPN=$sp timeout 400 python bench.py --no-cpu-baseline --steps 8 > gpurun_out/bench_sp$sp.json 2> gpurun_out/bench2.err; echo "spn $sp: $(cut -c1-160 gpurun_out/bench_sp$sp.json)"; done
