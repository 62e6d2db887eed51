mkdir -p gpurun_out
export BT_AFF=1 BT_XS=1
cd bayesian-torch_b200/csrc
for v in default wide default wide; do
cp libbt_$v.so libbt_b200.so; cd ../..
echo "== $v"
for a in "256 256 3 1 1 128 14" "512 2048 1 1 0 128 7"; do
timeout 120 python tools/one_layer.py $a 32 3 2>&1 | tail -1 | cut -c1-120
done
cd bayesian-torch_b200/csrc
done
cd ../..
timeout 300 python -m pytest tests/test_rng_mc_gpu.py -x -q 2>&1 | tail -2
