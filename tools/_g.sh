mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/t.log; cat gpurun_out/t.log
for pc in 0 1 0 1; do
BT_PACK_CACHE=$pc timeout 400 python bench.py --no-cpu-baseline --steps 8 > gpurun_out/bench_pc$pc.json 2> gpurun_out/bench2.err; echo "pack cache $pc: $(cut -c1-160 gpurun_out/bench_pc$pc.json)"
done
