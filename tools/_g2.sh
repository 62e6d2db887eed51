set -x
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 10000 --csv --log-file gpurun_out/launches_r1a.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python tools/one_layer.py 64 64 3 1 1 128 56 1 3 > gpurun_out/plain_l1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tc_fwd -s 2 -c 1 -o gpurun_out/prof_l1conv2_r1a python tools/one_layer.py 64 64 3 1 1 128 56 1 3 > gpurun_out/ncu_l1.log 2>&1
python tools/one_layer.py 512 512 3 1 1 128 7 1 3 > gpurun_out/plain_l4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tc_fwd -s 2 -c 1 -o gpurun_out/prof_l4conv2_r1a python tools/one_layer.py 512 512 3 1 1 128 7 1 3 > gpurun_out/ncu_l4.log 2>&1
cat gpurun_out/plain_l1.log gpurun_out/plain_l4.log
ls -la gpurun_out
