mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain_ncu.log 2>&1 &&
BT_NCU_RANGE=1 timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python tools/summarize_launches.py gpurun_out/launches.csv 30 gpurun_out/tc2_traffic.json
export BT_AFF=1 BT_XS=1
timeout 200 python tools/one_layer.py 256 256 3 1 1 128 14 32 2 > gpurun_out/plain.log 2>&1 &&
timeout 800 ncu --set full --clock-control none --import-source on -k regex:tc2_fwd_kernel -s 2 -c 1 -f -o gpurun_out/prof_l3conv2 python tools/one_layer.py 256 256 3 1 1 128 14 32 2 > gpurun_out/ncu_full.log 2>&1
cat gpurun_out/plain.log | tail -1
