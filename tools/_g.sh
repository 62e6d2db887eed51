mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
BT_FOLD=1 timeout 200 python tools/layer_table.py 8 > gpurun_out/table.txt 2>&1; head -3 gpurun_out/table.txt | cut -c1-100
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -1 gpurun_out/bench.json | cut -c1-300
BT_NCU_RANGE=1 timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
python tools/summarize_launches.py gpurun_out/launches.csv 12
