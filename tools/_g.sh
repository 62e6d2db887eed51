mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 ) > gpurun_out/t.log; cat gpurun_out/t.log
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/bench2.json 2> gpurun_out/bench2.err; cut -c1-200 gpurun_out/bench2.json; tail -2 gpurun_out/bench2.err
