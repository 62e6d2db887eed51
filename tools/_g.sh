mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 ) > gpurun_out/t.log; cat gpurun_out/t.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 400 python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err; cut -c1-220 gpurun_out/bench1.json; tail -2 gpurun_out/bench1.err
timeout 200 python tools/linear_sweep.py 4096 1 > gpurun_out/linear_sweep.txt 2>&1; timeout 300 python tools/linear_sweep.py 4096 8 >> gpurun_out/linear_sweep.txt 2>&1; cat gpurun_out/linear_sweep.txt
