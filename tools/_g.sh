mkdir -p gpurun_out
timeout 420 compute-sanitizer --tool memcheck --launch-timeout 120 --error-exitcode 7 python -m pytest tests/test_tc_gpu.py -q -x -k "shared_input or mc_batched or matches_oracle" > gpurun_out/memcheck.log 2>&1
echo "rc=$?"; grep -E "ERROR SUMMARY|passed|failed|Invalid|error" gpurun_out/memcheck.log | head -12
