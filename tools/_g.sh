mkdir -p gpurun_out
timeout 300 python tools/gap_profile.py 32 > gpurun_out/gaps32.txt 2>&1; tail -24 gpurun_out/gaps32.txt
