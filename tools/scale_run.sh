#!/bin/bash
# 1 -> 8 GPU scaling of both workloads on one box (gpurun --gpus 8 -- tools/scale_run.sh)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for N in ${NS:-1 2 4 8}; do
  if [ $N -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N))"; fi
  $L bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-eager --no-extras > gpurun_out/scale_mc_n$N.json 2> gpurun_out/scale_mc_n$N.err || tail -5 gpurun_out/scale_mc_n$N.err
  $L bench.py --gpus $N --steps 10 --warmup 3 --workload train > gpurun_out/scale_train_n$N.json 2> gpurun_out/scale_train_n$N.err || tail -5 gpurun_out/scale_train_n$N.err
done
python - <<'PY'
import json
for w in ("mc", "train"):
    for n in (1, 2, 4, 8):
        try:
            d = json.load(open(f"gpurun_out/scale_{w}_n{n}.json"))
            print(w, n, round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), "ms", round(d["ms_per_step"], 2), d.get("checksum", d.get("loss")))
        except Exception as e:
            print(w, n, "FAILED", e)
PY
