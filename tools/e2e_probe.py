"""Per-step times of bench.py's end-to-end arm (H2D of the next batch on a side stream during the step), several rounds."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
from bench import build_model  # noqa: E402

dev = torch.device("cuda:0")
model = build_model(dev)
x_host = torch.randn(128, 3, 224, 224).pin_memory()
x_dev = x_host.to(dev).bfloat16().contiguous(memory_format=torch.channels_last)
pred = bt.MCPredictor(model, 32, batch_samples=32)
pf = bt.HostBatchPrefetcher(dev, dtype=torch.bfloat16, memory_format=torch.channels_last)
mode = sys.argv[1] if len(sys.argv) > 1 else "e2e"


def step():
    if mode == "resident":
        r = pred(x_dev)
        return r.mean.cpu()
    if pf._pending is None:
        pf.submit(x_host)
    xd = pf.take()
    r = pred(xd)
    if mode != "nosubmit":
        pf.submit(x_host)
    return r.mean.cpu(), r.var.cpu()


for _ in range(3):
    pred(x_dev)
torch.cuda.synchronize()
for rnd in range(6):
    ts = []
    for i in range(6):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    print(mode, "round", rnd, " ".join(f"{t:6.1f}" for t in ts), flush=True)
    if pf._pending is not None and mode != "e2e":
        pf.take()
