"""Host (Python + ctypes + launch) time of one folded ResNet-50 forward vs its device time, per
samples-per-pass: is the 4-samples-per-GPU case of the 8-GPU run launch-bound?"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
import bench  # noqa: E402

dev = torch.device("cuda:0")
model = bench.build_model(dev)
x = torch.randn(128, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
for mc in (4, 8, 32):
    with torch.no_grad():
        for _ in range(3):
            bt.mc_predict(model, x, mc, batch_samples=mc)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(5):
            bt.mc_predict(model, x, mc, batch_samples=mc)
        e1.record()
        t_host = (time.perf_counter() - t0) / 5 * 1e3
        torch.cuda.synchronize()
        print(f"samples per pass {mc:2d}: host {t_host:6.2f} ms per forward (enqueue only), device {e0.elapsed_time(e1) / 5:6.2f} ms")
