"""BASELINE.json configs[4]: LinearReparameterization sweep, square 1024..16384, batch 4096, bf16
tensor-core path (in-kernel Philox, sampled W never in HBM) and the KL kernel's HBM GB/s."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
from bayesian_torch_b200 import layers as L  # noqa: E402

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
S = int(sys.argv[2]) if len(sys.argv) > 2 else 1
print(f"batch {B}, {S} weight sample(s) per call; peak = MEASURED_PEAKS.json sustained bf16 1377.7 TFLOP/s, HBM 6548.8 GB/s")
print(f"{'n':>6s} {'fwd ms':>8s} {'TFLOP/s':>8s} {'%peak':>6s} {'kl ms':>7s} {'KL GB/s':>8s} {'%hbm':>5s}")
for n in (1024, 2048, 4096, 8192, 16384):
    torch.manual_seed(0)
    layer = L.LinearReparameterization(n, n, bias=True).to(dev)
    layer.path = "tc"
    x = torch.randn(B, n, device=dev).bfloat16()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    with torch.no_grad():
        for i in range(7):
            if i == 2:
                ev[0].record()
            with bt.mc_batched(S, 0, shared_batch=B):
                y = layer(x, return_kl=False)
        ev[1].record()
        for i in range(7):
            if i == 2:
                ev[2].record()
            kl = layer.kl_loss()
        ev[3].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / 5
    kms = ev[2].elapsed_time(ev[3]) / 5
    tf = 2.0 * S * B * n * n / ms / 1e9
    gbs = 8.0 * (n * n + n) / kms / 1e6
    print(f"{n:6d} {ms:8.3f} {tf:8.1f} {100 * tf / 1377.7:6.1f} {kms:7.3f} {gbs:8.1f} {100 * gbs / 6548.8:5.1f}")
