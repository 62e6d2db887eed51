"""Conv2dFlipout / LinearFlipout forward on the tensor-core path vs the fp32 SIMT kernel, per ResNet-18 layer
shape (BASELINE.json configs[2]: batch 128, 224x224): ms, TFLOP/s (2 contractions = 4*M*N*K FLOPs).
usage: flipout_layers.py [batch] [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
from bayesian_torch_b200 import layers as L  # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 128
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda:0")
SHAPES = [  # cin, cout, k, stride, pad, hw   (ResNet-18 Bayesian layers after the stem)
    (64, 64, 3, 1, 1, 56), (64, 128, 3, 2, 1, 56), (128, 128, 3, 1, 1, 28), (64, 128, 1, 2, 0, 56),
    (128, 256, 3, 2, 1, 28), (256, 256, 3, 1, 1, 14), (128, 256, 1, 2, 0, 28),
    (256, 512, 3, 2, 1, 14), (512, 512, 3, 1, 1, 7), (256, 512, 1, 2, 0, 14),
]
paths = os.environ.get("BT_PATHS", "tc,simt").split(",")
print(f"batch {batch}; Flipout forward, bf16 channels-last; FLOPs = 2 contractions")
print(f"{'layer':28s} " + " ".join(f"{p + ' ms':>9s} {'TFLOP/s':>8s}" for p in paths))
for cin, cout, k, stride, pad, hw in SHAPES:
    torch.manual_seed(0)
    layer = L.Conv2dFlipout(cin, cout, k, stride=stride, padding=pad, bias=False).to(dev)
    layer.dnn_to_bnn_flag = True
    x = torch.randn(batch, cin, hw, hw, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    ho = (hw + 2 * pad - k) // stride + 1
    flops = 4.0 * batch * ho * ho * cout * cin * k * k
    cells = []
    for pth in paths:
        layer.path = pth
        with torch.no_grad():
            for i in range(iters + 2):
                if i == 2:
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                with bt.mc_sample(i):
                    layer(x)
            e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        cells.append(f"{ms:9.3f} {flops / ms / 1e9:8.1f}")
    print(f"{f'{cin}->{cout} {k}x{k}/{stride} @{hw}':28s} " + " ".join(cells))
