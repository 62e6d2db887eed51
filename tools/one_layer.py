"""Run one Conv2dReparameterization shape on the tensor-core path a few times (ncu target).
usage: one_layer.py cin cout k stride pad batch hw [samples] [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
from bayesian_torch_b200 import layers as L  # noqa: E402

cin, cout, k, stride, pad, batch, hw = (int(v) for v in sys.argv[1:8])
S = int(sys.argv[8]) if len(sys.argv) > 8 else 1
iters = int(sys.argv[9]) if len(sys.argv) > 9 else 5
dev = torch.device("cuda:0")
torch.manual_seed(0)
layer = L.Conv2dReparameterization(cin, cout, k, stride=stride, padding=pad, bias=False, prior_type="normal").to(dev)
layer.path = "tc"
layer.dnn_to_bnn_flag = True
if os.environ.get("BT_AFF", "0") == "1":  # folded BatchNorm + ReLU epilogue
    layer._ep_scale = torch.rand(cout, device=dev) + 0.5
    layer._ep_shift = torch.randn(cout, device=dev)
    layer._ep_relu = True
XS = S if os.environ.get("BT_XS", "0") == "1" else 1  # per-sample activations (any layer but the first)
x = torch.randn(XS * batch, cin, hw, hw, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
res = None
if os.environ.get("BT_RES", "0") == "1":  # residual input (the identity branch), read by the epilogue
    ho = (hw + 2 * pad - k) // stride + 1
    res = torch.randn(S * batch, cout, ho, ho, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
with torch.no_grad():
    for i in range(iters + 2):
        if i == 2:
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        with bt.mc_batched(S, 0, shared_batch=batch):
            y = layer(x, residual=res) if res is not None else layer(x)
    e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
ho = y.shape[2]
flops = 2.0 * S * batch * ho * ho * cout * cin * k * k
print(f"{sys.argv[1:]} {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s  out {tuple(y.shape)}")
