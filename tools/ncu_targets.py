"""Small programs for `ncu --set full` captures of each kernel family (one target per invocation).
usage: ncu_targets.py {kl|wgrad|flipout|stem|maxpool|tf32|hbm1x1|conv3x3|conv3x3_128|exp_res|lstm|pack}"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bayesian_torch_b200 as bt  # noqa: E402
from bayesian_torch_b200 import layers as L  # noqa: E402
from bayesian_torch_b200 import ops  # noqa: E402

what = sys.argv[1]
dev = torch.device("cuda:0")
torch.manual_seed(0)
cl = torch.channels_last


def conv(cin, cout, k, stride=1, pad=0, flip=False, relu=True):
    cls = L.Conv2dFlipout if flip else L.Conv2dReparameterization
    kw = {} if flip else {"prior_type": "normal"}
    lay = cls(cin, cout, k, stride=stride, padding=pad, bias=False, **kw).to(dev)
    lay.dnn_to_bnn_flag = True
    if not flip:
        lay._ep_scale, lay._ep_shift, lay._ep_relu = torch.rand(cout, device=dev) + 0.5, torch.randn(cout, device=dev), relu
    return lay


def run(fn, n=4):
    for _ in range(n):
        fn()
    torch.cuda.synchronize()


with torch.no_grad():
    if what == "kl":          # KL over one 16384 x 16384 tensor pair (2.1 GB of mu + rho) and over ResNet-50's table
        lin = L.LinearReparameterization(16384, 16384).to(dev)
        run(lambda: lin.kl_loss())
    elif what == "stem":      # 3 -> 64 7x7 / 2, 16 samples x 128 images (sample pairs on the N axis), + its patch kernel
        lay = conv(3, 64, 7, 2, 3)
        x = torch.randn(128, 3, 224, 224, device=dev).bfloat16().contiguous(memory_format=cl)
        def f():
            with bt.mc_batched(16, 0, shared_batch=128):
                lay(x)
        run(f)
    elif what == "maxpool":
        x = torch.randn(16 * 128, 64, 112, 112, device=dev).bfloat16().contiguous(memory_format=cl)
        run(lambda: ops.maxpool2d(x, 3, 2, 1))
    elif what == "flipout":   # ResNet-18 layer2 3x3 Flipout, batch 128
        lay = conv(128, 128, 3, 1, 1, flip=True)
        x = torch.randn(128, 128, 28, 28, device=dev).bfloat16().contiguous(memory_format=cl)
        def f():
            with bt.mc_sample(0):
                lay(x)
        run(f)
    elif what == "tf32":      # fp32 activations: 256 -> 256 3x3 at 14x14, 8 samples x 128 images
        lay = conv(256, 256, 3, 1, 1)
        x = torch.randn(8 * 128, 256, 14, 14, device=dev).contiguous(memory_format=cl)
        def f():
            with bt.mc_batched(8, 0):
                lay(x)
        run(f)
    elif what == "hbm1x1":    # 64 -> 256 1x1 at 56x56 with residual, 16 samples x 128 images: the HBM-bound class
        lay = conv(64, 256, 1)
        x = torch.randn(16 * 128, 64, 56, 56, device=dev).bfloat16().contiguous(memory_format=cl)
        r = torch.randn(16 * 128, 256, 56, 56, device=dev).bfloat16().contiguous(memory_format=cl)
        def f():
            with bt.mc_batched(16, 0):
                lay(x, residual=r)
        run(f)
    elif what == "conv3x3":   # 256 -> 256 3x3 at 14x14, 32 samples x 128 images: the compute-bound class (pre-sampled W, BN = 256 tiles)
        lay = conv(256, 256, 3, 1, 1)
        x = torch.randn(32 * 128, 256, 14, 14, device=dev).bfloat16().contiguous(memory_format=cl)
        def f():
            with bt.mc_batched(32, 0):
                lay(x)
        run(f)
    elif what == "conv3x3_128":   # 128 -> 128 3x3 at 28x28, 32 samples x 128 images: pre-sampled W, BN = 128 (one n-block)
        lay = conv(128, 128, 3, 1, 1)
        x = torch.randn(32 * 128, 128, 28, 28, device=dev).bfloat16().contiguous(memory_format=cl)
        def f():
            with bt.mc_batched(32, 0):
                lay(x)
        run(f)
    elif what == "exp_res":       # 1024 <- 256 1x1 at 14x14 with residual, 32 samples: the residual expansions of layer3
        lay = conv(256, 1024, 1)
        x = torch.randn(32 * 128, 256, 14, 14, device=dev).bfloat16().contiguous(memory_format=cl)
        r = torch.randn(32 * 128, 1024, 14, 14, device=dev).bfloat16().contiguous(memory_format=cl)
        def f():
            with bt.mc_batched(32, 0):
                lay(x, residual=r)
        run(f)
    elif what == "pack":      # parameter packing with the fused KL: Linear 8192 x 8192, forward(return_kl=True)
        lin = L.LinearReparameterization(8192, 8192).to(dev)
        x = torch.randn(4096, 8192, device=dev).bfloat16()
        run(lambda: lin(x))
if what == "wgrad":           # backward of a 128 -> 128 3x3 conv at 28x28, batch 128 (ResNet-18 layer2): dgrad + wgrad
    lay = L.Conv2dReparameterization(128, 128, 3, padding=1, bias=False, prior_type="normal").to(dev)
    x = torch.randn(128, 128, 28, 28, device=dev).bfloat16().contiguous(memory_format=cl).requires_grad_(True)
    def f():
        out, kl = lay(x)
        (out.float().square().mean() + kl).backward()
    run(f)
if what == "lstm":
    lstm = L.LSTMReparameterization(512, 512).to(dev)
    x = torch.randn(256, 8, 512, device=dev)
    with torch.no_grad():
        run(lambda: lstm(x))
print("ok", what)
