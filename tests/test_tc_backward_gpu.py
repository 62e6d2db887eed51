"""Backward on the tensor-core kernels (bf16 channels-last activations) against the fp32 SIMT backward
(itself pinned to the reference's autograd gradients by tests/test_parity_gpu.py) on the SAME Philox draw:
grad_x = conv_transpose(grad_out, W) runs as the forward kernel on grad_out (zeros inserted for strided
layers) with the transposed, tap-reversed, regenerated weights.  Differences left: bf16 rounding of the
weights / of grad_x; bound 2e-2 * max|grad| (observed ~5e-3)."""
import pytest
import torch

import bayesian_torch_b200 as bt
from bayesian_torch_b200 import layers as L

pytestmark = pytest.mark.gpu

CASES = [
    # kind, cin, cout, k, stride, pad, batch, hw
    ("reparam", 64, 64, 3, 1, 1, 2, 12),
    ("reparam", 64, 128, 3, 2, 1, 2, 14),     # strided: zeros inserted into grad_out
    ("reparam", 128, 64, 1, 1, 0, 3, 9),      # 1x1: a GEMM over pixels
    ("reparam", 64, 128, 1, 2, 0, 2, 16),     # 1x1 stride-2 downsample
    ("reparam", 64, 64, 3, 2, 1, 1, 15),      # odd input: the last row/column gets no gradient from the last tap
    ("flipout", 64, 64, 3, 1, 1, 2, 12),
    ("flipout", 64, 128, 3, 2, 1, 2, 14),
    ("flipout", 128, 256, 1, 2, 0, 2, 16),
    ("flipout", 256, 128, 3, 1, 1, 2, 7),
    ("reparam", 3, 64, 7, 2, 3, 2, 32),       # RGB stem: patch-matrix operand (no grad_x needed in practice; here SIMT dgrad)
    ("flipout", 3, 64, 7, 2, 3, 2, 32),
    ("reparam", 64, 320, 3, 1, 1, 2, 10),     # two n-tiles of 256, the second partly empty
    ("flipout", 128, 512, 3, 1, 1, 1, 9),     # Flipout at NT=256: all 512 TMEM columns
    ("reparam", 192, 64, 3, 1, 1, 3, 11),     # 27 k-units: an odd tail tile
]


def _layer(kind, cin, cout, k, stride, pad, dev):
    if kind == "reparam":
        return L.Conv2dReparameterization(cin, cout, k, stride=stride, padding=pad, prior_type="normal").to(dev)
    return L.Conv2dFlipout(cin, cout, k, stride=stride, padding=pad).to(dev)


def _grads(layer, x, go, path):
    layer.path, layer.sample_mode = path, "philox"
    for p in layer.parameters():
        p.grad = None
    xi = x.clone().requires_grad_(True)
    with bt.mc_sample(7):
        out, kl = layer(xi)
    (out.float() * go.float()).sum().backward()
    torch.cuda.synchronize()
    return xi.grad.float(), [p.grad.clone() for p in layer.parameters()]


@pytest.mark.parametrize("kind,cin,cout,k,stride,pad,batch,hw", CASES)
def test_conv_backward_tc_matches_simt(kind, cin, cout, k, stride, pad, batch, hw, cuda_device):
    torch.manual_seed(cin + cout + k + stride)
    bt.manual_seed(5)
    layer = _layer(kind, cin, cout, k, stride, pad, cuda_device)
    x = torch.randn(batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    ho = (hw + 2 * pad - k) // stride + 1
    go = torch.randn(batch, cout, ho, ho, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    dx_s, gp_s = _grads(layer, x, go, "simt")
    dx_t, gp_t = _grads(layer, x, go, "auto" if cin % 8 else "tc")   # 'tc' raises if a tensor-core kernel cannot run
    err = float((dx_t - dx_s).abs().max() / dx_s.abs().max())
    assert err < 2e-2, f"grad_x rel err {err}"
    for a, b in zip(gp_t, gp_s):
        e = float((a - b).abs().max() / b.abs().max().clamp_min(1e-6))
        assert e < 2e-2, f"parameter grad rel err {e}"


@pytest.mark.parametrize("kind,k,n,rows", [("reparam", 256, 128, 70), ("flipout", 512, 1000, 128), ("flipout", 128, 64, 33)])
def test_linear_backward_tc_matches_simt(kind, k, n, rows, cuda_device):
    torch.manual_seed(k + n)
    bt.manual_seed(6)
    layer = (L.LinearReparameterization(k, n) if kind == "reparam" else L.LinearFlipout(k, n)).to(cuda_device)
    x = torch.randn(rows, k, device=cuda_device).bfloat16()
    go = torch.randn(rows, n, device=cuda_device).bfloat16()
    dx_s, gp_s = _grads(layer, x, go, "simt")
    dx_t, gp_t = _grads(layer, x, go, "auto" if n % 64 else "tc")
    err = float((dx_t - dx_s).abs().max() / dx_s.abs().max())
    assert err < 2e-2, f"grad_x rel err {err}"
    for a, b in zip(gp_t, gp_s):
        e = float((a - b).abs().max() / b.abs().max().clamp_min(1e-6))
        assert e < 2e-2, f"parameter grad rel err {e}"
