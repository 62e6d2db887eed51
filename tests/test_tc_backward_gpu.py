"""Backward on the tensor-core kernels (bf16 channels-last activations), on the SAME Philox draw, against

  (a) the ORACLE's autograd (fp64 torch ops of the reference's op sequence, conv_variational.py:362-385 /
      conv_flipout.py:370-417) evaluated on the operands the kernels multiply -- x and grad_out in bf16,
      W = bf16(bf16(sigma) bf16(eps) + bf16(mu)) (Flipout: bf16(mu) and bf16(bf16(sigma) bf16(eps))) -- so what is
      left is fp32 accumulation order, the bf16 rounding of grad_x and the straight-through use of the unrounded
      eps in grad_rho.  Documented bounds: grad_x  <= 2^-7 max|grad_x| (bf16 output: half an ulp is 2^-9 of the
      value, sums of up to K = c_out * taps terms re-associate); parameter gradients (fp32 outputs, reduction
      over batch * pixels) <= 4e-3 max|grad|;
  (b) the fp32 SIMT backward (itself pinned to the reference's autograd gradients by tests/test_parity_gpu.py):
      regression bound 2e-2 * max|grad| -- it includes the bf16 rounding of the weights, which (a) factors out.

grad_x = conv_transpose(grad_out, W) runs as the forward kernel on grad_out (zeros inserted for strided
layers) with the transposed, tap-reversed, regenerated weights."""
import pytest
import torch
import torch.nn.functional as F

import bayesian_torch_b200 as bt
from bayesian_torch_b200 import _lib, ops
from bayesian_torch_b200 import layers as L

DX_TOL, DP_TOL = 2.0 ** -7, 4e-3


def bf16r(t):
    return t.to(torch.bfloat16).to(torch.float64)


def oracle_grads(layer, x, go, kind, seed, sample, conv_kw=None):
    """fp64 autograd of the reference's op sequence on the kernels' rounded operands.  Returns grad_x and the
    gradients in layer.parameters() order (mu_w, rho_w, mu_b, rho_b)."""
    dev = x.device
    mu, rho = (t.detach().cpu().double() for t in layer._w())
    lid = layer._layer_id
    eps = ops.philox_eps_weight(mu.shape, seed, sample, lid, dev).cpu().double()
    sig = F.softplus(rho)
    xr = x.detach().cpu().double().requires_grad_(True)
    gor = go.detach().cpu().double()
    fn = (lambda a, w: F.linear(a, w)) if conv_kw is None else (lambda a, w: F.conv2d(a, w, None, **conv_kw))
    if kind == "reparam":
        w = bf16r(bf16r(sig) * bf16r(eps) + bf16r(mu)).requires_grad_(True)
        out = fn(xr, w)
        (out * gor).sum().backward()
        d_mu, d_delta = w.grad, w.grad
        g_b = gor
    else:
        wm = bf16r(mu).requires_grad_(True)
        wd = bf16r(bf16r(sig) * bf16r(eps)).requires_grad_(True)
        s_in = ops.philox_sign_shape(x.shape, dev, seed, sample, lid, _lib.KIND_SIGN_IN).cpu().double()
        s_out = ops.philox_sign_shape(go.shape, dev, seed, sample, lid, _lib.KIND_SIGN_OUT).cpu().double()
        out = fn(xr, wm) + fn(xr * s_in, wd) * s_out
        (out * gor).sum().backward()
        d_mu, d_delta = wm.grad, wd.grad
        g_b = gor * s_out
    grads = [d_mu, d_delta * eps * torch.sigmoid(rho)]
    if layer.mu_bias is not None:
        red = [0] + list(range(2, gor.dim()))
        rho_b = layer.rho_bias.detach().cpu().double()
        eps_b = ops.philox_normal(rho_b.numel(), seed, sample, lid, _lib.KIND_EPS_B, dev).cpu().double()
        grads += [gor.sum(red), g_b.sum(red) * eps_b * torch.sigmoid(rho_b)]
    return xr.grad, grads


def check_vs_oracle(dx_t, gp_t, dx_o, gp_o, dx_on_tc=True):
    e = float((dx_t.cpu().double() - dx_o).abs().max() / dx_o.abs().max())
    assert e <= (DX_TOL if dx_on_tc else 2e-2), f"grad_x vs oracle autograd: rel err {e}"
    for i, (a, b) in enumerate(zip(gp_t, gp_o)):
        e = float((a.cpu().double() - b).abs().max() / b.abs().max().clamp_min(1e-9))
        assert e <= DP_TOL, f"parameter gradient {i} vs oracle autograd: rel err {e}"

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
    dx_o, gp_o = oracle_grads(layer, x, go, kind, 5, 7, dict(stride=stride, padding=pad))
    check_vs_oracle(dx_t, gp_t, dx_o, gp_o, dx_on_tc=cin % 8 == 0)   # the RGB stem's grad_x runs on the fp32 kernel (unrounded W)


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
    dx_o, gp_o = oracle_grads(layer, x, go, kind, 6, 7)
    check_vs_oracle(dx_t, gp_t, dx_o, gp_o, dx_on_tc=n % 64 == 0)



@pytest.mark.parametrize("kind,cin,cout,k,pad,batch,hw", [
    ("flipout", 128, 128, 3, 1, 32, 14),    # 49 m-tiles, K = 1152: streamed, Flipout delta pre-sampled forward and backward
    ("reparam", 256, 256, 3, 1, 16, 14),    # backward data through the pre-sampled pass (eps from the packed tensor)
    ("flipout", 512, 256, 1, 0, 24, 14),    # 1x1 Flipout, two n-blocks
])
def test_training_path_presampled_weights_are_bit_identical(kind, cin, cout, k, pad, batch, hw, cuda_device, monkeypatch):
    """Flipout forward and the backward-data run take the pre-sampled pass too (Flipout: delta = bf16(sigma * eps) only,
    mu sub-tiles by TMA from the packed parameters; dgrad: eps from the packed tensor): outputs and the input gradient are
    the bits of the fused-sampling schedule (BT_TC2_WPRE_TRAIN=0)."""
    torch.manual_seed(cin + cout + k)
    bt.manual_seed(9)
    layer = _layer(kind, cin, cout, k, 1, pad, cuda_device)
    x = torch.randn(batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    go = torch.randn(batch, cout, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    res = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("BT_TC2_WPRE_TRAIN", flag)
        monkeypatch.setenv("BT_TC2_WPRE", flag)          # (the Reparameterization forward of the second case as well)
        layer.path, layer.sample_mode = "tc", "philox"
        for p in layer.parameters():
            p.grad = None
        xi = x.clone().requires_grad_(True)
        with bt.mc_sample(3):
            out, kl = layer(xi)
        (out.float() * go.float()).sum().backward()
        torch.cuda.synchronize()
        res[flag] = (out.detach().clone(), xi.grad.clone(), [p.grad.clone() for p in layer.parameters()])
    assert float(res["0"][0].float().abs().max()) > 0
    assert torch.equal(res["1"][0], res["0"][0])
    assert torch.equal(res["1"][1], res["0"][1])
    # (the weight-gradient kernel is not touched by the switch; its split-pixel fp32 atomics make its sums run-to-run
    # reproducible to rounding only)
    for a, b in zip(res["1"][2], res["0"][2]):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5 * float(b.abs().max()))
