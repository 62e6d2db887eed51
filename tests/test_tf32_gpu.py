"""fp32 activations on the tensor cores: the tf32 instantiation of the fused sample-GEMM kernel (fp32 A tiles by
TMA, W = tf32(fmaf(sigma, eps, mu)) sampled in shared memory, tcgen05.mma.kind::tf32, fp32 accumulate / outputs).
This is what `dnn_to_bnn(model).cuda()(x)` -- the reference's README usage, fp32 NCHW -- runs on by default
(path 'auto'); path 'simt' stays the rtol-1e-5 parity mode.

Documented bound against the reference's fp32 arithmetic (linear_variational.py:181, conv_variational.py:384): tf32
keeps 10 mantissa bits of each operand (x truncated by the tensor core, W rounded to nearest), products and sums are
fp32: |err| <= TF32_TOL * max|out| with TF32_TOL = 2e-3 (observed printed below); KL is evaluated in fp32 (rtol 1e-5)."""
import pytest
import torch

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import _lib, ops
from bayesian_torch_b200 import layers as L

pytestmark = pytest.mark.gpu
TF32_TOL = 2e-3

CASES = [
    # cin, cout, k, stride, pad, batch, hw, bias, layout
    (64, 64, 1, 1, 0, 2, 14, True, "nhwc"),
    (64, 96, 3, 1, 1, 3, 17, False, "nchw"),     # NCHW input: re-laid out once by the host code
    (32, 128, 3, 2, 1, 2, 16, True, "nhwc"),     # C_in = 32: one k-block per tap
    (128, 136, 3, 1, 1, 2, 14, True, "nhwc"),    # streaming (K = 1152), N tail
    (3, 64, 7, 2, 3, 2, 32, True, "nchw"),       # RGB stem: fp32 patch matrix, padded K
    (256, 64, 1, 1, 0, 4, 28, False, "nhwc"),    # several k-blocks, stationary
    (48, 40, 3, 1, 1, 2, 9, True, "nhwc"),       # C_in % 32 != 0 -> patch matrix
]


@pytest.mark.parametrize("cin,cout,k,stride,pad,batch,hw,bias,layout", CASES)
def test_conv2d_tf32_matches_fp32_reference(cin, cout, k, stride, pad, batch, hw, bias, layout, cuda_device):
    torch.manual_seed(cin + cout + k)
    layer = L.Conv2dReparameterization(cin, cout, k, stride=stride, padding=pad, bias=bias, prior_type="normal").to(cuda_device)
    layer.sample_mode = "buffer"
    layer.eps_kernel.normal_()
    if bias:
        layer.eps_bias.normal_()
    x = torch.randn(batch, cin, hw, hw, device=cuda_device)
    if layout == "nhwc":
        x = x.contiguous(memory_format=torch.channels_last)
    g = ops._build_geom  # noqa: F841  (kept for debugging sessions)
    n0 = ops.PROFILE.launches
    out, kl = layer(x)                      # default path: auto
    assert out.dtype == torch.float32
    want, wkl = oracle.conv_reparam(x.cpu(), layer.mu_kernel.detach().cpu(), layer.rho_kernel.detach().cpu(),
                                    None if not bias else layer.mu_bias.detach().cpu(), None if not bias else layer.rho_bias.detach().cpu(),
                                    layer.eps_kernel.cpu(), None if not bias else layer.eps_bias.cpu(), stride, pad, 1, 1)
    err = float((out.cpu() - want).abs().max() / want.abs().max())
    print(f"[tf32] conv {cin}->{cout} k{k} s{stride}: rel err {err:.2e}")
    assert 1e-7 < err <= TF32_TOL, err       # (> 0: it really ran on tf32 tensor cores, not on the fp32 FFMA kernel)
    torch.testing.assert_close(kl.cpu(), wkl.detach(), rtol=1e-5, atol=1e-6)
    layer.path = "simt"
    exact, _ = layer(x)
    torch.testing.assert_close(exact.cpu(), want.detach(), rtol=1e-5, atol=1e-5 * float(want.abs().max()))


def test_linear_tf32_philox_and_mc_batched(cuda_device):
    torch.manual_seed(9)
    bt.manual_seed(17)
    lin = L.LinearReparameterization(256, 200).to(cuda_device)
    x = torch.randn(96, 256, device=cuda_device)
    with torch.no_grad():
        with bt.mc_sample(4):
            got, kl = lin(x)
        with bt.path("simt"), bt.mc_sample(4):
            want, wkl = lin(x)
        with bt.mc_batched(3, 3, shared_batch=96):
            folded = lin(x, return_kl=False)
    err = float((got - want).abs().max() / want.abs().max())
    print(f"[tf32] linear 256->200 philox: rel err {err:.2e}")
    assert 1e-7 < err <= TF32_TOL
    torch.testing.assert_close(kl, wkl, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(folded[96:192], got, rtol=0, atol=0)      # sample 4 of the folded call == the single call


def test_tf32_resnet_block_default_usage(cuda_device):
    """The reference's README flow on fp32 NCHW input: dnn_to_bnn(model); model.cuda(); model(x) -- every Bayesian conv
    lands on the tf32 kernel (one layout copy at the first layer), outputs within the tf32 bound of the fp32 kernels, and
    the training step's backward (fp32 kernels) still runs."""
    import torch.nn as nn
    torch.manual_seed(4)
    bt.manual_seed(23)
    net = nn.Sequential(nn.Conv2d(3, 32, 3, padding=1), nn.BatchNorm2d(32), nn.ReLU(), nn.Conv2d(32, 64, 3, stride=2, padding=1),
                        nn.BatchNorm2d(64), nn.ReLU(), nn.Conv2d(64, 64, 1), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(64, 10))
    bt.dnn_to_bnn(net, {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
                        "type": "Reparameterization", "moped_enable": False, "moped_delta": 0.5})
    net = net.to(cuda_device)
    x = torch.randn(8, 3, 32, 32, device=cuda_device)
    paths = []
    hook = ops.PROFILE
    hook.reset(); hook.timing = True
    with bt.mc_sample(1):
        out = net(x)
    hook.timing = False
    paths = [r[0] for r in hook.records]
    assert paths.count(_lib.PATH_TC_BF16) >= 3, paths          # the convolutions ran on tensor cores
    with bt.path("simt"), bt.mc_sample(1):
        want = net(x)
    err = float((out - want).abs().max() / want.abs().max())
    print(f"[tf32] small net: rel err {err:.2e}")
    assert err <= 5 * TF32_TOL
    loss = out.square().mean() + bt.get_kl_loss(net) / 8
    loss.backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in net.parameters())
