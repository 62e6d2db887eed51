"""ConvTranspose{1,2,3}d Reparameterization / Flipout beyond the reference fixtures (tests/test_parity_gpu.py
covers those): in-kernel Philox mode against the deterministic mode fed with the exported streams, the bf16
channels-last route (backward-data kernel of the equivalent convolution on tensor cores), and seeded larger
shapes against the CPU oracle."""
import pytest
import torch

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import _lib, ops
from bayesian_torch_b200 import layers as L

pytestmark = pytest.mark.gpu


def _feed_exported_streams(layer, x, out_shape, sample):
    """Fill eps / sign buffers with the streams the kernels generate for (seed, sample, layer)."""
    seed, lid = ops.get_seed(), layer._layer_id
    layer.eps_kernel.copy_(ops.philox_eps_weight(layer.mu_kernel.shape, seed, sample, lid, x.device))
    if layer.mu_bias is not None:
        layer.eps_bias.copy_(ops.philox_normal(layer.mu_bias.numel(), seed, sample, lid, _lib.KIND_EPS_B, x.device))
    if layer._flipout:
        # the equivalent convolution maps this layer's output to its input: our input signs are its SIGN_OUT stream
        layer.sign_input = ops.philox_sign_shape(x.shape, x.device, seed, sample, lid, _lib.KIND_SIGN_OUT)
        layer.sign_output = ops.philox_sign_shape(out_shape, x.device, seed, sample, lid, _lib.KIND_SIGN_IN)


@pytest.mark.parametrize("cls,rank", [("ConvTranspose1dReparameterization", 1), ("ConvTranspose2dReparameterization", 2),
                                      ("ConvTranspose3dReparameterization", 3), ("ConvTranspose1dFlipout", 1),
                                      ("ConvTranspose2dFlipout", 2), ("ConvTranspose3dFlipout", 3)])
def test_philox_mode_equals_buffer_mode_with_exported_streams(cuda_device, cls, rank):
    torch.manual_seed(3)
    layer = getattr(L, cls)(6, 10, 3, stride=2, padding=1, output_padding=1).to(cuda_device)
    layer.path = "simt"
    x = torch.randn(2, 6, *([5, 4, 3][:rank]), device=cuda_device, requires_grad=True)
    with bt.mc_sample(9):
        out_p, kl_p = layer(x)
    (out_p.square().sum() + kl_p).backward()
    grads_p = [x.grad.clone()] + [p.grad.clone() for p in layer.parameters()]
    with bt.mc_sample(9):
        again, _ = layer(x)
    assert torch.equal(again, out_p)                     # a sample index is a fixed draw
    with bt.mc_sample(10):
        other, _ = layer(x)
    assert not torch.allclose(other, out_p)
    x.grad = None
    layer.zero_grad()
    _feed_exported_streams(layer, x, out_p.shape, 9)
    layer.sample_mode = "buffer"
    out_b, kl_b = layer(x)
    torch.testing.assert_close(out_b, out_p, rtol=1e-6, atol=1e-6)
    (out_b.square().sum() + kl_b).backward()
    for a, b in zip([x.grad] + [p.grad for p in layer.parameters()], grads_p):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("flip", [False, True])
def test_seeded_larger_transpose_against_oracle(cuda_device, flip):
    g = torch.Generator().manual_seed(17)
    R = lambda *s: torch.randn(*s, generator=g)
    cls = L.ConvTranspose2dFlipout if flip else L.ConvTranspose2dReparameterization
    layer = cls(24, 40, (4, 3), stride=(2, 1), padding=(1, 2), dilation=(1, 2)).to(cuda_device)
    layer.sample_mode, layer.path = "buffer", "simt"
    layer.eps_kernel.copy_(R(24, 40, 4, 3)); layer.eps_bias.copy_(R(40))
    x = R(5, 24, 19, 23)
    args = [x, layer.mu_kernel, layer.rho_kernel, layer.mu_bias, layer.rho_bias, layer.eps_kernel, layer.eps_bias]
    kw = dict(stride=(2, 1), padding=(1, 2), dilation=(1, 2))
    if flip:
        s_in = torch.empty(5, 24, 19, 23).uniform_(-1, 1, generator=g).sign()
        want_shape = torch.nn.functional.conv_transpose2d(x, layer.mu_kernel.detach().cpu(), None, (2, 1), (1, 2), 0, 1, (1, 2)).shape
        s_out = torch.empty(want_shape).uniform_(-1, 1, generator=g).sign()
        layer.sign_input, layer.sign_output = s_in.to(cuda_device), s_out.to(cuda_device)
        want, wkl = oracle.conv_transpose_flipout(*args, s_in, s_out, **kw)
    else:
        want, wkl = oracle.conv_transpose_reparam(*args, **kw)
    out, kl = layer(x.to(cuda_device))
    scale = float(want.abs().max())
    torch.testing.assert_close(out.cpu(), want.detach(), rtol=1e-5, atol=1e-5 * scale)
    torch.testing.assert_close(kl.cpu(), wkl.detach(), rtol=1e-5, atol=1e-6)


def test_transpose_bf16_channels_last_route(cuda_device):
    """bf16 channels-last activations: the contraction goes through the tensor-core backward-data kernel
    when the equivalent convolution qualifies; compared with the fp32 parity path on the same draw."""
    torch.manual_seed(5)
    layer = L.ConvTranspose2dReparameterization(64, 64, 3, stride=2, padding=1, output_padding=1).to(cuda_device)
    x = torch.randn(4, 64, 14, 14, device=cuda_device)
    with bt.mc_sample(2), torch.no_grad():
        layer.path = "simt"
        want, _ = layer(x)
    with bt.mc_sample(2), torch.no_grad():   # a new block: a second call inside the first would draw afresh
        layer.path = "auto"
        got, _ = layer(x.bfloat16().contiguous(memory_format=torch.channels_last))
    assert got.dtype == torch.bfloat16 and got.shape == want.shape == (4, 64, 28, 28)
    err = float((got.float() - want).abs().max() / want.abs().max())
    assert err < 2e-2, err   # bf16 operands (x, sampled W) and bf16 output


def test_transpose_rejects_bad_output_padding_and_folding(cuda_device):
    layer = L.ConvTranspose2dReparameterization(4, 4, 3, stride=1, output_padding=1).to(cuda_device)
    with pytest.raises(RuntimeError, match="output_padding"):
        layer(torch.randn(1, 4, 5, 5, device=cuda_device))
    ok = L.ConvTranspose2dReparameterization(4, 4, 3).to(cuda_device)
    with pytest.raises(RuntimeError, match="mc_batched"), bt.mc_batched(2, 0):
        ok(torch.randn(2, 4, 5, 5, device=cuda_device))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ok.cpu()(torch.randn(1, 4, 5, 5))
