"""In-kernel Philox path: the exported stream equals the numpy restatement of the RNG
contract (oracle/philox.py), the fused kernels consume exactly that stream (Philox mode ==
deterministic mode fed with the exported eps / signs), moment tests on the sampled
weights, stream independence; and the Monte-Carlo reduction kernels against the oracle."""
import numpy as np
import pytest
import torch

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import _lib, ops
from bayesian_torch_b200 import layers as L
from oracle import philox

pytestmark = pytest.mark.gpu


def test_exported_stream_matches_contract(cuda_device):
    for n, seed, sample, layer, kind in [(1, 1, 0, 1, 0), (5, 2**40 + 17, 3, 9, 1), (4099, 77, 2**31 + 5, 123456, 0)]:
        got = ops.philox_normal(n, seed, sample, layer, kind, cuda_device).cpu().double().numpy()
        want = philox.normal(n, seed, sample, philox.stream_id(layer, kind))
        np.testing.assert_allclose(got, want, rtol=0, atol=2e-5)   # fast log / sincos intrinsics
    for n, kind in [(1, 2), (127, 2), (128, 3), (1000, 3)]:
        got = ops.philox_sign(n, 99, 4, 7, kind, cuda_device).cpu().numpy()
        np.testing.assert_array_equal(got, philox.sign(n, 99, 4, philox.stream_id(7, kind)))


def test_moments_of_sampled_noise(cuda_device):
    """(W - mu)/sigma must be N(0,1): mean, variance, skewness, excess kurtosis, KS distance."""
    n = 1 << 22
    z = ops.philox_normal(n, 1234, 0, 5, 0, cuda_device).double()
    assert abs(float(z.mean())) < 4 / n ** 0.5
    assert abs(float(z.var()) - 1) < 4 * (2 / n) ** 0.5
    assert abs(float((z ** 3).mean())) < 4 * (15 / n) ** 0.5
    assert abs(float((z ** 4).mean()) - 3) < 4 * (96 / n) ** 0.5
    zs = z.sort().values[:: n // 4096]
    cdf = 0.5 * (1 + torch.erf(zs / 2 ** 0.5))
    emp = torch.arange(zs.numel(), device=zs.device, dtype=torch.float64) * (n // 4096) / n
    assert float((cdf - emp).abs().max()) < 2e-3
    assert float(z.abs().max()) > 4.5          # tails are present
    s = ops.philox_sign(n, 1234, 0, 5, 2, cuda_device)
    assert abs(float(s.mean())) < 4 / n ** 0.5
    assert abs(float((s[:-1] * s[1:]).mean())) < 4 / n ** 0.5      # neighbouring signs uncorrelated


def test_streams_are_independent(cuda_device):
    n = 1 << 18
    a = ops.philox_normal(n, 1, 0, 1, 0, cuda_device)
    for other in [ops.philox_normal(n, 1, 1, 1, 0, cuda_device),     # next MC sample
                  ops.philox_normal(n, 1, 0, 2, 0, cuda_device),     # other layer
                  ops.philox_normal(n, 1, 0, 1, 1, cuda_device),     # bias stream
                  ops.philox_normal(n, 2, 0, 1, 0, cuda_device)]:    # other seed
        assert abs(float((a * other).mean())) < 5 / n ** 0.5
    assert torch.equal(a, ops.philox_normal(n, 1, 0, 1, 0, cuda_device))   # counter-based: reproducible


@pytest.mark.parametrize("kind", ["linear", "conv", "linear_fo", "conv_fo"])
def test_philox_mode_equals_deterministic_mode_on_exported_stream(kind, cuda_device):
    """The fused kernel's in-register eps/signs are the exported stream: feeding the export
    through the deterministic ('buffer') mode must reproduce the Philox-mode output, and
    the gradients (which REGENERATE eps/signs in backward) must agree too."""
    torch.manual_seed(0)
    bt.manual_seed(4242)
    if kind == "linear":
        layer, x = L.LinearReparameterization(40, 24), torch.randn(9, 40)
    elif kind == "conv":
        layer, x = L.Conv2dReparameterization(8, 12, 3, padding=1, stride=2, prior_type="normal"), torch.randn(2, 8, 9, 9)
    elif kind == "linear_fo":
        layer, x = L.LinearFlipout(40, 24), torch.randn(9, 40)
    else:
        layer, x = L.Conv2dFlipout(8, 12, 3, padding=1), torch.randn(2, 8, 9, 9)
    layer = layer.to(cuda_device)
    layer.path = "simt"
    x = x.to(cuda_device)

    def run(mode):
        layer.sample_mode = mode
        for p in layer.parameters():
            p.grad = None
        xi = x.clone().requires_grad_(True)
        with bt.mc_sample(11):
            out, kl = layer(xi)
        (out.square().sum() + kl).backward()
        return out.detach(), xi.grad, [p.grad.clone() for p in layer.parameters()]

    out_p, dx_p, g_p = run("philox")
    w = layer._w()[0]
    layer_id, seed = layer._layer_id, ops.get_seed()
    getattr(layer, "eps_" + layer._wname).copy_(ops.philox_eps_weight(w.shape, seed, 11, layer_id, cuda_device))
    layer.eps_bias.copy_(ops.philox_normal(w.shape[0], seed, 11, layer_id, _lib.KIND_EPS_B, cuda_device))
    if "fo" in kind:
        # the sign streams run over activations in channels-last order, whatever the memory layout
        layer.sign_input = ops.philox_sign_like(x, seed, 11, layer_id, _lib.KIND_SIGN_IN)
        layer.sign_output = ops.philox_sign_like(out_p, seed, 11, layer_id, _lib.KIND_SIGN_OUT)
    out_b, dx_b, g_b = run("buffer")
    torch.testing.assert_close(out_p, out_b, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dx_p, dx_b, rtol=1e-4, atol=1e-4)
    for a, b in zip(g_p, g_b):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-4)
    with bt.mc_sample(12):
        layer.sample_mode = "philox"
        other, _ = layer(x)
    assert not torch.allclose(other, out_p)        # a different MC sample is a different draw


def test_torch_mode_replays_torch_generator(cuda_device):
    """'torch' sample mode draws eps with torch's generator exactly like the reference
    (linear_variational.py:164,176), so the buffers hold the eps used and the oracle agrees."""
    layer = L.LinearReparameterization(16, 8).to(cuda_device)
    layer.sample_mode, layer.path = "torch", "simt"
    x = torch.randn(5, 16, device=cuda_device)
    torch.manual_seed(5)
    out, kl = layer(x)
    torch.manual_seed(5)
    e_w = torch.empty_like(layer.eps_weight).normal_(); e_b = torch.empty_like(layer.eps_bias).normal_()
    assert torch.equal(layer.eps_weight, e_w) and torch.equal(layer.eps_bias, e_b)
    want, _ = oracle.linear_reparam(x, layer.mu_weight, layer.rho_weight, layer.mu_bias, layer.rho_bias, e_w, e_b)
    torch.testing.assert_close(out.cpu(), want.detach(), rtol=1e-5, atol=1e-5)


def test_mc_batched_equals_sample_loop(cuda_device):
    """S samples folded into one launch == S launches with mc_sample(s)."""
    torch.manual_seed(1)
    net = torch.nn.Sequential(L.Conv2dReparameterization(3, 8, 3, padding=1, prior_type="normal"), torch.nn.ReLU(),
                              L.Conv2dReparameterization(8, 8, 3, stride=2, prior_type="normal"), torch.nn.Flatten(),
                              L.LinearReparameterization(8 * 3 * 3, 10)).to(cuda_device)
    for m in net:
        if hasattr(m, "_wname"):
            m.path = "simt"
            m.dnn_to_bnn_flag = True
    x = torch.randn(4, 3, 8, 8, device=cuda_device)
    loop = []
    for s in range(3):
        with bt.mc_sample(20 + s):
            loop.append(net(x))
    with bt.mc_batched(3, 20, shared_batch=4):
        folded = net(x)
    torch.testing.assert_close(folded, torch.cat(loop), rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_mc_accumulate_and_finalize(dtype, cuda_device):
    torch.manual_seed(0)
    S, B, Cc = 6, 37, 1000
    logits = (torch.randn(S, B, Cc) * 3).to(dtype)
    acc = bt.MCAccumulator(B, Cc, cuda_device)
    acc.add(logits[:2].reshape(2 * B, Cc).to(cuda_device), 2)     # folded samples
    for s in range(2, S):
        acc.add(logits[s].to(cuda_device))
    res = acc.finalize(S)
    p = torch.softmax(logits.double(), -1)
    mean, var = oracle.mc_moments(logits.double())
    torch.testing.assert_close(res.mean.cpu().double(), mean, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(res.var.cpu().double(), var, rtol=1e-3, atol=1e-7)
    torch.testing.assert_close(res.entropy.cpu().double(), oracle.predictive_entropy(p), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(res.mutual_info.cpu().double(), oracle.mutual_information(p), rtol=1e-3, atol=1e-5)


def test_mc_predict_single_rank(cuda_device):
    net = torch.nn.Sequential(torch.nn.Flatten(), L.LinearReparameterization(12, 5)).to(cuda_device)
    net[1].path = "simt"
    net[1].dnn_to_bnn_flag = True
    x = torch.randn(7, 3, 2, 2, device=cuda_device)
    res = bt.mc_predict(net, x, 8)
    res_b = bt.mc_predict(net, x, 8, batch_samples=4)
    probs = []
    for s in range(8):
        with bt.mc_sample(s):
            probs.append(torch.softmax(net(x).double(), -1))
    probs = torch.stack(probs).cpu()
    torch.testing.assert_close(res.mean.cpu().double(), probs.mean(0), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(res_b.mean, res.mean, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(res.var.cpu().double(), probs.var(0, unbiased=False), rtol=1e-3, atol=1e-6)
    torch.testing.assert_close(res.mutual_info.cpu().double(), oracle.mutual_information(probs), rtol=1e-3, atol=1e-5)


def test_full_size_mc_batching_is_consistent(cuda_device):
    """BASELINE config at full size (ResNet-50, batch 128, 32 samples): all 32 samples folded
    into ONE model pass (activations beyond 2^31 elements, CTA pairs, residual epilogues) must
    give the same predictive moments as four passes of 8 -- sample s is the same draw either way."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    model = bench.build_model(cuda_device)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(128, 3, 224, 224, generator=g).to(cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    bt.manual_seed(77)
    with torch.no_grad():
        a = bt.mc_predict(model, x, 32, batch_samples=8)
    bt.manual_seed(77)
    with torch.no_grad():
        b = bt.mc_predict(model, x, 32, batch_samples=32)
    assert torch.isfinite(a.mean).all() and float(a.var.max()) > 0
    # same weights and inputs; only the accumulation order over samples differs (fp32 sums)
    torch.testing.assert_close(b.mean, a.mean, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(b.var, a.var, rtol=1e-3, atol=1e-7)
    torch.testing.assert_close(b.entropy, a.entropy, rtol=1e-4, atol=1e-5)


def test_host_batch_prefetcher(cuda_device):
    """Double-buffered pinned-host -> device input pipeline: values, dtype, memory format, ordering."""
    pf = bt.HostBatchPrefetcher(cuda_device, dtype=torch.bfloat16, memory_format=torch.channels_last)
    batches = [torch.randn(4, 3, 16, 16).pin_memory() for _ in range(3)]
    pf.submit(batches[0])
    with pytest.raises(RuntimeError, match="take"):
        pf.submit(batches[1])
    for i in range(3):
        x = pf.take()
        if i + 1 < 3:
            pf.submit(batches[i + 1])
        assert x.dtype == torch.bfloat16 and x.is_contiguous(memory_format=torch.channels_last)
        assert torch.equal(x.float().cpu(), batches[i].bfloat16().float())
    with pytest.raises(RuntimeError, match="nothing submitted"):
        pf.take()
    with pytest.raises(RuntimeError, match="host tensor"):
        pf.submit(torch.zeros(1, device=cuda_device))


def test_mc_predictor_graph_replay_equals_eager(cuda_device):
    """bt.MCPredictor: the step captured into a CUDA graph (the C ABI allocates nothing, synchronises nothing and
    launches on the caller's stream) returns bit-for-bit what the eager mc_predict returns, also for a NEW input of
    the same shape on replay, and re-captures when a parameter changes."""
    import torch.nn as nn
    torch.manual_seed(11)
    bt.manual_seed(31)
    net = nn.Sequential(nn.Conv2d(64, 64, 3, padding=1, bias=False), nn.BatchNorm2d(64), nn.ReLU(), nn.Conv2d(64, 128, 1),
                        nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(128, 16))
    bt.dnn_to_bnn(net, {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
                        "type": "Reparameterization", "moped_enable": True, "moped_delta": 0.5})
    net = net.to(cuda_device).eval().to(memory_format=torch.channels_last).to(torch.bfloat16)
    net = bt.fold_batchnorm(net, inplace=True)
    x1 = torch.randn(8, 64, 12, 12, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    x2 = torch.randn(8, 64, 12, 12, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    pred = bt.MCPredictor(net, 6, batch_samples=3)
    with torch.no_grad():
        want1 = bt.mc_predict(net, x1, 6, batch_samples=3)
        want2 = bt.mc_predict(net, x2, 6, batch_samples=3)
    got1 = pred(x1)
    assert pred.graphed and pred.launches_per_step >= 2 * 3 + 2 + 1
    n0 = ops.PROFILE.launches
    got2 = pred(x2)
    assert ops.PROFILE.launches - n0 == 1          # replay: only the finalize kernel is launched from the host
    got1b = pred(x1)
    for got, want in ((got1, want1), (got2, want2), (got1b, want1)):
        for a, b in zip(got, want):
            torch.testing.assert_close(a, b, rtol=0, atol=0)
    with torch.no_grad():
        net.get_submodule("0").mu_kernel.mul_(1.5)     # parameter version changes -> new capture with re-packed weights
        want3 = bt.mc_predict(net, x1, 6, batch_samples=3)
    got3 = pred(x1)
    torch.testing.assert_close(got3.mean, want3.mean, rtol=0, atol=0)
    assert float((got3.mean - got1.mean).abs().max()) > 0
