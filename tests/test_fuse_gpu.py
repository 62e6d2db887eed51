"""fold_batchnorm(): the rewritten graph must reproduce the unfused model (same weight draws)
to output-rounding accuracy, on the fp32 parity path and on the bf16 tensor-core path."""
import pytest
import torch
import torch.nn as nn

import bayesian_torch_b200 as bt
from bayesian_torch_b200 import layers as L

pytestmark = pytest.mark.gpu


class Block(nn.Module):
    """torchvision-Bottleneck-shaped: conv-bn-relu, conv-bn-relu, conv-bn, += identity, relu."""

    def __init__(self, c, mid):
        super().__init__()
        self.conv1 = L.Conv2dReparameterization(c, mid, 1, bias=False, prior_type="normal")
        self.bn1 = nn.BatchNorm2d(mid)
        self.conv2 = L.Conv2dReparameterization(mid, mid, 3, padding=1, bias=True, prior_type="normal")
        self.bn2 = nn.BatchNorm2d(mid)
        self.conv3 = L.Conv2dReparameterization(mid, c, 1, bias=False, prior_type="normal")
        self.bn3 = nn.BatchNorm2d(c)
        self.relu = nn.ReLU(inplace=True)

    def forward(self, x):
        identity = x
        out = self.relu(self.bn1(self.conv1(x)))
        out = self.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        out += identity
        return self.relu(out)


def make(cuda_device, c=64, mid=64):
    torch.manual_seed(0)
    # a Bayesian stem first: under mc_batched(shared_batch=...) the un-folded model input may
    # only feed Bayesian layers (a residual taken straight from it would not broadcast)
    stem = L.Conv2dReparameterization(c, c, 1, bias=False, prior_type="normal")
    m = nn.Sequential(stem, nn.BatchNorm2d(c), nn.ReLU(), Block(c, mid), Block(c, mid)).to(cuda_device)
    for mod in m.modules():
        if isinstance(mod, nn.BatchNorm2d):
            mod.running_mean.normal_(0, 0.2); mod.running_var.uniform_(0.5, 1.5)
            mod.weight.data.uniform_(0.5, 1.5); mod.bias.data.normal_(0, 0.2)
        if hasattr(mod, "dnn_to_bnn_flag"):
            mod.dnn_to_bnn_flag = True
    return m.eval()


def test_fold_batchnorm_fp32_matches_unfused(cuda_device):
    m = make(cuda_device)
    x = torch.randn(3, 64, 9, 9, device=cuda_device)
    with torch.no_grad(), bt.path("simt"), bt.mc_sample(3):
        want = m(x)
    f = bt.fold_batchnorm(m)
    assert f.folded == {"bn": 7, "relu": 5, "join": 0, "residual": 2}
    f2 = bt.fold_batchnorm(m, join_in_epilogue=False)
    assert f2.folded == {"bn": 7, "relu": 5, "join": 2, "residual": 0}
    with torch.no_grad(), bt.path("simt"), bt.mc_sample(3):
        torch.testing.assert_close(f2(x), want, rtol=1e-5, atol=1e-5)
    assert not any(isinstance(mod, nn.BatchNorm2d) for n, mod in f.named_modules() if n in {nd.target for nd in f.graph.nodes})
    with torch.no_grad(), bt.path("simt"), bt.mc_sample(3):
        got = f(x)
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5)
    with torch.no_grad(), bt.path("simt"), bt.mc_sample(3):      # the input model is untouched
        torch.testing.assert_close(m(x), want, rtol=0, atol=0)


def test_fold_batchnorm_bf16_tensor_core(cuda_device):
    m = make(cuda_device).to(memory_format=torch.channels_last).to(torch.bfloat16)
    x = torch.randn(4, 64, 14, 14, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    with torch.no_grad(), bt.mc_sample(7):
        want = m(x).float()
    f = bt.fold_batchnorm(m)
    with torch.no_grad(), bt.mc_sample(7):
        got = f(x).float()
    # the unfused chain rounds to bf16 after conv, BN and the join; the fused one rounds once per layer
    assert float((got - want).abs().max()) <= 3 * 2.0 ** -8 * float(want.abs().max())
    with torch.no_grad(), bt.mc_batched(3, 7, shared_batch=4):
        folded3 = f(x).float()
    torch.testing.assert_close(folded3[:4], got, rtol=0, atol=0)


def test_folded_layer_refuses_backward(cuda_device):
    m = make(cuda_device)
    f = bt.fold_batchnorm(m)
    x = torch.randn(2, 64, 5, 5, device=cuda_device, requires_grad=True)
    with bt.path("simt"):
        y = f(x)
    with pytest.raises(RuntimeError, match="inference-only"):
        y.sum().backward()


def test_add_relu_inplace(cuda_device):
    for dt in (torch.float32, torch.bfloat16):
        y = torch.randn(5, 7, 3, 3, device=cuda_device).to(dt)
        r = torch.randn(5, 7, 3, 3, device=cuda_device).to(dt)
        want = torch.relu(y.float() + r.float()).to(dt)
        got = bt.ops.add_relu_(y, r)
        assert got.data_ptr() == y.data_ptr()
        torch.testing.assert_close(got, want, rtol=0, atol=0)


def test_maxpool_nhwc_matches_torch(cuda_device):
    for shape, k, st, pd in [((3, 64, 17, 23), 3, 2, 1), ((2, 16, 8, 8), 2, 2, 0), ((1, 8, 5, 5), 3, 1, 1)]:
        x = torch.randn(*shape, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
        got = bt.ops.maxpool2d(x, k, st, pd)
        want = torch.nn.functional.max_pool2d(x, k, st, pd)
        assert got.is_contiguous(memory_format=torch.channels_last)
        torch.testing.assert_close(got, want, rtol=0, atol=0)


def test_frozen_layers_keep_packed_parameters_until_they_change(cuda_device):
    """fold_batchnorm marks its layers inference-only: the tensor-core forward packs (mu, sigma) once and later
    calls skip that launch; an in-place parameter update (version counter) or invalidate_packed() re-packs."""
    from bayesian_torch_b200 import ops
    m = make(cuda_device).to(memory_format=torch.channels_last).to(torch.bfloat16)
    f = bt.fold_batchnorm(m)
    x = torch.randn(4, 64, 14, 14, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)

    def run():
        ops.PROFILE.reset()
        with torch.no_grad(), bt.mc_sample(5):
            y = f(x).float()
        return y, ops.PROFILE.launches
    y0, n0 = run()
    y1, n1 = run()
    layers = [mod for mod in f.modules() if getattr(mod, "_frozen", False)]
    assert len(layers) == 7 and n1 == n0 - len(layers)      # one packing launch per layer saved
    assert torch.equal(y1, y0)
    with torch.no_grad():
        layers[0].mu_kernel.mul_(1.5)                        # in-place update: version counter moves
    y2, n2 = run()
    assert n2 == n1 + 1 and not torch.equal(y2, y0)
    bt.invalidate_packed(f)
    y3, n3 = run()
    assert n3 == n0 and torch.equal(y3, y2)
    with bt.mc_sample(5):                                     # grad mode: never cached
        ops.PROFILE.reset()
        f(x)
        assert ops.PROFILE.launches == n0
