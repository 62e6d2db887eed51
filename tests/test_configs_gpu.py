"""Model-level parity for BASELINE.json configs[0..2] (the non-headline configs are parity cases, not bench
lines): the CUDA path runs the model in its default in-kernel Philox mode; a CPU twin of the same model has
every Bayesian layer replaced by the oracle restatement fed with the EXPORTED streams of the same
(seed, sample, layer) keys, so both sides see identical draws.

  configs[0]  Bayesian SimpleCNN (models/bayesian/simple_cnn_variational.py:13-89) on 1x28x28, batch 128:
              forward + KL + backward, fp32 parity path, rtol 1e-5 on outputs / KL, 1e-4 on gradients
  configs[1]  Bayesian ResNet-20 (models/bayesian/resnet_variational.py:41-183, option-A shortcuts) on 3x32x32,
              batch 256, S = 8 Monte-Carlo samples: predictive mean / variance; fp32 path at rtol 1e-4 and the
              bf16 tensor-core path at the documented bf16 bound
  configs[2]  torchvision ResNet-18 + dnn_to_bnn(type="Flipout"), one training step (forward, get_kl_loss,
              CE + kl / batch, backward) at a reduced batch / image size the CPU twin finishes in seconds

Gradients through ReLU / max-pool: both route gradients DISCONTINUOUSLY.  The two sides agree to ~1e-6 on every
activation, so an activation within ~1e-6 of zero (or of a pooling tie) routes differently, and one re-routed element
moves a weight gradient -- a random-sign sum over ~1e5 positions -- by ~1/sqrt(N) ~ 3e-3 of its size, on every layer
upstream.  With millions of activations a few such elements exist in most draws, so an element-wise 1e-4 check of the
true architecture passes or fails by luck.  Each training-step config is therefore checked twice: the TRUE
architecture with outputs / KL / loss at rtol 1e-5 and gradients by relative L2 error (<= 2e-2, a few re-routings),
and a SMOOTH twin of it (softplus for ReLU, average for max pooling: same layers, same shapes, same kernels) with
gradients element-wise at rtol 1e-4 (configs[0]) / 2e-3 (configs[2], 21 layers of fp32 re-association).
"""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import _lib, ops
from bayesian_torch_b200 import layers as L
from bayesian_torch_b200.layers.base_variational_layer import _BayesLayer

pytestmark = pytest.mark.gpu


class OracleFed(nn.Module):
    """CPU stand-in of one drop-in layer: oracle arithmetic on the streams the CUDA kernels generate."""

    def __init__(self, layer: _BayesLayer, dev):
        super().__init__()
        self.rank, self.flip, self.lid, self.dev = layer._rank, layer._flipout, layer._layer_id, dev
        self.flag = layer.dnn_to_bnn_flag
        mu, rho = layer._w()
        self.mu_w, self.rho_w = nn.Parameter(mu.detach().cpu().clone()), nn.Parameter(rho.detach().cpu().clone())
        self.mu_b = None if layer.mu_bias is None else nn.Parameter(layer.mu_bias.detach().cpu().clone())
        self.rho_b = None if layer.mu_bias is None else nn.Parameter(layer.rho_bias.detach().cpu().clone())
        self.conv = layer._conv_args() if self.rank else None
        self.sample = 0

    def kl_loss(self):
        return oracle.kl_normal_from_rho(self.mu_w, self.rho_w, self.mu_b, self.rho_b)

    def forward(self, x, return_kl=True):
        seed, s = ops.get_seed(), self.sample
        eps_w = ops.philox_eps_weight(self.mu_w.shape, seed, s, self.lid, self.dev).cpu()
        eps_b = None if self.mu_b is None else ops.philox_normal(self.mu_b.numel(), seed, s, self.lid, _lib.KIND_EPS_B, self.dev).cpu()
        args = [x, self.mu_w, self.rho_w, self.mu_b, self.rho_b, eps_w, eps_b]
        if self.flip:
            s_in = ops.philox_sign_shape(x.shape, self.dev, seed, s, self.lid, _lib.KIND_SIGN_IN).cpu()
            if self.rank:
                st, pd, dl, g = self.conv
                sp = [ops._conv_out_size(x.shape[2 + i], self.mu_w.shape[2 + i], st[i], pd[i], dl[i]) for i in range(self.rank)]
                oshape = (x.shape[0], self.mu_w.shape[0], *sp)
            else:
                oshape = (x.shape[0], self.mu_w.shape[0])
            s_out = ops.philox_sign_shape(oshape, self.dev, seed, s, self.lid, _lib.KIND_SIGN_OUT).cpu()
            args += [s_in, s_out]
        if self.rank:
            st, pd, dl, g = self.conv
            fn = oracle.conv_flipout if self.flip else oracle.conv_reparam
            out, kl = fn(*args, stride=st, padding=pd, dilation=dl, groups=g)
        else:
            out, kl = (oracle.linear_flipout if self.flip else oracle.linear_reparam)(*args)
        return out if (self.flag or not return_kl) else (out, kl)


def oracle_twin(model, dev):
    twin = copy.deepcopy(model).cpu()

    def swap(m, src):
        for name, child in list(m._modules.items()):
            if isinstance(child, _BayesLayer):
                setattr(m, name, OracleFed(src._modules[name], dev))
            elif child is not None:
                swap(child, src._modules[name])
    swap(twin, model)
    return twin


def pin_layer_ids(model, base):
    """Give the model's Bayesian layers fixed Philox stream ids.  The ids come from a process-wide counter, so the
    draws (and with them the whole numeric instance) would depend on which tests ran before.  That matters for the
    GRADIENT checks below: ReLU / max-pool route gradients discontinuously, an activation within ~1e-7 of a tie
    routes differently on the two sides, and one re-routed element moves a weight gradient (a random-sign sum over
    ~1e5 positions) by ~1/sqrt(N) ~ 3e-3 of its size.  Pinned ids make the instance -- and its absence of ties --
    reproducible."""
    for i, m in enumerate(mod for mod in model.modules() if isinstance(mod, _BayesLayer)):
        m._layer_id = base + i


def set_sample(twin, s):
    for m in twin.modules():
        if isinstance(m, OracleFed):
            m.sample = s


def close(a, b, rtol, scale_atol, msg):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    torch.testing.assert_close(a, b, rtol=rtol, atol=scale_atol * max(float(b.abs().max()), 1e-30), msg=lambda m: f"{msg}: {m}")


# ---------------------------------------------------------------------------------- configs[0]
class SCNN(nn.Module):
    """simple_cnn_variational.py:13-89 on the drop-in layers (dropout left out: its mask is torch's RNG)."""

    def __init__(self, smooth=False):
        super().__init__()
        self.act = F.softplus if smooth else F.relu
        self.pool = F.avg_pool2d if smooth else F.max_pool2d
        kw = dict(prior_mean=0.0, prior_variance=1.0, posterior_mu_init=0.0, posterior_rho_init=-3.0)
        self.conv1 = L.Conv2dReparameterization(1, 32, 3, stride=1, prior_type="normal", **kw)
        self.conv2 = L.Conv2dReparameterization(32, 64, 3, stride=1, prior_type="normal", **kw)
        self.fc1 = L.LinearReparameterization(9216, 128, **kw)
        self.fc2 = L.LinearReparameterization(128, 10, **kw)

    def forward(self, x):
        kl_sum = 0
        x, kl = self.conv1(x); kl_sum = kl_sum + kl
        x, kl = self.conv2(self.act(x)); kl_sum = kl_sum + kl
        x = torch.flatten(self.pool(self.act(x), 2), 1)
        x, kl = self.fc1(x); kl_sum = kl_sum + kl
        x, kl = self.fc2(self.act(x)); kl_sum = kl_sum + kl
        return F.log_softmax(x, dim=1), kl_sum


def compare_grads(model, twin, smooth, rtol, scale_atol):
    names = {"mu_kernel": "mu_w", "rho_kernel": "rho_w", "mu_weight": "mu_w", "rho_weight": "rho_w", "mu_bias": "mu_b", "rho_bias": "rho_b"}
    tw = dict(twin.named_parameters())
    checked = 0
    for n, p in model.named_parameters():
        lay, leaf = n.rsplit(".", 1)
        if leaf not in names:
            continue
        want = tw[f"{lay}.{names[leaf]}"].grad
        if smooth:
            close(p.grad, want, rtol, scale_atol, f"grad {n}")
        else:   # a few re-routed ReLU / max-pool elements (module docstring): relative L2 error
            err = float((p.grad.detach().cpu().double() - want.double()).norm() / want.double().norm().clamp_min(1e-30))
            assert err <= 2e-2, f"grad {n}: relative L2 error {err}"
        checked += 1
    return checked


@pytest.mark.parametrize("smooth", [False, True])
def test_config0_scnn_forward_kl_backward(cuda_device, smooth):
    torch.manual_seed(0)
    bt.manual_seed(77)
    model = SCNN(smooth).to(cuda_device)
    pin_layer_ids(model, 1)
    twin = oracle_twin(model, cuda_device)
    x = torch.randn(128, 1, 28, 28)
    y = torch.randint(0, 10, (128,))
    with bt.path("simt"), bt.mc_sample(3):
        out, kl = model(x.to(cuda_device))
    loss = F.nll_loss(out, y.to(cuda_device)) + kl / 128        # main_bayesian_mnist.py:27-29
    loss.backward()
    set_sample(twin, 3)
    wout, wkl = twin(x)
    wloss = F.nll_loss(wout, y) + wkl / 128
    wloss.backward()
    close(out, wout, 1e-5, 1e-5, "log-probabilities")
    close(kl, wkl, 1e-5, 1e-6, "kl")
    close(loss, wloss, 1e-5, 1e-6, "loss")
    assert compare_grads(model, twin, smooth, 1e-4, 2e-5) == 16


# ---------------------------------------------------------------------------------- configs[1]
class BasicBlock(nn.Module):
    def __init__(self, in_planes, planes, stride):
        super().__init__()
        kw = dict(prior_mean=0.0, prior_variance=1.0, prior_type="normal", posterior_mu_init=0.0, posterior_rho_init=-3.0, bias=False)
        self.conv1 = L.Conv2dReparameterization(in_planes, planes, 3, stride=stride, padding=1, **kw)
        self.bn1 = nn.BatchNorm2d(planes)       # SyncBatchNorm in the reference == BatchNorm in eval mode
        self.conv2 = L.Conv2dReparameterization(planes, planes, 3, stride=1, padding=1, **kw)
        self.bn2 = nn.BatchNorm2d(planes)
        self.pad = planes // 4 if (stride != 1 or in_planes != planes) else None   # option A (resnet_variational.py:75-81)

    def forward(self, x):
        kl_sum = 0
        out, kl = self.conv1(x); kl_sum = kl_sum + kl
        out, kl = self.conv2(F.relu(self.bn1(out))); kl_sum = kl_sum + kl
        sc = x if self.pad is None else F.pad(x[:, :, ::2, ::2], (0, 0, 0, 0, self.pad, self.pad), "constant", 0)
        return F.relu(self.bn2(out) + sc), kl_sum


class ResNet20(nn.Module):
    def __init__(self):
        super().__init__()
        kw = dict(prior_mean=0.0, prior_variance=1.0, prior_type="normal", posterior_mu_init=0.0, posterior_rho_init=-3.0)
        self.conv1 = L.Conv2dReparameterization(3, 16, 3, stride=1, padding=1, bias=False, **kw)
        self.bn1 = nn.BatchNorm2d(16)
        blocks, inp = [], 16
        for planes, stride in ((16, 1), (32, 2), (64, 2)):
            for s in [stride, 1, 1]:
                blocks.append(BasicBlock(inp, planes, s)); inp = planes
        self.blocks = nn.ModuleList(blocks)
        self.linear = L.LinearReparameterization(64, 10, **kw)

    def forward(self, x):
        kl_sum = 0
        out, kl = self.conv1(x); kl_sum = kl_sum + kl
        out = F.relu(self.bn1(out))
        for b in self.blocks:
            out, kl = b(out); kl_sum = kl_sum + kl
        out = F.avg_pool2d(out, out.size()[3]).flatten(1)
        out, kl = self.linear(out)
        return out, kl_sum + kl


def test_config1_resnet20_mc_inference(cuda_device):
    torch.manual_seed(1)
    bt.manual_seed(78)
    model = ResNet20().to(cuda_device).eval()
    pin_layer_ids(model, 1)
    for m in model.modules():
        if isinstance(m, nn.BatchNorm2d):
            m.running_mean.normal_(0, 0.1); m.running_var.uniform_(0.8, 1.2)
    twin = oracle_twin(model, cuda_device).eval()
    S, B = 8, 256
    x = torch.randn(B, 3, 32, 32)
    xg = x.to(cuda_device)
    assert sum(isinstance(m, _BayesLayer) for m in model.modules()) == 20
    with torch.no_grad():
        want, wkl = [], None
        for s in range(S):                      # evaluate() loop, main_bayesian_cifar_dnn2bnn.py:542-556
            set_sample(twin, s)
            o, wkl = twin(x)
            want.append(o)
        wmean, wvar = oracle.mc_moments(torch.stack(want))
        got32 = []
        with bt.path("simt"):
            for s in range(S):
                with bt.mc_sample(s):
                    o, kl = model(xg)
                got32.append(o)
        close(torch.stack(got32), torch.stack(want), 1e-4, 1e-4, "fp32 logits of the 8 samples")
        close(kl, wkl, 1e-5, 1e-6, "kl")
        acc = bt.MCAccumulator(B, 10, cuda_device)
        for o in got32:
            acc.add(o)
        res = acc.finalize(S)
        close(res.mean, wmean, 1e-4, 1e-4, "predictive mean")
        close(res.var, wvar, 1e-3, 1e-3, "predictive variance")
        # bf16 channels-last: every layer on the tensor-core kernels (stem through the patch matrix)
        mb = copy.deepcopy(model).to(memory_format=torch.channels_last).to(torch.bfloat16)
        for a, b in zip(mb.modules(), model.modules()):
            if isinstance(a, _BayesLayer):
                a._layer_id = b._layer_id        # same streams as the fp32 model
        xb = xg.bfloat16().contiguous(memory_format=torch.channels_last)
        got16 = []
        for s in range(S):
            with bt.mc_sample(s):
                got16.append(mb(xb)[0].float())
        err = float((torch.stack(got16).cpu() - torch.stack(want)).abs().max() / torch.stack(want).abs().max())
        assert err < 6e-2, f"bf16 ResNet-20 logits rel err {err}"   # 20 bf16 layers deep, rounding compounds (one layer: 2^-8)


# ---------------------------------------------------------------------------------- configs[2]
def smoothen(model):
    """ReLU -> Softplus, MaxPool2d -> AvgPool2d in place (same layers, shapes and kernels; smooth gradients)."""
    for m in list(model.modules()):
        for name, child in list(m._modules.items()):
            if isinstance(child, nn.ReLU):
                setattr(m, name, nn.Softplus())
            elif isinstance(child, nn.MaxPool2d):
                setattr(m, name, nn.AvgPool2d(child.kernel_size, child.stride, child.padding))
    return model


@pytest.mark.parametrize("smooth", [False, True])
def test_config2_resnet18_flipout_training_step(cuda_device, smooth):
    import torchvision
    torch.manual_seed(2)
    bt.manual_seed(79)
    model = torchvision.models.resnet18(num_classes=20)
    if smooth:
        smoothen(model)
    bt.dnn_to_bnn(model, {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
                          "type": "Flipout", "moped_enable": True, "moped_delta": 0.5})
    model = model.to(cuda_device).eval()    # eval: BatchNorm uses running statistics on both sides (batch 6 is too small for batch stats parity)
    pin_layer_ids(model, 1)
    twin = oracle_twin(model, cuda_device).eval()
    assert sum(isinstance(m, _BayesLayer) for m in model.modules()) == 21
    B = 6
    x, y = torch.randn(B, 3, 64, 64), torch.randint(0, 20, (B,))
    with bt.path("simt"), bt.mc_sample(1):
        out = model(x.to(cuda_device))
        kl = bt.get_kl_loss(model)
    loss = F.cross_entropy(out, y.to(cuda_device)) + kl / B       # main_bayesian_cifar_dnn2bnn.py:405-412
    loss.backward()
    set_sample(twin, 1)
    wout = twin(x)
    wkl = sum(m.kl_loss() for m in twin.modules() if isinstance(m, OracleFed))
    wloss = F.cross_entropy(wout, y) + wkl / B
    wloss.backward()
    close(out, wout, 1e-4, 1e-4, "logits")
    close(kl, wkl, 1e-5, 1e-6, "get_kl_loss")
    assert compare_grads(model, twin, smooth, 2e-3, 2e-4) == 2 * 21 + 2   # (element-wise: 21 layers of fp32 re-association)
