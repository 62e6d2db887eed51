"""Model-level parity for BASELINE.json configs[3] -- the configuration bench.py times -- and configs[4].

configs[3]: torchvision ResNet-50 -> dnn_to_bnn(Reparameterization, MOPED) -> bench.build_model() exactly (bf16,
channels-last, fold_batchnorm, maxpool kernel) -> ONE mc_batched(32, shared_batch=128) pass (sample pairs on the
stem, CTA-pair streaming, residual epilogues) -> bt_mc_accumulate / finalize.  The CPU twin is the SAME model before
the fold with every Bayesian layer replaced by the oracle restatement (fp32, unfused BatchNorm, fp32 input) fed the
exported Philox streams of the same (seed, sample, layer) keys: the S sequential forwards -> stack -> softmax -> mean
of examples/main_bayesian_cifar_dnn2bnn.py:527-560.  The network is per-image independent in eval mode, so the twin
evaluates a subset of the 128 images for all 32 samples.

Documented bf16 bounds (bf16 operands W = bf16(bf16(sigma) bf16(eps) + bf16(mu)), bf16
activations between 54 layers, fp32 accumulation) against the fp32 reference arithmetic:
    per-sample logits      max|err| <= LOGIT_TOL  * max|logit|
    predictive mean        max|err| <= MEAN_TOL            (probabilities)
    predictive variance    max|err| <= VAR_TOL
    entropy / MI           max|err| <= ENT_TOL nats
The measured values are printed (pytest -s) and recorded in DESIGN.md.
"""
import os
import sys

import pytest
import torch

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import ops
from test_configs_gpu import oracle_twin, set_sample

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

pytestmark = pytest.mark.gpu

LOGIT_TOL, MEAN_TOL, VAR_TOL, ENT_TOL = 6e-2, 4e-2, 2e-2, 8e-2


@pytest.fixture(scope="module")
def bench_model(cuda_device):
    holder = {}

    def prepare(model):   # fp32 eval-mode model on the device, before the bf16 conversion and the fold
        g = torch.Generator(device="cpu").manual_seed(11)
        for m in model.modules():
            if isinstance(m, torch.nn.BatchNorm2d):   # non-trivial running statistics: the fold must carry them
                m.running_mean.copy_(torch.randn(m.running_mean.shape, generator=g) * 0.05)
                m.running_var.copy_(torch.rand(m.running_var.shape, generator=g) * 0.4 + 0.8)
                m.bias.data.copy_(torch.randn(m.bias.shape, generator=g) * 0.05)
        holder["twin"] = oracle_twin(model, cuda_device).eval()
    model = bench.build_model(cuda_device, prepare=prepare)
    assert model.folded["bn"] == 53 and model.folded["residual"] == 16 and model.pools_rewritten == 1
    return model, holder["twin"]


def twin_logits(twin, x, samples):
    outs = []
    with torch.no_grad():
        for s in samples:
            set_sample(twin, s)
            outs.append(twin(x))
    return torch.stack(outs)


def report(tag, got, want, scale=None):
    err = (got.double() - want.double()).abs()
    sc = float(want.abs().max()) if scale is None else scale
    print(f"[config3] {tag}: max|err| {float(err.max()):.3e} mean|err| {float(err.mean()):.3e} (scale {sc:.3e}) "
          f"rel max {float(err.max()) / sc:.3e} rel mean {float(err.mean()) / sc:.3e}")
    return float(err.max()) / sc


def check_moments(res, want_logits, rows):
    probs = torch.softmax(want_logits.double(), dim=-1)
    wmean, wvar = probs.mean(0), probs.var(0, unbiased=False)
    went, wmi = oracle.predictive_entropy(probs), oracle.mutual_information(probs)
    assert report("predictive mean", res.mean[rows].cpu(), wmean, 1.0) <= MEAN_TOL
    assert report("predictive variance", res.var[rows].cpu(), wvar, 1.0) <= VAR_TOL
    assert report("entropy", res.entropy[rows].cpu(), went, 1.0) <= ENT_TOL
    assert report("mutual information", res.mutual_info[rows].cpu(), wmi, 1.0) <= ENT_TOL
    top_g, top_w = res.mean[rows].cpu().argmax(-1), wmean.argmax(-1)
    gap = wmean.sort(-1, descending=True).values
    decided = (gap[:, 0] - gap[:, 1]) > 2 * MEAN_TOL      # top-1 must agree wherever the margin exceeds the bound
    assert bool((top_g[decided] == top_w[decided]).all())


def test_config3_resnet50_mc_inference_bench_configuration(cuda_device, bench_model):
    model, twin = bench_model
    B, S, SUB = 128, 32, 4
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, 3, 224, 224, generator=g)
    xb = x.to(cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    bt.manual_seed(1234)
    with torch.no_grad():
        with ops.mc_batched(S, 0, shared_batch=B):
            logits = model(xb)
        res = bt.mc_predict(model, xb, S, batch_samples=S)     # the call bench.py times
    assert logits.shape == (S * B, 1000)
    got = logits.float().view(S, B, 1000)[:, :SUB].cpu()
    want = twin_logits(twin, x[:SUB], range(S))
    assert float(want.abs().max()) > 1.0 and float(torch.softmax(want, -1).max()) < 0.999, "degenerate test model"
    assert report("logits, 32 samples x 4 of 128 images @224", got, want) <= LOGIT_TOL
    check_moments(res, want, slice(0, SUB))
    # a second group of images from the far end of the batch, a few samples: tile / sample indexing at full size
    got2 = logits.float().view(S, B, 1000)[[0, 13, 31]][:, B - 2:].cpu()
    want2 = twin_logits(twin, x[B - 2:], [0, 13, 31])
    assert report("logits, samples 0/13/31 x last 2 images", got2, want2) <= LOGIT_TOL


@pytest.mark.parametrize("B,hw,S", [(4, 96, 4), (1, 224, 2), (3, 160, 5)])
def test_config3_resnet50_small_and_ragged(cuda_device, bench_model, B, hw, S):
    model, twin = bench_model
    g = torch.Generator().manual_seed(B * 1000 + hw)
    x = torch.randn(B, 3, hw, hw, generator=g)
    xb = x.to(cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    bt.manual_seed(4321)
    with torch.no_grad():
        with ops.mc_batched(S, 2, shared_batch=B):
            logits = model(xb)
        res = bt.mc_predict(model, xb, S, batch_samples=S, first_sample=2)
        loop = []
        for s in range(S):          # sample-at-a-time loop (the reference's driver shape) == the folded pass
            with ops.mc_sample(2 + s):
                loop.append(model(xb))
    got = logits.float().view(S, B, 1000).cpu()
    torch.testing.assert_close(torch.stack(loop).float().cpu(), got, rtol=0, atol=0)
    want = twin_logits(twin, x, range(2, 2 + S))
    assert report(f"logits B={B} {hw}x{hw} S={S}", got, want) <= LOGIT_TOL
    check_moments(res, want, slice(0, B))


# ---------------------------------------------------------------------------------- configs[4]
def _linear_case(N, rows, dtype, path, dev):
    from bayesian_torch_b200 import _lib
    from bayesian_torch_b200 import layers as L
    torch.manual_seed(N)
    bt.manual_seed(2024)
    layer = L.LinearReparameterization(N, N).to(dev)
    layer.path = path
    x = torch.randn(rows, N, device=dev).to(dtype)
    with torch.no_grad(), ops.mc_sample(3):
        out, kl = layer(x)
    eps_w = ops.philox_eps_weight(layer.mu_weight.shape, 2024, 3, layer._layer_id, dev).cpu()
    eps_b = ops.philox_normal(N, 2024, 3, layer._layer_id, _lib.KIND_EPS_B, dev).cpu()
    return layer, x, out, kl, eps_w, eps_b


def test_config4_linear_8192_bf16_at_size(cuda_device):
    """LinearReparameterization(8192, 8192), batch 4096, bf16 tensor-core path, in-kernel Philox: outputs of the
    first / last 128 rows against (a) the oracle on the operands the kernel multiplies (bound 2^-8 max|out|: fp32
    accumulation order + one bf16 rounding of the output) and (b) the reference's fp32 arithmetic on the unrounded
    parameters (linear_variational.py:160-181; bound 2^-7 max|out|: adds the bf16 roundings of W); KL rtol 1e-5."""
    N, ROWS = 8192, 4096
    layer, x, out, kl, eps_w, eps_b = _linear_case(N, ROWS, torch.bfloat16, "tc", cuda_device)
    mu, rho = layer.mu_weight.detach().cpu(), layer.rho_weight.detach().cpu()
    mu_b, rho_b = layer.mu_bias.detach().cpu(), layer.rho_bias.detach().cpu()
    rows = torch.cat([torch.arange(0, 128), torch.arange(ROWS - 128, ROWS)])
    xs = x[rows.to(cuda_device)].float().cpu().double()
    got = out[rows.to(cuda_device)].float().cpu().double()
    r = lambda t: t.to(torch.bfloat16).double()
    sig = torch.log1p(torch.exp(rho))
    w_kernel = r((r(sig) * r(eps_w) + r(mu)).float())
    b = (mu_b + torch.log1p(torch.exp(rho_b)) * eps_b).double()
    want_same = xs @ w_kernel.t() + b
    w_ref = (mu + sig * eps_w).double()
    want_ref = xs @ w_ref.t() + b
    e_same = report("Linear 8192 bf16 vs oracle on rounded operands", got, want_same)
    e_ref = report("Linear 8192 bf16 vs fp32 reference arithmetic", got, want_ref)
    assert e_same <= 2.0 ** -8 and e_ref <= 2.0 ** -7
    wkl = oracle.kl_normal_from_rho(mu, rho, mu_b, rho_b)
    torch.testing.assert_close(kl.cpu(), wkl.float(), rtol=1e-5, atol=1e-6)


def test_config4_linear_2048_fp32_at_size(cuda_device):
    """fp32 leg of the sweep: LinearReparameterization(2048, 2048), batch 4096, fp32 parity kernels, rtol 1e-5."""
    N, ROWS = 2048, 4096
    layer, x, out, kl, eps_w, eps_b = _linear_case(N, ROWS, torch.float32, "simt", cuda_device)
    want, wkl = oracle.linear_reparam(x.cpu(), layer.mu_weight.detach().cpu(), layer.rho_weight.detach().cpu(),
                                      layer.mu_bias.detach().cpu(), layer.rho_bias.detach().cpu(), eps_w, eps_b)
    torch.testing.assert_close(out.cpu(), want.detach(), rtol=1e-5, atol=1e-5 * float(want.abs().max()))
    torch.testing.assert_close(kl.cpu(), wkl.detach().float(), rtol=1e-5, atol=1e-6)
