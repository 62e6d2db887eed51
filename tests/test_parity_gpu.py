"""GPU parity tests proper: the CUDA path (through the Python layers -> C ABI) against the
golden fixtures produced by the reference itself, and against the CPU oracle on seeded
inputs.  fp32 deterministic mode: rtol 1e-5 (+ atol 1e-5 * max|out| for entries near 0),
the tolerance BASELINE.json's north_star states."""
import numpy as np
import pytest
import torch

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import layers as L
from bayesian_torch_b200 import ops
from conftest import cfg_of, golden_names, load_golden
from oracle.make_golden import KL_COEF

pytestmark = pytest.mark.gpu
LAYER_CASES = [n for n in golden_names() if not n.startswith(("dnn_to_bnn", "lstm_"))]
RTOL = 1e-5


def close(a, b, rtol=RTOL, scale_atol=1e-5, msg=""):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    atol = scale_atol * max(float(b.abs().max()), 1e-30)
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol, msg=lambda m: f"{msg}: {m}")


def build_layer(name, d, dev):
    """Instantiate the drop-in layer of a fixture and load the fixture's parameters / eps / signs."""
    c = cfg_of(d)
    w = d["mu_w"]
    bias = "mu_b" in d
    pm, ps = c["prior_mu"], c["prior_sigma"]
    conv_kw = {k: c[k] for k in ("stride", "padding", "dilation", "groups", "output_padding") if k in c}
    if name.startswith("convT"):   # kernel [c_in, c_out / groups, *k]
        cls = getattr(L, f"ConvTranspose{name[5]}d" + ("Reparameterization" if "_rp_" in name else "Flipout"))
        layer = cls(w.shape[0], w.shape[1] * conv_kw["groups"], tuple(w.shape[2:]), bias=bias, **conv_kw)
    elif name.startswith("conv1d_fo") or name.startswith("conv3d_fo"):
        cls = L.Conv1dFlipout if name.startswith("conv1d") else L.Conv3dFlipout
        layer = cls(w.shape[1] * conv_kw["groups"], w.shape[0], tuple(w.shape[2:]), bias=bias, **conv_kw)
    elif name.startswith("linear_rp"):
        layer = L.LinearReparameterization(w.shape[1], w.shape[0], prior_mean=pm, prior_variance=ps, bias=bias,
                                           prior_type="laplace" if "cfg_prior_type" in d else "normal")
    elif name.startswith("linear_fo"):
        layer = L.LinearFlipout(w.shape[1], w.shape[0], bias=bias)
    elif name.startswith("conv1d_rp"):
        layer = L.Conv1dReparameterization(w.shape[1] * conv_kw["groups"], w.shape[0], w.shape[2], bias=bias, **conv_kw)
    elif name.startswith("conv3d_rp"):
        layer = L.Conv3dReparameterization(w.shape[1], w.shape[0], tuple(w.shape[2:]), pm, ps, 0, -3.0, bias=bias, **conv_kw)
    elif name.startswith("conv2d_rp"):
        layer = L.Conv2dReparameterization(w.shape[1] * conv_kw["groups"], w.shape[0], tuple(w.shape[2:]), prior_mean=pm,
                                           prior_variance=ps, prior_type="laplace" if "cfg_prior_type" in d else "normal",
                                           bias=bias, **conv_kw)
    else:
        layer = L.Conv2dFlipout(w.shape[1], w.shape[0], tuple(w.shape[2:]), bias=bias, **conv_kw)
    layer = layer.to(dev)
    wn = layer._wname
    with torch.no_grad():
        getattr(layer, "mu_" + wn).copy_(d["mu_w"]); getattr(layer, "rho_" + wn).copy_(d["rho_w"])
        getattr(layer, "eps_" + wn).copy_(d["eps_w"])
        if bias:
            layer.mu_bias.copy_(d["mu_b"]); layer.rho_bias.copy_(d["rho_b"]); layer.eps_bias.copy_(d["eps_b"])
    if "sign_in" in d:
        layer.sign_input, layer.sign_output = d["sign_in"].to(dev), d["sign_out"].to(dev)
    if "prior_w_mu" in d:
        layer.prior_weight_mu = d["prior_w_mu"].to(dev)
        layer.prior_weight_sigma = d["prior_w_sigma"].to(dev)
    layer.sample_mode = "buffer"   # deterministic test mode: the reference's own eps / signs
    layer.path = "simt"            # fp32 parity path
    return layer


@pytest.mark.parametrize("name", LAYER_CASES)
def test_forward_and_kl_match_reference_fixture(name, cuda_device):
    d = load_golden(name)
    layer = build_layer(name, d, cuda_device)
    out, kl = layer(d["x"].to(cuda_device))
    assert out.shape == d["out"].shape
    close(out, d["out"], msg=f"{name} out")
    close(kl, d["kl"], rtol=RTOL, scale_atol=1e-6, msg=f"{name} kl")
    close(layer.kl_loss(), d["kl"], rtol=RTOL, scale_atol=1e-6, msg=f"{name} kl_loss")
    layer.dnn_to_bnn_flag = True      # converted models get `out` only (linear_variational.py:161-162)
    only = layer(d["x"].to(cuda_device))
    assert torch.is_tensor(only) and torch.equal(only, out)


@pytest.mark.parametrize("name", LAYER_CASES)
def test_backward_matches_reference_autograd(name, cuda_device):
    d = load_golden(name)
    layer = build_layer(name, d, cuda_device)
    x = d["x"].to(cuda_device).requires_grad_(True)
    out, kl = layer(x)
    loss = (out * d["g"].to(cuda_device)).sum() + KL_COEF * kl
    loss.backward()
    wn = layer._wname
    close(x.grad, d["dx"], rtol=1e-4, msg=f"{name} dx")
    for p in ("mu_" + wn, "rho_" + wn, "mu_bias", "rho_bias"):
        if getattr(layer, p, None) is not None:
            close(getattr(layer, p).grad, d["d_" + p], rtol=1e-4, scale_atol=2e-5, msg=f"{name} d_{p}")


@pytest.mark.parametrize("name", ["conv2d_rp_3x3_p1", "conv2d_rp_64ch", "conv2d_fo_3x3", "conv2d_rp_groups2"])
def test_channels_last_and_bf16_inputs(name, cuda_device):
    """Layout is consumed through strides (no copy); bf16 activations round only x and out."""
    d = load_golden(name)
    layer = build_layer(name, d, cuda_device)
    x = d["x"].to(cuda_device)
    ref, _ = layer(x)
    out_cl, _ = layer(x.contiguous(memory_format=torch.channels_last))
    assert out_cl.is_contiguous(memory_format=torch.channels_last)
    close(out_cl, ref, rtol=1e-6, scale_atol=1e-6, msg="channels_last")
    if "sign_in" not in d:
        xb = x.bfloat16()
        out_b, _ = layer(xb)
        assert out_b.dtype == torch.bfloat16
        want, _ = layer(xb.float())
        close(out_b, want, rtol=1e-2, scale_atol=1e-2, msg="bf16 activations")


def test_seeded_larger_shapes_against_oracle(cuda_device):
    """Sizes past the fixtures (multiple CTA tiles, ragged edges) checked against the CPU oracle."""
    g = torch.Generator().manual_seed(7)
    R = lambda *s: torch.randn(*s, generator=g)
    # linear 300 x 257 -> 129 (odd everything)
    lin = L.LinearReparameterization(257, 129).to(cuda_device)
    lin.sample_mode, lin.path = "buffer", "simt"
    lin.eps_weight.copy_(R(129, 257)); lin.eps_bias.copy_(R(129))
    x = R(300, 257)
    out, kl = lin(x.to(cuda_device))
    want, wkl = oracle.linear_reparam(x, lin.mu_weight, lin.rho_weight, lin.mu_bias, lin.rho_bias, lin.eps_weight,
                                      lin.eps_bias, dtype=torch.float64)
    close(out, want, rtol=RTOL, msg="linear 300x257x129"); close(kl, wkl, scale_atol=1e-6)
    # conv 5x5 stride 2 pad 2, 37 -> 70 channels, batch 5, 33x29 images
    conv = L.Conv2dReparameterization(37, 70, 5, stride=2, padding=2, prior_type="normal").to(cuda_device)
    conv.sample_mode, conv.path = "buffer", "simt"
    conv.eps_kernel.copy_(R(70, 37, 5, 5)); conv.eps_bias.copy_(R(70))
    x = R(5, 37, 33, 29)
    out, kl = conv(x.to(cuda_device))
    want, wkl = oracle.conv_reparam(x, conv.mu_kernel, conv.rho_kernel, conv.mu_bias, conv.rho_bias, conv.eps_kernel,
                                    conv.eps_bias, stride=2, padding=2, dtype=torch.float64)
    close(out, want, rtol=RTOL, msg="conv 5x5 s2"); close(kl, wkl, scale_atol=1e-6)


def test_empty_batch_and_error_behaviour(cuda_device):
    lin = L.LinearReparameterization(8, 4).to(cuda_device)
    out, kl = lin(torch.empty(0, 8, device=cuda_device))
    assert out.shape == (0, 4) and torch.isfinite(kl)
    with pytest.raises(RuntimeError, match="expects"):
        lin(torch.randn(2, 9, device=cuda_device))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lin(torch.randn(2, 8))
    conv = L.Conv2dReparameterization(4, 4, 5).to(cuda_device)
    with pytest.raises(RuntimeError, match="not positive"):
        conv(torch.randn(1, 4, 3, 3, device=cuda_device))
    with pytest.raises(ValueError, match="Unknown prior_type"):
        lin.kl_div(lin.mu_weight, torch.ones_like(lin.mu_weight), 0.0, 1.0, "cauchy")


def test_kl_closed_forms_and_multi_tensor(cuda_device):
    """Known answers and the one-launch get_kl_loss against the oracle."""
    base = L.LinearReparameterization(4, 4).to(cuda_device)
    mu = torch.full((1000,), 0.25, device=cuda_device)
    assert float(base.kl_div(mu, torch.full_like(mu, 0.7), 0.25, 0.7)) == pytest.approx(0.0, abs=1e-6)
    assert float(base.kl_div(mu, torch.ones_like(mu), -0.75, 1.0)) == pytest.approx(0.5, rel=1e-6)
    # ragged sizes: 1 element, < one vector, one chunk + tail, unaligned view, several chunks
    g = torch.Generator().manual_seed(3)
    sizes = [1, 3, 8191, 8192, 8193, 70001, 300000]
    mus = [torch.randn(n, generator=g) * 0.1 for n in sizes]
    rhos = [torch.randn(n, generator=g) * 0.5 - 3.0 for n in sizes]
    big = torch.randn(5001, generator=g)
    mus.append(big[1:] * 0.1); rhos.append(big[1:] * 0.2 - 2.0)     # 4-byte-aligned only
    want = sum(oracle.kl_normal_from_rho(m, r, prior_mu=0.1, prior_sigma=0.8, dtype=torch.float64) for m, r in zip(mus, rhos))
    dm = [m.to(cuda_device) for m in mus[:-1]] + [big.to(cuda_device)[1:] * 0.1]
    dr = [r.to(cuda_device) for r in rhos[:-1]] + [big.to(cuda_device)[1:] * 0.2 - 2.0]
    got = ops.kl_normal_multi(dm, dr, 0.1, 0.8)
    assert float(got) == pytest.approx(float(want), rel=1e-6)
    # gradients of the fused KL against autograd through the oracle
    ms = [m.clone().double().requires_grad_(True) for m in mus[:5]]
    rs = [r.clone().double().requires_grad_(True) for r in rhos[:5]]
    tot = sum(oracle.kl_normal(m, oracle.softplus_ref(r), 0.1, 0.8) for m, r in zip(ms, rs))
    (tot * 1.7).backward()
    gm = [m.to(cuda_device).requires_grad_(True) for m in mus[:5]]
    gr = [r.to(cuda_device).requires_grad_(True) for r in rhos[:5]]
    (ops.kl_normal_multi(gm, gr, 0.1, 0.8) * 1.7).backward()
    for a, b in zip(gm + gr, ms + rs):
        close(a.grad, b.grad, rtol=1e-5, scale_atol=1e-6, msg="kl grad")


def test_dnn_to_bnn_small_model_matches_reference(cuda_device):
    d = load_golden("dnn_to_bnn_small")
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(8, 8, 3, stride=2, bias=False),
                              torch.nn.Flatten(), torch.nn.Linear(8 * 3 * 3, 10))
    params = {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
              "type": "Reparameterization", "moped_enable": False, "moped_delta": 0.5}
    bt.dnn_to_bnn(net, params)
    net.load_state_dict({k[4:]: v for k, v in d.items() if k.startswith("bnn_")}, strict=True)
    net = net.to(cuda_device)
    for i, m in enumerate(net):
        if hasattr(m, "_wname"):
            m.sample_mode, m.path = "buffer", "simt"
            getattr(m, "eps_" + m._wname).copy_(d[f"eps_{i}_w"])
            if m.eps_bias is not None:
                m.eps_bias.copy_(d[f"eps_{i}_b"])
    y = net(d["x"].to(cuda_device))          # dnn_to_bnn_flag -> plain tensor
    close(y, d["out"], msg="dnn_to_bnn forward")
    close(bt.get_kl_loss(net), d["kl"], scale_atol=1e-6, msg="get_kl_loss")
    # MOPED on device uses the bt_moped_rho kernel
    w = d["dnn_0.weight"].to(cuda_device)
    close(bt.utils.get_rho(w, 0.5), oracle.get_rho(d["dnn_0.weight"], 0.5), rtol=1e-5, scale_atol=1e-6, msg="moped rho")


def test_laplace_prior_kl_div_and_get_kl_loss(cuda_device):
    """The fork's 'laplace' prior (base_variational_layer.py:74-97): kl_div called directly, kl_loss(),
    forward(return_kl=True) and get_kl_loss over a model mixing both priors; gradients against autograd
    through the oracle restatement."""
    torch.manual_seed(31)
    lin = L.LinearReparameterization(300, 70, prior_type="laplace", posterior_rho_init=-1.5).to(cuda_device)
    conv = L.Conv2dReparameterization(8, 16, 3, prior_type="normal").to(cuda_device)
    with torch.no_grad():
        lin.mu_weight.mul_(8.0)      # spread mu / sigma over both tails of the erf term
    mu, rho = lin.mu_weight.detach().cpu(), lin.rho_weight.detach().cpu()
    want = oracle.kl_laplace(mu.double(), oracle.softplus_ref(rho.double()))
    got = lin.kl_div(lin.mu_weight, torch.log1p(torch.exp(lin.rho_weight)), 0.3, 2.0, "laplace")   # prior args are ignored
    close(got, want.float(), rtol=RTOL, scale_atol=1e-6, msg="kl_div laplace")
    want_l = want + oracle.kl_laplace(lin.mu_bias.detach().cpu().double(), oracle.softplus_ref(lin.rho_bias.detach().cpu().double()))
    close(lin.kl_loss(), want_l.float(), rtol=RTOL, scale_atol=1e-6, msg="kl_loss laplace")
    lin.sample_mode, lin.path = "torch", "simt"
    out, kl = lin(torch.randn(5, 300, device=cuda_device))
    close(kl, want_l.float(), rtol=RTOL, scale_atol=1e-6, msg="forward kl laplace")
    net = torch.nn.Sequential(conv, lin)
    want_n = oracle.kl_normal_from_rho(conv.mu_kernel.detach().cpu(), conv.rho_kernel.detach().cpu(),
                                       conv.mu_bias.detach().cpu(), conv.rho_bias.detach().cpu())
    total = bt.get_kl_loss(net)
    close(total, (want_l + want_n).float(), rtol=RTOL, scale_atol=1e-6, msg="get_kl_loss mixed priors")
    total.backward()
    m = lin.mu_weight.detach().cpu().double().requires_grad_(True)
    r = lin.rho_weight.detach().cpu().double().requires_grad_(True)
    oracle.kl_laplace(m, oracle.softplus_ref(r)).backward()
    close(lin.mu_weight.grad, m.grad.float(), rtol=1e-4, scale_atol=1e-5, msg="d kl / d mu (laplace)")
    close(lin.rho_weight.grad, r.grad.float(), rtol=1e-4, scale_atol=1e-5, msg="d kl / d rho (laplace)")
    with pytest.raises(ValueError, match="Unknown prior_type"):
        lin.kl_div(lin.mu_weight, lin.mu_weight.abs() + 1, 0.0, 1.0, "cauchy")


@pytest.mark.parametrize("name", ["lstm_rp_T5", "lstm_fo_T4"])
def test_lstm_matches_reference_fixture(name, cuda_device):
    """LSTMReparameterization / LSTMFlipout against the reference run: the per-step draws of the fixture are
    loaded into the sub-layers' buffers by a pre-hook before every ih / hh call."""
    d = load_golden(name)
    flip = name.startswith("lstm_fo")
    lstm = (L.LSTMFlipout if flip else L.LSTMReparameterization)(6, 5).to(cuda_device)
    with torch.no_grad():
        for n, p in lstm.named_parameters():
            p.copy_(d["p_" + n])
    step = {"ih": 0, "hh": 0}

    def load(tag):
        def hook(mod, inp):
            t = step[tag]; step[tag] += 1
            mod.eps_weight.copy_(d[f"eps_{tag}_w_{t}"]); mod.eps_bias.copy_(d[f"eps_{tag}_b_{t}"])
            if flip:
                mod.sign_input, mod.sign_output = d[f"sin_{tag}_{t}"].to(cuda_device), d[f"sout_{tag}_{t}"].to(cuda_device)
        return hook
    for tag in ("ih", "hh"):
        sub = getattr(lstm, tag)
        sub.sample_mode, sub.path = "buffer", "simt"
        sub.register_forward_pre_hook(load(tag))
    x = d["x"].to(cuda_device).requires_grad_(True)
    hs, (hs2, cs), kl = lstm(x)
    assert hs2 is hs and hs.shape == d["hs"].shape
    close(hs, d["hs"], msg=f"{name} hidden"); close(cs, d["cs"], msg=f"{name} cell")
    close(kl, d["kl"], rtol=RTOL, scale_atol=1e-6, msg=f"{name} kl")
    ((hs * d["g"].to(cuda_device)).sum() + (cs * d["gc"].to(cuda_device)).sum() + KL_COEF * kl).backward()
    close(x.grad, d["dx"], rtol=1e-4, msg=f"{name} dx")
    # rho gradients: the reference's autograd reads the last step's eps for every step (oracle.lstm_unrolled,
    # alias_eps); this path differentiates each step's own draw -> checked against the oracle without aliasing
    po = {k[2:]: v.clone().requires_grad_(True) for k, v in d.items() if k.startswith("p_")}
    from test_oracle_golden import lstm_draws
    ohs, ocs, okl = oracle.lstm_unrolled(d["x"], po, lstm_draws(d, flip), flipout=flip)
    names = sorted(po)
    og = dict(zip(names, torch.autograd.grad((ohs * d["g"]).sum() + (ocs * d["gc"]).sum() + KL_COEF * okl, [po[k] for k in names])))
    for n, p in lstm.named_parameters():
        want = d["d_" + n] if ".mu_" in n else og[n]
        close(p.grad, want, rtol=1e-4, scale_atol=2e-5, msg=f"{name} d_{n}")
    lstm.dnn_to_bnn_flag = True
    step.update(ih=0, hh=0)
    assert len(lstm(d["x"].to(cuda_device))) == 2      # (hidden_seq, (hidden_seq, c_ts)) without kl
    assert float(bt.get_kl_loss(lstm)) == pytest.approx(2 * float(d["kl"]) / d["x"].shape[1], rel=1e-5)  # parent + children (reference double count)


@pytest.mark.parametrize("flip", [False, True])
def test_lstm_fused_step_equals_composed_step(flip, cuda_device):
    """The one-kernel time step (bt_lstm_step_fwd + its backward) against the same step run as two Bayesian linear
    launches and torch gate arithmetic, in-kernel Philox draws, ragged sizes, given initial state: outputs and every
    gradient at rtol 1e-5 / 1e-4."""
    torch.manual_seed(3)
    bt.manual_seed(41)
    lstm = (L.LSTMFlipout if flip else L.LSTMReparameterization)(19, 23).to(cuda_device)
    for sub in (lstm.ih, lstm.hh):
        sub.path = "simt"
    x = torch.randn(70, 3, 19, device=cuda_device)
    h0, c0 = torch.randn(70, 23, device=cuda_device), torch.randn(70, 23, device=cuda_device)
    g1, g2 = torch.randn(70, 3, 23, device=cuda_device), torch.randn(70, 3, 23, device=cuda_device)
    res = []
    for fused in (True, False):
        lstm.fused_step = fused
        for p in lstm.parameters():
            p.grad = None
        xi, hi, ci = (t.clone().requires_grad_(True) for t in (x, h0, c0))
        n0 = ops.PROFILE.launches
        with bt.mc_sample(5):
            hs, (_, cs), kl = lstm(xi, (hi, ci))
        fwd_launches = ops.PROFILE.launches - n0
        ((hs * g1).sum() + (cs * g2).sum() + 0.3 * kl).backward()
        res.append((hs, cs, kl, xi.grad, hi.grad, ci.grad, [p.grad.clone() for p in lstm.parameters()], fwd_launches))
    a, b = res
    assert a[7] == 3 + 2 and b[7] == 2 * 3 + 2          # one kernel per step (+ the two KL launches) against two per step
    for i in range(3):
        close(a[i], b[i], rtol=RTOL, scale_atol=1e-6, msg=f"lstm fused vs composed output {i}")
    for i in range(3, 6):
        close(a[i], b[i], rtol=1e-4, scale_atol=1e-5, msg=f"lstm fused vs composed input grad {i}")
    for ga, gb in zip(a[6], b[6]):
        close(ga, gb, rtol=1e-4, scale_atol=1e-5, msg="lstm fused vs composed parameter grad")
