"""Pin the CPU oracle (oracle/bt_oracle.py) against fixtures produced by the
reference itself (oracle/make_golden.py): outputs, KL and autograd gradients."""
import os

import numpy as np
import pytest
import torch

import oracle
from oracle.make_golden import KL_COEF
from conftest import cfg_of, golden_names, load_golden

LAYER_CASES = [n for n in golden_names() if not n.startswith(("dnn_to_bnn", "lstm_"))]


def run_oracle(name, d, dtype=torch.float32):
    c = cfg_of(d)
    prior = dict(prior_mu=c["prior_mu"], prior_sigma=c["prior_sigma"])
    if "prior_w_mu" in d:  # tensor-valued prior (utils/util.py:101-104)
        prior = None
    conv_kw = {k: c[k] for k in ("stride", "padding", "dilation", "groups", "output_padding") if k in c}
    leaves = {k: d[k].clone().to(dtype).requires_grad_(True) for k in ("x", "mu_w", "rho_w", "mu_b", "rho_b") if k in d}
    g = lambda k: leaves.get(k)
    eps_b = d.get("eps_b")
    if name.startswith("linear_rp"):
        out, kl = oracle.linear_reparam(g("x"), g("mu_w"), g("rho_w"), g("mu_b"), g("rho_b"), d["eps_w"], eps_b,
                                        dtype=dtype, **(prior or {}))
    elif name.startswith("convT") and "_rp_" in name:
        out, kl = oracle.conv_transpose_reparam(g("x"), g("mu_w"), g("rho_w"), g("mu_b"), g("rho_b"), d["eps_w"], eps_b,
                                                dtype=dtype, **conv_kw, **prior)
    elif name.startswith("convT"):
        out, kl = oracle.conv_transpose_flipout(g("x"), g("mu_w"), g("rho_w"), g("mu_b"), g("rho_b"), d["eps_w"], eps_b,
                                                d["sign_in"], d["sign_out"], dtype=dtype, **conv_kw, **prior)
    elif "_rp_" in name:
        out, kl = oracle.conv_reparam(g("x"), g("mu_w"), g("rho_w"), g("mu_b"), g("rho_b"), d["eps_w"], eps_b,
                                      dtype=dtype, **conv_kw, **prior)
    elif name.startswith("linear_fo"):
        out, kl = oracle.linear_flipout(g("x"), g("mu_w"), g("rho_w"), g("mu_b"), g("rho_b"), d["eps_w"], eps_b,
                                        d["sign_in"], d["sign_out"], dtype=dtype, **prior)
    else:
        out, kl = oracle.conv_flipout(g("x"), g("mu_w"), g("rho_w"), g("mu_b"), g("rho_b"), d["eps_w"], eps_b,
                                      d["sign_in"], d["sign_out"], dtype=dtype, **conv_kw, **prior)
    if "cfg_prior_type" in d:  # the fork's Laplace(0, 1) prior (base_variational_layer.py:74-97)
        kl = oracle.kl_laplace(g("mu_w"), oracle.softplus_ref(g("rho_w")))
        if g("mu_b") is not None:
            kl = kl + oracle.kl_laplace(g("mu_b"), oracle.softplus_ref(g("rho_b")))
    if prior is None:
        kl = oracle.kl_normal(g("mu_w"), oracle.softplus_ref(g("rho_w")), d["prior_w_mu"].to(dtype), d["prior_w_sigma"].to(dtype))
        kl = kl + oracle.kl_normal(g("mu_b"), oracle.softplus_ref(g("rho_b")), 0.0, 1.0)
    return out, kl, leaves


@pytest.mark.parametrize("name", LAYER_CASES)
def test_oracle_forward_matches_reference(name):
    d = load_golden(name)
    out, kl, _ = run_oracle(name, d)
    # same ATen ops in the same order => bit-exact on the machine that made the
    # fixtures; other CPUs / thread counts reorder the fp32 GEMM sums, so allow
    # round-off proportional to the output scale (cancelling sums land near 0).
    scale = max(float(d["out"].abs().max()), 1.0)
    torch.testing.assert_close(out.detach(), d["out"], rtol=1e-5, atol=4e-6 * scale)
    torch.testing.assert_close(kl.detach(), d["kl"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name", LAYER_CASES)
def test_oracle_backward_matches_reference(name):
    d = load_golden(name)
    out, kl, leaves = run_oracle(name, d)
    loss = (out * d["g"]).sum() + KL_COEF * kl
    keys = list(leaves)
    grads = torch.autograd.grad(loss, [leaves[k] for k in keys])
    wn = "weight" if name.startswith("linear") else "kernel"
    ref = {"x": "dx", "mu_w": f"d_mu_{wn}", "rho_w": f"d_rho_{wn}", "mu_b": "d_mu_bias", "rho_b": "d_rho_bias"}
    for k, gr in zip(keys, grads):
        torch.testing.assert_close(gr, d[ref[k]], rtol=1e-4, atol=1e-5, msg=lambda m: f"{name}:{k}: {m}")


@pytest.mark.parametrize("name", LAYER_CASES)
def test_oracle_fp64_agrees(name):
    """fp64 oracle vs the reference's fp32 result: bounds the fp32 round-off the
    CUDA parity tolerance (rtol 1e-5 + atol scaled by |out|max) has to absorb."""
    d = load_golden(name)
    out, kl, _ = run_oracle(name, d, torch.float64)
    scale = float(d["out"].abs().max())
    assert float((out.detach().float() - d["out"]).abs().max()) <= 2e-6 * max(scale, 1.0) * 8
    assert abs(float(kl) - float(d["kl"])) <= 1e-5 * abs(float(d["kl"]))


def lstm_draws(d, flip):
    def draws(tag, t):
        out = [d[f"eps_{tag}_w_{t}"], d[f"eps_{tag}_b_{t}"]]
        if flip:
            out += [d[f"sin_{tag}_{t}"], d[f"sout_{tag}_{t}"]]
        return out
    return draws


@pytest.mark.parametrize("name", ["lstm_rp_T5", "lstm_fo_T4"])
def test_oracle_lstm_matches_reference(name):
    """The unrolled cell with the per-step draws recovered from the reference run reproduces its hidden /
    cell sequences, the summed KL and the gradients."""
    d = load_golden(name)
    flip = name.startswith("lstm_fo")
    p = {k[2:]: v.clone().requires_grad_(True) for k, v in d.items() if k.startswith("p_")}
    x = d["x"].clone().requires_grad_(True)
    # alias_eps: the reference's backward reads the LAST step's eps for every step (see oracle.lstm_unrolled)
    hs, cs, kl = oracle.lstm_unrolled(x, p, lstm_draws(d, flip), flipout=flip, alias_eps=True)
    torch.testing.assert_close(hs.detach(), d["hs"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(cs.detach(), d["cs"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(kl.detach(), d["kl"], rtol=1e-6, atol=1e-6)
    loss = (hs * d["g"]).sum() + (cs * d["gc"]).sum() + KL_COEF * kl
    keys = sorted(p)
    grads = torch.autograd.grad(loss, [x] + [p[k] for k in keys])
    torch.testing.assert_close(grads[0], d["dx"], rtol=1e-4, atol=1e-5)
    for k, gr in zip(keys, grads[1:]):
        torch.testing.assert_close(gr, d["d_" + k], rtol=1e-4, atol=1e-5, msg=lambda m: f"{name}:{k}: {m}")
    # per-step draws kept apart (what the CUDA path differentiates): same outputs, dx and mu gradients; the
    # rho gradients are the ones that change
    p2 = {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
    x2 = d["x"].clone().requires_grad_(True)
    hs2, cs2, kl2 = oracle.lstm_unrolled(x2, p2, lstm_draws(d, flip), flipout=flip)
    assert torch.equal(hs2, hs) and torch.equal(cs2, cs)
    g2 = torch.autograd.grad((hs2 * d["g"]).sum() + (cs2 * d["gc"]).sum() + KL_COEF * kl2, [x2] + [p2[k] for k in keys])
    torch.testing.assert_close(g2[0], d["dx"], rtol=1e-4, atol=1e-5)
    for k, gr in zip(keys, g2[1:]):
        if ".mu_" in k:
            torch.testing.assert_close(gr, d["d_" + k], rtol=1e-4, atol=1e-5)
    assert not torch.allclose(g2[1 + keys.index("hh.rho_weight")], d["d_hh.rho_weight"], rtol=1e-2, atol=1e-4)


def test_moped_rho_and_closed_form_kl():
    d = load_golden("linear_rp_moped_tensorprior")
    torch.testing.assert_close(oracle.get_rho(d["moped_w"], float(d["moped_delta"])), d["moped_rho"], rtol=1e-6, atol=1e-7)
    # closed forms: q == p -> 0 ;  sigma equal, mean shift dm -> dm^2 / (2 sigma_p^2)
    mu = torch.full((64,), 0.25)
    sig = torch.full((64,), 0.7)
    assert float(oracle.kl_normal(mu, sig, 0.25, 0.7)) == pytest.approx(0.0, abs=1e-7)
    assert float(oracle.kl_normal(mu, torch.ones(64), -0.75, 1.0)) == pytest.approx(0.5, rel=1e-6)


def test_dnn_to_bnn_fixture_self_consistent():
    """The small converted model's KL equals the sum of the oracle's per-layer KLs."""
    d = load_golden("dnn_to_bnn_small")
    kl = 0.0
    for i, wn in ((0, "kernel"), (2, "kernel"), (4, "weight")):
        kl = kl + oracle.kl_normal_from_rho(d[f"bnn_{i}.mu_{wn}"], d[f"bnn_{i}.rho_{wn}"],
                                            d.get(f"bnn_{i}.mu_bias"), d.get(f"bnn_{i}.rho_bias"))
    torch.testing.assert_close(kl, d["kl"], rtol=1e-6, atol=1e-6)
    # MOPED copied the DNN weights into mu (dnn_to_bnn.py:95-101)
    torch.testing.assert_close(d["bnn_0.mu_kernel"], d["dnn_0.weight"])
    torch.testing.assert_close(d["bnn_0.rho_kernel"], oracle.get_rho(d["dnn_0.weight"], 0.5))


def test_mc_moments_and_uncertainty():
    torch.manual_seed(0)
    logits = torch.randn(6, 5, 7)
    mean, var = oracle.mc_moments(logits)
    p = torch.softmax(logits.double(), -1)
    torch.testing.assert_close(mean, p.mean(0))
    torch.testing.assert_close(var, p.var(0, unbiased=False))
    ent = -(p.mean(0) * torch.log(p.mean(0) + 1e-15)).sum(-1)
    torch.testing.assert_close(oracle.predictive_entropy(p), ent)
    assert (oracle.mutual_information(p) > -1e-9).all()


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="needs the reference checkout (build container only)")
def test_fixtures_regenerate_from_the_committed_recipe():
    """`python oracle/make_golden.py --check`: running the reference again reproduces every committed fixture
    (each case seeds its own layer construction and input, so fixtures do not depend on the case order)."""
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bt_make_golden", os.path.join(root, "oracle", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.check() == []
