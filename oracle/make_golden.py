"""Generate tests/golden/*.npz by RUNNING THE REFERENCE ITSELF on CPU.

Run in the build container only (needs /root/reference, read-only):
    python oracle/make_golden.py
The GPU box has no /root/reference; tests read the committed .npz files.

Harness-side shims, no edits to the reference:
  * `termcolor` is imported by conv_variational.py:55 but not installed -> stub;
  * the fork made `prior_type` a required kl_div argument and only wired it for
    Linear/Conv2d Reparameterization; for the other layers the forward is run
    with return_kl=False and KL is computed with the reference's own
    kl_div(..., 'normal') on the same tensors (= upstream behaviour).
  * ConvTranspose1dFlipout / ConvTranspose2dFlipout allocate prior_weight_sigma with the
    un-transposed kernel shape; the harness substitutes the same constant in the right shape
    (see ref_kl).
Each fixture stores inputs (x, mu/rho, eps, signs), outputs (out, kl) and the
autograd gradients of  L = sum(out*g) + c*kl  w.r.t. x and every parameter.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")
KL_COEF = 0.37


def import_reference():
    """The reference itself: the staged archive (oracle/build_ref.py) or /root/reference, with the termcolor stub."""
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from oracle.build_ref import import_reference as imp
    L, D, U, _ = imp()
    return L, D, U


def ref_kl(layer):
    """kl_loss() with 'normal' semantics for every layer class."""
    w = "mu_weight" if hasattr(layer, "mu_weight") else "mu_kernel"
    r = w.replace("mu", "rho")
    ps = layer.prior_weight_sigma
    if ps.shape != getattr(layer, w).shape:
        # ConvTranspose1dFlipout / ConvTranspose2dFlipout register prior_weight_sigma as [c_out, c_in/g, *k]
        # although the kernel is [c_in, c_out/g, *k] (conv_flipout.py:706-709, :904-908): kl_div cannot
        # broadcast it unless c_in == c_out.  The buffer is a constant fill, so the shape-correct constant is used.
        ps = torch.full_like(getattr(layer, w), float(layer.prior_variance))
    kl = layer.kl_div(getattr(layer, w), torch.log1p(torch.exp(getattr(layer, r))),
                      layer.prior_weight_mu, ps, "normal")
    if layer.mu_bias is not None:
        kl = kl + layer.kl_div(layer.mu_bias, torch.log1p(torch.exp(layer.rho_bias)),
                               layer.prior_bias_mu, layer.prior_bias_sigma, "normal")
    return kl


ONLY = None  # `python oracle/make_golden.py <prefix> ...`: write only the fixtures whose name starts with a prefix


def run_case(name, layer, x, flipout=False, seed=0, extra=None, laplace=False):
    if ONLY and not any(name.startswith(p) for p in ONLY):
        return
    # layer and input are built HERE under a per-case seed (callables), so a fixture depends on its own case only,
    # not on which cases ran before it
    torch.manual_seed(500 + seed)
    if isinstance(layer, types.FunctionType):
        layer = layer()
    if isinstance(x, types.FunctionType):
        x = x()
    torch.manual_seed(1000 + seed)
    x = x.clone().requires_grad_(True)
    torch.manual_seed(2000 + seed)
    out = layer(x, return_kl=False)
    if laplace:   # the fork's own kl_loss() with prior_type='laplace' (it prints on every call)
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            kl = layer.kl_loss()
    else:
        kl = ref_kl(layer)
    g = torch.randn(out.shape, generator=torch.Generator().manual_seed(3000 + seed))
    loss = (out * g).sum() + KL_COEF * kl
    params = [p for p in layer.parameters()]
    grads = torch.autograd.grad(loss, [x] + params)
    wn = "weight" if hasattr(layer, "mu_weight") else "kernel"
    d = {
        "x": x.detach(), "g": g, "out": out.detach(), "kl": kl.detach(),
        "mu_w": getattr(layer, "mu_" + wn).detach(), "rho_w": getattr(layer, "rho_" + wn).detach(),
        "eps_w": getattr(layer, "eps_" + wn).detach().clone(),
        "dx": grads[0],
    }
    names = [n for n, _ in layer.named_parameters()]
    for n, gr in zip(names, grads[1:]):
        d["d_" + n] = gr
    if layer.mu_bias is not None:
        d.update(mu_b=layer.mu_bias.detach(), rho_b=layer.rho_bias.detach(),
                 eps_b=layer.eps_bias.detach().clone())
    if flipout:
        # replay the reference's draw order to recover the sign tensors
        torch.manual_seed(2000 + seed)
        if x.dim() == 2 or hasattr(layer, "mu_weight"):
            # linear_flipout.py:150,162 then :169-170
            layer.eps_weight.clone().normal_()
            if layer.mu_bias is not None:
                layer.eps_bias.clone().normal_()
            s_in = x.detach().clone().uniform_(-1, 1).sign()
            s_out = out.detach().clone().uniform_(-1, 1).sign()
        else:
            # conv_flipout.py:385-386 then :390,401
            s_in = x.detach().clone().uniform_(-1, 1).sign()
            s_out = out.detach().clone().uniform_(-1, 1).sign()
        d.update(sign_in=s_in, sign_out=s_out)
    if extra:
        d.update(extra)
    np.savez_compressed(os.path.join(OUT, name + ".npz"),
                        **{k: (v.numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()})
    print(f"{name:34s} out{tuple(out.shape)} kl={float(kl):.6f}")


def cfg(**kw):
    return {"cfg_" + k: np.asarray(v) for k, v in kw.items()}


def main(out_dir=None):
    global OUT
    if out_dir is not None:
        OUT = out_dir
    os.makedirs(OUT, exist_ok=True)
    L, D, U = import_reference()
    torch.manual_seed(0)
    R = torch.randn

    # ---- LinearReparameterization (linear_variational.py:54) ----
    run_case("linear_rp_8x5x7_bias", lambda: L.LinearReparameterization(5, 7), lambda: R(8, 5), seed=1)
    run_case("linear_rp_33x70x37_nobias", lambda: L.LinearReparameterization(70, 37, bias=False), lambda: R(33, 70), seed=2)
    run_case("linear_rp_lead_2x3x16x24", lambda: L.LinearReparameterization(16, 24), lambda: R(2, 3, 16), seed=3)
    run_case("linear_rp_128x256x192_prior", lambda: L.LinearReparameterization(
        256, 192, prior_mean=0.3, prior_variance=0.5, posterior_rho_init=-2.0), lambda: R(128, 256), seed=4,
        extra=cfg(prior_mean=0.3, prior_variance=0.5))

    # ---- Conv2dReparameterization (conv_variational.py:234) ----
    def c2(name, x, seed, **kw):
        layer = lambda: L.Conv2dReparameterization(prior_type="normal", **kw)  # noqa: E731
        ex = cfg(stride=kw.get("stride", 1), padding=kw.get("padding", 0), dilation=kw.get("dilation", 1),
                 groups=kw.get("groups", 1), prior_mean=kw.get("prior_mean", 0), prior_variance=kw.get("prior_variance", 1))
        run_case(name, layer, x, seed=seed, extra=ex)
    c2("conv2d_rp_1x1", lambda: R(2, 16, 9, 9), 10, in_channels=16, out_channels=24, kernel_size=1)
    c2("conv2d_rp_3x3_p1", lambda: R(3, 8, 12, 12), 11, in_channels=8, out_channels=16, kernel_size=3, padding=1)
    c2("conv2d_rp_3x3_s2_p1", lambda: R(2, 8, 13, 11), 12, in_channels=8, out_channels=12, kernel_size=3, stride=2, padding=1)
    c2("conv2d_rp_7x7_s2_p3_c3", lambda: R(2, 3, 32, 32), 13, in_channels=3, out_channels=16, kernel_size=7, stride=2, padding=3, bias=False)
    c2("conv2d_rp_groups2", lambda: R(2, 8, 10, 10), 14, in_channels=8, out_channels=12, kernel_size=3, padding=1, groups=2)
    c2("conv2d_rp_dil2", lambda: R(2, 4, 14, 14), 15, in_channels=4, out_channels=6, kernel_size=3, padding=2, dilation=2)
    c2("conv2d_rp_prior", lambda: R(2, 4, 8, 8), 16, in_channels=4, out_channels=8, kernel_size=(3, 2), padding=(1, 0),
       stride=(1, 2), prior_mean=-0.2, prior_variance=2.0)
    c2("conv2d_rp_64ch", lambda: R(2, 64, 14, 14), 17, in_channels=64, out_channels=64, kernel_size=3, padding=1, bias=False)

    # ---- Conv1d / Conv3d Reparameterization (conv_variational.py:68, :650) ----
    run_case("conv1d_rp_k3", lambda: L.Conv1dReparameterization(6, 10, 3, stride=2, padding=1), lambda: R(3, 6, 25), seed=20,
             extra=cfg(stride=2, padding=1, dilation=1, groups=1))
    run_case("conv3d_rp_k3", lambda: L.Conv3dReparameterization(4, 6, 3, 0, 1, 0, -3.0, stride=1, padding=1), lambda: R(2, 4, 6, 7, 5),
             seed=21, extra=cfg(stride=1, padding=1, dilation=1, groups=1))

    # ---- Flipout (linear_flipout.py:49, conv_flipout.py:247) ----
    run_case("linear_fo_16x20x12", lambda: L.LinearFlipout(20, 12), lambda: R(16, 20), flipout=True, seed=30)
    run_case("linear_fo_nobias", lambda: L.LinearFlipout(33, 9, bias=False), lambda: R(5, 33), flipout=True, seed=31)
    run_case("conv2d_fo_3x3", lambda: L.Conv2dFlipout(8, 12, 3, padding=1), lambda: R(2, 8, 10, 10), flipout=True, seed=32,
             extra=cfg(stride=1, padding=1, dilation=1, groups=1))
    run_case("conv2d_fo_s2_nobias", lambda: L.Conv2dFlipout(6, 8, 3, stride=2, padding=1, bias=False), lambda: R(2, 6, 11, 11),
             flipout=True, seed=33, extra=cfg(stride=2, padding=1, dilation=1, groups=1))

    # ---- the fork's Laplace prior (base_variational_layer.py:74-97): Linear / Conv2d Reparameterization only ----
    run_case("linear_rp_laplace_24x40", lambda: L.LinearReparameterization(40, 24, prior_type="laplace", posterior_rho_init=-1.0), lambda: R(6, 40),
             seed=45, extra={"cfg_prior_type": np.asarray(1)}, laplace=True)
    run_case("conv2d_rp_laplace_3x3", lambda: L.Conv2dReparameterization(6, 8, 3, padding=1, prior_type="laplace"), lambda: R(2, 6, 9, 9), seed=46,
             extra=dict(cfg(stride=1, padding=1, dilation=1, groups=1), cfg_prior_type=np.asarray(1)), laplace=True)

    # ---- Conv1d / Conv3d Flipout (conv_flipout.py:57, :443) ----
    run_case("conv1d_fo_k3_s2", lambda: L.Conv1dFlipout(6, 10, 3, stride=2, padding=1), lambda: R(3, 6, 25), flipout=True, seed=34,
             extra=cfg(stride=2, padding=1, dilation=1, groups=1))
    run_case("conv3d_fo_k3", lambda: L.Conv3dFlipout(4, 6, 3, padding=1, bias=False), lambda: R(2, 4, 5, 6, 4), flipout=True, seed=35,
             extra=cfg(stride=1, padding=1, dilation=1, groups=1))

    # ---- ConvTranspose{1,2,3}d Reparameterization (conv_variational.py:822, :992, :1167) and Flipout
    #      (conv_flipout.py:640, :834, :1033); kernel [c_in, c_out / groups, *k] ----
    def ct(**kw):
        return cfg(stride=kw.get("stride", 1), padding=kw.get("padding", 0), dilation=kw.get("dilation", 1),
                   groups=kw.get("groups", 1), output_padding=kw.get("output_padding", 0))
    kw = dict(stride=2, padding=1, output_padding=1)
    run_case("convT1d_rp_k3_s2", lambda: L.ConvTranspose1dReparameterization(6, 10, 3, **kw), lambda: R(3, 6, 9), seed=60, extra=ct(**kw))
    kw = dict(stride=2, padding=1, output_padding=1)
    run_case("convT2d_rp_k3_s2", lambda: L.ConvTranspose2dReparameterization(8, 12, (3, 3), **kw), lambda: R(2, 8, 7, 6), seed=61, extra=ct(**kw))
    kw = dict(stride=1, padding=2, dilation=2, groups=2)
    run_case("convT2d_rp_dil2_groups2_nobias", lambda: L.ConvTranspose2dReparameterization(8, 12, (3, 3), bias=False, **kw),
             lambda: R(2, 8, 6, 6), seed=62, extra=ct(**kw))
    kw = dict(stride=2, padding=0)
    run_case("convT3d_rp_k2_s2", lambda: L.ConvTranspose3dReparameterization(4, 6, (2, 2, 2), **kw), lambda: R(2, 4, 3, 4, 3), seed=63, extra=ct(**kw))
    kw = dict(stride=2, padding=1, output_padding=1)
    run_case("convT1d_fo_k3_s2", lambda: L.ConvTranspose1dFlipout(6, 10, 3, **kw), lambda: R(3, 6, 9), flipout=True, seed=64, extra=ct(**kw))
    kw = dict(stride=2, padding=1, output_padding=1)
    run_case("convT2d_fo_k3_s2", lambda: L.ConvTranspose2dFlipout(8, 12, (3, 3), **kw), lambda: R(2, 8, 5, 6), flipout=True, seed=65, extra=ct(**kw))
    kw = dict(stride=1, padding=1)
    run_case("convT3d_fo_k3_nobias", lambda: L.ConvTranspose3dFlipout(4, 6, (3, 3, 3), bias=False, **kw), lambda: R(2, 4, 3, 4, 3), flipout=True,
             seed=66, extra=ct(**kw))

    # ---- MOPED rho (utils/util.py:63-69) and tensor-valued priors (utils/util.py:101-104) ----
    torch.manual_seed(540)   # (every case seeds its own construction: fixtures do not depend on the case order)
    w = R(40, 30) * 0.05
    lay = L.LinearReparameterization(30, 40)
    lay.mu_weight.data.copy_(w)
    lay.rho_weight.data.copy_(U.get_rho(w, 0.5))
    lay.prior_weight_mu = w.clone() * 0.9
    lay.prior_weight_sigma = torch.rand(40, 30) + 0.5
    run_case("linear_rp_moped_tensorprior", lay, lambda: R(4, 30), seed=40,
             extra={"moped_w": w, "moped_delta": np.asarray(0.5), "moped_rho": U.get_rho(w, 0.5),
                    "prior_w_mu": lay.prior_weight_mu, "prior_w_sigma": lay.prior_weight_sigma})

    # ---- LSTMReparameterization / LSTMFlipout (rnn_variational.py:46-153, rnn_flipout.py:46-155): a fresh draw per
    #      time step and per sub-layer; the draws are recovered by replaying torch's generator from the state
    #      recorded before each ih / hh call.  Harness shim: the fork's kl_div has no default prior_type and the
    #      Flipout linears call it without one -> default it to 'normal' (upstream behaviour).
    from bayesian_torch.layers.base_variational_layer import BaseVariationalLayer_ as RefBase
    RefBase.kl_div.__defaults__ = ("normal",)
    for name, cls, T, seed in (("lstm_rp_T5", L.LSTMReparameterization, 5, 70), ("lstm_fo_T4", L.LSTMFlipout, 4, 71)):
        if ONLY and not any(name.startswith(p) for p in ONLY):
            continue
        torch.manual_seed(1000 + seed)
        lstm = cls(6, 5)
        x = R(3, T, 6).requires_grad_(True)
        states = []
        hooks = [m.register_forward_pre_hook(lambda mod, inp, tag=tag: states.append((tag, torch.get_rng_state(), tuple(inp[0].shape))))
                 for tag, m in (("ih", lstm.ih), ("hh", lstm.hh))]
        torch.manual_seed(2000 + seed)
        hs, (_, cs), kl = lstm(x)
        for h in hooks:
            h.remove()
        g = torch.randn(hs.shape, generator=torch.Generator().manual_seed(3000 + seed))
        gc = torch.randn(cs.shape, generator=torch.Generator().manual_seed(3001 + seed))
        loss = (hs * g).sum() + (cs * gc).sum() + KL_COEF * kl
        names = [n for n, _ in lstm.named_parameters()]
        grads = torch.autograd.grad(loss, [x] + list(lstm.parameters()))
        d = {"x": x.detach(), "g": g, "gc": gc, "hs": hs.detach(), "cs": cs.detach(), "kl": kl.detach(), "dx": grads[0]}
        d.update({"p_" + n: p.detach() for n, p in lstm.named_parameters()})
        d.update({"d_" + n: gr for n, gr in zip(names, grads[1:])})
        keep = torch.get_rng_state()
        flip = cls is L.LSTMFlipout
        for i, (tag, st, xs) in enumerate(states):   # call i: step i // 2, layer ih / hh
            torch.set_rng_state(st)
            lay = getattr(lstm, tag)
            t = i // 2
            d[f"eps_{tag}_w_{t}"] = torch.empty_like(lay.eps_weight).normal_()     # linear_variational.py:164 / linear_flipout.py:150
            d[f"eps_{tag}_b_{t}"] = torch.empty_like(lay.eps_bias).normal_()       # :176 / :162
            if flip:                                                               # linear_flipout.py:169-170
                d[f"sin_{tag}_{t}"] = torch.empty(xs).uniform_(-1, 1).sign()
                d[f"sout_{tag}_{t}"] = torch.empty(xs[0], 20).uniform_(-1, 1).sign()
        torch.set_rng_state(keep)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **{k: v.numpy() for k, v in d.items()})
        print(f"{name:34s} hs{tuple(hs.shape)} kl={float(kl):.6f}")

    if ONLY and not any("dnn_to_bnn_small".startswith(p) for p in ONLY):
        return
    # ---- dnn_to_bnn + get_kl_loss on a small torch model (models/dnn_to_bnn.py:127-165) ----
    torch.manual_seed(50)
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(8, 8, 3, stride=2, bias=False),
                              torch.nn.Flatten(), torch.nn.Linear(8 * 3 * 3, 10))
    dnn_state = {k: v.clone() for k, v in net.state_dict().items()}
    params = {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
              "type": "Reparameterization", "moped_enable": True, "moped_delta": 0.5}
    D.dnn_to_bnn(net, params)
    for m in net.modules():
        if hasattr(m, "prior_type"):
            m.prior_type = "normal"
    torch.manual_seed(52)
    x = R(4, 3, 8, 8)
    torch.manual_seed(51)
    y = net(x)
    kl = D.get_kl_loss(net)
    d = {"x": x, "out": y.detach(), "kl": kl.detach()}
    d.update({"dnn_" + k: v for k, v in dnn_state.items()})
    d.update({"bnn_" + k: v.detach() for k, v in net.state_dict().items()})
    for i, m in enumerate(net):
        if hasattr(m, "eps_kernel"):
            d[f"eps_{i}_w"] = m.eps_kernel.clone()
        if hasattr(m, "eps_weight"):
            d[f"eps_{i}_w"] = m.eps_weight.clone()
        if getattr(m, "eps_bias", None) is not None:
            d[f"eps_{i}_b"] = m.eps_bias.clone()
    np.savez_compressed(os.path.join(OUT, "dnn_to_bnn_small.npz"), **{k: v.numpy() for k, v in d.items()})
    print("dnn_to_bnn_small", tuple(y.shape), float(kl), sorted(net.state_dict().keys()))


def check(tol=1e-6):
    """Regenerate every fixture into a scratch directory and compare with the committed files.  Returns the list of
    (fixture, key, max|delta|) that differ by more than `tol` * max(1, max|committed|)."""
    import tempfile
    committed = os.path.join(ROOT, "tests", "golden")
    bad = []
    with tempfile.TemporaryDirectory() as tmp:
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            main(tmp)
        names = sorted(f for f in os.listdir(tmp) if f.endswith(".npz"))
        have = sorted(f for f in os.listdir(committed) if f.endswith(".npz"))
        for f in sorted(set(names) ^ set(have)):
            bad.append((f, "<file>", float("inf")))
        for f in names:
            if f not in have:
                continue
            a, b = np.load(os.path.join(tmp, f)), np.load(os.path.join(committed, f))
            for k in sorted(set(a.files) | set(b.files)):
                if k not in a.files or k not in b.files or a[k].shape != b[k].shape:
                    bad.append((f, k, float("inf")))
                    continue
                d = float(np.max(np.abs(a[k].astype(np.float64) - b[k].astype(np.float64)))) if a[k].size else 0.0
                if d > tol * max(1.0, float(np.max(np.abs(b[k]))) if b[k].size else 1.0):
                    bad.append((f, k, d))
    return bad


if __name__ == "__main__":
    args = sys.argv[1:]
    if "--check" in args:
        diffs = check()
        for f, k, d in diffs:
            print(f"DIFF {f}:{k} max|delta| {d:.3e}")
        print("fixtures reproduce from the committed recipe" if not diffs else f"{len(diffs)} entries differ")
        sys.exit(1 if diffs else 0)
    ONLY = args or None
    main()
