"""Torch-CPU restatement of the reference's variational-layer arithmetic.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Every function cites the
reference lines (relative to /root/reference/bayesian_torch/) it follows.  All
functions are pure: eps / sign tensors are inputs, never drawn here, so the
same tensors can be fed to the CUDA path.  ``dtype=torch.float64`` gives a
higher-precision answer for tolerance studies; the reference computes in fp32.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

__all__ = [
    "softplus_ref", "kl_normal", "kl_laplace", "kl_normal_from_rho", "sample_weight",
    "linear_reparam", "conv_reparam", "linear_flipout", "conv_flipout",
    "conv_transpose_reparam", "conv_transpose_flipout", "lstm_unrolled",
    "get_rho", "mc_moments", "predictive_entropy", "mutual_information",
]


def _c(t, dtype):
    return None if t is None else t.to("cpu", dtype)


def softplus_ref(rho: torch.Tensor) -> torch.Tensor:
    """sigma = log1p(exp(rho)), un-stabilised exactly as the reference writes it
    (layers/variational_layers/linear_variational.py:163)."""
    return torch.log1p(torch.exp(rho))


def kl_normal(mu_q, sigma_q, mu_p, sigma_p) -> torch.Tensor:
    """Element-wise KL(N(mu_q,sigma_q) || N(mu_p,sigma_p)) then .mean()
    (layers/base_variational_layer.py:70-72).  mu_p / sigma_p may be python
    scalars or tensors broadcastable to mu_q."""
    mu_p = torch.as_tensor(mu_p, dtype=mu_q.dtype)
    sigma_p = torch.as_tensor(sigma_p, dtype=mu_q.dtype)
    kl = (torch.log(sigma_p) - torch.log(sigma_q)
          + (sigma_q ** 2 + (mu_q - mu_p) ** 2) / (2 * (sigma_p ** 2)) - 0.5)
    return kl.mean()


def kl_laplace(mu_q, sigma_q) -> torch.Tensor:
    """The fork's analytic KL(N(mu_q, sigma_q) || Laplace(0, 1)) then .mean()
    (layers/base_variational_layer.py:74-97: b_p = 1 and mu_p = 0 are hard-coded, the
    caller's prior arguments are ignored).  E|w| is the folded-normal mean."""
    std_normal = torch.distributions.Normal(torch.zeros((), dtype=mu_q.dtype), torch.ones((), dtype=mu_q.dtype))
    b_p = torch.tensor(1.0, dtype=mu_q.dtype)
    off = mu_q - torch.tensor(0.0, dtype=mu_q.dtype)
    exp_abs = (sigma_q * math.sqrt(2.0 / math.pi) * torch.exp(-off ** 2 / (2 * sigma_q ** 2))
               + off * (1 - 2 * std_normal.cdf(-off / sigma_q)))
    kl = torch.log(2 * b_p) - 0.5 * torch.log(2 * math.pi * sigma_q ** 2) - 0.5 + (1 / b_p) * exp_abs
    return kl.mean()


def kl_normal_from_rho(mu, rho, mu_b=None, rho_b=None, prior_mu=0.0,
                       prior_sigma=1.0, dtype=torch.float32) -> torch.Tensor:
    """kl_loss() of a layer: weight mean + bias mean
    (linear_variational.py:146-158, conv_variational.py:353-360)."""
    mu, rho, mu_b, rho_b = (_c(t, dtype) for t in (mu, rho, mu_b, rho_b))
    kl = kl_normal(mu, softplus_ref(rho), prior_mu, prior_sigma)
    if mu_b is not None:
        kl = kl + kl_normal(mu_b, softplus_ref(rho_b), prior_mu, prior_sigma)
    return kl


def sample_weight(mu, rho, eps):
    """w = mu + log1p(exp(rho)) * eps (linear_variational.py:163-166)."""
    return mu + softplus_ref(rho) * eps


def _conv(rank: int):
    return {1: F.conv1d, 2: F.conv2d, 3: F.conv3d}[rank]


def linear_reparam(x, mu_w, rho_w, mu_b, rho_b, eps_w, eps_b, prior_mu=0.0,
                   prior_sigma=1.0, dtype=torch.float32
                   ) -> Tuple[torch.Tensor, torch.Tensor]:
    """LinearReparameterization.forward (linear_variational.py:160-204)."""
    x, mu_w, rho_w, mu_b, rho_b, eps_w, eps_b = (
        _c(t, dtype) for t in (x, mu_w, rho_w, mu_b, rho_b, eps_w, eps_b))
    w = sample_weight(mu_w, rho_w, eps_w)
    b = None if mu_b is None else sample_weight(mu_b, rho_b, eps_b)
    out = F.linear(x, w, b)
    return out, kl_normal_from_rho(mu_w, rho_w, mu_b, rho_b, prior_mu, prior_sigma, dtype)


def conv_reparam(x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, stride=1, padding=0,
                 dilation=1, groups=1, prior_mu=0.0, prior_sigma=1.0,
                 dtype=torch.float32) -> Tuple[torch.Tensor, torch.Tensor]:
    """Conv{1,2,3}dReparameterization.forward (conv_variational.py:197-231,
    362-407, 785-819); the rank comes from the kernel tensor."""
    x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b = (
        _c(t, dtype) for t in (x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b))
    w = sample_weight(mu_k, rho_k, eps_k)
    b = None if mu_b is None else sample_weight(mu_b, rho_b, eps_b)
    out = _conv(mu_k.dim() - 2)(x, w, b, stride, padding, dilation, groups)
    return out, kl_normal_from_rho(mu_k, rho_k, mu_b, rho_b, prior_mu, prior_sigma, dtype)


def linear_flipout(x, mu_w, rho_w, mu_b, rho_b, eps_w, eps_b, sign_in, sign_out,
                   prior_mu=0.0, prior_sigma=1.0, dtype=torch.float32):
    """LinearFlipout.forward (flipout_layers/linear_flipout.py:145-197):
    out = x.mu^T + mu_b + ((x*s_in).(sigma*eps)^T + sigma_b*eps_b) * s_out."""
    x, mu_w, rho_w, mu_b, rho_b, eps_w, eps_b, sign_in, sign_out = (
        _c(t, dtype) for t in (x, mu_w, rho_w, mu_b, rho_b, eps_w, eps_b, sign_in, sign_out))
    delta_w = softplus_ref(rho_w) * eps_w
    delta_b = None if mu_b is None else softplus_ref(rho_b) * eps_b
    mean_path = F.linear(x, mu_w, mu_b)
    pert = F.linear(x * sign_in, delta_w, delta_b)
    out = mean_path + pert * sign_out
    return out, kl_normal_from_rho(mu_w, rho_w, mu_b, rho_b, prior_mu, prior_sigma, dtype)


def conv_flipout(x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, sign_in, sign_out,
                 stride=1, padding=0, dilation=1, groups=1, prior_mu=0.0,
                 prior_sigma=1.0, dtype=torch.float32):
    """Conv2dFlipout.forward (flipout_layers/conv_flipout.py:370-439)."""
    x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, sign_in, sign_out = (
        _c(t, dtype) for t in (x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, sign_in, sign_out))
    conv = _conv(mu_k.dim() - 2)
    delta_k = softplus_ref(rho_k) * eps_k
    delta_b = None if mu_b is None else softplus_ref(rho_b) * eps_b
    mean_path = conv(x, mu_k, mu_b, stride, padding, dilation, groups)
    pert = conv(x * sign_in, delta_k, delta_b, stride, padding, dilation, groups)
    out = mean_path + pert * sign_out
    return out, kl_normal_from_rho(mu_k, rho_k, mu_b, rho_b, prior_mu, prior_sigma, dtype)


def _conv_transpose(rank: int):
    return {1: F.conv_transpose1d, 2: F.conv_transpose2d, 3: F.conv_transpose3d}[rank]


def conv_transpose_reparam(x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, stride=1, padding=0,
                           output_padding=0, dilation=1, groups=1, prior_mu=0.0, prior_sigma=1.0,
                           dtype=torch.float32) -> Tuple[torch.Tensor, torch.Tensor]:
    """ConvTranspose{1,2,3}dReparameterization.forward (conv_variational.py:941-989,
    1116-1164, 1292-1340); kernel [c_in, c_out / groups, *k]."""
    x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b = (
        _c(t, dtype) for t in (x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b))
    w = sample_weight(mu_k, rho_k, eps_k)
    b = None if mu_b is None else sample_weight(mu_b, rho_b, eps_b)
    out = _conv_transpose(mu_k.dim() - 2)(x, w, b, stride, padding, output_padding, groups, dilation)
    return out, kl_normal_from_rho(mu_k, rho_k, mu_b, rho_b, prior_mu, prior_sigma, dtype)


def conv_transpose_flipout(x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, sign_in, sign_out,
                           stride=1, padding=0, output_padding=0, dilation=1, groups=1,
                           prior_mu=0.0, prior_sigma=1.0, dtype=torch.float32):
    """ConvTranspose{1,2,3}dFlipout.forward (flipout_layers/conv_flipout.py:757-831,
    956-1030, 1155-1228)."""
    x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, sign_in, sign_out = (
        _c(t, dtype) for t in (x, mu_k, rho_k, mu_b, rho_b, eps_k, eps_b, sign_in, sign_out))
    convt = _conv_transpose(mu_k.dim() - 2)
    delta_k = softplus_ref(rho_k) * eps_k
    delta_b = None if mu_b is None else softplus_ref(rho_b) * eps_b
    mean_path = convt(x, mu_k, mu_b, stride, padding, output_padding, groups, dilation)
    pert = convt(x * sign_in, delta_k, delta_b, stride, padding, output_padding, groups, dilation)
    out = mean_path + pert * sign_out
    return out, kl_normal_from_rho(mu_k, rho_k, mu_b, rho_b, prior_mu, prior_sigma, dtype)


def lstm_unrolled(x, p, draws, flipout=False, hidden_states=None, dtype=torch.float32, alias_eps=False):
    """LSTMReparameterization / LSTMFlipout.forward (rnn_variational.py:112-153, rnn_flipout.py:102-155).
    p: {"ih.mu_weight": ..., "hh.rho_bias": ...}; draws(tag, t) -> (eps_w, eps_b[, sign_in, sign_out]) of
    sub-layer `tag` at step t (a fresh draw per step and sub-layer).  Returns (hidden_seq, c_seq, kl) with
    kl summed over the steps as the reference does.

    alias_eps=True reproduces one more property of the reference: every sub-layer draws into ONE eps buffer
    through `.data.normal_()` (linear_variational.py:164,176; linear_flipout.py:150,162), the multiplication
    `sigma * eps` saves that buffer for backward, and the version counter does not see `.data` writes -- so after
    a multi-step forward the backward of EVERY step reads the LAST step's eps.  Outputs, KL, dx and the mu
    gradients are unaffected; the rho gradients are those of the wrong draw.  The CUDA path regenerates each
    call's own draw (the mathematically correct gradient, alias_eps=False); the aliasing mode exists so the
    restatement can be pinned against the reference's autograd output."""
    if alias_eps:
        bufs, inner = {}, draws

        def draws(tag, t):  # noqa: F811
            vals = inner(tag, t)
            if tag not in bufs:
                bufs[tag] = [torch.empty_like(vals[0]), torch.empty_like(vals[1])]
            bufs[tag][0].data.copy_(vals[0]); bufs[tag][1].data.copy_(vals[1])
            return [bufs[tag][0].data, bufs[tag][1].data] + list(vals[2:])
    B, T, _ = x.shape
    HS = p["hh.mu_weight"].shape[1]
    h_t = torch.zeros(B, HS, dtype=dtype) if hidden_states is None else hidden_states[0]
    c_t = torch.zeros(B, HS, dtype=dtype) if hidden_states is None else hidden_states[1]
    fn = linear_flipout if flipout else linear_reparam
    hs, cs, kl = [], [], 0
    for t in range(T):
        ff = []
        for tag, inp in (("ih", x[:, t, :]), ("hh", h_t)):
            out, k = fn(inp, p[tag + ".mu_weight"], p[tag + ".rho_weight"], p.get(tag + ".mu_bias"), p.get(tag + ".rho_bias"),
                        *draws(tag, t), dtype=dtype)
            ff.append(out)
            kl = kl + k
        gates = ff[0] + ff[1]
        i_t, f_t = torch.sigmoid(gates[:, :HS]), torch.sigmoid(gates[:, HS:HS * 2])
        g_t, o_t = torch.tanh(gates[:, HS * 2:HS * 3]), torch.sigmoid(gates[:, HS * 3:])
        c_t = f_t * c_t + i_t * g_t
        h_t = o_t * torch.tanh(c_t)
        hs.append(h_t.unsqueeze(0))
        cs.append(c_t.unsqueeze(0))
    return (torch.cat(hs, 0).transpose(0, 1).contiguous(), torch.cat(cs, 0).transpose(0, 1).contiguous(), kl)


def get_rho(w: torch.Tensor, delta: float) -> torch.Tensor:
    """MOPED rho init: log(expm1(delta*|w|) + 1e-20) (utils/util.py:63-69)."""
    return torch.log(torch.expm1(delta * torch.abs(w)) + 1e-20)


def mc_moments(logits: torch.Tensor, softmax: bool = True):
    """Predictive mean / variance over the S axis of [S,B,C] logits.  The mean
    of softmax is what evaluate() forms
    (examples/main_bayesian_cifar_dnn2bnn.py:545,554-556); variance is
    E[p^2]-E[p]^2 (population), the quantity the multi-GPU path all-reduces."""
    p = torch.softmax(logits.double(), dim=-1) if softmax else logits.double()
    mean = p.mean(0)
    var = (p * p).mean(0) - mean * mean
    return mean, var


def predictive_entropy(probs: torch.Tensor) -> torch.Tensor:
    """Entropy of the mean predictive distribution (utils/util.py:41-52);
    probs is [S,B,C]."""
    m = probs.double().mean(0)
    return -(m * torch.log(m + 1e-15)).sum(-1)


def mutual_information(probs: torch.Tensor) -> torch.Tensor:
    """H[mean p] - mean H[p] (utils/util.py:55-60)."""
    p = probs.double()
    per_sample = -(p * torch.log(p + 1e-15)).sum(-1)
    return predictive_entropy(p) - per_sample.mean(0)
