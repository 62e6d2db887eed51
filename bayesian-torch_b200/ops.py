"""Host-side operator layer: torch tensors in, C-ABI calls out.

PyTorch is plumbing here (device memory, streams, autograd bookkeeping); all
arithmetic of the hot path runs in libbt_b200.so.  CPU tensors raise -- there
is no CPU / eager fallback (BASELINE.json north_star).
"""
from __future__ import annotations

import contextlib
import os
import ctypes as C
import threading
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import Geom, LayerGrads, LayerParams

# --------------------------------------------------------------------------- RNG / mode state


class _State(threading.local):
    def __init__(self):
        self.seed = 0x5EEDB200
        self.sample_override: Optional[int] = None  # set by mc_sample()
        self.sample_mode = "philox"                  # "philox" | "torch" | "buffer"
        self.path = "auto"                           # "auto" | "simt" | "tc"
        self.n_samples = 1                           # >1 inside mc_batched()
        self.shared_batch: Optional[int] = None      # batch of the un-folded model input
        self.call_counts: Optional[dict] = None      # layer_id -> calls so far inside this mc_sample()/mc_batched() block


_STATE = _State()


class _Profile:
    """Launch counter + optional CUDA-event timing of every layer call (bench.py roofline)."""

    def __init__(self):
        self._base = 0
        self.timing = False      # record (path, flops, bytes, start, end) per layer call
        self.records = []

    @property
    def launches(self):
        """Kernels of libbt_b200.so launched since the last reset (counted inside the library)."""
        return int(_lib.lib().bt_launch_count()) - self._base

    def reset(self):
        self._base = int(_lib.lib().bt_launch_count())
        self.records = []


PROFILE = _Profile()
_layer_counter = [0]
_counter_lock = threading.Lock()

_PATHS = {"auto": _lib.PATH_AUTO, "simt": _lib.PATH_SIMT_F32, "tc": _lib.PATH_TC_BF16}


def next_layer_id() -> int:
    with _counter_lock:
        _layer_counter[0] += 1
        return _layer_counter[0] & 0x0FFFFFFF


def manual_seed(seed: int) -> None:
    """Seed of the in-kernel Philox streams (independent of torch's generator)."""
    _STATE.seed = int(seed) & 0xFFFFFFFFFFFFFFFF


def get_seed() -> int:
    return _STATE.seed


def set_sample_mode(mode: str) -> None:
    """'philox': eps/signs generated in registers (default, nothing materialised);
    'torch' : eps drawn with torch's generator into the layer's eps_* buffers exactly
              like the reference (linear_variational.py:164), then fed to the kernel;
    'buffer': use whatever the eps_* buffers / sign tensors hold (parity tests)."""
    if mode not in ("philox", "torch", "buffer"):
        raise ValueError(f"unknown sample mode {mode!r}")
    _STATE.sample_mode = mode


def get_sample_mode() -> str:
    return _STATE.sample_mode


def set_path(path: str) -> None:
    """'auto' (tensor cores when the problem qualifies), 'simt' (fp32 parity path), 'tc'."""
    if path not in _PATHS:
        raise ValueError(f"unknown path {path!r}")
    _STATE.path = path


def get_path() -> str:
    return _STATE.path


@contextlib.contextmanager
def sample_mode(mode: str):
    old = _STATE.sample_mode
    set_sample_mode(mode)
    try:
        yield
    finally:
        _STATE.sample_mode = old


@contextlib.contextmanager
def path(p: str):
    old = _STATE.path
    set_path(p)
    try:
        yield
    finally:
        _STATE.path = old


@contextlib.contextmanager
def mc_sample(sample_id: int):
    """Pin the Monte-Carlo sample index used by every layer inside the block: sample s
    of a model is the same weight draw on any rank / any GPU count.  One model forward per
    block: a layer called AGAIN inside the same block (LSTM time steps, weight sharing, a
    second forward) draws afresh, as the reference does on every call (call_seed below)."""
    old = (_STATE.sample_override, _STATE.call_counts)
    _STATE.sample_override = int(sample_id) & 0xFFFFFFFF
    _STATE.call_counts = {}
    try:
        yield
    finally:
        _STATE.sample_override, _STATE.call_counts = old


def call_seed(seed: int, layer_id: int) -> int:
    """Seed of this call of layer `layer_id`.  Inside mc_sample() / mc_batched() the sample id is pinned, so a
    layer that runs several times in one forward (LSTM steps, weight sharing) would repeat its draw; the
    reference redraws on every call (linear_variational.py:164).  The k-th call (k >= 1) of a layer inside the
    block therefore uses seed + k * 0x9E3779B97F4A7C15 (mod 2^64): a different Philox key, i.e. an independent
    stream; the first call uses the seed itself.  Outside such a block the per-layer call counter already
    advances the sample id."""
    counts = _STATE.call_counts
    if counts is None:
        return seed
    k = counts.get(layer_id, 0)
    counts[layer_id] = k + 1
    return derive_seed(seed, k)


def derive_seed(seed: int, k: int) -> int:
    return (int(seed) + int(k) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF


@contextlib.contextmanager
def mc_batched(n_samples: int, first_sample: int = 0, shared_batch: Optional[int] = None):
    """Run S weight samples in ONE pass: inside the block the leading batch axis of every
    Bayesian layer input is read as [S*B, ...] (sample-major) and sample s of the batch
    uses weight draw first_sample + s.  Deterministic torch layers (BN, ReLU, pooling)
    work unchanged on the folded batch.  shared_batch=B: a layer input whose batch is B
    (the model input) is shared by all samples and only its OUTPUT is folded to [S*B, ...]."""
    old = (_STATE.n_samples, _STATE.sample_override, _STATE.shared_batch, _STATE.call_counts)
    _STATE.n_samples = int(n_samples)
    _STATE.sample_override = int(first_sample) & 0xFFFFFFFF
    _STATE.shared_batch = shared_batch
    _STATE.call_counts = {}
    try:
        yield
    finally:
        _STATE.n_samples, _STATE.sample_override, _STATE.shared_batch, _STATE.call_counts = old


def current_n_samples() -> int:
    return _STATE.n_samples


# --------------------------------------------------------------------------- helpers

_DT = {torch.float32: _lib.BT_F32, torch.bfloat16: _lib.BT_BF16}


def _require_cuda(*tensors) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError(
                "bayesian_torch_b200 runs on CUDA (sm_100a) only: got a CPU tensor; there is no CPU fallback")
        dev = dev or t.device
        if t.device != dev:
            raise RuntimeError(f"tensors on different devices: {t.device} vs {dev}")
    return dev


def _f32c(t: Optional[torch.Tensor], name: str) -> Optional[torch.Tensor]:
    if t is None:
        return None
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32 (got {t.dtype}); variational parameters stay fp32 masters")
    return t if t.is_contiguous() else t.contiguous()


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr_array(tensors: Sequence[Optional[torch.Tensor]]):
    arr = (C.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = _ptr(t)
    return arr


# --------------------------------------------------------------------------- KL


def kl_workspace(n: int, device) -> torch.Tensor:
    nbytes = _lib.lib().bt_workspace_bytes(b"kl", None, n)
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def _kl_call(mus, rhos, q_is_sigma, prior_mu, prior_sigma, pmu_t, psig_t, want_each, d_mu=None, d_rho=None,
             grad_scale=None, prior_type="normal"):
    dev = _require_cuda(*mus, *rhos)
    n = len(mus)
    numel = (C.c_int64 * n)(*[m.numel() for m in mus])
    each = torch.empty(n, dtype=torch.float32, device=dev) if want_each else None
    total = torch.empty((), dtype=torch.float32, device=dev)
    ws = kl_workspace(n, dev)
    any_pt = any(t is not None for t in pmu_t) or any(t is not None for t in psig_t)
    if prior_type == "laplace":
        with torch.cuda.device(dev):
            rc = _lib.lib().bt_kl_laplace_multi(
                n, _ptr_array(mus), _ptr_array(rhos), numel, int(q_is_sigma), _ptr(each), _ptr(total),
                _ptr_array(d_mu) if d_mu is not None else None, _ptr_array(d_rho) if d_rho is not None else None,
                _ptr(grad_scale), _ptr(ws), _stream())
        _lib.check(rc, "bt_kl_laplace_multi")
        return total, each
    with torch.cuda.device(dev):
        rc = _lib.lib().bt_kl_normal_multi(
            n, _ptr_array(mus), _ptr_array(rhos), numel, int(q_is_sigma), float(prior_mu), float(prior_sigma),
            _ptr_array(pmu_t) if any_pt else None, _ptr_array(psig_t) if any_pt else None,
            _ptr(each), _ptr(total), _ptr_array(d_mu) if d_mu is not None else None,
            _ptr_array(d_rho) if d_rho is not None else None, _ptr(grad_scale), _ptr(ws), _stream())
    _lib.check(rc, "bt_kl_normal_multi")
    return total, each


class _KLMulti(torch.autograd.Function):
    """sum_i mean(KL(N(mu_i, softplus(rho_i)) || prior_i)) in one launch, with analytic gradients."""

    @staticmethod
    def forward(ctx, cfg, *params):
        n = len(params) // 2
        mus = [_f32c(p.detach(), "mu") for p in params[:n]]
        rhos = [_f32c(p.detach(), "rho") for p in params[n:]]
        total, _ = _kl_call(mus, rhos, cfg["q_is_sigma"], cfg["prior_mu"], cfg["prior_sigma"], cfg["pmu_t"],
                            cfg["psig_t"], False, prior_type=cfg.get("prior_type", "normal"))
        ctx.cfg = cfg
        ctx.save_for_backward(*params)
        return total

    @staticmethod
    def backward(ctx, g):
        params = ctx.saved_tensors
        cfg = ctx.cfg
        n = len(params) // 2
        mus = [_f32c(p.detach(), "mu") for p in params[:n]]
        rhos = [_f32c(p.detach(), "rho") for p in params[n:]]
        d_mu = [torch.empty_like(m) for m in mus]
        d_rho = [torch.empty_like(r) for r in rhos]
        gs = g.detach().to(torch.float32).contiguous()
        _kl_call(mus, rhos, cfg["q_is_sigma"], cfg["prior_mu"], cfg["prior_sigma"], cfg["pmu_t"], cfg["psig_t"], False,
                 d_mu, d_rho, gs, prior_type=cfg.get("prior_type", "normal"))
        grads = [d.view_as(p) for d, p in zip(d_mu + d_rho, params)]
        return (None, *grads)


def kl_normal_multi(mus: Sequence[torch.Tensor], rhos: Sequence[torch.Tensor], prior_mu: float = 0.0,
                    prior_sigma: float = 1.0, prior_mu_t: Optional[Sequence] = None,
                    prior_sigma_t: Optional[Sequence] = None, q_is_sigma: bool = False,
                    prior_type: str = "normal") -> torch.Tensor:
    """Sum over tensors of the mean element-wise Gaussian KL (base_variational_layer.py:70-72).
    One kernel launch for the whole table; differentiable w.r.t. every mu and rho.
    prior_type 'laplace': the fork's KL(N || Laplace(0, 1)) (base_variational_layer.py:74-97), which
    ignores the prior arguments."""
    n = len(mus)
    if n == 0:
        raise ValueError("kl_normal_multi: empty tensor list")
    if prior_type not in ("normal", "laplace"):
        raise ValueError(f"Unknown prior_type: {prior_type}")
    cfg = {
        "prior_type": prior_type,
        "q_is_sigma": bool(q_is_sigma), "prior_mu": float(prior_mu), "prior_sigma": float(prior_sigma),
        "pmu_t": [None if t is None else _f32c(t.detach(), "prior_mu") for t in (prior_mu_t or [None] * n)],
        "psig_t": [None if t is None else _f32c(t.detach(), "prior_sigma") for t in (prior_sigma_t or [None] * n)],
    }
    out = None
    for i in range(0, n, 192):  # KL_MAX_TENSORS per launch
        sub = dict(cfg, pmu_t=cfg["pmu_t"][i:i + 192], psig_t=cfg["psig_t"][i:i + 192])
        part = _KLMulti.apply(sub, *mus[i:i + 192], *rhos[i:i + 192])
        out = part if out is None else out + part
    return out


# --------------------------------------------------------------------------- layer op


@dataclass
class LayerCall:
    """Everything one forward/backward of a Bayesian layer needs besides x and the parameters."""
    rank: int                                   # 0 linear, 1..3 conv
    flipout: bool
    stride: Tuple[int, ...] = ()
    padding: Tuple[int, ...] = ()
    dilation: Tuple[int, ...] = ()
    groups: int = 1
    prior_mu: float = 0.0
    prior_sigma: float = 1.0
    prior_t: Tuple[Optional[torch.Tensor], ...] = (None, None, None, None)  # w_mu, w_sigma, b_mu, b_sigma
    eps_w: Optional[torch.Tensor] = None        # None -> Philox
    eps_b: Optional[torch.Tensor] = None
    sign_in: Optional[torch.Tensor] = None
    sign_out: Optional[torch.Tensor] = None
    seed: int = 0
    sample_id: int = 0
    layer_id: int = 0
    n_samples: int = 1
    x_shared: bool = False                      # mc_batched: x is the un-folded model input
    want_kl: bool = True
    path: int = _lib.PATH_AUTO
    out_scale: Optional[torch.Tensor] = None    # folded eval-mode BatchNorm (inference only)
    out_shift: Optional[torch.Tensor] = None
    out_relu: bool = False
    residual: Optional[torch.Tensor] = None     # added after the affine, before the ReLU (inference only)
    output_padding: Tuple[int, ...] = ()        # ConvTranspose layers only
    pack_cache: Optional[dict] = None           # inference-only layer (fold_batchnorm): where its packed parameters live


def _conv_out_size(i, k, s, p, d):
    return (i + 2 * p - d * (k - 1) - 1) // s + 1


def _fill3(dst, vals):
    for i, v in enumerate(vals):
        dst[i] = int(v)


def _build_geom(call: LayerCall, x: torch.Tensor, w_shape, out: torch.Tensor) -> Geom:
    """x: [rows, c_in] (linear) or [B*, C, *spatial]; out likewise. Strides in elements."""
    g = Geom()
    g.rank = call.rank
    g.groups = call.groups
    g.n_samples = call.n_samples
    g.x_dtype = _DT[x.dtype]
    g.out_dtype = _DT[out.dtype]
    S = call.n_samples
    if call.x_shared:
        g.batch = x.shape[0]
    else:
        if x.shape[0] % S != 0:
            raise RuntimeError(f"mc_batched({S}): leading batch {x.shape[0]} is not a multiple of the sample count")
        g.batch = x.shape[0] // S
    g.c_in = x.shape[1]
    g.c_out = w_shape[0]
    xs, os_ = x.stride(), out.stride()
    g.x_batch_stride, g.x_chan_stride = xs[0], xs[1]
    g.out_batch_stride, g.out_chan_stride = os_[0], os_[1]
    g.x_sample_stride = xs[0] * g.batch if (S > 1 and not call.x_shared) else 0
    g.out_sample_stride = os_[0] * g.batch if S > 1 else 0
    r = call.rank
    _fill3(g.in_size, x.shape[2:2 + r]); _fill3(g.out_size, out.shape[2:2 + r]); _fill3(g.k_size, w_shape[2:2 + r])
    _fill3(g.stride, call.stride); _fill3(g.pad, call.padding); _fill3(g.dil, call.dilation)
    _fill3(g.x_sp_stride, xs[2:2 + r]); _fill3(g.out_sp_stride, os_[2:2 + r])
    return g


def _build_params(call: LayerCall, mu_w, rho_w, mu_b, rho_b) -> Tuple[LayerParams, list]:
    p = LayerParams()
    keep = [mu_w, rho_w, mu_b, rho_b]
    p.mu_w, p.rho_w, p.mu_b, p.rho_b = _ptr(mu_w), _ptr(rho_w), _ptr(mu_b), _ptr(rho_b)
    eps_w, eps_b = _f32c(call.eps_w, "eps_w"), _f32c(call.eps_b, "eps_b")
    keep += [eps_w, eps_b]
    p.eps_w, p.eps_b = _ptr(eps_w), _ptr(eps_b)
    p.sign_in, p.sign_out = _ptr(call.sign_in), _ptr(call.sign_out)
    pts = [None if t is None else _f32c(t, "prior") for t in call.prior_t]
    keep += pts
    p.prior_mu_w_t, p.prior_sigma_w_t, p.prior_mu_b_t, p.prior_sigma_b_t = (_ptr(t) for t in pts)
    p.prior_mu, p.prior_sigma = float(call.prior_mu), float(call.prior_sigma)
    p.seed, p.sample_id, p.layer_id = call.seed, call.sample_id, call.layer_id
    sc, sh = _f32c(call.out_scale, "out_scale"), _f32c(call.out_shift, "out_shift")
    keep += [sc, sh]
    p.out_scale, p.out_shift, p.out_relu = _ptr(sc), _ptr(sh), int(bool(call.out_relu))
    p.residual = _ptr(call.residual)
    keep.append(call.residual)
    p.flags = 0
    return p, keep


def _entry(call: LayerCall, bwd: bool):
    L = _lib.lib()
    if call.flipout:
        return (L.bt_linear_flipout_bwd if bwd else L.bt_linear_flipout_fwd) if call.rank == 0 else \
               (L.bt_conv2d_flipout_bwd if bwd else L.bt_conv2d_flipout_fwd)
    return (L.bt_linear_reparam_bwd if bwd else L.bt_linear_reparam_fwd) if call.rank == 0 else \
           (L.bt_conv_reparam_bwd if bwd else L.bt_conv_reparam_fwd)


def _like_x_format(x: torch.Tensor, shape) -> torch.Tensor:
    """Output allocated in x's memory format (channels-last stays channels-last).  A layout
    that reads as both (one channel, or 1x1 images) goes channels-last for bf16 activations so
    the next layer can stay on the tensor-core path."""
    ambiguous_cl = x.dtype == torch.bfloat16
    if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last) and (not x.is_contiguous() or ambiguous_cl):
        return torch.empty(shape, dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    if x.dim() == 5 and x.is_contiguous(memory_format=torch.channels_last_3d) and (not x.is_contiguous() or ambiguous_cl):
        return torch.empty(shape, dtype=x.dtype, device=x.device, memory_format=torch.channels_last_3d)
    if x.dim() == 3 and not x.is_contiguous() and x.stride(1) == 1:
        return torch.empty((shape[0], shape[2], shape[1]), dtype=x.dtype, device=x.device).permute(0, 2, 1)
    return torch.empty(shape, dtype=x.dtype, device=x.device)


def _dense(x: torch.Tensor) -> torch.Tensor:
    """Any dense layout is consumed through strides; only overlapping / exotic views are copied."""
    if x.is_contiguous():
        return x
    if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last):
        return x
    if x.dim() == 5 and x.is_contiguous(memory_format=torch.channels_last_3d):
        return x
    if x.dim() == 3 and x.stride(1) == 1 and x.stride(2) == x.shape[1] and x.stride(0) == x.shape[1] * x.shape[2]:
        return x   # [B, C, L] view of a dense [B, L, C] buffer: channels-last for Conv1d
    return x.contiguous()


def _to_channels_last(call: "LayerCall", x: torch.Tensor) -> torch.Tensor:
    """The reference feeds contiguous NCHW activations (torch's default); the tensor-core kernels read pixels as GEMM
    rows, i.e. channels-last.  A 4-d NCHW input of a layer that may run on tensor cores is re-laid out ONCE here (one
    transposing copy); the layer's output is then channels-last, and BatchNorm / ReLU / pooling keep that format, so
    the layers downstream get channels-last inputs without further copies.  The fp32 kernels (path 'simt') read any
    strides and are left alone."""
    if (call.rank == 2 and call.path != _lib.PATH_SIMT_F32 and call.groups == 1 and x.dim() == 4 and x.shape[1] > 1
            and x.is_contiguous() and not x.is_contiguous(memory_format=torch.channels_last)):
        return x.contiguous(memory_format=torch.channels_last)
    return x


def _match_strides(t: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    """fp32 copy of t laid out exactly like ref (the kernels index signs with x's / out's offsets)."""
    t = t.detach().reshape(ref.shape)
    if t.dtype == torch.float32 and t.stride() == ref.stride():
        return t
    return torch.empty_like(ref, dtype=torch.float32).copy_(t)


_ZERO = {}
_PACK_CACHE_ON = os.environ.get("BT_PACK_CACHE", "1") != "0"


def invalidate_packed(model) -> None:
    """Drop the cached packed parameters of every frozen layer of `model` (see _BayesLayerFn.forward).  Needed only
    after writing a fold_batchnorm'd model's parameters through `.data`, which no version counter sees."""
    for m in model.modules():
        cache = getattr(m, "_pack_cache", None)
        if cache is not None:
            cache.clear()



def _zero_scalar(dev) -> torch.Tensor:
    """fp32 zero on `dev`, created once: what a layer returns for kl when return_kl is off."""
    key = (dev.type, dev.index)
    z = _ZERO.get(key)
    if z is None:
        z = _ZERO[key] = torch.zeros((), dtype=torch.float32, device=dev)
    return z


class _BayesLayerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, call: LayerCall, x, mu_w, rho_w, mu_b, rho_b):
        dev = _require_cuda(x, mu_w, rho_w, mu_b, rho_b)
        if x.dtype not in _DT:
            raise RuntimeError(f"activations must be float32 or bfloat16, got {x.dtype}")
        mu_wc, rho_wc = _f32c(mu_w.detach(), "mu"), _f32c(rho_w.detach(), "rho")
        mu_bc = None if mu_b is None else _f32c(mu_b.detach(), "mu_bias")
        rho_bc = None if rho_b is None else _f32c(rho_b.detach(), "rho_bias")
        xd = _to_channels_last(call, _dense(x.detach()))
        r = call.rank
        nb = xd.shape[0] * (call.n_samples if call.x_shared else 1)
        if r == 0:
            out_shape = (nb, mu_wc.shape[0])
        else:
            sp = [_conv_out_size(xd.shape[2 + i], mu_wc.shape[2 + i], call.stride[i], call.padding[i], call.dilation[i])
                  for i in range(r)]
            if min(sp) <= 0:
                raise RuntimeError(f"conv output size {sp} is not positive for input {tuple(xd.shape)}")
            out_shape = (nb, mu_wc.shape[0], *sp)
        if xd.shape[1] != mu_wc.shape[1] * call.groups:
            raise RuntimeError(f"input has {xd.shape[1]} channels/features, layer expects {mu_wc.shape[1] * call.groups}")
        out = _like_x_format(xd, out_shape)
        if call.sign_in is not None:
            call.sign_in = _match_strides(call.sign_in, xd)
        if call.sign_out is not None:
            call.sign_out = _match_strides(call.sign_out, out)
        if call.residual is not None:
            res = call.residual.detach()
            if tuple(res.shape) != tuple(out.shape):
                raise RuntimeError(f"residual has shape {tuple(res.shape)}, the layer output is {tuple(out.shape)}")
            if res.dtype != out.dtype or res.stride() != out.stride() or res.device != out.device:
                res = torch.empty_like(out).copy_(res)   # the kernels index it with out's offsets
            call.residual = res
        g = _build_geom(call, xd, mu_wc.shape, out)
        p, keep = _build_params(call, mu_wc, rho_wc, mu_bc, rho_bc)
        L = _lib.lib()
        op = b"flipout_fwd" if call.flipout else b"reparam_fwd"
        ws_bytes = max(int(L.bt_workspace_bytes(op, C.byref(g), 2)), 256)
        ws = None
        if call.pack_cache is not None and call.eps_w is None and _PACK_CACHE_ON:
            # The tensor-core forward packs (mu, softplus(rho)) * out_scale into the workspace before its main
            # launch.  A frozen layer keeps its workspace between calls and skips that launch while the
            # parameters (storage and version), the epilogue scale and the geometry are what they were.
            key = (mu_wc.data_ptr(), mu_w._version, rho_wc.data_ptr(), rho_w._version,
                   _ptr(call.out_scale), None if call.out_scale is None else call.out_scale._version,
                   bytes(g), ws_bytes, dev.index, _stream(), call.path, call.residual is None)
            hit = call.pack_cache.get("ws")
            if hit is not None and hit[0] == key:
                ws = hit[1]
                p.flags |= _lib.FLAG_PACKED_VALID
            else:
                ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
                call.pack_cache["ws"] = (key, ws)   # lives and dies with the layer object
        if ws is None:
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        p.workspace_bytes = ws.numel()
        kl = torch.empty((), dtype=torch.float32, device=dev) if call.want_kl else None
        with torch.cuda.device(dev):
            used = L.bt_select_path(C.byref(g), C.byref(p), int(call.flipout), call.path)
            if PROFILE.timing:
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record()
            rc = _entry(call, False)(C.byref(g), C.byref(p), _ptr(xd), _ptr(out), _ptr(kl), _ptr(ws), call.path, _stream())
            if PROFILE.timing:
                ev1.record()
                m = out.numel() // out.shape[1]
                kred = mu_wc.numel() // mu_wc.shape[0]
                flops = 2.0 * m * out.shape[1] * kred * (2 if call.flipout else 1)
                nbytes = xd.numel() * xd.element_size() + out.numel() * out.element_size() + 8 * mu_wc.numel()
                if call.residual is not None:
                    nbytes += call.residual.numel() * call.residual.element_size()
                PROFILE.records.append((used, flops, nbytes, ev0, ev1))
        _lib.check(rc, "layer forward")
        ctx.call = call
        ctx.out_meta = (tuple(out.shape), tuple(out.stride()))
        ctx.save_for_backward(x, mu_w, rho_w, mu_b, rho_b)
        if kl is None:
            kl = _zero_scalar(dev)   # a cached constant: no fill kernel per layer call
            ctx.mark_non_differentiable(kl)
        return out, kl

    @staticmethod
    def backward(ctx, g_out, g_kl):
        call: LayerCall = ctx.call
        if call.out_scale is not None or call.out_shift is not None or call.out_relu or call.residual is not None:
            raise RuntimeError("the folded BatchNorm / ReLU / residual epilogue is inference-only; unfuse the model to train")
        x, mu_w, rho_w, mu_b, rho_b = ctx.saved_tensors
        dev = x.device
        mu_wc, rho_wc = _f32c(mu_w.detach(), "mu"), _f32c(rho_w.detach(), "rho")
        mu_bc = None if mu_b is None else _f32c(mu_b.detach(), "mu_bias")
        rho_bc = None if rho_b is None else _f32c(rho_b.detach(), "rho_bias")
        xd = _to_channels_last(call, _dense(x.detach()))
        if g_out is None:  # only kl was used downstream
            g_out = torch.empty_strided(ctx.out_meta[0], ctx.out_meta[1], dtype=x.dtype, device=dev).zero_()
        S = call.n_samples
        if S != 1:
            raise RuntimeError("backward through mc_batched() is not supported; loop over mc_sample(s) for training")
        # grad_out must share dtype with x and have a dense layout; a gradient that arrives in another memory format
        # than the forward output had (autograd materialises unused outputs as contiguous zeros; some torch ops drop
        # channels-last) is re-laid out like the output, so the tensor-core backward still applies
        go = g_out.detach().to(x.dtype)
        if tuple(go.shape) == ctx.out_meta[0] and tuple(go.stride()) != ctx.out_meta[1]:
            go = torch.empty_strided(ctx.out_meta[0], ctx.out_meta[1], dtype=x.dtype, device=dev).copy_(go)
        go = _dense(go)
        geom = _build_geom(call, xd, mu_wc.shape, go)
        p, keep = _build_params(call, mu_wc, rho_wc, mu_bc, rho_bc)
        grads = LayerGrads()
        gx = torch.empty_like(xd) if ctx.needs_input_grad[1] else None
        need_w = ctx.needs_input_grad[2] or ctx.needs_input_grad[3]
        gmu = torch.empty_like(mu_wc) if need_w else None
        grho = torch.empty_like(rho_wc) if need_w else None
        need_b = mu_b is not None and (ctx.needs_input_grad[4] or ctx.needs_input_grad[5])
        gmb = torch.empty_like(mu_bc) if need_b else None
        grb = torch.empty_like(rho_bc) if need_b else None
        grads.grad_x, grads.grad_mu_w, grads.grad_rho_w = _ptr(gx), _ptr(gmu), _ptr(grho)
        grads.grad_mu_b, grads.grad_rho_b = _ptr(gmb), _ptr(grb)
        gk = None
        if call.want_kl and g_kl is not None:
            gk = g_kl.detach().to(torch.float32).contiguous()
        op = b"flipout_bwd" if call.flipout else b"reparam_bwd"
        ws = torch.empty(max(int(_lib.lib().bt_workspace_bytes(op, C.byref(geom), 2)), 256), dtype=torch.uint8, device=dev)
        p.workspace_bytes = ws.numel()
        with torch.cuda.device(dev):
            rc = _entry(call, True)(C.byref(geom), C.byref(p), _ptr(xd), _ptr(go), _ptr(gk), C.byref(grads), _ptr(ws),
                                    call.path, _stream())
        _lib.check(rc, "layer backward")
        if gx is not None and gx.shape != x.shape:
            gx = gx.view_as(x)
        return None, gx, gmu, grho, gmb, grb


class _BayesTransposeFn(torch.autograd.Function):
    """conv_transpose(x, W) + b with W = mu + softplus(rho) * eps and b = mu_b + softplus(rho_b) * eps_b sampled in
    the kernels (Flipout: the mean path plus the sign-flipped perturbation path, bias perturbation inside the sign
    product).  With E the convolution whose weight has W's shape [c_in, c_out / groups, *k] (E maps this layer's
    OUTPUT space to its INPUT space):
        forward      = E's backward-data  on grad_out := x, the sampled bias added in its epilogue
                       (bt_conv_*_bwd, grad_x only, BT_FLAG_TRANSPOSE_BIAS)
        grad_x       = E's forward        on x_E := grad_out           (bt_conv_*_fwd, no bias)
        grad_mu/rho  = E's backward-weights with x_E := grad_out, grad_out_E := x; the bias gradients reduce x_E
                       (BT_FLAG_BIAS_GRAD_X_SIDE)
    all three regenerate the same eps (and signs) from the (seed, sample, layer) key or read the same buffers."""

    @staticmethod
    def forward(ctx, call: LayerCall, x, mu_w, rho_w, mu_b, rho_b, oshape):
        dev = _require_cuda(x, mu_w, rho_w, mu_b, rho_b)
        if x.dtype not in _DT:
            raise RuntimeError(f"activations must be float32 or bfloat16, got {x.dtype}")
        mu_wc, rho_wc = _f32c(mu_w.detach(), "mu"), _f32c(rho_w.detach(), "rho")
        mu_bc = None if mu_b is None else _f32c(mu_b.detach(), "mu_bias")
        rho_bc = None if rho_b is None else _f32c(rho_b.detach(), "rho_bias")
        xd = _dense(x.detach())
        if xd.shape[1] != mu_wc.shape[0]:
            raise RuntimeError(f"input has {xd.shape[1]} channels, layer expects {mu_wc.shape[0]}")
        if min(oshape[2:]) <= 0:
            raise RuntimeError(f"conv_transpose output size {tuple(oshape[2:])} is not positive for input {tuple(xd.shape)}")
        for i in range(call.rank):  # E's forward arithmetic must map the output back onto x exactly
            if _conv_out_size(oshape[2 + i], mu_wc.shape[2 + i], call.stride[i], call.padding[i], call.dilation[i]) != xd.shape[2 + i]:
                raise RuntimeError("output_padding must be smaller than either stride or dilation")
        out = _like_x_format(xd, tuple(oshape))
        if call.sign_in is not None:    # E's input is our output, E's output is our input
            call.sign_in = _match_strides(call.sign_in, out)
        if call.sign_out is not None:
            call.sign_out = _match_strides(call.sign_out, xd)
        geom = _build_geom(call, out, mu_wc.shape, xd)
        p, keep = _build_params(call, mu_wc, rho_wc, mu_bc, rho_bc)
        p.flags |= _lib.FLAG_TRANSPOSE_BIAS
        grads = LayerGrads()
        grads.grad_x = _ptr(out)
        op = b"flipout_bwd" if call.flipout else b"reparam_bwd"
        ws = torch.empty(max(int(_lib.lib().bt_workspace_bytes(op, C.byref(geom), 2)), 256), dtype=torch.uint8, device=dev)
        p.workspace_bytes = ws.numel()
        with torch.cuda.device(dev):
            rc = _entry(call, True)(C.byref(geom), C.byref(p), _ptr(out), _ptr(xd), None, C.byref(grads), _ptr(ws),
                                    call.path, _stream())
        _lib.check(rc, "conv_transpose forward")
        ctx.call = call
        ctx.save_for_backward(x, mu_w, rho_w, mu_b, rho_b)
        return out

    @staticmethod
    def backward(ctx, g_out):
        call: LayerCall = ctx.call
        x, mu_w, rho_w, mu_b, rho_b = ctx.saved_tensors
        dev = x.device
        mu_wc, rho_wc = _f32c(mu_w.detach(), "mu"), _f32c(rho_w.detach(), "rho")
        mu_bc = None if mu_b is None else _f32c(mu_b.detach(), "mu_bias")
        rho_bc = None if rho_b is None else _f32c(rho_b.detach(), "rho_bias")
        xd = _dense(x.detach())
        go = _dense(g_out.detach().to(x.dtype))
        L = _lib.lib()
        gx = gmu = grho = gmb = grb = None
        with torch.cuda.device(dev):
            if ctx.needs_input_grad[1]:
                p, keep = _build_params(call, mu_wc, rho_wc, None, None)
                gx = torch.empty_like(xd)
                geom = _build_geom(call, go, mu_wc.shape, gx)
                op = b"flipout_fwd" if call.flipout else b"reparam_fwd"
                ws = torch.empty(max(int(L.bt_workspace_bytes(op, C.byref(geom), 2)), 256), dtype=torch.uint8, device=dev)
                p.workspace_bytes = ws.numel()
                rc = _entry(call, False)(C.byref(geom), C.byref(p), _ptr(go), _ptr(gx), None, _ptr(ws), call.path, _stream())
                _lib.check(rc, "conv_transpose backward (data)")
            need_w = ctx.needs_input_grad[2] or ctx.needs_input_grad[3]
            need_b = mu_b is not None and (ctx.needs_input_grad[4] or ctx.needs_input_grad[5])
            if need_w or need_b:
                p, keep = _build_params(call, mu_wc, rho_wc, mu_bc, rho_bc)
                p.flags |= _lib.FLAG_BIAS_GRAD_X_SIDE
                geom = _build_geom(call, go, mu_wc.shape, xd)
                grads = LayerGrads()
                if need_w:
                    gmu, grho = torch.empty_like(mu_wc), torch.empty_like(rho_wc)
                    grads.grad_mu_w, grads.grad_rho_w = _ptr(gmu), _ptr(grho)
                if need_b:
                    gmb, grb = torch.empty_like(mu_bc), torch.empty_like(rho_bc)
                    grads.grad_mu_b, grads.grad_rho_b = _ptr(gmb), _ptr(grb)
                op = b"flipout_bwd" if call.flipout else b"reparam_bwd"
                ws = torch.empty(max(int(L.bt_workspace_bytes(op, C.byref(geom), 2)), 256), dtype=torch.uint8, device=dev)
                p.workspace_bytes = ws.numel()
                rc = _entry(call, True)(C.byref(geom), C.byref(p), _ptr(go), _ptr(xd), None, C.byref(grads), _ptr(ws),
                                        call.path, _stream())
                _lib.check(rc, "conv_transpose backward (weights)")
        if gx is not None and gx.shape != x.shape:
            gx = gx.view_as(x)
        return None, gx, gmu, grho, gmb, grb, None


def bayes_transpose_layer(call: LayerCall, x, mu_w, rho_w, mu_b, rho_b, oshape):
    """conv_transpose(x, sampled W) + sampled bias of a ConvTranspose{1,2,3}d Bayesian layer, all in the kernels."""
    return _BayesTransposeFn.apply(call, x, mu_w, rho_w, mu_b, rho_b, tuple(int(v) for v in oshape))


def bayes_layer(call: LayerCall, x, mu_w, rho_w, mu_b, rho_b):
    """(out, kl) of one Bayesian layer call; kl is a zero constant when call.want_kl is False."""
    return _BayesLayerFn.apply(call, x, mu_w, rho_w, mu_b, rho_b)


# --------------------------------------------------------------------------- fused LSTM step


def _fill_lstm_layer(dst, call: LayerCall, mu_w, rho_w, mu_b, rho_b, keep):
    dst.mu_w, dst.rho_w, dst.mu_b, dst.rho_b = _ptr(mu_w), _ptr(rho_w), _ptr(mu_b), _ptr(rho_b)
    eps_w, eps_b = _f32c(call.eps_w, "eps_w"), _f32c(call.eps_b, "eps_b")
    s_in = None if call.sign_in is None else call.sign_in.detach().to(torch.float32).contiguous()
    s_out = None if call.sign_out is None else call.sign_out.detach().to(torch.float32).contiguous()
    keep += [eps_w, eps_b, s_in, s_out]
    dst.eps_w, dst.eps_b, dst.sign_in, dst.sign_out = _ptr(eps_w), _ptr(eps_b), _ptr(s_in), _ptr(s_out)
    dst.seed, dst.sample_id, dst.layer_id = call.seed, call.sample_id, call.layer_id


def _rows(t: torch.Tensor) -> torch.Tensor:
    """2-d fp32 view with unit column stride (rows may be strided: a time slice of [batch, seq, features])."""
    if t.dtype != torch.float32:
        raise RuntimeError(f"the fused LSTM step computes in float32, got {t.dtype}")
    return t if t.stride(1) == 1 else t.contiguous()


class _LSTMStepFn(torch.autograd.Function):
    """One time step of LSTMReparameterization / LSTMFlipout in ONE kernel (bt_lstm_step_fwd): both sampled
    contractions, the gate non-linearities and the cell update.  Backward: bt_lstm_step_bwd_gates (element-wise)
    and the two linear layers' own backward kernels on the pre-activation gate gradient, which regenerate the
    step's weights from the calls' keys (or read the calls' eps / sign tensors)."""

    @staticmethod
    def forward(ctx, calls, x, h_prev, c_prev, ih_mu_w, ih_rho_w, ih_mu_b, ih_rho_b, hh_mu_w, hh_rho_w, hh_mu_b, hh_rho_b):
        ih_call, hh_call = calls
        dev = _require_cuda(x, h_prev, c_prev, ih_mu_w, ih_rho_w, hh_mu_w, hh_rho_w)
        xd, hd, cd = _rows(x.detach()), _rows(h_prev.detach()), _rows(c_prev.detach())
        B, H = hd.shape
        params = [None if t is None else _f32c(t.detach(), "parameter") for t in
                  (ih_mu_w, ih_rho_w, ih_mu_b, ih_rho_b, hh_mu_w, hh_rho_w, hh_mu_b, hh_rho_b)]
        if params[0].shape != (4 * H, xd.shape[1]) or params[4].shape != (4 * H, H):
            raise RuntimeError(f"LSTM step: weights {tuple(params[0].shape)} / {tuple(params[4].shape)} do not match x "
                               f"{tuple(xd.shape)} and hidden {H}")
        h_t = torch.empty(B, H, dtype=torch.float32, device=dev)
        c_t = torch.empty(B, H, dtype=torch.float32, device=dev)
        need_grad = any(ctx.needs_input_grad)   # (grad mode is off inside forward; this is what autograd will ask for)
        gates = torch.empty(B, 4 * H, dtype=torch.float32, device=dev) if need_grad else None
        st = _lib.LstmStep()
        keep = []
        st.batch, st.in_features, st.hidden, st.flipout = B, xd.shape[1], H, int(ih_call.flipout)
        _fill_lstm_layer(st.ih, ih_call, *params[:4], keep)
        _fill_lstm_layer(st.hh, hh_call, *params[4:], keep)
        st.x, st.x_row_stride = _ptr(xd), xd.stride(0)
        st.h_prev, st.h_row_stride = _ptr(hd), hd.stride(0)
        st.c_prev, st.c_row_stride = _ptr(cd), cd.stride(0)
        st.h_out, st.h_out_row_stride, st.c_out, st.c_out_row_stride = _ptr(h_t), H, _ptr(c_t), H
        st.gates_out = _ptr(gates)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().bt_lstm_step_fwd(C.byref(st), _stream()), "bt_lstm_step_fwd")
        ctx.calls = calls
        ctx.gates = gates
        ctx.save_for_backward(x, h_prev, c_prev, c_t, ih_mu_w, ih_rho_w, ih_mu_b, ih_rho_b, hh_mu_w, hh_rho_w, hh_mu_b, hh_rho_b)
        return h_t, c_t

    @staticmethod
    def backward(ctx, g_h, g_c):
        x, h_prev, c_prev, c_t, *ps = ctx.saved_tensors
        dev = x.device
        B, H = c_t.shape
        gh = None if g_h is None else _rows(g_h.detach().to(torch.float32))
        gc = None if g_c is None else _rows(g_c.detach().to(torch.float32))
        cd = _rows(c_prev.detach())
        g_pre = torch.empty(B, 4 * H, dtype=torch.float32, device=dev)
        g_cprev = torch.empty(B, H, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = _lib.lib().bt_lstm_step_bwd_gates(B, H, _ptr(ctx.gates), _ptr(cd), cd.stride(0), _ptr(c_t), c_t.stride(0),
                                                   _ptr(gh), 0 if gh is None else gh.stride(0), _ptr(gc), 0 if gc is None else gc.stride(0),
                                                   _ptr(g_pre), _ptr(g_cprev), _stream())
        _lib.check(rc, "bt_lstm_step_bwd_gates")
        outs = []
        for li, (call, inp) in enumerate(zip(ctx.calls, (x, h_prev))):
            mu_w, rho_w, mu_b, rho_b = ps[4 * li:4 * li + 4]
            mu_wc, rho_wc = _f32c(mu_w.detach(), "mu"), _f32c(rho_w.detach(), "rho")
            mu_bc = None if mu_b is None else _f32c(mu_b.detach(), "mu_bias")
            rho_bc = None if rho_b is None else _f32c(rho_b.detach(), "rho_bias")
            xd = inp.detach().contiguous()   # (grad_x is written with x's strides: a dense copy of the time slice)
            call = LayerCall(rank=0, flipout=call.flipout, eps_w=call.eps_w, eps_b=call.eps_b, sign_in=call.sign_in, sign_out=call.sign_out,
                             seed=call.seed, sample_id=call.sample_id, layer_id=call.layer_id, want_kl=False, path=_lib.PATH_SIMT_F32)
            if call.sign_in is not None:
                call.sign_in = _match_strides(call.sign_in, xd)
                call.sign_out = _match_strides(call.sign_out, g_pre)
            geom = _build_geom(call, xd, mu_wc.shape, g_pre)
            p, keep = _build_params(call, mu_wc, rho_wc, mu_bc, rho_bc)
            grads = LayerGrads()
            base = 1 if li == 0 else 2                       # needs_input_grad index of x / h_prev
            gx = torch.empty_like(xd) if ctx.needs_input_grad[base] else None
            gmu, grho = torch.empty_like(mu_wc), torch.empty_like(rho_wc)
            gmb = None if mu_bc is None else torch.empty_like(mu_bc)
            grb = None if rho_bc is None else torch.empty_like(rho_bc)
            grads.grad_x, grads.grad_mu_w, grads.grad_rho_w = _ptr(gx), _ptr(gmu), _ptr(grho)
            grads.grad_mu_b, grads.grad_rho_b = _ptr(gmb), _ptr(grb)
            op = b"flipout_bwd" if call.flipout else b"reparam_bwd"
            ws = torch.empty(max(int(_lib.lib().bt_workspace_bytes(op, C.byref(geom), 2)), 256), dtype=torch.uint8, device=dev)
            p.workspace_bytes = ws.numel()
            with torch.cuda.device(dev):
                rc = _entry(call, True)(C.byref(geom), C.byref(p), _ptr(xd), _ptr(g_pre), None, C.byref(grads), _ptr(ws),
                                        _lib.PATH_SIMT_F32, _stream())
            _lib.check(rc, "LSTM step backward")
            outs.append((gx, gmu, grho, gmb, grb))
        (gx, *ih_g), (ghp, *hh_g) = outs
        return (None, gx, ghp, g_cprev, *ih_g, *hh_g)


def lstm_step(ih_call: LayerCall, hh_call: LayerCall, x, h_prev, c_prev, ih_params, hh_params):
    """(h_t, c_t) of one fused LSTM time step; *_params = (mu_w, rho_w, mu_b, rho_b) of the ih / hh sub-layers."""
    return _LSTMStepFn.apply((ih_call, hh_call), x, h_prev, c_prev, *ih_params, *hh_params)


# --------------------------------------------------------------------------- misc exports


def philox_normal(n: int, seed: int, sample_id: int, layer_id: int, kind: int, device="cuda") -> torch.Tensor:
    """The exact N(0,1) stream the fused kernels generate in registers (bt_philox_normal_fill)."""
    out = torch.empty(n, dtype=torch.float32, device=device)
    with torch.cuda.device(out.device):
        _lib.check(_lib.lib().bt_philox_normal_fill(_ptr(out), n, seed, sample_id, layer_id, kind, _stream()), "philox_normal")
    return out


def philox_sign(n: int, seed: int, sample_id: int, layer_id: int, kind: int, device="cuda") -> torch.Tensor:
    out = torch.empty(n, dtype=torch.float32, device=device)
    with torch.cuda.device(out.device):
        _lib.check(_lib.lib().bt_philox_sign_fill(_ptr(out), n, seed, sample_id, layer_id, kind, _stream()), "philox_sign")
    return out


def philox_sign_shape(shape, device, seed: int, sample_id: int, layer_id: int, kind: int) -> torch.Tensor:
    """+-1 tensor of logical shape [B, C, *spatial] holding the signs the kernels generate for each
    element: the sign streams run over an activation tensor in CHANNELS-LAST order ((b, spatial...), c),
    whatever its memory layout (csrc/bt_common.cuh: bt_sign_index)."""
    shape = tuple(int(v) for v in shape)
    n = 1
    for v in shape:
        n *= v
    flat = philox_sign(n, seed, sample_id, layer_id, kind, device)
    if len(shape) <= 2:
        return flat.view(shape)
    return flat.view((shape[0],) + shape[2:] + (shape[1],)).movedim(-1, 1)


def philox_sign_like(t: torch.Tensor, seed: int, sample_id: int, layer_id: int, kind: int) -> torch.Tensor:
    return philox_sign_shape(t.shape, t.device, seed, sample_id, layer_id, kind)


def philox_eps_weight(w_shape, seed: int, sample_id: int, layer_id: int, device="cuda") -> torch.Tensor:
    """eps tensor in the weight's logical layout [c_out, c_in/g, *k] holding the values the
    kernels use: the stream is indexed in packed (n, tap, channel) order (shared with the
    tensor-core path, which stages weights channel-fastest)."""
    n, c = int(w_shape[0]), int(w_shape[1])
    taps = 1
    for k in w_shape[2:]:
        taps *= int(k)
    flat = philox_normal(n * c * taps, seed, sample_id, layer_id, _lib.KIND_EPS_W, device)
    return flat.view(n, taps, c).permute(0, 2, 1).reshape(tuple(w_shape)).contiguous()


def add_relu_(y: torch.Tensor, r: torch.Tensor) -> torch.Tensor:
    """y <- relu(y + r) in one pass (the residual join of a ResNet block), in place."""
    _require_cuda(y, r)
    if torch.is_grad_enabled() and (y.requires_grad or r.requires_grad):
        return torch.relu(y + r)   # the raw kernel has no autograd node: differentiable graphs go through torch
    if y.shape != r.shape or y.dtype != r.dtype or y.stride() != r.stride() or y.dtype not in _DT:
        return torch.relu_(y.add_(r))
    if not (y.is_contiguous() or (y.dim() == 4 and y.is_contiguous(memory_format=torch.channels_last))):
        return torch.relu_(y.add_(r))
    with torch.cuda.device(y.device):
        _lib.check(_lib.lib().bt_add_relu(_ptr(y), _ptr(r), y.numel(), _DT[y.dtype], _stream()), "add_relu")
    return y


def maxpool2d(x: torch.Tensor, kernel_size: int, stride: int, padding: int) -> torch.Tensor:
    """MaxPool2d (floor mode, square window) on a bf16 channels-last tensor in one 16-byte-per-thread
    pass; anything else goes to torch."""
    if not (x.is_cuda and x.dim() == 4 and x.dtype == torch.bfloat16 and x.shape[1] % 8 == 0
            and x.is_contiguous(memory_format=torch.channels_last) and 2 * padding <= kernel_size) \
            or (torch.is_grad_enabled() and x.requires_grad):   # no autograd node on the raw kernel
        return torch.nn.functional.max_pool2d(x, kernel_size, stride, padding)
    B, Cc, H, W = x.shape
    oh, ow = (H + 2 * padding - kernel_size) // stride + 1, (W + 2 * padding - kernel_size) // stride + 1
    y = torch.empty((B, Cc, oh, ow), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().bt_maxpool2d_nhwc(_ptr(x), _ptr(y), B, Cc, H, W, kernel_size, stride, padding, _lib.BT_BF16,
                                                _stream()), "maxpool2d")
    return y


def moped_rho(w: torch.Tensor, delta: float) -> torch.Tensor:
    """rho = log(expm1(delta*|w|) + 1e-20)  (utils/util.py:63-69) as one kernel."""
    _require_cuda(w)
    wc = _f32c(w.detach(), "w")
    rho = torch.empty_like(wc)
    with torch.cuda.device(wc.device):
        _lib.check(_lib.lib().bt_moped_rho(_ptr(wc), _ptr(rho), wc.numel(), float(delta), _stream()), "moped_rho")
    return rho
