"""Base class of the B200-native Bayesian layers.

Mirrors the interface of the reference's BaseVariationalLayer_
(bayesian_torch/layers/base_variational_layer.py:42-100): the ``dnn_to_bnn_flag``
property, ``kl_div`` and ``get_kernel_size``.  The arithmetic runs in
libbt_b200.so; one generic implementation serves every Reparameterization /
Flipout class, the public classes only pin the reference's constructor
signatures and attribute names.
"""
from __future__ import annotations

import collections.abc
from itertools import repeat
from typing import Optional

import torch
import torch.nn as nn
from torch.nn import Parameter

from .. import _lib, ops


def get_kernel_size(x, n):
    """int -> n-tuple, iterable -> tuple (base_variational_layer.py:37-40)."""
    if isinstance(x, collections.abc.Iterable):
        return tuple(x)
    return tuple(repeat(x, n))


def _ntuple(v, n):
    t = get_kernel_size(v, n)
    if len(t) != n:
        raise ValueError(f"expected {n} values, got {t}")
    return tuple(int(i) for i in t)


class BaseVariationalLayer_(nn.Module):
    def __init__(self):
        super().__init__()
        self._dnn_to_bnn_flag = False

    @property
    def dnn_to_bnn_flag(self):
        return self._dnn_to_bnn_flag

    @dnn_to_bnn_flag.setter
    def dnn_to_bnn_flag(self, value):
        self._dnn_to_bnn_flag = value

    def kl_div(self, mu_q, sigma_q, mu_p, sigma_p, prior_type="normal"):
        """mean over elements of KL(N(mu_q, sigma_q) || N(mu_p, sigma_p))
        (base_variational_layer.py:55-72).  mu_p / sigma_p: python scalars or tensors shaped
        like mu_q.  prior_type None is read as 'normal' (upstream behaviour; the fork's
        Conv2d default of None raises there)."""
        if prior_type is None:
            prior_type = "normal"
        if prior_type not in ("normal", "laplace"):
            raise ValueError(f"Unknown prior_type: {prior_type}")
        if prior_type == "laplace":   # location 0 / scale 1 hard-coded by the fork (base_variational_layer.py:80-81)
            return ops.kl_normal_multi([mu_q], [sigma_q], q_is_sigma=True, prior_type="laplace")

        def split(p):
            if torch.is_tensor(p) and p.numel() > 1:
                return 0.0 if p is mu_p else 1.0, p.expand_as(mu_q)
            return float(p), None
        pm, pm_t = split(mu_p)
        ps, ps_t = split(sigma_p)
        return ops.kl_normal_multi([mu_q], [sigma_q], pm, ps, [pm_t], [ps_t], q_is_sigma=True)


class _BayesLayer(BaseVariationalLayer_):
    """Generic variational layer: rank 0 = Linear, 1..3 = ConvNd; Reparameterization or Flipout."""

    _rank = 0
    _flipout = False
    _transpose = False  # ConvTransposeNd: weight [c_in, c_out / groups, *k], the layer is the data-gradient of a conv
    _wname = "weight"  # parameter suffix: mu_weight / mu_kernel

    # per-layer overrides of the process-wide settings in ops (None = follow the global)
    sample_mode: Optional[str] = None
    path: Optional[str] = None

    def _setup(self, w_shape, bias: bool, prior_mean, prior_variance, mu_init, rho_init, bias_n=None):
        w = self._wname
        self.register_parameter("mu_" + w, Parameter(torch.empty(w_shape)))
        self.register_parameter("rho_" + w, Parameter(torch.empty(w_shape)))
        self.register_buffer("eps_" + w, torch.empty(w_shape), persistent=False)
        self.register_buffer("prior_weight_mu", torch.empty(w_shape), persistent=False)
        self.register_buffer("prior_weight_sigma", torch.empty(w_shape), persistent=False)
        n = w_shape[0] if bias_n is None else bias_n
        if bias:
            self.mu_bias = Parameter(torch.empty(n))
            self.rho_bias = Parameter(torch.empty(n))
            self.register_buffer("eps_bias", torch.empty(n), persistent=False)
            self.register_buffer("prior_bias_mu", torch.empty(n), persistent=False)
            self.register_buffer("prior_bias_sigma", torch.empty(n), persistent=False)
        else:
            self.register_parameter("mu_bias", None)
            self.register_parameter("rho_bias", None)
            self.register_buffer("eps_bias", None, persistent=False)
            self.register_buffer("prior_bias_mu", None, persistent=False)
            self.register_buffer("prior_bias_sigma", None, persistent=False)
        self._mu0, self._rho0 = float(mu_init), float(rho_init)
        self._layer_id = ops.next_layer_id()
        self._calls = 0
        self.sign_input = None   # 'buffer' mode: Flipout sign tensors supplied by the caller
        self.sign_output = None
        self.quant_prepare = False
        # inference-time epilogue folded in by models.fuse.fold_batchnorm (None / False = plain layer)
        self.register_buffer("_ep_scale", None, persistent=False)
        self.register_buffer("_ep_shift", None, persistent=False)
        self._ep_relu = False
        self.init_parameters()

    # -- reference API ---------------------------------------------------------------------
    def init_parameters(self):
        """Constant priors, mu ~ N(mu_init, 0.1), rho ~ N(rho_init, 0.1)
        (linear_variational.py:133-144)."""
        w = self._wname
        self.prior_weight_mu.fill_(self.prior_mean)
        self.prior_weight_sigma.fill_(self.prior_variance)
        getattr(self, "mu_" + w).data.normal_(mean=self._mu0, std=0.1)
        getattr(self, "rho_" + w).data.normal_(mean=self._rho0, std=0.1)
        if self.mu_bias is not None:
            self.prior_bias_mu.fill_(self.prior_mean)
            self.prior_bias_sigma.fill_(self.prior_variance)
            self.mu_bias.data.normal_(mean=self._mu0, std=0.1)
            self.rho_bias.data.normal_(mean=self._rho0, std=0.1)
        self._mark_constant_priors()

    def prepare(self):
        raise NotImplementedError("INT8 post-training quantisation stubs are outside this build's scope (SURVEY.md row 9)")

    @staticmethod
    def _mark_of(b):
        # `layer.prior_bias_mu.data = t` (utils/util.py:MOPED, Linear branch) changes neither id() nor _version,
        # only the storage: the mark carries data_ptr and shape too
        return (id(b), b._version, b.data_ptr(), tuple(b.shape))

    def _mark_constant_priors(self):
        self._prior_marks = {n: self._mark_of(b) for n, b in self._prior_buffers().items() if b is not None}

    def __getstate__(self):
        # (runs on the SOURCE of a deepcopy / pickle) remember which priors are still the constant fill: the
        # marks hold id() / data_ptr() of this object's buffers and mean nothing for the clone's
        state = self.__dict__.copy()
        marks = state.get("_prior_marks")
        if marks is not None:
            state["_prior_const_at_copy"] = [n for n, b in self._prior_buffers().items()
                                             if b is not None and marks.get(n) == self._mark_of(b)]
        return state

    def __setstate__(self, state):
        # copy.deepcopy / pickle: the clone is a NEW layer with its own Philox stream (clones sharing a
        # layer id would draw bit-identical eps, e.g. ModuleList([deepcopy(block) for ...]))
        const = state.pop("_prior_const_at_copy", None)
        super().__setstate__(state)
        if "_layer_id" in self.__dict__:
            self._layer_id = ops.next_layer_id()
            self._calls = 0
            self._pack_cache = None
        if const is not None:
            self._prior_marks = {n: self._mark_of(getattr(self, n)) for n in const}

    def _prior_buffers(self):
        return {n: getattr(self, n) for n in ("prior_weight_mu", "prior_weight_sigma", "prior_bias_mu", "prior_bias_sigma")}

    def _prior_tensors(self):
        """(w_mu, w_sigma, b_mu, b_sigma): a buffer is passed to the kernel only if someone
        replaced or modified it after init (utils/util.py:MOPED assigns tensor priors,
        :101-104); untouched constant buffers are folded into the two scalars."""
        out = []
        for n, b in self._prior_buffers().items():
            mark = self._prior_marks.get(n)
            changed = b is not None and (mark is None or mark != self._mark_of(b))
            out.append(b if changed else None)
        return tuple(out)

    def _apply(self, fn, *args, **kwargs):
        # .half()/.bfloat16()/.to(dtype) keep mu/rho/eps/prior in fp32: the kernels read fp32
        # masters and round the SAMPLED weight once (bf16 operands, fp32 accumulate).
        def keep_fp32(t):
            r = fn(t)
            if t.dtype == torch.float32 and r.is_floating_point() and r.dtype != torch.float32:
                r = r.to(torch.float32)
            # .to(memory_format=channels_last) would re-stride the 4-d / 5-d kernels; the CUDA path reads them
            # in the reference's [c_out, c_in/g, *k] order, so a re-strided kernel cost a copy on every call
            if r.dim() >= 4 and not r.is_contiguous():
                r = r.contiguous()
            return r
        const = {n: b is not None and self._prior_marks.get(n) == self._mark_of(b)
                 for n, b in self._prior_buffers().items()} if hasattr(self, "_prior_marks") else {}
        super()._apply(keep_fp32, *args, **kwargs)
        for n, was_const in const.items():  # moving devices re-creates buffers; constants stay constants
            if was_const:
                b = getattr(self, n)
                self._prior_marks[n] = self._mark_of(b)
        return self

    def _w(self):
        w = self._wname
        return getattr(self, "mu_" + w), getattr(self, "rho_" + w)

    def kl_loss(self):
        """kl(weight) + kl(bias), each a mean over elements (linear_variational.py:146-158)."""
        mu_w, rho_w = self._w()
        pt = self._prior_tensors()
        mus, rhos, pm, ps = [mu_w], [rho_w], [pt[0]], [pt[1]]
        if self.mu_bias is not None:
            mus.append(self.mu_bias); rhos.append(self.rho_bias); pm.append(pt[2]); ps.append(pt[3])
        if self._prior_kind() == "laplace":
            return ops.kl_normal_multi(mus, rhos, prior_type="laplace")
        return ops.kl_normal_multi(mus, rhos, float(self.prior_mean), float(self.prior_variance), pm, ps)

    def _prior_kind(self):
        """'normal' or 'laplace': the fork's prior_type argument (Linear / Conv2d Reparameterization only;
        None reads as 'normal')."""
        pt = getattr(self, "prior_type", None)
        if pt is None:
            return "normal"
        if pt not in ("normal", "laplace"):
            raise ValueError(f"Unknown prior_type: {pt}")
        return pt

    # -- forward ----------------------------------------------------------------------------
    def _conv_args(self):
        r = self._rank
        return (_ntuple(self.stride, r), _ntuple(self.padding, r), _ntuple(self.dilation, r), int(self.groups))

    def forward(self, input, return_kl=True, residual=None):
        """forward(input, return_kl) as in the reference.  residual (an extension used by
        fold_batchnorm; inference only) is an output-shaped tensor added after the folded
        BatchNorm and before the folded ReLU: the identity branch of a residual block."""
        if self.dnn_to_bnn_flag:
            return_kl = False
        if self._transpose:
            if residual is not None:
                raise RuntimeError("residual epilogue is not available on ConvTranspose layers")
            return self._forward_transpose(input, return_kl)
        mu_w, rho_w = self._w()
        r = self._rank
        mode = self.sample_mode or ops.get_sample_mode()
        x = input
        lead = None
        if r == 0:
            lead = x.shape[:-1]
            x = x.reshape(-1, x.shape[-1])
        elif x.dim() != r + 2:
            raise RuntimeError(f"{type(self).__name__} expects a {r + 2}-d input, got {tuple(x.shape)}")
        call = ops.LayerCall(rank=r, flipout=self._flipout)
        if r > 0:
            call.stride, call.padding, call.dilation, call.groups = self._conv_args()
        call.prior_mu, call.prior_sigma = float(self.prior_mean), float(self.prior_variance)
        call.prior_t = self._prior_tensors()
        call.seed, call.layer_id = ops.get_seed(), self._layer_id
        call.n_samples = ops.current_n_samples()
        # the leading dim BEFORE Linear flattens [B, T, F] -> [B*T, F]: a shared model input keeps batch B there
        call.x_shared = call.n_samples > 1 and ops._STATE.shared_batch is not None and input.shape[0] == ops._STATE.shared_batch
        laplace = self._prior_kind() == "laplace"   # its KL is a separate launch (kl_loss), not the fused normal-prior one
        call.want_kl = bool(return_kl) and not laplace
        call.path = ops._PATHS[self.path or ops.get_path()]
        call.out_scale, call.out_shift, call.out_relu = self._ep_scale, self._ep_shift, self._ep_relu
        call.residual = residual
        if getattr(self, "_frozen", False) and not torch.is_grad_enabled():
            if getattr(self, "_pack_cache", None) is None:
                self._pack_cache = {}
            call.pack_cache = self._pack_cache
        self._set_noise(call, x, mode)
        out, kl = ops.bayes_layer(call, x, mu_w, rho_w, self.mu_bias, self.rho_bias)
        if lead is not None:
            out = out.reshape(-1, *lead[1:], out.shape[-1])   # leading axis may be folded S*B
        if return_kl:
            return out, (self.kl_loss() if laplace else kl)
        return out

    def _set_noise(self, call, x, mode):
        """Fill the call's noise source for THIS call of the layer: Philox keys (sample id from the enclosing
        mc_sample / mc_batched block or the layer's own call counter) or, in the tensor-fed modes, the eps / sign
        tensors -- drawn with torch's generator in the reference's order ('torch') or taken from the buffers
        ('buffer').  x: the (flattened) input of the call."""
        mu_w, _ = self._w()
        r = self._rank
        if ops._STATE.sample_override is not None:
            call.sample_id = ops._STATE.sample_override
            call.seed = ops.call_seed(call.seed, self._layer_id)   # repeated calls inside the block: fresh streams
        else:
            call.sample_id = self._calls & 0xFFFFFFFF
            self._calls += call.n_samples
        eps_w = getattr(self, "eps_" + self._wname)
        if mode == "torch":
            ops._require_cuda(x)
            if self._flipout and r > 0:   # conv_flipout.py:385-386,390,401 draw order
                call.sign_in, call.sign_out = self._draw_signs(x, mu_w)
            call.eps_w = eps_w.normal_()   # linear_variational.py:164
            if self.mu_bias is not None:
                call.eps_b = self.eps_bias.normal_()
            if self._flipout and r == 0:  # linear_flipout.py:150,162 then :169-170
                call.sign_in, call.sign_out = self._draw_signs(x, mu_w)
        elif mode == "buffer":
            call.eps_w = eps_w
            call.eps_b = self.eps_bias
            if self._flipout:
                if self.sign_input is None or self.sign_output is None:
                    raise RuntimeError("sample mode 'buffer' on a Flipout layer needs layer.sign_input / layer.sign_output")
                call.sign_in, call.sign_out = self.sign_input, self.sign_output
        if torch.is_grad_enabled() and mode != "philox":
            # backward regenerates W (and the rho gradient) from eps: keep THIS call's draw, not the buffer a later
            # call of the same layer (weight sharing, LSTM steps) overwrites -- the reference's autograd saves the
            # buffer itself and reads the last draw there (oracle.lstm_unrolled, alias_eps)
            call.eps_w = None if call.eps_w is None else call.eps_w.clone()
            call.eps_b = None if call.eps_b is None else call.eps_b.clone()

    def _forward_transpose(self, x, return_kl):
        """ConvTranspose{1,2,3}d{Reparameterization,Flipout}.forward (conv_variational.py:941-989,
        conv_flipout.py:757-831): out = conv_transpose(x, mu + sigma * eps) + (mu_b + sigma_b * eps_b), Flipout
        likewise with the two paths.  The contraction is the DATA GRADIENT of the convolution whose weight has
        this layer's [c_in, c_out / groups, *k] shape, so it runs on the backward-data kernel of that
        convolution, which also adds the sampled bias in its epilogue (ops.bayes_transpose_layer); KL is the shared
        multi-tensor kernel."""
        r = self._rank
        if x.dim() != r + 2:
            raise RuntimeError(f"{type(self).__name__} expects a {r + 2}-d input, got {tuple(x.shape)}")
        if ops.current_n_samples() != 1:
            raise RuntimeError("ConvTranspose layers do not fold samples (mc_batched): loop over mc_sample(s)")
        mu_w, rho_w = self._w()
        mode = self.sample_mode or ops.get_sample_mode()
        call = ops.LayerCall(rank=r, flipout=self._flipout)
        call.stride, call.padding, call.dilation, call.groups = self._conv_args()
        call.output_padding = _ntuple(self.output_padding, r)
        call.seed, call.layer_id = ops.get_seed(), self._layer_id
        call.want_kl = False
        call.path = ops._PATHS[self.path or ops.get_path()]
        if ops._STATE.sample_override is not None:
            call.sample_id = ops._STATE.sample_override
            call.seed = ops.call_seed(call.seed, self._layer_id)
        else:
            call.sample_id = self._calls & 0xFFFFFFFF
            self._calls += 1
        eps_w = getattr(self, "eps_" + self._wname)
        ops._require_cuda(x)
        out_sp = [(x.shape[2 + i] - 1) * call.stride[i] - 2 * call.padding[i] + call.dilation[i] * (mu_w.shape[2 + i] - 1)
                  + call.output_padding[i] + 1 for i in range(r)]
        oshape = (x.shape[0], mu_w.shape[1] * call.groups, *out_sp)
        s_in = s_out = None
        if mode == "torch":
            if self._flipout:   # conv_flipout.py:776-777 then :781,:791 draw order
                s_in = torch.empty_like(x, dtype=torch.float32, memory_format=torch.contiguous_format).uniform_(-1, 1).sign()
                s_out = torch.empty(oshape, dtype=torch.float32, device=x.device).uniform_(-1, 1).sign()
            call.eps_w = eps_w.normal_()
            if self.mu_bias is not None:
                call.eps_b = self.eps_bias.normal_()
        elif mode == "buffer":
            call.eps_w = eps_w
            call.eps_b = self.eps_bias
            if self._flipout:
                if self.sign_input is None or self.sign_output is None:
                    raise RuntimeError("sample mode 'buffer' on a Flipout layer needs layer.sign_input / layer.sign_output")
                s_in, s_out = self.sign_input, self.sign_output
        # (in-kernel Philox: weights, bias draw and signs are all generated by the kernels; the equivalent convolution
        # reads THIS layer's output as its input, so our output signs are its SIGN_IN stream)
        if self._flipout and mode != "philox":
            call.sign_in, call.sign_out = s_out, s_in   # (equivalent convolution's view, see above)
        if torch.is_grad_enabled() and mode != "philox":   # this call's draw, not the buffer a later call overwrites
            call.eps_w = call.eps_w.clone()
            call.eps_b = None if call.eps_b is None else call.eps_b.clone()
        out = ops.bayes_transpose_layer(call, x, mu_w, rho_w, self.mu_bias, self.rho_bias, oshape)
        if return_kl:
            return out, self.kl_loss()
        return out

    def _draw_signs(self, x, mu_w):
        """sign(U(-1,1)) per element of x and of the output, drawn with torch's generator in
        the reference's order so a seeded run replays the reference's stream."""
        s_in = torch.empty_like(x, dtype=torch.float32, memory_format=torch.contiguous_format).uniform_(-1, 1).sign()
        if self._rank == 0:
            oshape = (x.shape[0], mu_w.shape[0])
        else:
            st, pd, dl, _ = self._conv_args()
            sp = [ops._conv_out_size(x.shape[2 + i], mu_w.shape[2 + i], st[i], pd[i], dl[i]) for i in range(self._rank)]
            oshape = (x.shape[0], mu_w.shape[0], *sp)
        s_out = torch.empty(oshape, dtype=torch.float32, device=x.device).uniform_(-1, 1).sign()
        return s_in, s_out
