"""LSTMReparameterization on the B200-native path.

Drop-in for bayesian_torch/layers/variational_layers/rnn_variational.py:46-153: an LSTM cell unrolled
in Python over two Bayesian linear layers, `ih` (in_features -> 4*out_features) and `hh`
(out_features -> 4*out_features), with a FRESH weight draw at every time step (each `self.ih(x_t)` call
samples, :126-127) and the per-step KL of both layers accumulated over the sequence (:131).

A time step is ONE kernel here (ops.lstm_step -> bt_lstm_step_fwd, csrc/bt_lstm.cu): both contractions with
their weights sampled while they are staged, the gate non-linearities and the cell update; its backward is an
element-wise gate kernel plus the two linear layers' own backward kernels.  fp32 activations, one weight sample
per call; bf16 activations or sample-folded calls (mc_batched) run the step as two sample-GEMM launches of the
Bayesian linear layers plus element-wise gate arithmetic.  KL does not depend on the draw, so it is evaluated
once per call (one multi-tensor launch) and multiplied by the sequence length instead of T * 2 launches.
"""
import torch

from ... import ops
from ..base_variational_layer import BaseVariationalLayer_
from .linear_variational import LinearReparameterization

__all__ = ["LSTMReparameterization"]


class _BayesLSTM(BaseVariationalLayer_):
    _linear = None  # the Bayesian linear class of ih / hh
    fused_step = True   # False: always run a step as two linear-layer launches + element-wise gate ops

    def __init__(self, in_features, out_features, prior_mean=0, prior_variance=1, posterior_mu_init=0,
                 posterior_rho_init=-3.0, bias=True):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.prior_mean = prior_mean
        self.prior_variance = prior_variance
        self.posterior_mu_init = (posterior_mu_init,)    # 1-tuples, as the reference stores them (:78,:80)
        self.posterior_rho_init = (posterior_rho_init,)
        self.bias = bias
        self.kl = 0
        kw = dict(prior_mean=prior_mean, prior_variance=prior_variance, posterior_mu_init=posterior_mu_init,
                  posterior_rho_init=posterior_rho_init, bias=bias)
        self.ih = self._linear(in_features=in_features, out_features=out_features * 4, **kw)
        self.hh = self._linear(in_features=out_features, out_features=out_features * 4, **kw)

    def kl_loss(self):
        return self.ih.kl_loss() + self.hh.kl_loss()

    def _step_calls(self, x_t, h_t):
        """The noise sources of this step's ih / hh calls, set up exactly as the sub-layers' own forward would
        (their forward pre-hooks run too: a step IS a call of each)."""
        calls = []
        for sub, inp in ((self.ih, x_t), (self.hh, h_t)):
            for hook in list(sub._forward_pre_hooks.values()):
                hook(sub, (inp,))
            call = ops.LayerCall(rank=0, flipout=sub._flipout)
            call.seed, call.layer_id, call.n_samples = ops.get_seed(), sub._layer_id, 1
            sub._set_noise(call, inp, sub.sample_mode or ops.get_sample_mode())
            calls.append(call)
        return calls

    def forward(self, X, hidden_states=None, return_kl=True):
        if self.dnn_to_bnn_flag:
            return_kl = False
        batch_size, seq_size, _ = X.size()
        if hidden_states is None:
            h_t = torch.zeros(batch_size, self.out_features, device=X.device, dtype=X.dtype)
            c_t = torch.zeros(batch_size, self.out_features, device=X.device, dtype=X.dtype)
        else:
            h_t, c_t = hidden_states
        hs, cs = [], []
        fused = (self.fused_step and X.is_cuda and X.dtype == torch.float32 and h_t.dtype == torch.float32 and c_t.dtype == torch.float32
                 and ops.current_n_samples() == 1 and (self.ih.path or ops.get_path()) != "tc")
        for t in range(seq_size):
            x_t = X[:, t, :]
            if fused:
                h_t, c_t = ops.lstm_step(*self._step_calls(x_t, h_t), x_t, h_t, c_t,
                                         (self.ih.mu_weight, self.ih.rho_weight, self.ih.mu_bias, self.ih.rho_bias),
                                         (self.hh.mu_weight, self.hh.rho_weight, self.hh.mu_bias, self.hh.rho_bias))
            else:
                # two sample-GEMM launches + element-wise gates (rows of the [4H] axis: input, forget, cell, output)
                pre_i, pre_f, pre_g, pre_o = (self.ih(x_t, return_kl=False) + self.hh(h_t, return_kl=False)).chunk(4, dim=1)
                c_t = torch.sigmoid(pre_f) * c_t + torch.sigmoid(pre_i) * torch.tanh(pre_g)
                h_t = torch.sigmoid(pre_o) * torch.tanh(c_t)
            hs.append(h_t)
            cs.append(c_t)
        hidden_seq, c_ts = torch.stack(hs, dim=1), torch.stack(cs, dim=1)      # [batch, seq, hidden]
        if return_kl:
            kl = self.kl_loss() * seq_size   # sum over the steps of kl_i + kl_h (:131); the terms are equal
            self.kl = kl
            return hidden_seq, (hidden_seq, c_ts), kl
        return hidden_seq, (hidden_seq, c_ts)


class LSTMReparameterization(_BayesLSTM):
    _linear = LinearReparameterization
