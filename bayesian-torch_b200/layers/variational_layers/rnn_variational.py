"""LSTMReparameterization on the B200-native path.

Drop-in for bayesian_torch/layers/variational_layers/rnn_variational.py:46-153: an LSTM cell unrolled
in Python over two Bayesian linear layers, `ih` (in_features -> 4*out_features) and `hh`
(out_features -> 4*out_features), with a FRESH weight draw at every time step (each `self.ih(x_t)` call
samples, :126-127) and the per-step KL of both layers accumulated over the sequence (:131).  The two
GEMMs of a step run on the fused sample-GEMM kernels of LinearReparameterization; the gate
arithmetic is torch element-wise ops.  KL does not depend on the draw, so it is evaluated once per call
(one multi-tensor launch) and multiplied by the sequence length instead of T * 2 launches.
"""
import torch

from ..base_variational_layer import BaseVariationalLayer_
from .linear_variational import LinearReparameterization

__all__ = ["LSTMReparameterization"]


class _BayesLSTM(BaseVariationalLayer_):
    _linear = None  # the Bayesian linear class of ih / hh

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

    def forward(self, X, hidden_states=None, return_kl=True):
        if self.dnn_to_bnn_flag:
            return_kl = False
        batch_size, seq_size, _ = X.size()
        if hidden_states is None:
            h_t = torch.zeros(batch_size, self.out_features, device=X.device, dtype=X.dtype)
            c_t = torch.zeros(batch_size, self.out_features, device=X.device, dtype=X.dtype)
        else:
            h_t, c_t = hidden_states
        HS = self.out_features
        hidden_seq, c_ts = [], []
        for t in range(seq_size):
            gates = self.ih(X[:, t, :], return_kl=False) + self.hh(h_t, return_kl=False)
            i_t = torch.sigmoid(gates[:, :HS])
            f_t = torch.sigmoid(gates[:, HS:HS * 2])
            g_t = torch.tanh(gates[:, HS * 2:HS * 3])
            o_t = torch.sigmoid(gates[:, HS * 3:])
            c_t = f_t * c_t + i_t * g_t
            h_t = o_t * torch.tanh(c_t)
            hidden_seq.append(h_t.unsqueeze(0))
            c_ts.append(c_t.unsqueeze(0))
        # (sequence, batch, feature) -> (batch, sequence, feature)
        hidden_seq = torch.cat(hidden_seq, dim=0).transpose(0, 1).contiguous()
        c_ts = torch.cat(c_ts, dim=0).transpose(0, 1).contiguous()
        if return_kl:
            kl = self.kl_loss() * seq_size   # sum over the steps of kl_i + kl_h (:131); the terms are equal
            self.kl = kl
            return hidden_seq, (hidden_seq, c_ts), kl
        return hidden_seq, (hidden_seq, c_ts)


class LSTMReparameterization(_BayesLSTM):
    _linear = LinearReparameterization
