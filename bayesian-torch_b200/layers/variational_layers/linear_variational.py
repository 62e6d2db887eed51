"""LinearReparameterization on the B200-native path.

Drop-in for bayesian_torch/layers/variational_layers/linear_variational.py:54-204:
same constructor arguments, parameter / buffer names (mu_weight, rho_weight, mu_bias,
rho_bias, eps_weight, eps_bias, prior_*), forward(x, return_kl=True) -> (out, kl) | out
and kl_loss().  w = mu + log1p(exp(rho)) * eps is formed tile by tile inside the GEMM
kernel and never written to HBM.
"""
from ..base_variational_layer import _BayesLayer

__all__ = ["LinearReparameterization"]


class LinearReparameterization(_BayesLayer):
    _rank, _flipout, _wname = 0, False, "weight"

    def __init__(self, in_features, out_features, prior_mean=0, prior_variance=1, posterior_mu_init=0,
                 posterior_rho_init=-3.0, bias=True, prior_type='normal'):
        super().__init__()
        assert prior_type is not None, "prior_type must be specified for LinearReparameterization layer"
        self.in_features = in_features
        self.out_features = out_features
        self.prior_mean = prior_mean
        self.prior_variance = prior_variance   # used as sigma, like the reference (:135)
        self.prior_type = prior_type
        # 1-tuples, as the reference stores them (trailing commas at :84,:87)
        self.posterior_mu_init = (posterior_mu_init,)
        self.posterior_rho_init = (posterior_rho_init,)
        self.bias = bias
        self._setup((out_features, in_features), bias, prior_mean, prior_variance, posterior_mu_init, posterior_rho_init)
