"""LinearFlipout on the B200-native path.

Drop-in for bayesian_torch/layers/flipout_layers/linear_flipout.py:49-197:
out = x.mu^T + mu_b + ((x*s_in).(sigma*eps)^T + sigma_b*eps_b) * s_out with per-element
Rademacher signs; both contractions run in one kernel.
"""
from ..base_variational_layer import _BayesLayer

__all__ = ["LinearFlipout"]


class LinearFlipout(_BayesLayer):
    _rank, _flipout, _wname = 0, True, "weight"

    def __init__(self, in_features, out_features, prior_mean=0, prior_variance=1, posterior_mu_init=0,
                 posterior_rho_init=-3.0, bias=True):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.prior_mean = prior_mean
        self.prior_variance = prior_variance
        self.posterior_mu_init = posterior_mu_init     # scalars here, unlike Reparameterization (:80-81)
        self.posterior_rho_init = posterior_rho_init
        self._setup((out_features, in_features), bias, prior_mean, prior_variance, posterior_mu_init, posterior_rho_init)
