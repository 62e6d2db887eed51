"""Conv2dFlipout on the B200-native path.

Drop-in for bayesian_torch/layers/flipout_layers/conv_flipout.py:247-439 (constructor
:248-260, forward :370-439): two convolutions sharing x -- the mean path and the
sign-flipped perturbation path -- fused into one implicit-GEMM kernel.
"""
from ..base_variational_layer import _BayesLayer, get_kernel_size

__all__ = ["Conv2dFlipout"]


class Conv2dFlipout(_BayesLayer):
    _rank, _flipout, _wname = 2, True, "kernel"

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 prior_mean=0, prior_variance=1, posterior_mu_init=0, posterior_rho_init=-3.0, bias=True):
        super().__init__()
        if in_channels % groups != 0 or out_channels % groups != 0:
            raise ValueError('invalid in_channels size')
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.kernel_size = kernel_size
        self.stride = stride
        self.padding = padding
        self.dilation = dilation
        self.groups = groups
        self.prior_mean = prior_mean
        self.prior_variance = prior_variance
        self.posterior_mu_init = posterior_mu_init
        self.posterior_rho_init = posterior_rho_init
        self.bias = bias
        self.kl = 0
        k = get_kernel_size(kernel_size, 2)
        self._setup((out_channels, in_channels // groups, *k), bias, prior_mean, prior_variance, posterior_mu_init,
                    posterior_rho_init)
