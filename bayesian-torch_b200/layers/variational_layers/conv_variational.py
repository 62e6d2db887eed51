"""Conv{1,2,3}dReparameterization on the B200-native path.

Drop-ins for bayesian_torch/layers/variational_layers/conv_variational.py (Conv1d :68-231,
Conv2d :234-407, Conv3d :650-819): same constructor argument order -- including the
fork's `prior_type` between `prior_variance` and `posterior_mu_init` on Conv2d and the
positional prior/posterior arguments of Conv3d -- same parameter names (mu_kernel,
rho_kernel, mu_bias, rho_bias) and the same ValueError on bad groups.  `prior_type`
None is read as 'normal' (SURVEY.md section 0).  The convolution is an implicit GEMM
whose weight tile is sampled in shared memory.

ConvTranspose{1,2,3}dReparameterization (conv_variational.py:822-1340; weight
[in_channels, out_channels / groups, *k], `output_padding` after `groups`) run on the backward-data
kernel of the convolution with the same weight (base_variational_layer._forward_transpose).
"""
from ..base_variational_layer import _BayesLayer, get_kernel_size

__all__ = ["Conv1dReparameterization", "Conv2dReparameterization", "Conv3dReparameterization",
           "ConvTranspose1dReparameterization", "ConvTranspose2dReparameterization", "ConvTranspose3dReparameterization"]


class _ConvReparam(_BayesLayer):
    _flipout, _wname = False, "kernel"

    def _init_conv(self, in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                   prior_variance, posterior_mu_init, posterior_rho_init, bias, tuple_inits=True):
        if in_channels % groups != 0:
            raise ValueError('invalid in_channels size')
        if out_channels % groups != 0:
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
        self.posterior_mu_init = (posterior_mu_init,) if tuple_inits else posterior_mu_init
        self.posterior_rho_init = (posterior_rho_init,) if tuple_inits else posterior_rho_init
        self.bias = bias
        k = get_kernel_size(kernel_size, self._rank)
        if self._transpose:
            self._setup((in_channels, out_channels // groups, *k), bias, prior_mean, prior_variance, posterior_mu_init,
                        posterior_rho_init, bias_n=out_channels)
        else:
            self._setup((out_channels, in_channels // groups, *k), bias, prior_mean, prior_variance, posterior_mu_init,
                        posterior_rho_init)


class Conv1dReparameterization(_ConvReparam):
    _rank = 1

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 prior_mean=0, prior_variance=1, posterior_mu_init=0, posterior_rho_init=-3.0, bias=True):
        super().__init__()
        self._init_conv(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                        prior_variance, posterior_mu_init, posterior_rho_init, bias)


class Conv2dReparameterization(_ConvReparam):
    _rank = 2

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 prior_mean=0, prior_variance=1, prior_type=None, posterior_mu_init=0, posterior_rho_init=-3.0,
                 bias=True):
        super().__init__()
        self.prior_type = prior_type
        self._init_conv(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                        prior_variance, posterior_mu_init, posterior_rho_init, bias)


class Conv3dReparameterization(_ConvReparam):
    _rank = 3

    def __init__(self, in_channels, out_channels, kernel_size, prior_mean, prior_variance, posterior_mu_init,
                 posterior_rho_init, stride=1, padding=0, dilation=1, groups=1, bias=True):
        super().__init__()
        self._init_conv(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                        prior_variance, posterior_mu_init, posterior_rho_init, bias)


class _ConvTransposeReparam(_ConvReparam):
    _transpose = True

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 output_padding=0, prior_mean=0, prior_variance=1, posterior_mu_init=0, posterior_rho_init=-3.0,
                 bias=True):
        super().__init__()
        self.output_padding = output_padding
        self._init_conv(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                        prior_variance, posterior_mu_init, posterior_rho_init, bias)


class ConvTranspose1dReparameterization(_ConvTransposeReparam):
    _rank = 1


class ConvTranspose2dReparameterization(_ConvTransposeReparam):
    _rank = 2


class ConvTranspose3dReparameterization(_ConvTransposeReparam):
    _rank = 3
