"""Conv{1,2,3}dFlipout and ConvTranspose{1,2,3}dFlipout on the B200-native path.

Drop-ins for bayesian_torch/layers/flipout_layers/conv_flipout.py (Conv1d :57-244, Conv2d :247-439,
Conv3d :443-637, ConvTranspose1d :640-831, ConvTranspose2d :834-1030, ConvTranspose3d :1033-1228):
same constructor argument order (note `output_padding` sits after `groups` on ConvTranspose1dFlipout and
after `padding` on the 2-d / 3-d classes, as in the reference), same parameter names, forward(x, return_kl).
Two convolutions sharing x -- the mean path and the sign-flipped perturbation path -- fused into one
implicit-GEMM kernel; the transposed classes run on the backward-data kernel of the convolution with the
same weight (base_variational_layer._forward_transpose).
"""
from ..base_variational_layer import _BayesLayer, get_kernel_size

__all__ = ["Conv1dFlipout", "Conv2dFlipout", "Conv3dFlipout",
           "ConvTranspose1dFlipout", "ConvTranspose2dFlipout", "ConvTranspose3dFlipout"]


class _ConvFlipout(_BayesLayer):
    _flipout, _wname = True, "kernel"

    def _init_conv(self, in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                   prior_variance, posterior_mu_init, posterior_rho_init, bias, output_padding=None):
        if in_channels % groups != 0 or out_channels % groups != 0:
            raise ValueError('invalid in_channels size')
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.kernel_size = kernel_size
        self.stride = stride
        self.padding = padding
        self.dilation = dilation
        self.groups = groups
        if output_padding is not None:
            self.output_padding = output_padding
        self.prior_mean = prior_mean
        self.prior_variance = prior_variance
        self.posterior_mu_init = posterior_mu_init
        self.posterior_rho_init = posterior_rho_init
        self.bias = bias
        self.kl = 0
        k = get_kernel_size(kernel_size, self._rank)
        if self._transpose:
            self._setup((in_channels, out_channels // groups, *k), bias, prior_mean, prior_variance, posterior_mu_init,
                        posterior_rho_init, bias_n=out_channels)
        else:
            self._setup((out_channels, in_channels // groups, *k), bias, prior_mean, prior_variance, posterior_mu_init,
                        posterior_rho_init)


class _ConvNdFlipout(_ConvFlipout):
    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 prior_mean=0, prior_variance=1, posterior_mu_init=0, posterior_rho_init=-3.0, bias=True):
        super().__init__()
        self._init_conv(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                        prior_variance, posterior_mu_init, posterior_rho_init, bias)


class Conv1dFlipout(_ConvNdFlipout):
    _rank = 1


class Conv2dFlipout(_ConvNdFlipout):
    _rank = 2


class Conv3dFlipout(_ConvNdFlipout):
    _rank = 3


class ConvTranspose1dFlipout(_ConvFlipout):
    _rank, _transpose = 1, True

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1,
                 output_padding=0, prior_mean=0, prior_variance=1, posterior_mu_init=0, posterior_rho_init=-3.0,
                 bias=True):
        super().__init__()
        self._init_conv(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                        prior_variance, posterior_mu_init, posterior_rho_init, bias, output_padding=output_padding)


class _ConvTransposeNdFlipout(_ConvFlipout):
    _transpose = True

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, output_padding=0, dilation=1,
                 groups=1, prior_mean=0, prior_variance=1, posterior_mu_init=0, posterior_rho_init=-3.0, bias=True):
        super().__init__()
        self._init_conv(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, prior_mean,
                        prior_variance, posterior_mu_init, posterior_rho_init, bias, output_padding=output_padding)


class ConvTranspose2dFlipout(_ConvTransposeNdFlipout):
    _rank = 2


class ConvTranspose3dFlipout(_ConvTransposeNdFlipout):
    _rank = 3
