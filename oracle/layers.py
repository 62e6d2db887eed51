"""CPU nn.Module wrappers around the oracle functions, used ONLY as the timed CPU baseline
(bench.py cpu_baseline / --impl reference) and by tests.  TEST INFRASTRUCTURE, not product.

They restate the reference modules' forward on CPU with torch ops: eps ~ N(0,1) drawn per
call and shared by the batch (linear_variational.py:164, conv_variational.py:367), weight
materialised, F.linear / F.conv2d (linear_variational.py:181, conv_variational.py:384)."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import bt_oracle as O


class OracleLinearReparam(nn.Module):
    def __init__(self, d: nn.Linear, rho0: float = -3.0):
        super().__init__()
        self.mu_weight = nn.Parameter(d.weight.detach().clone())
        self.rho_weight = nn.Parameter(torch.full_like(d.weight, rho0))
        self.mu_bias = None if d.bias is None else nn.Parameter(d.bias.detach().clone())
        self.rho_bias = None if d.bias is None else nn.Parameter(torch.full_like(d.bias, rho0))

    def forward(self, x):
        eps_w = torch.randn_like(self.mu_weight)
        eps_b = None if self.mu_bias is None else torch.randn_like(self.mu_bias)
        out, _ = O.linear_reparam(x, self.mu_weight, self.rho_weight, self.mu_bias, self.rho_bias, eps_w, eps_b)
        return out

    def kl_loss(self):
        return O.kl_normal_from_rho(self.mu_weight, self.rho_weight, self.mu_bias, self.rho_bias)


class OracleConv2dReparam(nn.Module):
    def __init__(self, d: nn.Conv2d, rho0: float = -3.0):
        super().__init__()
        self.stride, self.padding, self.dilation, self.groups = d.stride, d.padding, d.dilation, d.groups
        self.mu_kernel = nn.Parameter(d.weight.detach().clone())
        self.rho_kernel = nn.Parameter(torch.full_like(d.weight, rho0))
        self.mu_bias = None if d.bias is None else nn.Parameter(d.bias.detach().clone())
        self.rho_bias = None if d.bias is None else nn.Parameter(torch.full_like(d.bias, rho0))

    def forward(self, x):
        eps_k = torch.randn_like(self.mu_kernel)
        eps_b = None if self.mu_bias is None else torch.randn_like(self.mu_bias)
        out, _ = O.conv_reparam(x, self.mu_kernel, self.rho_kernel, self.mu_bias, self.rho_bias, eps_k, eps_b,
                                self.stride, self.padding, self.dilation, self.groups)
        return out

    def kl_loss(self):
        return O.kl_normal_from_rho(self.mu_kernel, self.rho_kernel, self.mu_bias, self.rho_bias)


def oracle_dnn_to_bnn(m: nn.Module, rho0: float = -3.0) -> nn.Module:
    """Recursive in-place swap of Conv2d / Linear leaves (the walk of models/dnn_to_bnn.py:127-154)."""
    for name, child in list(m._modules.items()):
        if child is None:
            continue
        if child._modules:
            oracle_dnn_to_bnn(child, rho0)
        elif isinstance(child, nn.Conv2d):
            setattr(m, name, OracleConv2dReparam(child, rho0))
        elif isinstance(child, nn.Linear):
            setattr(m, name, OracleLinearReparam(child, rho0))
    return m


@torch.no_grad()
def oracle_mc_predict(model: nn.Module, x: torch.Tensor, num_mc: int):
    """The reference's evaluate() loop (examples/main_bayesian_cifar_dnn2bnn.py:542-556)."""
    outs = torch.stack([model(x) for _ in range(num_mc)])
    return O.mc_moments(outs)
