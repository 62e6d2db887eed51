"""dnn_to_bnn / get_kl_loss against the B200-native layers.

Same behaviour and parameter dict as bayesian_torch/models/dnn_to_bnn.py:52-165
({"prior_mu","prior_sigma","posterior_mu_init","posterior_rho_init","type",
"moped_enable","moped_delta"}): modules whose class name contains Conv / Linear are swapped
in place for "<TorchClass><type>" looked up BY NAME in the layers package, MOPED copies the
DNN weights into mu and get_rho(w, delta) into rho, and dnn_to_bnn_flag is set.
get_kl_loss returns the same sum of per-layer kl_loss() values, computed by ONE
multi-tensor kernel launch instead of a Python loop of ~12 launches per layer.
"""
from __future__ import annotations

from .. import layers as bayesian_layers
from .. import ops
from ..utils.util import get_rho


def _moped(params, d, bnn_layer, wname):
    if params.get("moped_enable", False):
        delta = params["moped_delta"]
        getattr(bnn_layer, "mu_" + wname).data.copy_(d.weight.data)
        getattr(bnn_layer, "rho_" + wname).data.copy_(get_rho(d.weight.data, delta))
        if bnn_layer.mu_bias is not None:
            bnn_layer.mu_bias.data.copy_(d.bias.data)
            bnn_layer.rho_bias.data.copy_(get_rho(d.bias.data, delta))


def _layer_class(d, params):
    name = d.__class__.__name__ + params["type"]
    if not hasattr(bayesian_layers, name):
        raise AttributeError(f"no B200-native layer named {name} (SURVEY.md section 8a lists the supported classes)")
    return getattr(bayesian_layers, name)


def bnn_linear_layer(params, d):
    bnn_layer = _layer_class(d, params)(
        in_features=d.in_features, out_features=d.out_features, prior_mean=params["prior_mu"],
        prior_variance=params["prior_sigma"], posterior_mu_init=params["posterior_mu_init"],
        posterior_rho_init=params["posterior_rho_init"], bias=d.bias is not None)
    _moped(params, d, bnn_layer, "weight")
    bnn_layer.dnn_to_bnn_flag = True
    return bnn_layer


def bnn_conv_layer(params, d):
    bnn_layer = _layer_class(d, params)(
        in_channels=d.in_channels, out_channels=d.out_channels, kernel_size=d.kernel_size, stride=d.stride,
        padding=d.padding, dilation=d.dilation, groups=d.groups, prior_mean=params["prior_mu"],
        prior_variance=params["prior_sigma"], posterior_mu_init=params["posterior_mu_init"],
        posterior_rho_init=params["posterior_rho_init"], bias=d.bias is not None)
    _moped(params, d, bnn_layer, "kernel")
    bnn_layer.dnn_to_bnn_flag = True
    return bnn_layer


def bnn_lstm_layer(params, d):
    """nn.LSTM -> LSTM<type> (dnn_to_bnn.py:103-121): input_size / hidden_size only; MOPED is not applied
    to LSTM layers there either."""
    bnn_layer = _layer_class(d, params)(
        in_features=d.input_size, out_features=d.hidden_size, prior_mean=params["prior_mu"],
        prior_variance=params["prior_sigma"], posterior_mu_init=params["posterior_mu_init"],
        posterior_rho_init=params["posterior_rho_init"], bias=d.bias is not None)
    if params.get("moped_enable", False):
        print("WARNING: MOPED method is not supported for LSTM layers!!!")
    bnn_layer.dnn_to_bnn_flag = True
    return bnn_layer


def dnn_to_bnn(m, bnn_prior_parameters):
    """In-place recursive swap (dnn_to_bnn.py:127-154)."""
    for name, child in list(m._modules.items()):
        if child is None:
            continue
        cls = child.__class__.__name__
        if child._modules:
            dnn_to_bnn(child, bnn_prior_parameters)
        elif "Conv" in cls:
            setattr(m, name, bnn_conv_layer(bnn_prior_parameters, child))
        elif "Linear" in cls:
            setattr(m, name, bnn_linear_layer(bnn_prior_parameters, child))
        elif "LSTM" in cls:
            setattr(m, name, bnn_lstm_layer(bnn_prior_parameters, child))
    return


def get_kl_loss(m):
    """Sum over modules having kl_loss() of weight-KL-mean + bias-KL-mean (dnn_to_bnn.py:157-165).
    Native layers are batched into one launch per distinct (prior_mu, prior_sigma, prior_type)."""
    groups = {}
    other = None
    for layer in m.modules():
        if not hasattr(layer, "kl_loss"):
            continue
        if isinstance(layer, bayesian_layers.base_variational_layer._BayesLayer):
            key = (float(layer.prior_mean), float(layer.prior_variance), layer._prior_kind())
            g = groups.setdefault(key, ([], [], [], []))
            mu_w, rho_w = layer._w()
            pt = layer._prior_tensors()
            g[0].append(mu_w); g[1].append(rho_w); g[2].append(pt[0]); g[3].append(pt[1])
            if layer.mu_bias is not None:
                g[0].append(layer.mu_bias); g[1].append(layer.rho_bias); g[2].append(pt[2]); g[3].append(pt[3])
        else:
            kl = layer.kl_loss()
            other = kl if other is None else other + kl
    total = other
    for (pm, ps, kind), (mus, rhos, pmt, pst) in groups.items():
        kl = ops.kl_normal_multi(mus, rhos, pm, ps, pmt, pst, prior_type=kind)
        total = kl if total is None else total + kl
    return total
