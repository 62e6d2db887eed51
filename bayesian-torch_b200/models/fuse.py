"""Inference-time graph rewrite: fold eval-mode BatchNorm, ReLU and the residual join into
the Bayesian layers' kernels.

The reference's ResNets follow every Bayesian convolution with BatchNorm (+ ReLU)
(models/bayesian/resnet_variational_large.py:97-119; torchvision.models.resnet after
dnn_to_bnn) and today run them as separate element-wise passes over the activations, which on
a B200 cost more than the convolutions themselves (30 of ResNet-50's 54 Bayesian layers are
bound by activation bandwidth).  fold_batchnorm() traces the model with
torch.fx (Bayesian layers are leaves) and rewrites

    y = relu(bn(conv(x)))            ->  y = conv(x)            [scale/shift/ReLU in the kernel epilogue]
    y = relu(bn(conv(x)) + r)        ->  y = conv(x, residual=r)   [r is read tile by tile by the conv's epilogue]
                                         (join_in_epilogue=False: y = add_relu_(conv(x), r), one separate join pass)

It is exact up to rounding: scale rides on the sampled weights, shift enters the fp32
accumulator, one rounding to the output dtype; the residual sum is rounded like the eager
graph rounds it.  Eval mode only; the result is a torch.fx.GraphModule
sharing no state with the input model unless inplace=True.
"""
from __future__ import annotations

import copy
import operator

import torch
import torch.fx as fx
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from ..layers.base_variational_layer import _BayesLayer

_BN = (nn.BatchNorm1d, nn.BatchNorm2d, nn.BatchNorm3d)


class _Tracer(fx.Tracer):
    def is_leaf_module(self, m, qualname):
        return isinstance(m, _BayesLayer) or super().is_leaf_module(m, qualname)


def _is_relu(node, modules):
    if node.op == "call_module":
        return isinstance(modules[node.target], nn.ReLU)
    if node.op == "call_function":
        return node.target in (F.relu, torch.relu, torch.relu_)
    if node.op == "call_method":
        return node.target in ("relu", "relu_")
    return False


def _is_add(node):
    if node.op == "call_function":
        return node.target in (operator.add, operator.iadd, torch.add) and len(node.args) == 2 and not node.kwargs
    if node.op == "call_method":
        return node.target in ("add", "add_") and len(node.args) == 2 and not node.kwargs
    return False


def bn_scale_shift(bn):
    """(scale, shift) with bn(y) == y * scale + shift in eval mode."""
    if bn.training or not bn.track_running_stats or bn.running_mean is None:
        raise RuntimeError("fold_batchnorm needs eval-mode BatchNorm with running statistics")
    var = bn.running_var.detach().float()
    mean = bn.running_mean.detach().float()
    w = bn.weight.detach().float() if bn.weight is not None else torch.ones_like(var)
    b = bn.bias.detach().float() if bn.bias is not None else torch.zeros_like(var)
    scale = w * torch.rsqrt(var + bn.eps)
    return scale.contiguous(), (b - mean * scale).contiguous()


def fold_batchnorm(model: nn.Module, inplace: bool = False, join_in_epilogue: bool = True) -> fx.GraphModule:
    if model.training:
        raise RuntimeError("fold_batchnorm is an inference-time rewrite: call model.eval() first")
    if not inplace:
        src = model
        model = copy.deepcopy(model)
        # the folded model REPLACES the original (same weights, same draws): its layers keep the Philox
        # stream ids a plain deepcopy would renew (_BayesLayer.__setstate__)
        for a, b in zip(model.modules(), src.modules()):
            if isinstance(a, _BayesLayer):
                a._layer_id, a._calls = b._layer_id, b._calls
    graph = _Tracer().trace(model)
    gm = fx.GraphModule(model, graph)
    modules = dict(gm.named_modules())
    calls = {}
    for n in graph.nodes:
        if n.op == "call_module":
            calls[n.target] = calls.get(n.target, 0) + 1
    folded = {"bn": 0, "relu": 0, "join": 0, "residual": 0}
    pools = 0
    for node in list(graph.nodes):   # square, floor-mode, undilated MaxPool2d -> the channels-last kernel
        if node.op == "call_module" and isinstance(modules.get(node.target), nn.MaxPool2d):
            mp = modules[node.target]
            k, st, pd = mp.kernel_size, mp.stride if mp.stride is not None else mp.kernel_size, mp.padding
            if all(isinstance(v, int) for v in (k, st, pd)) and mp.dilation == 1 and not mp.ceil_mode and not mp.return_indices:
                with graph.inserting_after(node):
                    new = graph.call_function(ops.maxpool2d, (node.args[0], k, st, pd))
                node.replace_all_uses_with(new)
                graph.erase_node(node)
                pools += 1
    for node in list(graph.nodes):
        if node.op != "call_module" or not isinstance(modules.get(node.target), _BayesLayer):
            continue
        layer = modules[node.target]
        if calls[node.target] != 1 or not layer.dnn_to_bnn_flag or layer._flipout:
            continue  # shared module, (out, kl) tuple output, or Flipout (its epilogue has the sign product)
        layer._frozen = True   # inference-only from here on: the packed parameters are kept between calls (ops._PACKED)
        users = list(node.users)
        if len(users) == 1 and users[0].op == "call_module" and isinstance(modules[users[0].target], _BN) \
                and calls[users[0].target] == 1 and layer._ep_scale is None:
            bn_node = users[0]
            scale, shift = bn_scale_shift(modules[bn_node.target])
            dev = layer._w()[0].device
            layer._ep_scale, layer._ep_shift = scale.to(dev), shift.to(dev)
            bn_node.replace_all_uses_with(node)
            graph.erase_node(bn_node)
            folded["bn"] += 1
        users = list(node.users)
        if len(users) != 1:
            continue
        u = users[0]
        if _is_relu(u, modules):
            layer._ep_relu = True
            u.replace_all_uses_with(node)
            graph.erase_node(u)
            folded["relu"] += 1
        elif _is_add(u) and len(u.users) == 1 and _is_relu(next(iter(u.users)), modules):
            other = u.args[1] if u.args[0] is node else u.args[0]
            if not isinstance(other, fx.Node) or other is node:
                continue
            relu = next(iter(u.users))
            if join_in_epilogue and not layer._ep_relu:
                # out = relu(bn(conv(x)) + identity): the identity tile is read by the conv's own
                # epilogue.  The conv must run after the identity branch (a downsample conv).
                order = {n: i for i, n in enumerate(graph.nodes)}
                if order[other] > order[node]:
                    other.append(node)
                node.kwargs = {**node.kwargs, "residual": other}
                layer._ep_relu = True
                relu.replace_all_uses_with(node)
                graph.erase_node(relu)
                graph.erase_node(u)
                folded["residual"] += 1
                continue
            with graph.inserting_after(u):
                joined = graph.call_function(ops.add_relu_, (node, other))
            relu.replace_all_uses_with(joined)
            graph.erase_node(relu)
            graph.erase_node(u)
            folded["join"] += 1
    graph.lint()
    gm.recompile()
    gm.folded = folded
    gm.pools_rewritten = pools
    return gm
