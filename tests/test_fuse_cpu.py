"""fold_batchnorm() graph surgery on CPU (no kernels run): which nodes disappear, what the
layers carry afterwards, and that training-mode models are refused."""
import pytest
import torch
import torch.nn as nn
import torchvision

import bayesian_torch_b200 as bt
from bayesian_torch_b200.models.fuse import bn_scale_shift


def test_resnet18_graph_is_folded():
    torch.manual_seed(0)
    m = torchvision.models.resnet18()
    bt.dnn_to_bnn(m, {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
                      "type": "Reparameterization", "moped_enable": False, "moped_delta": 0.5})
    with pytest.raises(RuntimeError, match="eval"):
        bt.fold_batchnorm(m)
    f = bt.fold_batchnorm(m.eval())
    # 20 convs each followed by a BatchNorm; conv1 of the stem and of every block also by a ReLU;
    # 8 residual joins, each read by the epilogue of the block's last conv
    assert f.folded == {"bn": 20, "relu": 9, "join": 0, "residual": 8}
    last = getattr(f.layer2, "0").conv2
    assert last._ep_relu                                      # ReLU after the residual sum
    calls = {n.target: n for n in f.graph.nodes if n.op == "call_module"}
    order = {n: i for i, n in enumerate(f.graph.nodes)}
    res = calls["layer2.0.conv2"].kwargs["residual"]
    assert res.target == "layer2.0.downsample.0" and order[res] < order[calls["layer2.0.conv2"]]
    g = bt.fold_batchnorm(m, join_in_epilogue=False)
    assert g.folded == {"bn": 20, "relu": 9, "join": 8, "residual": 0}
    targets = {n.target for n in f.graph.nodes if n.op == "call_module"}
    assert not any("bn" in t or "downsample.1" in t for t in targets)
    assert f.conv1._ep_relu and f.conv1._ep_scale.shape == (64,)
    assert not getattr(g.layer1, "0").conv2._ep_relu          # separate join pass: its ReLU comes after the join
    assert m.conv1._ep_scale is None                          # the input model is untouched


def test_bn_scale_shift_is_the_eval_affine():
    bn = nn.BatchNorm2d(5).eval()
    bn.running_mean.normal_(); bn.running_var.uniform_(0.5, 2); bn.weight.data.normal_(); bn.bias.data.normal_()
    x = torch.randn(2, 5, 3, 3)
    sc, sh = bn_scale_shift(bn)
    torch.testing.assert_close(bn(x), x * sc.view(1, -1, 1, 1) + sh.view(1, -1, 1, 1), rtol=1e-5, atol=1e-6)
