"""Host-side logic of the drop-in layers that needs no GPU: constructor signatures,
parameter / buffer names, state_dict keys, dnn_to_bnn swapping and MOPED init."""
import inspect

import pytest
import torch

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import layers as L
from conftest import load_golden

# the reference's signatures, argument by argument
SIGS = {
    "LinearReparameterization": ["in_features", "out_features", "prior_mean", "prior_variance", "posterior_mu_init",
                                 "posterior_rho_init", "bias", "prior_type"],
    "Conv1dReparameterization": ["in_channels", "out_channels", "kernel_size", "stride", "padding", "dilation",
                                 "groups", "prior_mean", "prior_variance", "posterior_mu_init", "posterior_rho_init", "bias"],
    "Conv2dReparameterization": ["in_channels", "out_channels", "kernel_size", "stride", "padding", "dilation",
                                 "groups", "prior_mean", "prior_variance", "prior_type", "posterior_mu_init",
                                 "posterior_rho_init", "bias"],
    "Conv3dReparameterization": ["in_channels", "out_channels", "kernel_size", "prior_mean", "prior_variance",
                                 "posterior_mu_init", "posterior_rho_init", "stride", "padding", "dilation", "groups", "bias"],
    "LinearFlipout": ["in_features", "out_features", "prior_mean", "prior_variance", "posterior_mu_init",
                      "posterior_rho_init", "bias"],
    "Conv2dFlipout": ["in_channels", "out_channels", "kernel_size", "stride", "padding", "dilation", "groups",
                      "prior_mean", "prior_variance", "posterior_mu_init", "posterior_rho_init", "bias"],
}
_CONV_FO = SIGS["Conv2dFlipout"]
SIGS["Conv1dFlipout"] = SIGS["Conv3dFlipout"] = _CONV_FO                                  # conv_flipout.py:58-70, :444-456
_CT_A = _CONV_FO[:7] + ["output_padding"] + _CONV_FO[7:]                                     # output_padding after groups
_CT_B = _CONV_FO[:5] + ["output_padding"] + _CONV_FO[5:]                                     # output_padding after padding
for _r in "123":                                                                             # conv_variational.py:822-1340
    SIGS[f"ConvTranspose{_r}dReparameterization"] = _CT_A
SIGS["ConvTranspose1dFlipout"] = _CT_A                                                       # conv_flipout.py:641-654
SIGS["ConvTranspose2dFlipout"] = SIGS["ConvTranspose3dFlipout"] = _CT_B                      # conv_flipout.py:835-848, :1034-1047


@pytest.mark.parametrize("name", sorted(SIGS))
def test_constructor_signature(name):
    params = list(inspect.signature(getattr(L, name).__init__).parameters)[1:]
    assert params == SIGS[name]


def test_names_shapes_and_state_dict():
    lin = L.LinearReparameterization(5, 7)
    assert sorted(lin.state_dict()) == ["mu_bias", "mu_weight", "rho_bias", "rho_weight"]  # buffers non-persistent
    assert lin.eps_weight.shape == lin.prior_weight_mu.shape == lin.mu_weight.shape == (7, 5)
    assert lin.posterior_mu_init == (0,) and lin.posterior_rho_init == (-3.0,)
    assert repr(lin) == "LinearReparameterization()"
    conv = L.Conv2dReparameterization(8, 12, (3, 2), groups=2, bias=False)
    assert conv.mu_kernel.shape == (12, 4, 3, 2) and conv.mu_bias is None and conv.eps_bias is None
    assert sorted(conv.state_dict()) == ["mu_kernel", "rho_kernel"]
    assert L.Conv1dReparameterization(4, 6, 3).mu_kernel.shape == (6, 4, 3)
    assert L.Conv1dReparameterization(4, 6, (3,)).mu_kernel.shape == (6, 4, 3)   # tuple from dnn_to_bnn works here
    assert L.Conv3dReparameterization(4, 6, 3, 0, 1, 0, -3.0).mu_kernel.shape == (6, 4, 3, 3, 3)
    fo = L.LinearFlipout(5, 7)
    assert fo.posterior_rho_init == -3.0   # scalar in Flipout layers
    assert float(lin.prior_weight_sigma[0, 0]) == 1.0 and float(lin.rho_weight.mean()) == pytest.approx(-3.0, abs=0.1)
    with pytest.raises(ValueError, match="invalid in_channels size"):
        L.Conv2dReparameterization(6, 8, 3, groups=4)
    lin.dnn_to_bnn_flag = True
    assert lin.dnn_to_bnn_flag is True


def test_transposed_and_flipout_conv_shapes():
    """ConvTranspose kernels are [c_in, c_out / groups, *k] with a c_out-sized bias (conv_variational.py:876-905)."""
    for r, k in ((1, (3,)), (2, (3, 2)), (3, (2, 3, 2))):
        for kind in ("Reparameterization", "Flipout"):
            t = getattr(L, f"ConvTranspose{r}d{kind}")(8, 12, k, groups=2)
            assert t.mu_kernel.shape == t.eps_kernel.shape == t.prior_weight_sigma.shape == (8, 6, *k)
            assert t.mu_bias.shape == t.prior_bias_mu.shape == (12,)
            assert t.output_padding == 0 and repr(t) == f"ConvTranspose{r}d{kind}()"
            assert sorted(t.state_dict()) == ["mu_bias", "mu_kernel", "rho_bias", "rho_kernel"]
    assert L.Conv1dFlipout(4, 6, 3).mu_kernel.shape == (6, 4, 3)
    assert L.Conv3dFlipout(4, 6, (3, 2, 1), bias=False).mu_kernel.shape == (6, 4, 3, 2, 1)
    with pytest.raises(ValueError, match="invalid in_channels size"):
        L.ConvTranspose2dFlipout(6, 8, 3, groups=4)
    net = torch.nn.Sequential(torch.nn.ConvTranspose2d(4, 6, 3, stride=2), torch.nn.Conv1d(3, 5, 3), torch.nn.Conv3d(2, 4, 3))
    params = {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
              "type": "Flipout", "moped_enable": True, "moped_delta": 0.5}
    w = net[0].weight.detach().clone()
    bt.dnn_to_bnn(net, params)   # dispatches on any class name containing "Conv" (dnn_to_bnn.py:131-139)
    assert [type(m).__name__ for m in net] == ["ConvTranspose2dFlipout", "Conv1dFlipout", "Conv3dFlipout"]
    torch.testing.assert_close(net[0].mu_kernel.detach(), w)   # MOPED copies the [c_in, c_out, k, k] weight as is


def test_dtype_casts_keep_fp32_masters():
    m = torch.nn.Sequential(L.Conv2dReparameterization(4, 4, 3), torch.nn.BatchNorm2d(4)).bfloat16()
    assert m[0].mu_kernel.dtype == torch.float32 and m[0].eps_kernel.dtype == torch.float32
    assert m[1].weight.dtype == torch.bfloat16
    assert m[0]._prior_tensors() == (None, None, None, None)      # constants stay scalars after _apply
    m = m.to(memory_format=torch.channels_last)                    # kernels keep the reference's [c_out, c_in, kh, kw] order
    assert m[0].mu_kernel.is_contiguous() and m[0].rho_kernel.is_contiguous() and m[0].eps_kernel.is_contiguous()
    m[0].prior_weight_mu = torch.zeros_like(m[0].mu_kernel)        # utils.util.MOPED style assignment
    assert m[0]._prior_tensors()[0] is m[0].prior_weight_mu


def test_dnn_to_bnn_swaps_by_name_and_moped_matches_reference():
    d = load_golden("dnn_to_bnn_small")
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.ReLU(), torch.nn.Conv2d(8, 8, 3, stride=2, bias=False),
                              torch.nn.Flatten(), torch.nn.Linear(8 * 3 * 3, 10))
    net.load_state_dict({k[4:]: v for k, v in d.items() if k.startswith("dnn_")})
    params = {"prior_mu": 0.0, "prior_sigma": 1.0, "posterior_mu_init": 0.0, "posterior_rho_init": -3.0,
              "type": "Reparameterization", "moped_enable": True, "moped_delta": 0.5}
    bt.dnn_to_bnn(net, params)
    assert [type(m).__name__ for m in net] == ["Conv2dReparameterization", "ReLU", "Conv2dReparameterization",
                                               "Flatten", "LinearReparameterization"]
    assert all(m.dnn_to_bnn_flag for m in net if hasattr(m, "dnn_to_bnn_flag"))
    sd = net.state_dict()
    assert sorted(sd) == sorted(k[4:] for k in d if k.startswith("bnn_"))
    for k, v in sd.items():  # same MOPED init as the reference's dnn_to_bnn (dnn_to_bnn.py:95-101)
        torch.testing.assert_close(v, d["bnn_" + k], rtol=1e-6, atol=1e-7)
    # and a reference checkpoint loads strictly
    net.load_state_dict({k[4:]: v for k, v in d.items() if k.startswith("bnn_")}, strict=True)
    fo = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3), torch.nn.Linear(4, 2))
    bt.dnn_to_bnn(fo, dict(params, type="Flipout", moped_enable=False))
    assert [type(m).__name__ for m in fo] == ["Conv2dFlipout", "LinearFlipout"]


def test_shard_samples_partition():
    for S in (1, 7, 32, 50):
        for G in (1, 2, 3, 4, 8):
            spans = [bt.shard_samples(S, r, G) for r in range(G)]
            assert spans[0][0] == 0 and spans[-1][1] == S
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        bt.shard_samples(4, 4, 4)
