"""Regression tests for the round-1 review findings: workspace sizing from the launch plan,
tensor priors assigned through `.data`, Philox stream identity of cloned / repeatedly-called layers, the
shared-input tag of mc_batched, and autograd behaviour of the raw pooling / join kernels."""
import copy
import ctypes as C
import pickle

import pytest
import torch
import torch.nn as nn

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import _lib, ops
from bayesian_torch_b200 import layers as L


# ------------------------------------------------------------------------------------------ CPU
def test_prior_assigned_through_data_is_seen_as_a_tensor_prior():
    """utils/util.py MOPED (Linear branch) writes `layer.prior_bias_mu.data = det_layer.bias`: neither id() nor
    _version of the buffer changes, only its storage."""
    lin = L.LinearReparameterization(6, 4)
    assert lin._prior_tensors() == (None, None, None, None)
    lin.prior_bias_mu.data = torch.arange(4.0)
    pt = lin._prior_tensors()
    assert pt[0] is None and pt[1] is None and pt[3] is None
    assert pt[2] is lin.prior_bias_mu
    lin.prior_weight_mu = torch.ones(4, 6)          # attribute assignment (utils/util.py:101-104)
    assert lin._prior_tensors()[0] is lin.prior_weight_mu
    lin.prior_weight_sigma.mul_(2.0)                # in-place edit
    assert lin._prior_tensors()[1] is lin.prior_weight_sigma


def test_clones_get_their_own_stream_and_keep_constant_priors():
    conv = L.Conv2dReparameterization(4, 8, 3, prior_type="normal")
    conv.prior_bias_sigma.data = torch.full((8,), 0.5)
    blocks = nn.ModuleList([copy.deepcopy(conv) for _ in range(3)])
    ids = {conv._layer_id} | {b._layer_id for b in blocks}
    assert len(ids) == 4, "deep copies must not share the (seed, sample, layer) Philox key"
    for b in blocks:
        pt = b._prior_tensors()
        assert pt[0] is None and pt[1] is None and pt[2] is None      # still the constant fill: scalar-prior kernel
        assert pt[3] is b.prior_bias_sigma                             # the tensor prior survives the copy as a tensor prior
        assert b._calls == 0
    clone = pickle.loads(pickle.dumps(conv))
    assert clone._layer_id not in ids
    assert torch.equal(clone.mu_kernel, conv.mu_kernel)


def test_call_seed_counts_calls_inside_mc_blocks_only():
    seed = 0x1234
    assert ops.call_seed(seed, 7) == seed                   # outside a block: the per-layer call counter moves the sample id
    with bt.mc_sample(3):
        a0, a1, b0, a2 = ops.call_seed(seed, 7), ops.call_seed(seed, 7), ops.call_seed(seed, 8), ops.call_seed(seed, 7)
    assert a0 == seed and b0 == seed
    assert a1 == ops.derive_seed(seed, 1) and a2 == ops.derive_seed(seed, 2) and len({a0, a1, a2}) == 3
    with bt.mc_batched(4, 0, shared_batch=2):               # a new block starts counting again
        assert ops.call_seed(seed, 7) == seed
        with bt.mc_sample(1):                               # nested blocks keep their own counters
            assert ops.call_seed(seed, 7) == seed
        assert ops.call_seed(seed, 7) == ops.derive_seed(seed, 1)
    assert ops.derive_seed(2 ** 64 - 1, 1) < 2 ** 64


def test_workspace_size_follows_the_launch_plan_without_a_gpu():
    """bt_workspace_bytes answers on a box without a GPU too (the plan is rejected there: packed area only)."""
    g = _lib.Geom()
    g.rank, g.groups, g.n_samples, g.x_dtype, g.out_dtype = 2, 1, 1, _lib.BT_BF16, _lib.BT_BF16
    g.batch, g.c_in, g.c_out = 2, 64, 64
    for i, v in enumerate((64, 64)):
        g.in_size[i], g.out_size[i], g.k_size[i], g.stride[i], g.pad[i], g.dil[i] = v, 4, 16, 16, 0, 1
    g.x_chan_stride, g.x_sp_stride[1], g.x_sp_stride[0], g.x_batch_stride = 1, 64, 64 * 64, 64 * 64 * 64
    g.out_chan_stride, g.out_sp_stride[1], g.out_sp_stride[0], g.out_batch_stride = 1, 64, 64 * 4, 64 * 16
    n = _lib.lib().bt_workspace_bytes(b"reparam_fwd", C.byref(g), 2)
    packed = 3 * 64 * (64 * 256) * 4
    assert n >= 256 + packed
    if torch.cuda.is_available():   # stride 16 > 8: the im2col map cannot describe it -> patch matrix must be counted
        assert n >= 256 + packed + 2 * 16 * (64 * 256) * 2


# ------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_stride16_conv_patch_matrix_fits_its_workspace(cuda_device):
    """Conv2d(64, 64, 16, stride=16), bf16 channels-last, path auto: c_in % 64 == 0 looks im2col-able to the
    old size function, but the plan falls back to the patch matrix (stride > 8)."""
    torch.manual_seed(5)
    layer = L.Conv2dReparameterization(64, 64, 16, stride=16, prior_type="normal").to(cuda_device)
    layer.sample_mode = "buffer"
    layer.eps_kernel.normal_(); layer.eps_bias.normal_()
    x = torch.randn(3, 64, 64, 64, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    guard = torch.zeros(1 << 20, device=cuda_device)      # neighbours of the workspace in the caching allocator
    layer.path = "tc"
    out, kl = layer(x)
    torch.cuda.synchronize()
    assert float(guard.abs().sum()) == 0.0
    layer.path = "simt"
    ref, _ = layer(x.float())
    err = float((out.float() - ref).abs().max() / ref.abs().max())
    assert err < 2e-2, err
    # an understated workspace is refused, not overrun
    g = ops._build_geom(ops.LayerCall(rank=2, flipout=False, stride=(16, 16), padding=(0, 0), dilation=(1, 1)), x,
                        layer.mu_kernel.shape, out)
    p, keep = ops._build_params(ops.LayerCall(rank=2, flipout=False), layer.mu_kernel.detach(), layer.rho_kernel.detach(), None, None)
    need = int(_lib.lib().bt_workspace_bytes(b"reparam_fwd", C.byref(g), 2))
    ws = torch.empty(need, dtype=torch.uint8, device=cuda_device)
    p.workspace_bytes = need // 2
    rc = _lib.lib().bt_conv_reparam_fwd(C.byref(g), C.byref(p), x.data_ptr(), out.data_ptr(), None, ws.data_ptr(), _lib.PATH_TC_BF16,
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == -1 and b"needed" in _lib.lib().bt_last_error()


@pytest.mark.gpu
def test_kl_uses_a_prior_assigned_through_data(cuda_device):
    torch.manual_seed(6)
    lin = L.LinearReparameterization(12, 9).to(cuda_device)
    pm = torch.randn(9, device=cuda_device)
    lin.prior_bias_mu.data = pm
    got = lin.kl_loss()
    sig = lambda r: torch.log1p(torch.exp(r))
    want = oracle.kl_normal(lin.mu_weight.cpu(), sig(lin.rho_weight.cpu()), torch.zeros(9, 12), torch.ones(9, 12)) + \
        oracle.kl_normal(lin.mu_bias.cpu(), sig(lin.rho_bias.cpu()), pm.cpu(), torch.ones(9))
    torch.testing.assert_close(got.cpu(), want.detach(), rtol=1e-5, atol=1e-6)
    out, kl = lin(torch.randn(4, 12, device=cuda_device))
    torch.testing.assert_close(kl.cpu(), want.detach(), rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
def test_lstm_steps_draw_fresh_weights_inside_mc_sample(cuda_device):
    """Inside mc_sample() the sample id is pinned; step t of the LSTM must still redraw (rnn_variational.py:126-127):
    call k of a layer uses the stream of derive_seed(seed, k)."""
    torch.manual_seed(7)
    bt.manual_seed(99)
    lstm = L.LSTMReparameterization(6, 5).to(cuda_device)
    for sub in (lstm.ih, lstm.hh):
        sub.path = "simt"
    x = torch.randn(3, 4, 6, device=cuda_device)
    with torch.no_grad(), bt.mc_sample(2):
        hs, _ = lstm(x, return_kl=False)
    # oracle twin fed with the per-step streams
    draws = []
    for t in range(4):
        d = {}
        for tag, sub in (("ih", lstm.ih), ("hh", lstm.hh)):
            sd = ops.derive_seed(99, t)
            d[tag] = (ops.philox_eps_weight(sub.mu_weight.shape, sd, 2, sub._layer_id, cuda_device).cpu(),
                      ops.philox_normal(sub.mu_bias.numel(), sd, 2, sub._layer_id, _lib.KIND_EPS_B, cuda_device).cpu())
        draws.append(d)
    assert not torch.equal(draws[0]["ih"][0], draws[1]["ih"][0])
    h = torch.zeros(3, 5); c = torch.zeros(3, 5)
    want = []
    xc = x.cpu()
    for t in range(4):
        gi, _ = oracle.linear_reparam(xc[:, t], lstm.ih.mu_weight.cpu(), lstm.ih.rho_weight.cpu(), lstm.ih.mu_bias.cpu(),
                                      lstm.ih.rho_bias.cpu(), *draws[t]["ih"])
        gh, _ = oracle.linear_reparam(h, lstm.hh.mu_weight.cpu(), lstm.hh.rho_weight.cpu(), lstm.hh.mu_bias.cpu(),
                                      lstm.hh.rho_bias.cpu(), *draws[t]["hh"])
        g = (gi + gh).detach()
        i_t, f_t, g_t, o_t = torch.sigmoid(g[:, :5]), torch.sigmoid(g[:, 5:10]), torch.tanh(g[:, 10:15]), torch.sigmoid(g[:, 15:])
        c = f_t * c + i_t * g_t
        h = o_t * torch.tanh(c)
        want.append(h)
    torch.testing.assert_close(hs.cpu(), torch.stack(want, 1), rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
def test_mc_batched_shared_sequence_input_through_linear(cuda_device):
    """A shared model input [B, T, F] is flattened to [B*T, F] by Linear: it must still be recognised as shared."""
    torch.manual_seed(8)
    bt.manual_seed(5)
    lin = L.LinearReparameterization(16, 8).to(cuda_device)
    lin.path = "simt"
    B, T, S = 3, 4, 4      # B*T = 12 is a multiple of S: the old size test silently read x as sample-folded
    x = torch.randn(B, T, 16, device=cuda_device)
    with torch.no_grad():
        with bt.mc_batched(S, 10, shared_batch=B):
            got = lin(x, return_kl=False)
        assert got.shape == (S * B, T, 8)
        for s in range(S):
            with bt.mc_sample(10 + s):
                one = lin(x, return_kl=False)
            torch.testing.assert_close(got[s * B:(s + 1) * B], one, rtol=1e-6, atol=1e-6)


@pytest.mark.gpu
def test_raw_pool_and_join_kernels_do_not_cut_autograd(cuda_device):
    x = torch.randn(2, 8, 9, 9, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = ops.maxpool2d(x, 3, 2, 1)
    y.float().sum().backward()
    assert x.grad is not None and float(x.grad.float().abs().sum()) > 0
    a = torch.randn(2, 8, 4, 4, device=cuda_device, requires_grad=True)
    r = torch.randn(2, 8, 4, 4, device=cuda_device)
    z = ops.add_relu_(a * 1.0, r)
    z.sum().backward()
    assert a.grad is not None
    with torch.no_grad():   # the raw kernels still serve inference
        y2 = ops.maxpool2d(x.detach(), 3, 2, 1)
    torch.testing.assert_close(y2, torch.nn.functional.max_pool2d(x.detach(), 3, 2, 1))


@pytest.mark.gpu
def test_workspace_without_room_for_presampled_weights_still_computes_the_same(cuda_device):
    """bt_workspace_bytes counts the pre-sampled weights of a streamed layer; a caller that states only the packed
    area (what the ABI-4 size function returned) gets the fused-sampling schedule and the same bits."""
    torch.manual_seed(8)
    bt.manual_seed(77)
    layer = L.Conv2dReparameterization(256, 256, 3, padding=1, bias=False, prior_type="normal").to(cuda_device)
    layer.path = "tc"
    layer.dnn_to_bnn_flag = True
    x = torch.randn(16, 256, 14, 14, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    with torch.no_grad(), bt.mc_sample(4):
        want = layer(x)                                       # 25 m-tiles, K = 2304: pre-sampled W, 256-column tiles
    call = ops.LayerCall(rank=2, flipout=False, stride=(1, 1), padding=(1, 1), dilation=(1, 1), seed=77, sample_id=4,
                         layer_id=layer._layer_id)
    out = torch.empty_like(want)
    g = ops._build_geom(call, x, layer.mu_kernel.shape, out)
    p, keep = ops._build_params(call, layer.mu_kernel.detach(), layer.rho_kernel.detach(), None, None)
    full = int(_lib.lib().bt_workspace_bytes(b"reparam_fwd", C.byref(g), 2))
    packed = 256 + ((3 * 256 * 2304 * 4 + 255) // 256) * 256  # KL scratch + (mu, sigma, eps) area
    assert full >= packed + 256 * 2304 * 2                    # the size function counts one sample of bf16 weights
    ws = torch.empty(full, dtype=torch.uint8, device=cuda_device)
    outs = []
    for stated in (full, packed):
        p.workspace_bytes = stated
        out = torch.empty_like(want)
        rc = _lib.lib().bt_conv_reparam_fwd(C.byref(g), C.byref(p), x.data_ptr(), out.data_ptr(), None, ws.data_ptr(), _lib.PATH_TC_BF16,
                                            torch.cuda.current_stream().cuda_stream)
        assert rc == 0, _lib.lib().bt_last_error()
        outs.append(out)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[1])
    assert torch.equal(outs[0], want)
