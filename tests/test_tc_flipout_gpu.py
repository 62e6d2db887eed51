"""Flipout on the tcgen05 / TMEM path: one kernel, two accumulators per output tile
(x . mu^T and (x * s_in) . (sigma * eps)^T), activations sign-flipped in shared memory,
s_out applied when the accumulators are joined in the epilogue.

The oracle (linear_flipout.py:145-174, conv_flipout.py:370-417 restated in oracle/bt_oracle.py)
is evaluated on the SAME rounded operands the kernel multiplies: x in bf16, mu in bf16,
delta_w = bf16(bf16(softplus(rho)) * bf16(eps)); what is left is fp32 accumulation order and the
bf16 rounding of the output: |err| <= 2^-8 * max|out| (same bound as the Reparameterization path)."""
import pytest
import torch

import bayesian_torch_b200 as bt
import oracle
from bayesian_torch_b200 import _lib, ops
from bayesian_torch_b200 import layers as L

pytestmark = pytest.mark.gpu
BF16_OUT_TOL = 2.0 ** -8


def bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)


def oracle_flipout_bf16(layer, x, conv_kw=None):
    mu, rho = (t.detach().cpu() for t in layer._w())
    eps = getattr(layer, "eps_" + layer._wname).cpu()
    sigma = oracle.sample_weight(torch.zeros_like(mu), rho, torch.ones_like(mu))
    mu_r = bf16_round(mu).double()
    dw = bf16_round((bf16_round(sigma).double() * bf16_round(eps).double()).float()).double()
    mb = db = None
    if layer.mu_bias is not None:
        mb = layer.mu_bias.detach().cpu().double()
        db = oracle.sample_weight(torch.zeros_like(layer.mu_bias).cpu(), layer.rho_bias.detach().cpu(), layer.eps_bias.cpu()).double()
    xr = bf16_round(x.cpu().float()).double()
    s_in, s_out = layer.sign_input.cpu().double(), layer.sign_output.cpu().double()
    if conv_kw is None:
        f = torch.nn.functional.linear
        return f(xr, mu_r, mb) + f(xr * s_in, dw, db) * s_out
    f = torch.nn.functional.conv2d
    return f(xr, mu_r, mb, **conv_kw) + f(xr * s_in, dw, db, **conv_kw) * s_out


def rand_signs(shape, device):
    return torch.empty(shape, device=device).uniform_(-1, 1).sign()


def run_tc(layer, x):
    layer.sample_mode, layer.path = "buffer", "tc"
    out, kl = layer(x)
    torch.cuda.synchronize()
    return out, kl


CONV_CASES = [
    # cin, cout, k, stride, pad, batch, hw, bias
    (64, 64, 1, 1, 0, 2, 14, True),     # 1x1 as a GEMM over pixels, BN=64
    (64, 64, 3, 1, 1, 3, 17, False),    # 3x3 im2col, padding rows (zero-filled taps need no signs), ragged M
    (64, 128, 1, 2, 0, 2, 16, True),    # strided 1x1 downsample (im2col mode), BN=128
    (128, 128, 3, 1, 1, 4, 14, True),   # K=1152: streaming W ring, two tiles per item
    (256, 160, 3, 2, 1, 2, 15, True),   # N tail (160 = 128 + 32), stride 2
    (512, 64, 3, 1, 1, 1, 7, False),    # deep K (72 k-blocks), tiny M
    (128, 256, 1, 1, 0, 4, 28, True),   # two n-blocks, many m-tiles per CTA
]


@pytest.mark.parametrize("cin,cout,k,stride,pad,batch,hw,bias", CONV_CASES)
def test_conv2d_flipout_tc_matches_oracle(cin, cout, k, stride, pad, batch, hw, bias, cuda_device):
    torch.manual_seed(cin * 1000 + cout + k)
    layer = L.Conv2dFlipout(cin, cout, k, stride=stride, padding=pad, bias=bias).to(cuda_device)
    layer.eps_kernel.normal_()
    if bias:
        layer.eps_bias.normal_()
    x = torch.randn(batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    ho = (hw + 2 * pad - k) // stride + 1
    layer.sign_input = rand_signs(x.shape, cuda_device)
    layer.sign_output = rand_signs((batch, cout, ho, ho), cuda_device)
    out, kl = run_tc(layer, x)
    assert out.dtype == torch.bfloat16 and out.is_contiguous(memory_format=torch.channels_last)
    want = oracle_flipout_bf16(layer, x, dict(stride=stride, padding=pad))
    err = float((out.cpu().double() - want).abs().max() / want.abs().max())
    assert err <= BF16_OUT_TOL, f"rel err {err}"
    wkl = oracle.kl_normal_from_rho(layer.mu_kernel.detach(), layer.rho_kernel.detach(),
                                    None if not bias else layer.mu_bias.detach(), None if not bias else layer.rho_bias.detach())
    assert float(kl) == pytest.approx(float(wkl), rel=1e-5)


@pytest.mark.parametrize("rows,k,n", [(300, 256, 192), (128, 512, 992), (2048, 1024, 1024), (5, 64, 32)])
def test_linear_flipout_tc_matches_oracle(rows, k, n, cuda_device):
    torch.manual_seed(rows + k + n)
    layer = L.LinearFlipout(k, n).to(cuda_device)
    layer.eps_weight.normal_(); layer.eps_bias.normal_()
    x = torch.randn(rows, k, device=cuda_device).bfloat16()
    layer.sign_input, layer.sign_output = rand_signs((rows, k), cuda_device), rand_signs((rows, n), cuda_device)
    out, _ = run_tc(layer, x)
    want = oracle_flipout_bf16(layer, x)
    err = float((out.cpu().double() - want).abs().max() / want.abs().max())
    assert err <= BF16_OUT_TOL, f"rel err {err}"


@pytest.mark.parametrize("kind", ["conv3x3", "conv1x1", "linear"])
def test_flipout_tc_philox_equals_exported_stream_and_simt(kind, cuda_device):
    """In-kernel Rademacher bits and normals on the tensor-core path == the exported streams fed through the
    deterministic mode == the same draw on the fp32 SIMT path (one RNG contract for both kernel families)."""
    torch.manual_seed(3)
    bt.manual_seed(77)
    if kind == "linear":
        layer = L.LinearFlipout(128, 96).to(cuda_device)
        x = torch.randn(70, 128, device=cuda_device).bfloat16()
    else:
        k = 3 if kind == "conv3x3" else 1
        layer = L.Conv2dFlipout(64, 128, k, padding=k // 2).to(cuda_device)
        x = torch.randn(2, 64, 12, 12, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    layer.sample_mode, layer.path = "philox", "tc"
    with bt.mc_sample(5):
        out_p, _ = layer(x)
    seed, lid = ops.get_seed(), layer._layer_id
    w = layer._w()[0]
    getattr(layer, "eps_" + layer._wname).copy_(ops.philox_eps_weight(w.shape, seed, 5, lid, cuda_device))
    layer.eps_bias.copy_(ops.philox_normal(w.shape[0], seed, 5, lid, _lib.KIND_EPS_B, cuda_device))
    layer.sign_input = ops.philox_sign_like(x, seed, 5, lid, _lib.KIND_SIGN_IN)
    layer.sign_output = ops.philox_sign_like(out_p, seed, 5, lid, _lib.KIND_SIGN_OUT)
    out_b, _ = run_tc(layer, x)
    assert float((out_p.float() - out_b.float()).abs().max()) <= 2.0 ** -7 * float(out_b.float().abs().max())
    layer.sample_mode, layer.path = "philox", "simt"
    with bt.mc_sample(5):
        out_s, _ = layer(x)
    err = float((out_p.float() - out_s.float()).abs().max() / out_s.float().abs().max())
    assert err < 2e-2          # bf16 weight rounding vs fp32 weights
    with bt.mc_sample(6):
        layer.path = "tc"
        other, _ = layer(x)
    assert not torch.allclose(other.float(), out_p.float(), atol=1e-2)


def test_flipout_tc_mc_batched_equals_sample_loop(cuda_device):
    torch.manual_seed(4)
    layer = L.Conv2dFlipout(64, 64, 3, padding=1).to(cuda_device)
    layer.path = "tc"
    layer.dnn_to_bnn_flag = True
    x = torch.randn(3, 64, 10, 10, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    loop = []
    for s in range(3):
        with bt.mc_sample(40 + s):
            loop.append(layer(x))
    with bt.mc_batched(3, 40, shared_batch=3):
        folded = layer(x)
    torch.cuda.synchronize()
    assert torch.equal(folded, torch.cat(loop))


@pytest.mark.parametrize("cin,cout,k,pad,batch,hw", [(128, 128, 3, 1, 8, 14), (256, 256, 3, 1, 16, 14), (512, 96, 3, 1, 6, 9)])
def test_flipout_cluster_pair_matches_single_cta(cin, cout, k, pad, batch, hw, cuda_device, monkeypatch):
    torch.manual_seed(cin + cout)
    layer = L.Conv2dFlipout(cin, cout, k, padding=pad).to(cuda_device)
    layer.dnn_to_bnn_flag = True
    layer.path = "tc"
    x = torch.randn(batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    outs = {}
    for cl in ("1", "2"):
        monkeypatch.setenv("BT_TC2_CL", cl)
        with torch.no_grad(), bt.mc_sample(3):
            outs[cl] = layer(x)
        torch.cuda.synchronize()
    assert float(outs["1"].float().abs().max()) > 0
    torch.testing.assert_close(outs["2"], outs["1"], rtol=0, atol=0)


@pytest.mark.parametrize("kind", ["stem", "linear_odd", "head1000"])
def test_flipout_tc_unaligned_shapes_match_oracle_on_exported_streams(kind, cuda_device):
    """Shapes whose tile rows are not aligned runs of the sign streams: C_in % 64 != 0 goes through two patch
    matrices (x and x * s_in, the RGB stem), c_out % 32 != 0 gathers s_out from two stream words.  The oracle
    is fed the exported Philox streams."""
    torch.manual_seed(8)
    bt.manual_seed(31)
    if kind == "stem":
        layer = L.Conv2dFlipout(3, 64, 7, stride=2, padding=3).to(cuda_device)
        x = torch.randn(2, 3, 32, 32, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
        kw = dict(stride=2, padding=3)
    elif kind == "linear_odd":
        layer, kw = L.LinearFlipout(100, 40).to(cuda_device), None
        x = torch.randn(37, 100, device=cuda_device).bfloat16()
    else:
        layer, kw = L.LinearFlipout(512, 1000).to(cuda_device), None
        x = torch.randn(128, 512, device=cuda_device).bfloat16()
    layer.sample_mode, layer.path = "philox", "tc"
    with bt.mc_sample(9):
        out, _ = layer(x)
    torch.cuda.synchronize()
    seed, lid = ops.get_seed(), layer._layer_id
    w = layer._w()[0]
    getattr(layer, "eps_" + layer._wname).copy_(ops.philox_eps_weight(w.shape, seed, 9, lid, cuda_device))
    layer.eps_bias.copy_(ops.philox_normal(w.shape[0], seed, 9, lid, _lib.KIND_EPS_B, cuda_device))
    layer.sign_input = ops.philox_sign_like(x, seed, 9, lid, _lib.KIND_SIGN_IN)
    layer.sign_output = ops.philox_sign_like(out, seed, 9, lid, _lib.KIND_SIGN_OUT)
    want = oracle_flipout_bf16(layer, x, kw)
    err = float((out.detach().cpu().double() - want).abs().max() / want.abs().max())
    assert err <= 2.0 ** -7, f"rel err {err}"     # + bf16 rounding of the in-kernel eps (exported stream is fp32)


def test_flipout_tc_rejects_what_it_cannot_run(cuda_device):
    layer = L.Conv2dFlipout(64, 64, 3, padding=1, groups=2).to(cuda_device)
    x = torch.randn(1, 64, 8, 8, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    layer.path = "tc"
    with pytest.raises(RuntimeError, match="tensor-core path unavailable"):
        layer(x)
    layer.path = "auto"
    out, _ = layer(x)
    assert out.shape == (1, 64, 8, 8)
