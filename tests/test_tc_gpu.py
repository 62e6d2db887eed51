"""tcgen05 / TMEM path (bf16 operands, fp32 accumulate, weight sampled in shared memory)
against the CPU oracle.  The oracle is evaluated on the SAME rounded operands the kernel
multiplies -- x in bf16 and W = bf16(bf16(softplus(rho))*bf16(eps) + bf16(mu)) -- so the only differences
left are fp32 accumulation order and the bf16 rounding of the output: documented bound
|err| <= 2^-8 * max|out| for bf16 outputs (observed ~2e-3)."""
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


def oracle_bf16(layer, x, conv_kw=None):
    mu, rho = layer._w()
    eps = getattr(layer, "eps_" + layer._wname)
    # the tensor-core path's weight contract: mu, sigma = softplus(rho) and eps are rounded to bf16 (what
    # model.bfloat16() does to the reference's parameters), w = bf16(sigma * eps + mu) with one rounding
    mu_c, rho_c = mu.detach().cpu(), rho.detach().cpu()
    sigma = oracle.sample_weight(torch.zeros_like(mu_c), rho_c, torch.ones_like(mu_c))
    w = bf16_round((bf16_round(sigma).double() * bf16_round(eps.cpu()).double() + bf16_round(mu_c).double()).float())
    b = None
    if layer.mu_bias is not None:
        b = oracle.sample_weight(layer.mu_bias.detach().cpu(), layer.rho_bias.detach().cpu(), layer.eps_bias.cpu())
    xr = bf16_round(x.cpu().float())
    if conv_kw is None:
        return torch.nn.functional.linear(xr.double(), w.double(), None if b is None else b.double())
    return torch.nn.functional.conv2d(xr.double(), w.double(), None if b is None else b.double(), **conv_kw)


def run_tc(layer, x):
    layer.sample_mode, layer.path = "buffer", "tc"
    out, kl = layer(x)
    torch.cuda.synchronize()
    return out, kl


CONV_CASES = [
    # cin, cout, k, stride, pad, batch, hw, bias          what it exercises
    (64, 64, 1, 1, 0, 2, 14, True),     # 1x1, BN=64, stationary
    (64, 64, 3, 1, 1, 3, 17, False),    # 3x3 stationary panel (K=576), ragged M
    (64, 128, 1, 2, 0, 2, 16, True),    # strided 1x1 downsample, BN=128
    (128, 128, 3, 1, 1, 2, 14, True),   # K=1152 > smem panel -> streaming, multi-accumulator
    (256, 136, 3, 2, 1, 2, 15, True),   # N tail (136 = 128 + 8), stride 2, K=2304
    (16, 32, 3, 1, 1, 2, 9, True),      # C_in < 64: several taps per k-block, K tail (144 = 2*64 + 16)
    (8, 8, 5, 1, 2, 1, 11, False),      # 25 taps of 8 channels (patch-matrix path)
    (3, 64, 7, 2, 3, 2, 32, True),      # the RGB stem: K = 147 padded to 152, RNG stream pitch != GEMM pitch
    (1, 32, 3, 1, 0, 4, 28, True),      # SCNN conv1: a single input channel
    (128, 256, 1, 1, 0, 4, 28, True),   # 1x1, two n-blocks, several m-tiles per CTA (TMA path)
    (512, 64, 3, 1, 1, 1, 7, False),    # deep K (72 k-blocks) streaming, tiny M
    (64, 72, 3, 1, 1, 2, 20, True),     # N = 72: second n-block has 8 live columns
]


@pytest.mark.parametrize("cin,cout,k,stride,pad,batch,hw,bias", CONV_CASES)
def test_conv2d_tc_matches_oracle(cin, cout, k, stride, pad, batch, hw, bias, cuda_device):
    torch.manual_seed(cin * 1000 + cout + k)
    layer = L.Conv2dReparameterization(cin, cout, k, stride=stride, padding=pad, bias=bias, prior_type="normal").to(cuda_device)
    layer.eps_kernel.normal_()
    if bias:
        layer.eps_bias.normal_()
    x = torch.randn(batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    out, kl = run_tc(layer, x)
    assert out.dtype == torch.bfloat16 and out.is_contiguous(memory_format=torch.channels_last)
    want = oracle_bf16(layer, x, dict(stride=stride, padding=pad))
    err = float((out.cpu().double() - want).abs().max() / want.abs().max())
    assert err <= BF16_OUT_TOL, f"rel err {err}"
    wkl = oracle.kl_normal_from_rho(layer.mu_kernel.detach(), layer.rho_kernel.detach(),
                                    None if not bias else layer.mu_bias.detach(), None if not bias else layer.rho_bias.detach())
    assert float(kl) == pytest.approx(float(wkl), rel=1e-5)     # KL stays fp32 whatever the GEMM precision


def _conv_generic(cuda_device, cin, cout, ksz, stride, pad, dil, batch, hw, out_f32=False):
    torch.manual_seed(cin + cout)
    layer = L.Conv2dReparameterization(cin, cout, ksz, stride=stride, padding=pad, dilation=dil, bias=True,
                                       prior_type="normal").to(cuda_device)
    layer.eps_kernel.normal_(); layer.eps_bias.normal_()
    x = torch.randn(batch, cin, *hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    out, _ = run_tc(layer, x)
    want = oracle_bf16(layer, x, dict(stride=stride, padding=pad, dilation=dil))
    err = float((out.cpu().double() - want).abs().max() / want.abs().max())
    assert err <= BF16_OUT_TOL, f"rel err {err}"


@pytest.mark.parametrize("ksz,stride,pad,dil,hw", [
    ((3, 1), (1, 1), (1, 0), (1, 1), (13, 10)),   # H-only filter: catches a swapped (W,H) corner / offset order
    ((1, 5), (1, 2), (0, 2), (1, 1), (9, 21)),    # W-only filter, W stride 2
    ((3, 3), (2, 1), (1, 1), (1, 1), (16, 11)),   # anisotropic stride
    ((3, 3), (1, 1), (2, 2), (2, 2), (12, 12)),   # dilation 2
    ((3, 2), (1, 1), (0, 1), (1, 2), (8, 9)),     # even filter width, mixed padding / dilation
])
def test_conv2d_tma_im2col_geometry(ksz, stride, pad, dil, hw, cuda_device):
    _conv_generic(cuda_device, 64, 64, ksz, stride, pad, dil, 3, hw)


def test_conv1d_tc_matches_oracle(cuda_device):
    torch.manual_seed(11)
    layer = L.Conv1dReparameterization(64, 128, 3, stride=2, padding=1).to(cuda_device)
    layer.eps_kernel.normal_(); layer.eps_bias.normal_()
    x = torch.randn(4, 37, 64, device=cuda_device).bfloat16().permute(0, 2, 1)   # [B, C, L] with channels fastest
    out, _ = run_tc(layer, x)
    w = bf16_round(oracle.sample_weight(layer.mu_kernel.detach().cpu(), layer.rho_kernel.detach().cpu(), layer.eps_kernel.cpu()))
    b = oracle.sample_weight(layer.mu_bias.detach().cpu(), layer.rho_bias.detach().cpu(), layer.eps_bias.cpu())
    want = torch.nn.functional.conv1d(bf16_round(x.cpu().float()).double(), w.double(), b.double(), stride=2, padding=1)
    err = float((out.cpu().double() - want).abs().max() / want.abs().max())
    assert err <= BF16_OUT_TOL, f"rel err {err}"


@pytest.mark.parametrize("rows,k,n", [(300, 256, 192), (128, 2048, 1000), (4096, 1024, 1024), (5, 64, 8), (77, 72, 40),
                                      (33, 100, 16)])
def test_linear_tc_matches_oracle(rows, k, n, cuda_device):
    torch.manual_seed(rows + k + n)
    layer = L.LinearReparameterization(k, n).to(cuda_device)
    layer.eps_weight.normal_(); layer.eps_bias.normal_()
    x = torch.randn(rows, k, device=cuda_device).bfloat16()
    out, _ = run_tc(layer, x)
    want = oracle_bf16(layer, x)
    err = float((out.cpu().double() - want).abs().max() / want.abs().max())
    assert err <= BF16_OUT_TOL, f"rel err {err}"


@pytest.mark.parametrize("cin", [64, 3])
def test_tc_philox_equals_exported_stream_and_simt(cin, cuda_device):
    """In-kernel Philox on the tensor-core path == the exported stream fed through the
    deterministic mode == the same draw on the fp32 SIMT path (one RNG contract).  cin=3 runs
    the patch-matrix path, whose padded K makes weight chunks straddle Philox blocks."""
    torch.manual_seed(3)
    bt.manual_seed(99)
    layer = L.Conv2dReparameterization(cin, 128, 3, padding=1, prior_type="normal").to(cuda_device)
    x = torch.randn(2, cin, 12, 12, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    layer.sample_mode, layer.path = "philox", "tc"
    with bt.mc_sample(5):
        out_p, _ = layer(x)
    layer.eps_kernel.copy_(ops.philox_eps_weight(layer.mu_kernel.shape, ops.get_seed(), 5, layer._layer_id, cuda_device))
    layer.eps_bias.copy_(ops.philox_normal(128, ops.get_seed(), 5, layer._layer_id, _lib.KIND_EPS_B, cuda_device))
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


def test_tc_mc_batched_equals_sample_loop(cuda_device):
    torch.manual_seed(4)
    layer = L.Conv2dReparameterization(64, 64, 3, padding=1, bias=False, prior_type="normal").to(cuda_device)
    layer.path = "tc"
    layer.dnn_to_bnn_flag = True
    x = torch.randn(3, 64, 10, 10, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    loop = []
    for s in range(4):
        with bt.mc_sample(40 + s):
            loop.append(layer(x))
    with bt.mc_batched(4, 40, shared_batch=3):
        folded = layer(x)
    torch.cuda.synchronize()
    assert torch.equal(folded, torch.cat(loop))
    with bt.mc_batched(4, 40):                 # per-sample inputs, as deeper layers see them
        again = layer(torch.cat([x] * 4).contiguous(memory_format=torch.channels_last))
    assert torch.equal(again, folded)


@pytest.mark.parametrize("cin,cout,k,stride,pad,batch,hw,S", [
    (3, 64, 7, 2, 3, 3, 36, 40),      # RGB stem through the patch matrix: 3 k-blocks, ragged last tile
    (64, 96, 1, 1, 0, 5, 20, 24),     # 1x1 (plain GEMM map), one k-block, two tiles kept per item
    (64, 64, 1, 1, 0, 5, 20, 6),      # 64 output channels: two samples stacked on the N axis of one tile (sampled bias per sample)
    (64, 64, 3, 1, 1, 3, 17, 7),      # odd sample count: no stacking
    (128, 64, 3, 2, 1, 2, 30, 8),     # K = 1152: too long for the kept ring -> the usual order
    (3, 64, 7, 2, 3, 8, 64, 4),       # the RGB stem (patch matrix, K padded 147 -> 152): 64 m-tiles per sample -> pre-sampled W by default
    (64, 128, 3, 1, 1, 6, 20, 3),     # 19 m-tiles per sample, odd sample count, N = 128
])
def test_tc_shared_input_sample_fastest_order(cuda_device, cin, cout, k, stride, pad, batch, hw, S, monkeypatch):
    """x shared by all samples: the (m-group, n-block, sample) order with the A tiles kept in shared
    memory gives bit-identical outputs to the per-sample loop and to the (sample, ...) order."""
    torch.manual_seed(21)
    layer = L.Conv2dReparameterization(cin, cout, k, stride=stride, padding=pad, bias=True, prior_type="normal").to(cuda_device)
    layer.path = "tc"
    layer.dnn_to_bnn_flag = True
    x = torch.randn(batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        loop = []
        for s in range(S):
            with bt.mc_sample(11 + s):
                loop.append(layer(x))
        monkeypatch.setenv("BT_TC2_SINNER", "1")
        with bt.mc_batched(S, 11, shared_batch=batch):
            folded = layer(x)
        monkeypatch.setenv("BT_TC2_SINNER", "0")
        with bt.mc_batched(S, 11, shared_batch=batch):
            usual = layer(x)
        monkeypatch.delenv("BT_TC2_SINNER")     # the default order
        with bt.mc_batched(S, 11, shared_batch=batch):
            auto = layer(x)
        monkeypatch.setenv("BT_TC2_SINNER", "1")
        monkeypatch.setenv("BT_TC2_WPRE", "0")  # sample-fastest order with the fused sampling (the first run pre-samples where it can)
        with bt.mc_batched(S, 11, shared_batch=batch):
            fused = layer(x)
    torch.cuda.synchronize()
    assert torch.equal(folded, torch.cat(loop))
    assert torch.equal(folded, usual)
    assert torch.equal(folded, auto)
    assert torch.equal(folded, fused)


def test_conv3d_tc_matches_oracle(cuda_device):
    torch.manual_seed(12)
    layer = L.Conv3dReparameterization(16, 24, 3, 0, 1, 0, -3.0, stride=1, padding=1).to(cuda_device)
    layer.eps_kernel.normal_(); layer.eps_bias.normal_()
    x = torch.randn(2, 16, 5, 6, 7, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last_3d)
    out, _ = run_tc(layer, x)
    assert out.is_contiguous(memory_format=torch.channels_last_3d)
    w = bf16_round(oracle.sample_weight(layer.mu_kernel.detach().cpu(), layer.rho_kernel.detach().cpu(), layer.eps_kernel.cpu()))
    b = oracle.sample_weight(layer.mu_bias.detach().cpu(), layer.rho_bias.detach().cpu(), layer.eps_bias.cpu())
    want = torch.nn.functional.conv3d(bf16_round(x.cpu().float()).double(), w.double(), b.double(), padding=1)
    err = float((out.detach().cpu().double() - want).abs().max() / want.abs().max())
    assert err <= BF16_OUT_TOL, f"rel err {err}"


def test_tc_rejects_what_it_cannot_run(cuda_device):
    layer = L.Conv2dReparameterization(16, 16, 3, padding=1, groups=2, prior_type="normal").to(cuda_device)
    layer.path = "tc"
    with pytest.raises(RuntimeError, match="tensor-core path unavailable"):
        layer(torch.randn(1, 16, 8, 8, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last))
    fo = L.Conv2dFlipout(64, 64, 1).to(cuda_device)     # fp32 activations: tf32 tensor cores for Reparameterization only
    fo.path = "tc"
    with pytest.raises(RuntimeError, match="tensor-core path unavailable"):
        fo(torch.randn(1, 64, 8, 8, device=cuda_device).contiguous(memory_format=torch.channels_last))
    layer.path = "auto"                        # auto falls back to the fp32 kernel, documented in DESIGN.md
    out, _ = layer(torch.randn(1, 16, 8, 8, device=cuda_device))
    assert out.shape == (1, 16, 8, 8)


RESIDUAL_CASES = [
    # cin, cout, k, pad, batch, hw, S, relu
    (64, 256, 1, 0, 4, 14, 1, True),    # bottleneck conv3: BN=128, two n-blocks, stationary
    (128, 64, 3, 1, 3, 7, 1, True),     # BN=64, ragged M (147 rows): residual rows past Ms are never fetched
    (64, 192, 1, 0, 2, 9, 3, True),     # N tail (128 + 64), three stacked samples
    (256, 256, 3, 1, 2, 14, 2, False),  # streaming W ring, MT accumulators, no ReLU
]


@pytest.mark.parametrize("cin,cout,k,pad,batch,hw,S,relu", RESIDUAL_CASES)
def test_conv2d_tc_residual_epilogue(cin, cout, k, pad, batch, hw, S, relu, cuda_device):
    """out = relu(bf16(affine(conv)) + residual), rounded like the eager bf16 graph: bit-exact
    against the same kernel without the residual followed by torch's bf16 add and relu."""
    torch.manual_seed(cin + cout + k)
    layer = L.Conv2dReparameterization(cin, cout, k, padding=pad, bias=True, prior_type="normal").to(cuda_device)
    layer.dnn_to_bnn_flag = True
    layer.path = "tc"
    layer._ep_scale = torch.rand(cout, device=cuda_device) + 0.5
    layer._ep_shift = torch.randn(cout, device=cuda_device)
    x = torch.randn(S * batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    res = torch.randn(S * batch, cout, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        with bt.mc_batched(S, 5):
            plain = layer(x)
        layer._ep_relu = relu
        with bt.mc_batched(S, 5):      # (one call per block: a second call inside a block draws afresh, ops.call_seed)
            fused = layer(x, residual=res)
        with bt.mc_batched(S, 5):
            fused_nchw = layer(x, residual=res.contiguous())      # another layout: copied to out's strides
    want = plain + res
    if relu:
        want = torch.relu(want)
    torch.testing.assert_close(fused, want, rtol=0, atol=0)
    torch.testing.assert_close(fused_nchw, want, rtol=0, atol=0)
    with pytest.raises(RuntimeError, match="residual has shape"):
        layer(x, residual=res[:, :1])


CLUSTER_CASES = [
    # cin, cout, k, pad, batch, hw, S     streaming layers with >= 2 m-groups
    (128, 128, 3, 1, 8, 14, 2),    # K=1152, 13 tiles per sample: odd tail of a CTA pair
    (256, 256, 3, 1, 16, 14, 1),   # MT=4 groups, two n-blocks
    (1024, 256, 1, 0, 8, 14, 3),   # K=1024 1x1 (tiled map), three samples
    (512, 72, 3, 1, 6, 9, 1),      # N tail, BN=128 with 72 live rows split 64 + 8 between the CTAs
]


@pytest.mark.parametrize("cin,cout,k,pad,batch,hw,S", CLUSTER_CASES)
def test_cluster_pair_matches_single_cta(cin, cout, k, pad, batch, hw, S, cuda_device, monkeypatch):
    """CTA pairs that split the sampling of every W k-block (DSMEM stores into both rings) must
    give bit-identical outputs to the single-CTA schedule: same Philox stream, same MMA order."""
    torch.manual_seed(cin + cout)
    layer = L.Conv2dReparameterization(cin, cout, k, padding=pad, bias=True, prior_type="normal").to(cuda_device)
    layer.dnn_to_bnn_flag = True
    layer.path = "tc"
    layer._ep_scale = torch.rand(cout, device=cuda_device) + 0.5
    layer._ep_shift = torch.randn(cout, device=cuda_device)
    layer._ep_relu = True
    x = torch.randn(S * batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    outs = {}
    for cl in ("1", "2"):
        monkeypatch.setenv("BT_TC2_CL", cl)
        with torch.no_grad(), bt.mc_batched(S, 3):
            outs[cl] = layer(x)
        torch.cuda.synchronize()
    assert float(outs["1"].float().abs().max()) > 0
    torch.testing.assert_close(outs["2"], outs["1"], rtol=0, atol=0)


PRESAMPLE_CASES = [
    # cin, cout, k, pad, batch, hw, S, residual     streamed layers with >= 16 m-tiles per sample
    (128, 128, 3, 1, 16, 14, 2, False),   # one n-block: pre-sampled W, contiguous item ranges
    (256, 256, 3, 1, 16, 14, 2, False),   # two n-blocks: round-robin items
    (1024, 512, 1, 0, 16, 14, 3, False),  # long 1x1 reduction, four n-blocks, three samples
    (256, 520, 1, 0, 16, 14, 2, True),    # N tail (520 = 4 x 128 + 8) with a residual input
    (512, 256, 1, 0, 24, 14, 2, False),   # the panel would fit (stationary): switched to streaming when W is pre-sampled
    (512, 512, 1, 0, 16, 14, 2, True),    # BN = 256 tiles (N % 256 == 0, >= 8 k-blocks) with a residual input
]


@pytest.mark.parametrize("cin,cout,k,pad,batch,hw,S,res", PRESAMPLE_CASES)
def test_presampled_weights_and_round_robin_items_match_the_fused_schedule(cin, cout, k, pad, batch, hw, S, res, cuda_device, monkeypatch):
    """tc2_presample_kernel writes the weights the in-kernel sampling loop would draw (same Philox positions, same
    bf16 arithmetic) and the round-robin item order changes WHICH SM computes a tile, not how: outputs are
    bit-identical to the fused, contiguous-range schedule."""
    torch.manual_seed(cin + cout + k)
    layer = L.Conv2dReparameterization(cin, cout, k, padding=pad, bias=True, prior_type="normal").to(cuda_device)
    layer.dnn_to_bnn_flag = True
    layer.path = "tc"
    layer._ep_scale = torch.rand(cout, device=cuda_device) + 0.5
    layer._ep_shift = torch.randn(cout, device=cuda_device)
    layer._ep_relu = True
    x = torch.randn(S * batch, cin, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    r = torch.randn(S * batch, cout, hw, hw, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last) if res else None
    outs = {}
    for wpre, rr, bn256 in (("0", "0", "0"), ("1", "1", "1"), ("0", "1", "0"), ("1", "0", "1"), ("1", "1", "0")):
        monkeypatch.setenv("BT_TC2_WPRE", wpre)
        monkeypatch.setenv("BT_TC2_RR", rr)
        monkeypatch.setenv("BT_TC2_BN256", bn256)   # 256-column tiles where the plan allows them (cases 2, 3, 5, 6)
        n0 = ops.PROFILE.launches
        with torch.no_grad(), bt.mc_batched(S, 5):
            outs[wpre + rr + bn256] = layer(x, residual=r) if res else layer(x)
        torch.cuda.synchronize()
        if wpre == "1":
            assert ops.PROFILE.launches - n0 >= 2      # the pre-sampling pass ran (and the GEMM)
    assert float(outs["000"].float().abs().max()) > 0
    for key in ("111", "010", "101", "110"):
        torch.testing.assert_close(outs[key], outs["000"], rtol=0, atol=0)


def test_tc_forward_kl_is_fused_into_the_packing_pass(cuda_device):
    """forward(return_kl=True) on the tensor-core path: the KL term comes out of the parameter-packing launch (mu / rho
    read once for sampling and KL), with scalar and with tensor-valued priors, bias included; rtol 1e-5 vs the oracle,
    and no separate KL kernel is launched."""
    torch.manual_seed(21)
    layer = L.Conv2dReparameterization(64, 136, 3, padding=1, bias=True, prior_mean=0.1, prior_variance=0.7, prior_type="normal").to(cuda_device)
    layer.eps_kernel.normal_(); layer.eps_bias.normal_()
    x = torch.randn(2, 64, 9, 9, device=cuda_device).bfloat16().contiguous(memory_format=torch.channels_last)
    n0 = ops.PROFILE.launches
    out, kl = run_tc(layer, x)
    assert ops.PROFILE.launches - n0 == 2          # the packing pass (with the KL) and the GEMM, nothing else
    sig = lambda r: torch.log1p(torch.exp(r))
    mu, rho, mb, rb = (t.detach().cpu() for t in (layer.mu_kernel, layer.rho_kernel, layer.mu_bias, layer.rho_bias))
    want = oracle.kl_normal(mu, sig(rho), torch.full_like(mu, 0.1), torch.full_like(mu, 0.7)) + \
        oracle.kl_normal(mb, sig(rb), torch.full_like(mb, 0.1), torch.full_like(mb, 0.7))
    torch.testing.assert_close(kl.cpu(), want, rtol=1e-5, atol=1e-6)
    # tensor-valued priors (utils/util.py:101-104 MOPED): weight prior tensors, bias mean tensor
    pm, ps = torch.randn_like(mu) * 0.1, torch.rand_like(mu) + 0.5
    pbm = torch.randn_like(mb) * 0.1
    layer.prior_weight_mu, layer.prior_weight_sigma = pm.to(cuda_device), ps.to(cuda_device)
    layer.prior_bias_mu.data = pbm.to(cuda_device)
    out2, kl2 = run_tc(layer, x)
    want2 = oracle.kl_normal(mu, sig(rho), pm, ps) + oracle.kl_normal(mb, sig(rb), pbm, torch.full_like(mb, 0.7))
    torch.testing.assert_close(kl2.cpu(), want2, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(out2, out, rtol=0, atol=0)
    # gradients of the fused KL flow through the same autograd node as before
    (kl2 * 3.0).backward()
    m = mu.double().requires_grad_(True); r = rho.double().requires_grad_(True)
    (3.0 * oracle.kl_normal(m, torch.log1p(torch.exp(r)), pm.double(), ps.double())).backward()
    torch.testing.assert_close(layer.mu_kernel.grad.cpu().double(), m.grad, rtol=1e-4, atol=1e-8)
    torch.testing.assert_close(layer.rho_kernel.grad.cpu().double(), r.grad, rtol=1e-4, atol=1e-8)

