// fp32 FFMA implicit-GEMM forward with the weight sampled while it is staged.
//
// This is the deterministic PARITY path (bt_path BT_PATH_SIMT_F32): any rank (Linear,
// Conv1/2/3d), stride, padding, dilation, groups and memory layout, fp32 accumulation in
// k order.  It restates, in one launch, the op sequence of
// LinearReparameterization.forward (linear_variational.py:160-181),
// Conv*dReparameterization.forward (conv_variational.py:362-385) and the Flipout
// forwards (linear_flipout.py:145-174, conv_flipout.py:370-417):
//   W tile  = mu + softplus(rho) * eps        (never written to global memory)
//   out     = A(x) . W^T + (mu_b + softplus(rho_b) * eps_b)
//   flipout = A(x).mu^T + mu_b + (A(x*s_in).(softplus(rho)*eps)^T + softplus(rho_b)*eps_b) * s_out
// GEMM view: M = batch*prod(out), N = c_out/groups, K = (c_in/groups)*taps; the k index
// runs (channel, tap) with the tap fastest, i.e. the weight tensor's own memory order.
// CTA tile 128x64x16, 256 threads, 8x4 outputs per thread, register-prefetched operands.
#include "bt_simt.cuh"

namespace {

constexpr int BM = 128, BN = 64, BK = 16, NT = 256;
constexpr int A_PER_THREAD = BM * BK / NT;  // 8
constexpr int W_PER_THREAD = BN * BK / NT;  // 4

__device__ __forceinline__ void fill_taps(int4* taps, const SGeo& g) {
  for (int t = threadIdx.x; t < g.T; t += blockDim.x) {
    const int kw = t % g.KW;
    const int kh = (t / g.KW) % g.KH;
    const int kd = t / (g.KW * g.KH);
    taps[t] = make_int4(kd * g.dd, kh * g.dh, kw * g.dw, 0);
  }
}

template <typename TX, typename TO, bool kFlip>
__global__ void __launch_bounds__(NT) simt_fwd_kernel(const SGeo g, const SParams p, const TX* __restrict__ x,
                                                      TO* __restrict__ out) {
  constexpr int NA = kFlip ? 2 : 1;
  __shared__ __align__(16) float As[NA][BK][BM];
  __shared__ __align__(16) float Bs[NA][BK][BN + 4];
  __shared__ int4 taps[BT_MAX_TAPS];

  const int tid = threadIdx.x;
  const int s = blockIdx.z / g.groups;
  const int grp = blockIdx.z % g.groups;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  fill_taps(taps, g);

  const BtRngKey key_w = bt_rng_key(p.seed, p.sample_id + s, p.layer_id, BT_KIND_EPS_W);
  const BtRngKey key_b = bt_rng_key(p.seed, p.sample_id + s, p.layer_id, BT_KIND_EPS_B);
  const BtRngKey key_si = bt_rng_key(p.seed, p.sample_id + s, p.layer_id, BT_KIND_SIGN_IN);
  const BtRngKey key_so = bt_rng_key(p.seed, p.sample_id + s, p.layer_id, BT_KIND_SIGN_OUT);

  // ---- A-operand loader state: one output position per thread, k advances by 2 per load
  const int a_m = tid % BM;
  const int64_t am = m0 + a_m;
  const bool a_valid = am < g.M;
  int ab = 0, aod = 0, aoh = 0, aow = 0;
  if (a_valid) {
    int64_t t = am;
    aow = (int)(t % g.OW); t /= g.OW;
    aoh = (int)(t % g.OH); t /= g.OH;
    aod = (int)(t % g.OD); t /= g.OD;
    ab = (int)t;
  }
  const int id0 = aod * g.sd - g.pd, ih0 = aoh * g.sh - g.ph, iw0 = aow * g.sw - g.pw;
  const int64_t x_off0 = (int64_t)s * g.xss + (int64_t)ab * g.xbs + (int64_t)grp * g.Cg * g.xcs;
  const int a_step = NT / BM;  // 2
  int a_c = (tid / BM) / g.T, a_t = (tid / BM) % g.T;
  const int a_adv_c = a_step / g.T, a_adv_t = a_step % g.T;

  // ---- W-operand loader state: fixed k lane, 4 output channels per thread
  const int w_k = tid % BK;
  const int w_n = tid / BK;  // + 16*i
  int w_c = w_k / g.T, w_t = w_k % g.T;
  const int w_adv_c = BK / g.T, w_adv_t = BK % g.T;

  float a_reg[NA][A_PER_THREAD];
  float w_reg[NA][W_PER_THREAD];

  auto load_tiles = [&](int64_t k0) {
#pragma unroll
    for (int i = 0; i < A_PER_THREAD; ++i) {
      float v = 0.f, v2 = 0.f;
      if (a_valid && a_c < g.Cg) {
        const int4 tp = taps[a_t];
        const int id = id0 + tp.x, ih = ih0 + tp.y, iw = iw0 + tp.z;
        if ((unsigned)id < (unsigned)g.ID && (unsigned)ih < (unsigned)g.IH && (unsigned)iw < (unsigned)g.IW) {
          const int64_t off = x_off0 + (int64_t)a_c * g.xcs + (int64_t)id * g.xs0 + (int64_t)ih * g.xs1 +
                              (int64_t)iw * g.xs2;
          v = bt_to_float<TX>(x[off]);
          if (kFlip) {
            float sg;
            if (p.sign_in) {
              sg = p.sign_in[off];
            } else {
              const int64_t cg = (int64_t)grp * g.Cg + a_c;
              const uint64_t li = bt_sign_index((uint64_t)ab, (int64_t)g.Cg * g.groups, cg, g.ID, id, g.IH, ih, g.IW, iw);
              sg = bt_sign1(li, key_si);
            }
            v2 = v * sg;
          }
        }
      }
      a_reg[0][i] = v;
      if (kFlip) a_reg[NA - 1][i] = v2;
      a_c += a_adv_c; a_t += a_adv_t;
      if (a_t >= g.T) { a_t -= g.T; ++a_c; }
    }
    const int64_t k = k0 + w_k;
#pragma unroll
    for (int i = 0; i < W_PER_THREAD; ++i) {
      const int n = n0 + w_n + (NT / BK) * i;
      float w = 0.f, w2 = 0.f;
      if (n < g.Ng && k < g.K) {
        const int64_t ng = (int64_t)grp * g.Ng + n;
        const int64_t widx = ng * g.K + k;
        const float mu = __ldg(p.mu_w + widx);
        const float sig = bt_softplus(__ldg(p.rho_w + widx));
        // RNG position: packed (n, tap, channel) order shared with the tensor-core path
        const float eps = p.eps_w ? __ldg(p.eps_w + widx)
                                  : bt_normal1(((uint64_t)ng * g.T + w_t) * g.Cg + w_c, key_w);
        if (kFlip) { w = mu; w2 = sig * eps; } else { w = fmaf(sig, eps, mu); }
      }
      w_reg[0][i] = w;
      if (kFlip) w_reg[NA - 1][i] = w2;
    }
    w_c += w_adv_c; w_t += w_adv_t;
    if (w_t >= g.T) { w_t -= g.T; ++w_c; }
  };

  float acc[NA][8][4];
#pragma unroll
  for (int a = 0; a < NA; ++a)
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[a][i][j] = 0.f;

  const int tx = tid % 16, ty = tid / 16;
  __syncthreads();  // taps visible
  const int64_t nk = (g.K + BK - 1) / BK;
  load_tiles(0);
  for (int64_t kt = 0; kt < nk; ++kt) {
#pragma unroll
    for (int a = 0; a < NA; ++a) {
#pragma unroll
      for (int i = 0; i < A_PER_THREAD; ++i) As[a][tid / BM + a_step * i][a_m] = a_reg[a][i];
#pragma unroll
      for (int i = 0; i < W_PER_THREAD; ++i) Bs[a][w_k][w_n + (NT / BK) * i] = w_reg[a][i];
    }
    __syncthreads();
    if (kt + 1 < nk) load_tiles((kt + 1) * BK);  // global loads overlap the FFMA block below
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
#pragma unroll
      for (int a = 0; a < NA; ++a) {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[a][kk][tx * 4]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[a][kk][64 + tx * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[a][kk][ty * 4]);
        const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[a][i][j] = fmaf(av[i], bv[j], acc[a][i][j]);
      }
    }
    __syncthreads();
  }

  // ---- epilogue: sampled bias, flipout recombination, strided store
  float b_mean[4], b_pert[4];
  int64_t ngl[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int n = n0 + ty * 4 + j;
    ngl[j] = (int64_t)grp * g.Ng + n;
    b_mean[j] = 0.f; b_pert[j] = 0.f;
    if (n < g.Ng && p.mu_b) {
      const float eps = p.eps_b ? p.eps_b[ngl[j]] : bt_normal1((uint64_t)ngl[j], key_b);
      const float pert = bt_softplus(p.rho_b[ngl[j]]) * eps;
      if (kFlip) { b_mean[j] = p.mu_b[ngl[j]]; b_pert[j] = pert; } else { b_mean[j] = p.mu_b[ngl[j]] + pert; }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t m = m0 + (i < 4 ? tx * 4 + i : 64 + tx * 4 + (i - 4));
    if (m >= g.M) continue;
    int64_t t = m;
    const int ow = (int)(t % g.OW); t /= g.OW;
    const int oh = (int)(t % g.OH); t /= g.OH;
    const int od = (int)(t % g.OD); t /= g.OD;
    const int64_t b = t;
    const int64_t obase = (int64_t)s * g.oss + b * g.obs + (int64_t)od * g.os0 + (int64_t)oh * g.os1 +
                          (int64_t)ow * g.os2;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (n0 + ty * 4 + j >= g.Ng) continue;
      const int64_t off = obase + ngl[j] * g.ocs;
      float v = acc[0][i][j] + b_mean[j];
      if (kFlip) {
        float sg;
        if (p.sign_out) {
          sg = p.sign_out[off];
        } else {
          const uint64_t li = bt_sign_index((uint64_t)b, (int64_t)g.Ng * g.groups, ngl[j], g.OD, od, g.OH, oh, g.OW, ow);
          sg = bt_sign1(li, key_so);
        }
        v += (acc[NA - 1][i][j] + b_pert[j]) * sg;
      }
      if (p.out_scale) v *= p.out_scale[ngl[j]];
      if (p.out_shift) v += p.out_shift[ngl[j]];
      if (p.residual)  // eager-graph rounding: the affine result is rounded to out's dtype before the add
        v = bt_to_float<TO>(bt_from_float<TO>(v)) + bt_to_float<TO>(static_cast<const TO*>(p.residual)[off]);
      if (p.out_relu) v = fmaxf(v, 0.f);
      out[off] = bt_from_float<TO>(v);
    }
  }
}

template <typename TX, typename TO>
int launch_fwd(const SGeo& g, const SParams& p, const void* x, void* out, bool flip, cudaStream_t stream) {
  const int64_t mt = (g.M + BM - 1) / BM;
  const int nt = (g.Ng + BN - 1) / BN;
  const int64_t z = (int64_t)g.n_samples * g.groups;
  BT_REQUIRE(mt <= 0x7fffffffll && nt <= 65535 && z <= 65535, BT_ERR_UNSUPPORTED,
             "simt forward: grid (%lld,%d,%lld) exceeds CUDA limits", (long long)mt, nt, (long long)z);
  dim3 grid((unsigned)mt, (unsigned)nt, (unsigned)z);
  if (flip)
    simt_fwd_kernel<TX, TO, true><<<grid, NT, 0, stream>>>(g, p, (const TX*)x, (TO*)out);
  else
    simt_fwd_kernel<TX, TO, false><<<grid, NT, 0, stream>>>(g, p, (const TX*)x, (TO*)out);
  BT_LAUNCH_OK("simt_fwd_kernel");
  return BT_OK;
}

}  // namespace

int bt_simt_forward(const bt_geom* geom, const bt_layer_params* lp, const void* x, void* out, bool flipout,
                    cudaStream_t stream) {
  SGeo g;
  int rc = bt_sgeo_from(geom, &g, "simt forward");
  if (rc != BT_OK) return rc;
  BT_REQUIRE(lp && lp->mu_w && lp->rho_w && ((x && out) || g.M == 0), BT_ERR_INVALID,
             "simt forward: null tensor pointer");
  BT_REQUIRE((lp->mu_b == nullptr) == (lp->rho_b == nullptr), BT_ERR_INVALID,
             "simt forward: mu_b and rho_b must both be given or both be null");
  BT_REQUIRE(!(flipout && geom->n_samples > 1 && (lp->sign_in || lp->sign_out)), BT_ERR_UNSUPPORTED,
             "simt forward: tensor-fed signs with n_samples > 1");
  BT_REQUIRE(!(geom->n_samples > 1 && (lp->eps_w || lp->eps_b)), BT_ERR_UNSUPPORTED,
             "simt forward: tensor-fed eps with n_samples > 1 (every sample would reuse the same draw)");
  if (g.M == 0) return BT_OK;  // empty batch: nothing to launch
  const SParams p = bt_sparams_from(lp);
  const int xd = geom->x_dtype, od = geom->out_dtype;
  if (xd == BT_F32 && od == BT_F32) return launch_fwd<float, float>(g, p, x, out, flipout, stream);
  if (xd == BT_BF16 && od == BT_BF16) return launch_fwd<__nv_bfloat16, __nv_bfloat16>(g, p, x, out, flipout, stream);
  if (xd == BT_BF16 && od == BT_F32) return launch_fwd<__nv_bfloat16, float>(g, p, x, out, flipout, stream);
  if (xd == BT_F32 && od == BT_BF16) return launch_fwd<float, __nv_bfloat16>(g, p, x, out, flipout, stream);
  bt_set_error("simt forward: unsupported dtype pair (%d,%d)", xd, od);
  return BT_ERR_UNSUPPORTED;
}
