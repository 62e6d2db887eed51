// fp32 SIMT weight-gradient, bias gradient and the (mu, rho) epilogue of backward.
//   dW[n,k]     = sum_m grad_out[m,n] * A(x)[m,k]                      (split over m, fp32 atomics)
//   grad_mu_w   = dW (+ grad_kl * dKL/dmu)
//   grad_rho_w  = dW * eps * sigmoid(rho) (+ grad_kl * dKL/drho)        eps re-read or REGENERATED
//   Flipout     : grad_mu_w = g^T x ; dDelta = (g*s_out)^T (x*s_in) ; grad_rho_w = dDelta*eps*sigmoid(rho)
// which is what autograd derives from the reference op sequence (SURVEY.md 3.2;
// linear_variational.py:163-181, linear_flipout.py:149-174, base_variational_layer.py:70-72).
#include "bt_simt.cuh"

template <typename TX, typename TO>
int bt_launch_dgrad(const SGeo& g, const SParams& p, const void* gout, void* gx, bool flip, cudaStream_t stream);

namespace {

constexpr int WN = 64, WK = 64, WR = 16, NT = 256;

template <typename TX, typename TO, bool kFlip>
__global__ void __launch_bounds__(NT) simt_wgrad_kernel(const SGeo g, const SParams p, const TX* __restrict__ x,
                                                        const TO* __restrict__ gout, float* __restrict__ dst1,
                                                        float* __restrict__ dst2, int splits, int64_t rows_per_split) {
  constexpr int NA = kFlip ? 2 : 1;
  __shared__ __align__(16) float Ys[NA][WR][WN + 4];
  __shared__ __align__(16) float Xs[NA][WR][WK + 4];
  const int tid = threadIdx.x;
  const int grp = blockIdx.z / splits, split = blockIdx.z % splits;
  const int n0 = blockIdx.y * WN;
  const int64_t k0 = (int64_t)blockIdx.x * WK;
  const int64_t m_begin = (int64_t)split * rows_per_split;
  const int64_t m_end = min(g.M, m_begin + rows_per_split);
  const BtRngKey key_si = bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_SIGN_IN);
  const BtRngKey key_so = bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_SIGN_OUT);
  const int64_t Cout = (int64_t)g.Ng * g.groups, Cin = (int64_t)g.Cg * g.groups;

  // loader lanes: a thread always reads output position (m_step + tid%16); 4 channels n and 4 taps k
  const int l_m = tid % WR, l_j = tid / WR;  // l_j in 0..15, + 16*i
  int kc[4], kdd[4], kdh[4], kdw[4];
  bool kv[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t k = k0 + l_j + 16 * i;
    kv[i] = k < g.K;
    const int tap = kv[i] ? (int)(k % g.T) : 0;
    kc[i] = kv[i] ? (int)(k / g.T) : 0;
    kdw[i] = (tap % g.KW) * g.dw;
    kdh[i] = ((tap / g.KW) % g.KH) * g.dh;
    kdd[i] = (tap / (g.KW * g.KH)) * g.dd;
  }

  float acc[NA][4][4];
#pragma unroll
  for (int a = 0; a < NA; ++a)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[a][i][j] = 0.f;
  const int tn = tid / 16, tk = tid % 16;

  for (int64_t ms = m_begin; ms < m_end; ms += WR) {
    const int64_t m = ms + l_m;
    float yv[NA][4], xv[NA][4];
#pragma unroll
    for (int a = 0; a < NA; ++a)
#pragma unroll
      for (int i = 0; i < 4; ++i) { yv[a][i] = 0.f; xv[a][i] = 0.f; }
    if (m < m_end) {
      int64_t t = m;
      const int ow = (int)(t % g.OW); t /= g.OW;
      const int oh = (int)(t % g.OH); t /= g.OH;
      const int od = (int)(t % g.OD); t /= g.OD;
      const int64_t b = t;
      const int64_t obase = b * g.obs + (int64_t)od * g.os0 + (int64_t)oh * g.os1 + (int64_t)ow * g.os2;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int n = n0 + l_j + 16 * i;
        if (n < g.Ng) {
          const int64_t ng = (int64_t)grp * g.Ng + n;
          const int64_t off = obase + ng * g.ocs;
          const float v = bt_to_float<TO>(gout[off]);
          yv[0][i] = v;
          if (kFlip) {
            const float sg = p.sign_out ? p.sign_out[off]
                                        : bt_sign1(bt_sign_index((uint64_t)b, Cout, ng, g.OD, od, g.OH, oh, g.OW, ow), key_so);
            yv[NA - 1][i] = v * sg;
          }
        }
      }
      const int id0 = od * g.sd - g.pd, ih0 = oh * g.sh - g.ph, iw0 = ow * g.sw - g.pw;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int id = id0 + kdd[i], ih = ih0 + kdh[i], iw = iw0 + kdw[i];
        if (kv[i] && (unsigned)id < (unsigned)g.ID && (unsigned)ih < (unsigned)g.IH && (unsigned)iw < (unsigned)g.IW) {
          const int64_t cg = (int64_t)grp * g.Cg + kc[i];
          const int64_t off = b * g.xbs + cg * g.xcs + (int64_t)id * g.xs0 + (int64_t)ih * g.xs1 + (int64_t)iw * g.xs2;
          const float v = bt_to_float<TX>(x[off]);
          xv[0][i] = v;
          if (kFlip) {
            const float sg = p.sign_in ? p.sign_in[off]
                                       : bt_sign1(bt_sign_index((uint64_t)b, Cin, cg, g.ID, id, g.IH, ih, g.IW, iw), key_si);
            xv[NA - 1][i] = v * sg;
          }
        }
      }
    }
    __syncthreads();  // previous step's reads done
#pragma unroll
    for (int a = 0; a < NA; ++a)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        Ys[a][l_m][l_j + 16 * i] = yv[a][i];
        Xs[a][l_m][l_j + 16 * i] = xv[a][i];
      }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < WR; ++r) {
#pragma unroll
      for (int a = 0; a < NA; ++a) {
        const float4 y4 = *reinterpret_cast<const float4*>(&Ys[a][r][tn * 4]);
        const float4 x4 = *reinterpret_cast<const float4*>(&Xs[a][r][tk * 4]);
        const float yy[4] = {y4.x, y4.y, y4.z, y4.w};
        const float xx[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[a][i][j] = fmaf(yy[i], xx[j], acc[a][i][j]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int n = n0 + tn * 4 + i;
    if (n >= g.Ng) continue;
    const int64_t ng = (int64_t)grp * g.Ng + n;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t k = k0 + tk * 4 + j;
      if (k >= g.K) continue;
      atomicAdd(dst1 + ng * g.K + k, acc[0][i][j]);
      if (kFlip) atomicAdd(dst2 + ng * g.K + k, acc[NA - 1][i][j]);
    }
  }
}

struct KlGradArgs {
  const float* grad_kl;  // device scalar or NULL
  const float* pmu_t;
  const float* psig_t;
  float prior_mu, prior_sigma;
};

// Element-wise epilogue over a parameter tensor: turns dW (in gmu) [and dDelta (in grho) for
// Flipout] into grad_mu / grad_rho and adds the KL gradient.
template <bool kFlip>
__global__ void param_grad_finalize_kernel(float* __restrict__ gmu, float* __restrict__ grho,
                                           const float* __restrict__ mu, const float* __restrict__ rho,
                                           const float* __restrict__ eps, int64_t numel, int T, int Cg, bool is_bias,
                                           BtRngKey key, KlGradArgs kl) {
  const float gk = kl.grad_kl ? __ldg(kl.grad_kl) / (float)numel : 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += (int64_t)gridDim.x * blockDim.x) {
    const float r = rho[i];
    const float sig = bt_softplus(r);
    const float dsig = r > 20.f ? 1.f : bt_sigmoid(r);
    float e;
    if (eps) {
      e = eps[i];
    } else if (is_bias) {
      e = bt_normal1((uint64_t)i, key);
    } else {  // logical (n, c, tap) -> packed RNG position (n, tap, c)
      const int64_t n = i / ((int64_t)Cg * T);
      const int64_t rem = i - n * (int64_t)Cg * T;
      const int64_t c = rem / T, tap = rem - c * T;
      e = bt_normal1((uint64_t)((n * T + tap) * Cg + c), key);
    }
    const float dW = gmu[i];
    const float dD = kFlip ? grho[i] : dW;
    float om = dW, orr = dD * e * dsig;
    if (kl.grad_kl) {
      const float pm = kl.pmu_t ? kl.pmu_t[i] : kl.prior_mu;
      const float ps = kl.psig_t ? kl.psig_t[i] : kl.prior_sigma;
      const float inv2 = 1.f / (ps * ps);
      om += gk * (mu[i] - pm) * inv2;
      orr += gk * (sig * inv2 - 1.f / sig) * dsig;
    }
    gmu[i] = om;
    grho[i] = orr;
  }
}

// One CTA per channel: s1 = sum grad_out, s2 = sum grad_out*s_out (Flipout), then the
// bias (mu_b, rho_b) gradients including the KL term.  XSIDE: the reduced tensor is the one passed as `x`
// (the backward of a ConvTranspose layer, whose output gradient sits on the x side of the equivalent convolution;
// its output signs are that convolution's SIGN_IN stream).
template <typename TO, bool kFlip, bool XSIDE>
__global__ void __launch_bounds__(256) bias_grad_kernel(const SGeo g, const SParams p, const TO* __restrict__ gout,
                                                        float* __restrict__ gmu_b, float* __restrict__ grho_b,
                                                        KlGradArgs kl) {
  __shared__ float red[2][8];
  const int n = blockIdx.x;
  const int64_t Cout = XSIDE ? (int64_t)g.Cg * g.groups : (int64_t)g.Ng * g.groups;
  const BtRngKey key_so = bt_rng_key(p.seed, p.sample_id, p.layer_id, XSIDE ? BT_KIND_SIGN_IN : BT_KIND_SIGN_OUT);
  const float* __restrict__ sign_t = XSIDE ? p.sign_in : p.sign_out;
  const int D = XSIDE ? g.ID : g.OD, H = XSIDE ? g.IH : g.OH, W = XSIDE ? g.IW : g.OW;
  const int64_t bs = XSIDE ? g.xbs : g.obs, cs = XSIDE ? g.xcs : g.ocs, s0 = XSIDE ? g.xs0 : g.os0, s1s = XSIDE ? g.xs1 : g.os1,
                s2s = XSIDE ? g.xs2 : g.os2;
  const int64_t M = g.batch * D * H * W;
  float s1 = 0.f, s2 = 0.f;
  for (int64_t m = threadIdx.x; m < M; m += blockDim.x) {
    int64_t t = m;
    const int ow = (int)(t % W); t /= W;
    const int oh = (int)(t % H); t /= H;
    const int od = (int)(t % D); t /= D;
    const int64_t b = t;
    const int64_t off = b * bs + (int64_t)n * cs + (int64_t)od * s0 + (int64_t)oh * s1s + (int64_t)ow * s2s;
    const float v = bt_to_float<TO>(gout[off]);
    s1 += v;
    if (kFlip) {
      const float sg = sign_t ? sign_t[off] : bt_sign1(bt_sign_index((uint64_t)b, Cout, n, D, od, H, oh, W, ow), key_so);
      s2 += v * sg;
    }
  }
  s1 = bt_warp_sum(s1);
  s2 = bt_warp_sum(s2);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s1; red[1][threadIdx.x >> 5] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    s1 = 0.f; s2 = 0.f;
    for (int w = 0; w < 8; ++w) { s1 += red[0][w]; s2 += red[1][w]; }
    const float r = p.rho_b[n];
    const float sig = bt_softplus(r);
    const float dsig = r > 20.f ? 1.f : bt_sigmoid(r);
    const float e = p.eps_b ? p.eps_b[n]
                            : bt_normal1((uint64_t)n, bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_EPS_B));
    float om = s1, orr = (kFlip ? s2 : s1) * e * dsig;
    if (kl.grad_kl) {
      const float gk = __ldg(kl.grad_kl) / (float)Cout;
      const float pm = kl.pmu_t ? kl.pmu_t[n] : kl.prior_mu;
      const float ps = kl.psig_t ? kl.psig_t[n] : kl.prior_sigma;
      const float inv2 = 1.f / (ps * ps);
      om += gk * (p.mu_b[n] - pm) * inv2;
      orr += gk * (sig * inv2 - 1.f / sig) * dsig;
    }
    if (gmu_b) gmu_b[n] = om;
    if (grho_b) grho_b[n] = orr;
  }
}

template <typename TX, typename TO>
int backward_typed(const SGeo& g, const SParams& p, const bt_layer_params* lp, const void* x, const void* gout,
                   const float* grad_kl, const bt_layer_grads* gr, bool flip, cudaStream_t stream) {
  int rc;
  if (gr->grad_x) {
    rc = bt_launch_dgrad<TX, TO>(g, p, gout, gr->grad_x, flip, stream);
    if (rc != BT_OK) return rc;
  }
  const int64_t wn = (int64_t)g.Ng * g.groups * g.K;
  if (gr->grad_mu_w || gr->grad_rho_w) {
    BT_REQUIRE(gr->grad_mu_w && gr->grad_rho_w, BT_ERR_INVALID,
               "backward: grad_mu_w and grad_rho_w must be requested together");
    BT_CUDA_OK(cudaMemsetAsync(gr->grad_mu_w, 0, (size_t)wn * 4, stream));
    if (flip) BT_CUDA_OK(cudaMemsetAsync(gr->grad_rho_w, 0, (size_t)wn * 4, stream));
    if (g.M > 0) {
      const int kt = (int)((g.K + WK - 1) / WK), nt = (g.Ng + WN - 1) / WN;
      // split the long reduction over m so that ~4 CTAs per SM are in flight
      int64_t want = ((int64_t)bt_num_sms() * 4 + (int64_t)kt * nt * g.groups - 1) / ((int64_t)kt * nt * g.groups);
      const int64_t max_splits = (g.M + WR * 8 - 1) / (WR * 8);
      if (want > max_splits) want = max_splits;
      if (want < 1) want = 1;
      if (want * g.groups > 65535) want = 65535 / g.groups;
      int64_t rows = (g.M + want - 1) / want;
      rows = (rows + WR - 1) / WR * WR;
      const int splits = (int)((g.M + rows - 1) / rows);
      dim3 grid((unsigned)kt, (unsigned)nt, (unsigned)(g.groups * splits));
      if (flip)
        simt_wgrad_kernel<TX, TO, true><<<grid, NT, 0, stream>>>(g, p, (const TX*)x, (const TO*)gout, gr->grad_mu_w,
                                                                  gr->grad_rho_w, splits, rows);
      else
        simt_wgrad_kernel<TX, TO, false><<<grid, NT, 0, stream>>>(g, p, (const TX*)x, (const TO*)gout, gr->grad_mu_w,
                                                                   gr->grad_rho_w, splits, rows);
      BT_LAUNCH_OK("simt_wgrad_kernel");
    }
    KlGradArgs kl = {grad_kl, lp->prior_mu_w_t, lp->prior_sigma_w_t, lp->prior_mu, lp->prior_sigma};
    int64_t grid64 = (wn + 255) / 256; if (grid64 > (int64_t)bt_num_sms() * 8) grid64 = (int64_t)bt_num_sms() * 8; if (grid64 < 1) grid64 = 1; const int grid = (int)grid64;
    const BtRngKey key = bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_EPS_W);
    if (flip)
      param_grad_finalize_kernel<true><<<grid, 256, 0, stream>>>(gr->grad_mu_w, gr->grad_rho_w, p.mu_w, p.rho_w,
                                                                 p.eps_w, wn, g.T, g.Cg, false, key, kl);
    else
      param_grad_finalize_kernel<false><<<grid, 256, 0, stream>>>(gr->grad_mu_w, gr->grad_rho_w, p.mu_w, p.rho_w,
                                                                  p.eps_w, wn, g.T, g.Cg, false, key, kl);
    BT_LAUNCH_OK("param_grad_finalize_kernel");
  }
  if (p.mu_b && (gr->grad_mu_b || gr->grad_rho_b)) {
    KlGradArgs kl = {grad_kl, lp->prior_mu_b_t, lp->prior_sigma_b_t, lp->prior_mu, lp->prior_sigma};
    if (lp->flags & BT_FLAG_BIAS_GRAD_X_SIDE) {   // backward of a ConvTranspose layer: its output gradient is `x` here
      const int cin = g.Cg * g.groups;
      if (flip)
        bias_grad_kernel<TX, true, true><<<cin, 256, 0, stream>>>(g, p, (const TX*)x, gr->grad_mu_b, gr->grad_rho_b, kl);
      else
        bias_grad_kernel<TX, false, true><<<cin, 256, 0, stream>>>(g, p, (const TX*)x, gr->grad_mu_b, gr->grad_rho_b, kl);
    } else {
      const int cout = g.Ng * g.groups;
      if (flip)
        bias_grad_kernel<TO, true, false><<<cout, 256, 0, stream>>>(g, p, (const TO*)gout, gr->grad_mu_b, gr->grad_rho_b, kl);
      else
        bias_grad_kernel<TO, false, false><<<cout, 256, 0, stream>>>(g, p, (const TO*)gout, gr->grad_mu_b, gr->grad_rho_b, kl);
    }
    BT_LAUNCH_OK("bias_grad_kernel");
  }
  return BT_OK;
}

}  // namespace

int bt_simt_backward(const bt_geom* geom, const bt_layer_params* lp, const void* x, const void* grad_out,
                     const float* grad_kl, const bt_layer_grads* grads, void* workspace, bool flipout,
                     cudaStream_t stream) {
  (void)workspace;
  SGeo g;
  int rc = bt_sgeo_from(geom, &g, "simt backward");
  if (rc != BT_OK) return rc;
  BT_REQUIRE(geom->n_samples == 1, BT_ERR_UNSUPPORTED, "simt backward: n_samples must be 1 (loop over samples)");
  BT_REQUIRE(geom->x_dtype == geom->out_dtype, BT_ERR_UNSUPPORTED, "simt backward: x and grad_out dtypes must match");
  const SParams p = bt_sparams_from(lp);
  if (geom->x_dtype == BT_F32)
    return backward_typed<float, float>(g, p, lp, x, grad_out, grad_kl, grads, flipout, stream);
  if (geom->x_dtype == BT_BF16)
    return backward_typed<__nv_bfloat16, __nv_bfloat16>(g, p, lp, x, grad_out, grad_kl, grads, flipout, stream);
  bt_set_error("simt backward: unsupported dtype %d", geom->x_dtype);
  return BT_ERR_UNSUPPORTED;
}
