// fp32 SIMT data-gradient: grad_x = conv_transpose(grad_out, W) with W re-sampled on the fly
// (same eps tensor or same Philox key as forward -- forward saves nothing).
// This is what autograd derives for `input` from F.linear / F.conv*d in
// linear_variational.py:181 and conv_variational.py:384; for Flipout
// (linear_flipout.py:168-174) grad_x = dgrad(g, mu) + s_in * dgrad(g * s_out, sigma*eps).
// GEMM view: M' = batch*prod(in), N' = c_in/groups, K' = (c_out/groups)*taps (tap fastest).
#include "bt_simt.cuh"

namespace {

constexpr int BM = 128, BN = 64, BK = 16, NT = 256;
constexpr int A_PER_THREAD = BM * BK / NT;
constexpr int W_PER_THREAD = BN * BK / NT;

template <typename TX, typename TO, bool kFlip>
__global__ void __launch_bounds__(NT) simt_dgrad_kernel(const SGeo g, const SParams p, const TO* __restrict__ gout,
                                                        TX* __restrict__ gx) {
  constexpr int NA = kFlip ? 2 : 1;
  __shared__ __align__(16) float As[NA][BK][BM];
  __shared__ __align__(16) float Bs[NA][BK][BN + 4];
  __shared__ int4 taps[BT_MAX_TAPS];

  const int tid = threadIdx.x;
  const int grp = blockIdx.z;
  const int64_t MI = g.batch * g.ID * g.IH * g.IW;
  const int64_t KK = (int64_t)g.Ng * g.T;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int c0 = blockIdx.y * BN;
  for (int t = tid; t < g.T; t += NT) {
    const int kw = t % g.KW, kh = (t / g.KW) % g.KH, kd = t / (g.KW * g.KH);
    taps[t] = make_int4(kd * g.dd, kh * g.dh, kw * g.dw, 0);
  }
  const BtRngKey key_w = bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_EPS_W);
  const BtRngKey key_si = bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_SIGN_IN);
  const BtRngKey key_so = bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_SIGN_OUT);
  const int64_t Cout = (int64_t)g.Ng * g.groups, Cin = (int64_t)g.Cg * g.groups;

  // A' loader: one input position per thread
  const int a_m = tid % BM;
  const int64_t am = m0 + a_m;
  const bool a_valid = am < MI;
  int ab = 0, aid = 0, aih = 0, aiw = 0;
  if (a_valid) {
    int64_t t = am;
    aiw = (int)(t % g.IW); t /= g.IW;
    aih = (int)(t % g.IH); t /= g.IH;
    aid = (int)(t % g.ID); t /= g.ID;
    ab = (int)t;
  }
  const int a_step = NT / BM;
  int a_n = (tid / BM) / g.T, a_t = (tid / BM) % g.T;
  const int a_adv_n = a_step / g.T, a_adv_t = a_step % g.T;
  // W' loader: fixed k' lane
  const int w_k = tid % BK, w_c = tid / BK;
  int w_n = w_k / g.T, w_t = w_k % g.T;
  const int w_adv_n = BK / g.T, w_adv_t = BK % g.T;

  float a_reg[NA][A_PER_THREAD], w_reg[NA][W_PER_THREAD];

  auto load_tiles = [&]() {
#pragma unroll
    for (int i = 0; i < A_PER_THREAD; ++i) {
      float v = 0.f, v2 = 0.f;
      if (a_valid && a_n < g.Ng) {
        const int4 tp = taps[a_t];
        const int td = aid + g.pd - tp.x, th = aih + g.ph - tp.y, tw = aiw + g.pw - tp.z;
        if (td >= 0 && th >= 0 && tw >= 0) {
          const int od = td / g.sd, oh = th / g.sh, ow = tw / g.sw;
          if (od * g.sd == td && oh * g.sh == th && ow * g.sw == tw && od < g.OD && oh < g.OH && ow < g.OW) {
            const int64_t ng = (int64_t)grp * g.Ng + a_n;
            const int64_t off = (int64_t)ab * g.obs + ng * g.ocs + (int64_t)od * g.os0 + (int64_t)oh * g.os1 +
                                (int64_t)ow * g.os2;
            v = bt_to_float<TO>(gout[off]);
            if (kFlip) {
              float sg;
              if (p.sign_out) {
                sg = p.sign_out[off];
              } else {
                const uint64_t li = bt_sign_index((uint64_t)ab, Cout, ng, g.OD, od, g.OH, oh, g.OW, ow);
                sg = bt_sign1(li, key_so);
              }
              v2 = v * sg;
            }
          }
        }
      }
      a_reg[0][i] = v;
      if (kFlip) a_reg[NA - 1][i] = v2;
      a_n += a_adv_n; a_t += a_adv_t;
      if (a_t >= g.T) { a_t -= g.T; ++a_n; }
    }
#pragma unroll
    for (int i = 0; i < W_PER_THREAD; ++i) {
      const int c = c0 + w_c + (NT / BK) * i;
      float w = 0.f, w2 = 0.f;
      if (c < g.Cg && w_n < g.Ng) {
        const int64_t ng = (int64_t)grp * g.Ng + w_n;
        const int64_t widx = ng * g.K + (int64_t)c * g.T + w_t;
        const float mu = __ldg(p.mu_w + widx);
        const float sig = bt_softplus(__ldg(p.rho_w + widx));
        const float eps = p.eps_w ? __ldg(p.eps_w + widx) : bt_normal1(((uint64_t)ng * g.T + w_t) * g.Cg + c, key_w);
        if (kFlip) { w = mu; w2 = sig * eps; } else { w = fmaf(sig, eps, mu); }
      }
      w_reg[0][i] = w;
      if (kFlip) w_reg[NA - 1][i] = w2;
    }
    w_n += w_adv_n; w_t += w_adv_t;
    if (w_t >= g.T) { w_t -= g.T; ++w_n; }
  };

  float acc[NA][8][4];
#pragma unroll
  for (int a = 0; a < NA; ++a)
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[a][i][j] = 0.f;

  const int tx = tid % 16, ty = tid / 16;
  __syncthreads();
  const int64_t nk = (KK + BK - 1) / BK;
  load_tiles();
  for (int64_t kt = 0; kt < nk; ++kt) {
#pragma unroll
    for (int a = 0; a < NA; ++a) {
#pragma unroll
      for (int i = 0; i < A_PER_THREAD; ++i) As[a][tid / BM + a_step * i][a_m] = a_reg[a][i];
#pragma unroll
      for (int i = 0; i < W_PER_THREAD; ++i) Bs[a][w_k][w_c + (NT / BK) * i] = w_reg[a][i];
    }
    __syncthreads();
    if (kt + 1 < nk) load_tiles();
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

  // ConvTranspose forward (this kernel's grad_x IS that layer's output): the layer's sampled bias per output
  // channel, mu_b + softplus(rho_b) eps_b (conv_variational.py:941-989); Flipout keeps the perturbation apart,
  // it is multiplied by the output sign below (conv_flipout.py:757-831)
  float b_mean[4] = {0.f, 0.f, 0.f, 0.f}, b_pert[4] = {0.f, 0.f, 0.f, 0.f};
  if (p.dgrad_bias) {
    const BtRngKey key_b = bt_rng_key(p.seed, p.sample_id, p.layer_id, BT_KIND_EPS_B);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c0 + ty * 4 + j;
      if (c >= g.Cg) continue;
      const int64_t cg = (int64_t)grp * g.Cg + c;
      const float e = p.eps_b ? p.eps_b[cg] : bt_normal1((uint64_t)cg, key_b);
      b_mean[j] = p.mu_b[cg];
      b_pert[j] = bt_softplus(p.rho_b[cg]) * e;
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t m = m0 + (i < 4 ? tx * 4 + i : 64 + tx * 4 + (i - 4));
    if (m >= MI) continue;
    int64_t t = m;
    const int iw = (int)(t % g.IW); t /= g.IW;
    const int ih = (int)(t % g.IH); t /= g.IH;
    const int id = (int)(t % g.ID); t /= g.ID;
    const int64_t b = t;
    const int64_t xbase = b * g.xbs + (int64_t)id * g.xs0 + (int64_t)ih * g.xs1 + (int64_t)iw * g.xs2;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c0 + ty * 4 + j;
      if (c >= g.Cg) continue;
      const int64_t cg = (int64_t)grp * g.Cg + c;
      const int64_t off = xbase + cg * g.xcs;
      float v = acc[0][i][j];
      if (kFlip) {
        float sg;
        if (p.sign_in) {
          sg = p.sign_in[off];
        } else {
          const uint64_t li = bt_sign_index((uint64_t)b, Cin, cg, g.ID, id, g.IH, ih, g.IW, iw);
          sg = bt_sign1(li, key_si);
        }
        v += (acc[NA - 1][i][j] + b_pert[j]) * sg + b_mean[j];
      } else {
        v += b_mean[j] + b_pert[j];
      }
      gx[off] = bt_from_float<TX>(v);
    }
  }
}

}  // namespace

template <typename TX, typename TO>
int bt_launch_dgrad(const SGeo& g, const SParams& p, const void* gout, void* gx, bool flip, cudaStream_t stream) {
  const int64_t MI = g.batch * g.ID * g.IH * g.IW;
  if (MI == 0) return BT_OK;
  const int64_t mt = (MI + BM - 1) / BM;
  const int nt = (g.Cg + BN - 1) / BN;
  BT_REQUIRE(mt <= 0x7fffffffll && nt <= 65535 && g.groups <= 65535, BT_ERR_UNSUPPORTED, "simt dgrad: grid too large");
  dim3 grid((unsigned)mt, (unsigned)nt, (unsigned)g.groups);
  if (flip)
    simt_dgrad_kernel<TX, TO, true><<<grid, NT, 0, stream>>>(g, p, (const TO*)gout, (TX*)gx);
  else
    simt_dgrad_kernel<TX, TO, false><<<grid, NT, 0, stream>>>(g, p, (const TO*)gout, (TX*)gx);
  BT_LAUNCH_OK("simt_dgrad_kernel");
  return BT_OK;
}

template int bt_launch_dgrad<float, float>(const SGeo&, const SParams&, const void*, void*, bool, cudaStream_t);
template int bt_launch_dgrad<__nv_bfloat16, __nv_bfloat16>(const SGeo&, const SParams&, const void*, void*, bool,
                                                           cudaStream_t);
