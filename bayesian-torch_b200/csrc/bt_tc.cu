// tcgen05 / TMEM implicit GEMM with the weight tile SAMPLED IN SHARED MEMORY (sm_100a).
//
//   out[s][m][n] = sum_k A(x_s)[m][k] * bf16(mu[n][k] + sigma[n][k] * eps_s[n][k]) + (mu_b + sigma_b * eps_b)[n]
//
// replaces, for bf16 activations, the reference's per-call sequence  log1p(exp(rho)) ->
// normal_() -> sigma*eps -> mu+tmp -> F.linear / F.conv2d  (linear_variational.py:163-181,
// conv_variational.py:366-385).  The sampled weight never exists in global memory: producer
// warps turn (mu, sigma, Philox eps) into bf16 tiles written straight into the 128B-swizzled
// shared-memory layout tcgen05.mma consumes.
//
// Persistent, warp-specialised CTA (1 per SM, 544 threads):
//   warps  0-3  epilogue   : TMEM -> registers (tcgen05.ld) -> +sampled bias -> global
//   warps  4-7  A loaders  : im2col rows of x (NHWC, bf16) by 16-byte cp.async into swizzled smem
//   warps  8-15 W producers: mu + sigma*eps -> bf16 -> swizzled smem (in-register Philox4x32-10)
//   warp   16   MMA issuer : one thread issues tcgen05.mma (M=128, N=BN, K=16), accumulators in TMEM
// Pipelines (all mbarrier based): A ring (a_full/a_empty), W ring (w_full/w_empty), accumulator
// sets in TMEM (acc_full/acc_empty).  Two ways to amortise the sampling cost over the batch:
//   stationary: the whole [BN x K] sampled panel fits in shared memory -> generated once per
//               (sample, n-block) and reused for every m-tile the CTA owns;
//   streaming : MT (<= 512/BN) accumulators stay resident in TMEM so each generated W k-block
//               feeds MT m-tiles before it is overwritten.
// Work item = (sample s, n-block, group of MT m-tiles); CTAs take contiguous item ranges.
#include <stdlib.h>

#include "bt_common.cuh"
#include "bt_tc.cuh"

using namespace btc;

namespace {

constexpr int BM = 128;          // UMMA M
constexpr int BK = 64;           // k-block: 64 bf16 = one 128-byte swizzle row
constexpr int A_SLOT = BM * BK * 2;  // 16 KB
constexpr int CTRL_BYTES = 3072;
constexpr int MAX_NA = 12, MAX_RW = 24, MAX_SETS = 8, MAX_MT = 4, MAX_TAPS = 64;
constexpr int NUM_THREADS = 17 * 32;
constexpr int A_LAG = 2;         // cp.async groups in flight per loader thread
constexpr int SMEM_LIMIT = 232448;

struct TcArgs {
  const __nv_bfloat16* x;
  void* out;
  const float *mu_p, *sig_p, *eps_p;   // packed [N][K], k = tap*Cin + c
  const float *mu_b, *rho_b, *eps_b;
  int64_t M, total_items;
  int64_t xss, xbs, xs0, xs1, xs2, oss, obs, os0, os1, os2;
  int N, K, Cin, C8, T, S, NB;
  int ID, IH, IW, OD, OH, OW, KH, KW;
  int sd, sh, sw, pd, ph, pw, dd, dh, dw;
  int num_mt, num_mg, MT, NSETS, nkb, NA, RW, stationary;
  uint64_t seed;
  uint32_t sample_id, layer_id;
};

struct Ctrl {  // lives in the first CTRL_BYTES of dynamic shared memory
  uint64_t a_full[MAX_NA], a_empty[MAX_NA];
  uint64_t w_full[MAX_RW], w_empty[MAX_RW];
  uint64_t acc_full[MAX_SETS], acc_empty[MAX_SETS];
  uint32_t tmem_base;
  uint32_t pad[3];
  float bias[256];
  int4 taps[MAX_TAPS];
};
static_assert(sizeof(Ctrl) <= CTRL_BYTES, "control block too large");

// ---------------------------------------------------------------- parameter packing ----
// mu[n][c][tap] -> mu_p[n][tap*Cin + c], sigma_p = softplus(rho) likewise (sigma does not depend on
// the sample, so it is hoisted out of the per-sample loop), optional eps likewise.
__global__ void tc_pack_kernel(const float* __restrict__ mu, const float* __restrict__ rho,
                               const float* __restrict__ eps, float* __restrict__ mu_p, float* __restrict__ sig_p,
                               float* __restrict__ eps_p, int64_t total, int Cin, int T) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    // i indexes the packed layout: (n, tap, c)
    const int64_t K = (int64_t)Cin * T;
    const int64_t n = i / K;
    const int64_t r = i - n * K;
    const int64_t tap = r / Cin, c = r - tap * Cin;
    const int64_t src = n * K + c * T + tap;
    mu_p[i] = mu[src];
    sig_p[i] = bt_softplus(rho[src]);
    if (eps_p) eps_p[i] = eps[src];
  }
}

template <int BN, typename TO>
__global__ void __launch_bounds__(NUM_THREADS, 1) tc_fwd_kernel(const __grid_constant__ TcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw);
  Ctrl* ctrl = reinterpret_cast<Ctrl*>(smem);
  constexpr int W_SLOT = BN * BK * 2;
  const uint32_t a_base = base + CTRL_BYTES;
  const uint32_t w_base = a_base + a.NA * A_SLOT;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- one-time setup
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.NA; ++i) { mbar_init(smem_u32(&ctrl->a_full[i]), 128); mbar_init(smem_u32(&ctrl->a_empty[i]), 1); }
    for (int i = 0; i < a.RW; ++i) { mbar_init(smem_u32(&ctrl->w_full[i]), 256); mbar_init(smem_u32(&ctrl->w_empty[i]), 1); }
    for (int i = 0; i < a.NSETS; ++i) { mbar_init(smem_u32(&ctrl->acc_full[i]), 1); mbar_init(smem_u32(&ctrl->acc_empty[i]), 128); }
    mbar_fence_init();
  }
  for (int t = threadIdx.x; t < a.T; t += NUM_THREADS) {
    const int kw = t % a.KW, kh = (t / a.KW) % a.KH, kd = t / (a.KW * a.KH);
    ctrl->taps[t] = make_int4(kd * a.dd, kh * a.dh, kw * a.dw, 0);
  }
  if (warp == 16) tmem_alloc(smem_u32(&ctrl->tmem_base), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = ctrl->tmem_base;

  // contiguous, balanced item range of this CTA
  const int64_t i0 = a.total_items * blockIdx.x / gridDim.x;
  const int64_t i1 = a.total_items * (blockIdx.x + 1) / gridDim.x;

  if (warp < 4) {
    // =========================== EPILOGUE ===========================
    const int q = warp;  // TMEM lane quarter this warp may read
    int64_t itcount = 0;
    for (int64_t it = i0; it < i1; ++it, ++itcount) {
      const int64_t key = it / a.num_mg;
      const int mg = (int)(it - key * a.num_mg);
      const int s = (int)(key / a.NB), nb = (int)(key % a.NB);
      const int set = (int)(itcount % a.NSETS);
      const uint32_t ph = (uint32_t)((itcount / a.NSETS) & 1);
      // sampled bias of this (sample, n-block): mu_b + softplus(rho_b) * eps_b
      named_bar_sync(1, 128);
      for (int j = threadIdx.x; j < BN; j += 128) {
        const int n = nb * BN + j;
        float b = 0.f;
        if (a.mu_b && n < a.N) {
          const float e = a.eps_b ? a.eps_b[n]
                                  : bt_normal1((uint64_t)n, bt_rng_key(a.seed, a.sample_id + s, a.layer_id, BT_KIND_EPS_B));
          b = a.mu_b[n] + bt_softplus(a.rho_b[n]) * e;
        }
        ctrl->bias[j] = b;
      }
      named_bar_sync(1, 128);
      mbar_wait(smem_u32(&ctrl->acc_full[set]), ph);
      tc_fence_after();
      const int mtv = min(a.MT, a.num_mt - mg * a.MT);
      for (int t = 0; t < mtv; ++t) {
        const int64_t m = ((int64_t)(mg * a.MT + t)) * BM + q * 32 + lane;
        const bool valid = m < a.M;
        TO* orow = nullptr;
        if (valid) {
          int64_t r = m;
          const int ow = (int)(r % a.OW); r /= a.OW;
          const int oh = (int)(r % a.OH); r /= a.OH;
          const int od = (int)(r % a.OD); r /= a.OD;
          orow = reinterpret_cast<TO*>(a.out) + (int64_t)s * a.oss + r * a.obs + (int64_t)od * a.os0 +
                 (int64_t)oh * a.os1 + (int64_t)ow * a.os2 + (int64_t)nb * BN;
        }
        const uint32_t tacc = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)((set * a.MT + t) * BN);
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          tmem_ld32(tacc + c * 32, v);
          tmem_ld_wait();
          if (valid) {
            const int ncol0 = nb * BN + c * 32;
#pragma unroll
            for (int j8 = 0; j8 < 4; ++j8) {
              if (ncol0 + j8 * 8 >= a.N) break;  // N % 8 == 0: an 8-column chunk is all in or all out
              float f[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[j8 * 8 + j]) + ctrl->bias[c * 32 + j8 * 8 + j];
              if constexpr (sizeof(TO) == 2) {
                uint4 pk;
                __nv_bfloat162 h;
                h = __floats2bfloat162_rn(f[0], f[1]); pk.x = *reinterpret_cast<uint32_t*>(&h);
                h = __floats2bfloat162_rn(f[2], f[3]); pk.y = *reinterpret_cast<uint32_t*>(&h);
                h = __floats2bfloat162_rn(f[4], f[5]); pk.z = *reinterpret_cast<uint32_t*>(&h);
                h = __floats2bfloat162_rn(f[6], f[7]); pk.w = *reinterpret_cast<uint32_t*>(&h);
                *reinterpret_cast<uint4*>(orow + c * 32 + j8 * 8) = pk;
              } else {
                *reinterpret_cast<float4*>(orow + c * 32 + j8 * 8) = make_float4(f[0], f[1], f[2], f[3]);
                *reinterpret_cast<float4*>(orow + c * 32 + j8 * 8 + 4) = make_float4(f[4], f[5], f[6], f[7]);
              }
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&ctrl->acc_empty[set]));
    }
  } else if (warp < 8) {
    // =========================== A LOADERS (im2col, cp.async) ===========================
    const int r = threadIdx.x - 128;  // tile row owned by this thread
    const uint32_t row_off = (uint32_t)r * 128u;
    const uint32_t sw = (uint32_t)(r & 7);
    int64_t acount = 0;
    for (int64_t it = i0; it < i1; ++it) {
      const int64_t key = it / a.num_mg;
      const int mg = (int)(it - key * a.num_mg);
      const int s = (int)(key / a.NB);
      const int mtv = min(a.MT, a.num_mt - mg * a.MT);
      // decode this thread's row of every m-tile of the group once
      int64_t xoff[MAX_MT];
      int id0[MAX_MT], ih0[MAX_MT], iw0[MAX_MT];
#pragma unroll
      for (int t = 0; t < MAX_MT; ++t) {
        xoff[t] = -1; id0[t] = 0; ih0[t] = 0; iw0[t] = 0;
        if (t < mtv) {
          const int64_t m = ((int64_t)(mg * a.MT + t)) * BM + r;
          if (m < a.M) {
            int64_t rr = m;
            const int ow = (int)(rr % a.OW); rr /= a.OW;
            const int oh = (int)(rr % a.OH); rr /= a.OH;
            const int od = (int)(rr % a.OD); rr /= a.OD;
            xoff[t] = (int64_t)s * a.xss + rr * a.xbs;
            id0[t] = od * a.sd - a.pd; ih0[t] = oh * a.sh - a.ph; iw0[t] = ow * a.sw - a.pw;
          }
        }
      }
      for (int kb = 0; kb < a.nkb; ++kb) {
        const int q0 = kb * 8;  // first 16-byte chunk (8 channels) of this k-block
        const int tap0 = q0 / a.C8, c80 = q0 - tap0 * a.C8;
#pragma unroll
        for (int t = 0; t < MAX_MT; ++t) {
          if (t >= mtv) break;
          const int slot = (int)(acount % a.NA);
          mbar_wait(smem_u32(&ctrl->a_empty[slot]), (uint32_t)(((acount / a.NA) & 1) ^ 1));
          const uint32_t dst = a_base + slot * A_SLOT + row_off;
          int tap = tap0, c8 = c80;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const void* src = a.x;
            uint32_t nbytes = 0;
            if (xoff[t] >= 0 && tap < a.T) {
              const int4 tp = ctrl->taps[tap];
              const int id = id0[t] + tp.x, ih = ih0[t] + tp.y, iw = iw0[t] + tp.z;
              if ((unsigned)id < (unsigned)a.ID && (unsigned)ih < (unsigned)a.IH && (unsigned)iw < (unsigned)a.IW) {
                src = a.x + xoff[t] + (int64_t)id * a.xs0 + (int64_t)ih * a.xs1 + (int64_t)iw * a.xs2 + c8 * 8;
                nbytes = 16;
              }
            }
            cp_async_16(dst + (((uint32_t)j ^ sw) << 4), src, nbytes);
            if (++c8 == a.C8) { c8 = 0; ++tap; }
          }
          cp_async_commit();
          if (acount >= A_LAG) {
            cp_async_wait<A_LAG>();
            fence_proxy_async();
            mbar_arrive(smem_u32(&ctrl->a_full[(acount - A_LAG) % a.NA]));
          }
          ++acount;
        }
      }
    }
    cp_async_wait<0>();
    fence_proxy_async();
    for (int64_t j = (acount > A_LAG ? acount - A_LAG : 0); j < acount; ++j) mbar_arrive(smem_u32(&ctrl->a_full[j % a.NA]));
  } else if (warp < 16) {
    // =========================== W PRODUCERS (sample in smem) ===========================
    const int p = threadIdx.x - 256;
    const int col4 = p & 15;        // float4 (4 consecutive k) within the 64-wide k-block
    const int row_base = p >> 4;    // 0..15, rows row_base + 16*i
    const uint32_t col_off = (uint32_t)(col4 & 1) * 8u;
    const uint32_t chunk16 = (uint32_t)(col4 >> 1);
    int64_t wcount = 0, prev_key = -1;
    for (int64_t it = i0; it < i1; ++it) {
      const int64_t key = it / a.num_mg;
      const bool new_gen = !a.stationary || key != prev_key;
      prev_key = key;
      if (!new_gen) continue;
      const int s = (int)(key / a.NB), nb = (int)(key % a.NB);
      const BtRngKey rk = bt_rng_key(a.seed, a.sample_id + s, a.layer_id, BT_KIND_EPS_W);
      for (int kb = 0; kb < a.nkb; ++kb, ++wcount) {
        const int slot = (int)(wcount % a.RW);
        const int k = kb * BK + col4 * 4;
        const bool kin = k < a.K;  // K % 8 == 0: a float4 is all in or all out
        constexpr int HALF = BN / 32;  // rows per thread per half-tile (register budget: 96/thread)
        const uint32_t wdst = w_base + slot * W_SLOT;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float4 mu4[HALF], sg4[HALF];
#pragma unroll
          for (int i = 0; i < HALF; ++i) {  // loads first (L2-resident parameters)
            const int n = nb * BN + row_base + 16 * (h * HALF + i);
            mu4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            sg4[i] = mu4[i];
            if (kin && n < a.N) {
              const int64_t e = (int64_t)n * a.K + k;
              mu4[i] = __ldg(reinterpret_cast<const float4*>(a.mu_p + e));
              sg4[i] = __ldg(reinterpret_cast<const float4*>(a.sig_p + e));
            }
          }
          if (h == 0) mbar_wait(smem_u32(&ctrl->w_empty[slot]), (uint32_t)(((wcount / a.RW) & 1) ^ 1));
#pragma unroll
          for (int i = 0; i < HALF; ++i) {
            const int row = row_base + 16 * (h * HALF + i);
            const int n = nb * BN + row;
            float4 e4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (kin && n < a.N) {
              const int64_t e = (int64_t)n * a.K + k;
              e4 = a.eps_p ? __ldg(reinterpret_cast<const float4*>(a.eps_p + e)) : bt_normal4((uint64_t)e >> 2, rk);
            }
            const __nv_bfloat162 lo = __floats2bfloat162_rn(fmaf(sg4[i].x, e4.x, mu4[i].x), fmaf(sg4[i].y, e4.y, mu4[i].y));
            const __nv_bfloat162 hi = __floats2bfloat162_rn(fmaf(sg4[i].z, e4.z, mu4[i].z), fmaf(sg4[i].w, e4.w, mu4[i].w));
            const uint32_t addr = wdst + (uint32_t)row * 128u + ((chunk16 ^ (uint32_t)(row & 7)) << 4) + col_off;
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(*reinterpret_cast<const uint32_t*>(&lo)),
                         "r"(*reinterpret_cast<const uint32_t*>(&hi))
                         : "memory");
          }
        }
        fence_proxy_async();
        mbar_arrive(smem_u32(&ctrl->w_full[slot]));
      }
    }
  } else {
    // =========================== MMA ISSUER ===========================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
      int64_t acount = 0, wuse = 0, gen_base = 0, itcount = 0, prev_key = -1;
      for (int64_t it = i0; it < i1; ++it, ++itcount) {
        const int64_t key = it / a.num_mg;
        const int mg = (int)(it - key * a.num_mg);
        const bool first = !a.stationary || key != prev_key;
        const bool last = !a.stationary || it + 1 == i1 || (it + 1) / a.num_mg != key;
        prev_key = key;
        if (first) { gen_base = wuse; wuse += a.nkb; }
        const int set = (int)(itcount % a.NSETS);
        mbar_wait(smem_u32(&ctrl->acc_empty[set]), (uint32_t)(((itcount / a.NSETS) & 1) ^ 1));
        tc_fence_after();
        const int mtv = min(a.MT, a.num_mt - mg * a.MT);
        for (int kb = 0; kb < a.nkb; ++kb) {
          const int64_t widx = gen_base + kb;
          const int wslot = (int)(widx % a.RW);
          if (first) {
            mbar_wait(smem_u32(&ctrl->w_full[wslot]), (uint32_t)((widx / a.RW) & 1));
            tc_fence_after();
          }
          const uint64_t bdesc = umma_desc_sw128(w_base + wslot * W_SLOT);
          for (int t = 0; t < mtv; ++t, ++acount) {
            const int aslot = (int)(acount % a.NA);
            mbar_wait(smem_u32(&ctrl->a_full[aslot]), (uint32_t)((acount / a.NA) & 1));
            tc_fence_after();
            const uint64_t adesc = umma_desc_sw128(a_base + aslot * A_SLOT);
            const uint32_t d = tmem + (uint32_t)((set * a.MT + t) * BN);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)  // +32 bytes (>>4 = 2) per K=16 step inside the swizzle atom
              umma_bf16(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
            umma_commit(smem_u32(&ctrl->a_empty[aslot]));
          }
          if (last) umma_commit(smem_u32(&ctrl->w_empty[wslot]));
        }
        umma_commit(smem_u32(&ctrl->acc_full[set]));
      }
    }
    __syncwarp();
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == 16) tmem_dealloc(tmem, 512);
}

}  // namespace

// ------------------------------------------------------------------------ host side ----
static int64_t weight_elems(const bt_geom* g) {
  int64_t t = 1;
  for (int i = 0; i < g->rank; ++i) t *= g->k_size[i];
  return g->c_out * (g->c_in / g->groups) * t;
}

// (mu, softplus(rho), eps) -> packed [N][tap][Cin] fp32 arrays shared by both tensor-core kernels.
int bt_tc_pack(const bt_layer_params* p, int64_t wn, int Cin, int T, float* mu_p, float* sig_p, float* eps_p,
               cudaStream_t stream) {
  int64_t blocks = (wn + 255) / 256;
  const int64_t cap = (int64_t)bt_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  tc_pack_kernel<<<(int)blocks, 256, 0, stream>>>(p->mu_w, p->rho_w, p->eps_w, mu_p, sig_p, eps_p, wn, Cin, T);
  BT_LAUNCH_OK("tc_pack_kernel");
  return BT_OK;
}

bool bt_tc1_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why) {
#define TC_NO(cond, msg) if (cond) { *why = msg; return false; }
  TC_NO(flipout, "Flipout runs on the fp32 path in this build");
  TC_NO(g->groups != 1, "groups != 1");
  TC_NO(g->x_dtype != BT_BF16, "activations are not bf16");
  TC_NO(g->x_chan_stride != 1 || g->out_chan_stride != 1, "channels are not the fastest axis (need channels-last)");
  TC_NO(g->c_in % 8 != 0 || g->c_out % 8 != 0, "c_in / c_out not multiples of 8");
  int64_t taps = 1;
  for (int i = 0; i < g->rank; ++i) {
    taps *= g->k_size[i];
    TC_NO(g->x_sp_stride[i] % 8 != 0 || g->out_sp_stride[i] % 8 != 0, "spatial strides not multiples of 8 elements");
  }
  TC_NO(taps > MAX_TAPS, "kernel volume > 64 taps");
  TC_NO(g->x_batch_stride % 8 != 0 || g->out_batch_stride % 8 != 0 || g->x_sample_stride % 8 != 0 ||
            g->out_sample_stride % 8 != 0, "batch strides not multiples of 8 elements");
  TC_NO(g->n_samples > 1 && (p->eps_w || p->eps_b), "tensor-fed eps with n_samples > 1");
  TC_NO(g->batch <= 0, "empty batch");
#undef TC_NO
  return true;
}

int64_t bt_tc_workspace_bytes(const bt_geom* g) {
  const int64_t wn = weight_elems(g);
  return ((3 * wn * 4 + 255) / 256) * 256;  // mu_p, sigma_p, eps_p
}

template <int BN, typename TO>
static int tc_launch(const TcArgs& a, int grid, int smem_bytes, cudaStream_t stream) {
  BT_CUDA_OK(cudaFuncSetAttribute(tc_fwd_kernel<BN, TO>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  tc_fwd_kernel<BN, TO><<<grid, NUM_THREADS, smem_bytes, stream>>>(a);
  BT_LAUNCH_OK("tc_fwd_kernel");
  return BT_OK;
}

int bt_tc1_forward(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, void* workspace,
                   bool flipout, cudaStream_t stream) {
  const char* why = "";
  BT_REQUIRE(bt_tc1_supported(g, p, flipout, &why), BT_ERR_UNSUPPORTED, "tensor-core forward: %s", why);
  BT_REQUIRE(workspace, BT_ERR_INVALID, "tensor-core forward: workspace required (bt_workspace_bytes)");
  BT_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0 && ((uintptr_t)workspace & 15) == 0, BT_ERR_ALIGN,
             "tensor-core forward: x / out / workspace must be 16-byte aligned");
  TcArgs a;
  memset(&a, 0, sizeof(a));
  int64_t in[3] = {1, 1, 1}, o[3] = {1, 1, 1}, k[3] = {1, 1, 1}, s[3] = {1, 1, 1}, pd[3] = {0, 0, 0}, d[3] = {1, 1, 1},
          xs[3] = {0, 0, 0}, os[3] = {0, 0, 0};
  for (int i = 0; i < g->rank; ++i) {
    const int j = 3 - g->rank + i;
    in[j] = g->in_size[i]; o[j] = g->out_size[i]; k[j] = g->k_size[i]; s[j] = g->stride[i]; pd[j] = g->pad[i];
    d[j] = g->dil[i]; xs[j] = g->x_sp_stride[i]; os[j] = g->out_sp_stride[i];
    const int64_t expect = (in[j] + 2 * pd[j] - d[j] * (k[j] - 1) - 1) / s[j] + 1;
    BT_REQUIRE(o[j] == expect && o[j] > 0, BT_ERR_INVALID, "tensor-core forward: inconsistent out_size[%d]", i);
  }
  a.x = (const __nv_bfloat16*)x;
  a.out = out;
  a.mu_b = p->mu_b; a.rho_b = p->rho_b; a.eps_b = p->eps_b;
  a.M = g->batch * o[0] * o[1] * o[2];
  a.N = (int)g->c_out; a.Cin = (int)g->c_in; a.C8 = a.Cin / 8;
  a.T = (int)(k[0] * k[1] * k[2]);
  a.K = a.Cin * a.T;
  a.S = g->n_samples;
  a.ID = (int)in[0]; a.IH = (int)in[1]; a.IW = (int)in[2];
  a.OD = (int)o[0]; a.OH = (int)o[1]; a.OW = (int)o[2];
  a.KH = (int)k[1]; a.KW = (int)k[2];
  a.sd = (int)s[0]; a.sh = (int)s[1]; a.sw = (int)s[2];
  a.pd = (int)pd[0]; a.ph = (int)pd[1]; a.pw = (int)pd[2];
  a.dd = (int)d[0]; a.dh = (int)d[1]; a.dw = (int)d[2];
  a.xss = g->x_sample_stride; a.xbs = g->x_batch_stride; a.xs0 = xs[0]; a.xs1 = xs[1]; a.xs2 = xs[2];
  a.oss = g->out_sample_stride; a.obs = g->out_batch_stride; a.os0 = os[0]; a.os1 = os[1]; a.os2 = os[2];
  a.seed = p->seed; a.sample_id = p->sample_id; a.layer_id = p->layer_id;

  // ---- pack (mu, softplus(rho), eps) into the kernel's k order
  const int64_t wn = (int64_t)a.N * a.K;
  float* mu_p = (float*)workspace;
  float* sig_p = mu_p + wn;
  float* eps_p = p->eps_w ? sig_p + wn : nullptr;
  {
    const int rc = bt_tc_pack(p, wn, a.Cin, a.T, mu_p, sig_p, eps_p, stream);
    if (rc != BT_OK) return rc;
  }
  a.mu_p = mu_p; a.sig_p = sig_p; a.eps_p = eps_p;

  // ---- tiling / pipeline configuration
  const int BN = a.N <= 64 ? 64 : 128;
  const int NSLOTS = 512 / BN;
  const int WB = BN * BK * 2;
  a.NB = (a.N + BN - 1) / BN;
  a.num_mt = (int)((a.M + BM - 1) / BM);
  a.nkb = (a.K + BK - 1) / BK;
  const int avail = SMEM_LIMIT - 1024 - CTRL_BYTES;
  const int sms = bt_num_sms();
  a.stationary = (a.nkb <= MAX_RW && (int64_t)a.nkb * WB + 4 * A_SLOT <= avail) ? 1 : 0;
  if (a.stationary) {
    a.RW = a.nkb;
    a.MT = 1;
  } else {
    a.RW = 3;
    a.MT = 1;
    for (int mt = NSLOTS < MAX_MT ? NSLOTS : MAX_MT; mt >= 1; mt >>= 1) {
      const int64_t items = (int64_t)a.S * a.NB * ((a.num_mt + mt - 1) / mt);
      if (items >= sms || mt == 1) { a.MT = mt; break; }
    }
  }
  a.NSETS = NSLOTS / a.MT;
  if (a.NSETS > MAX_SETS) a.NSETS = MAX_SETS;
  a.num_mg = (a.num_mt + a.MT - 1) / a.MT;
  a.NA = (avail - a.RW * WB) / A_SLOT;
  if (a.NA > MAX_NA) a.NA = MAX_NA;
  BT_REQUIRE(a.NA >= A_LAG + 1, BT_ERR_UNSUPPORTED, "tensor-core forward: no room for the A pipeline");
  a.total_items = (int64_t)a.S * a.NB * a.num_mg;
  const int grid = (int)(a.total_items < sms ? a.total_items : sms);
  const int smem_bytes = 1024 + CTRL_BYTES + a.NA * A_SLOT + a.RW * WB;
  const bool o32 = g->out_dtype == BT_F32;
  if (BN == 64) return o32 ? tc_launch<64, float>(a, grid, smem_bytes, stream) : tc_launch<64, __nv_bfloat16>(a, grid, smem_bytes, stream);
  return o32 ? tc_launch<128, float>(a, grid, smem_bytes, stream) : tc_launch<128, __nv_bfloat16>(a, grid, smem_bytes, stream);
}

// ---- dispatch between the two tensor-core kernels: the TMA-fed one (bt_tc2.cu) when the
// tensors are dense channels-last with c_in % 64 == 0 (or Linear), else the cp.async one above.
// BT_TC_IMPL=1|2 pins one of them (tests / profiling).
static int tc_impl_env() {
  const char* v = getenv("BT_TC_IMPL");
  return v && *v ? atoi(v) : 0;
}

bool bt_tc_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why) {
  const int pin = tc_impl_env();
  if (pin != 1 && bt_tc2_supported(g, p, flipout, why)) return true;
  if (pin == 2) return false;
  return bt_tc1_supported(g, p, flipout, why);
}

int bt_tc_forward(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, void* workspace,
                  bool flipout, cudaStream_t stream) {
  const char* why = "";
  if (tc_impl_env() != 1 && bt_tc2_supported(g, p, flipout, &why)) return bt_tc2_forward(g, p, x, out, workspace, flipout, stream);
  return bt_tc1_forward(g, p, x, out, workspace, flipout, stream);
}
