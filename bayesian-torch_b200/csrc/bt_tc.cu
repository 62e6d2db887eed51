// TMA-fed tcgen05 / TMEM implicit GEMM over sampled weights (sm_100a).
//
//   out[s][m][n] = act( sum_k A(x_s)[m][k] * bf16(mu'[n][k] + sigma'[n][k] * eps_s[n][k]) + shift_s[n] + residual[s][m][n] )
//
// replaces, for bf16 channels-last (or fp32: tf32 instantiation) activations, the reference's per-call sequence
// log1p(exp(rho)) -> normal_() -> sigma*eps -> mu+tmp -> F.linear / F.conv2d
// (linear_variational.py:163-181, conv_variational.py:366-385), Flipout's dual contraction
// (linear_flipout.py:145-197, conv_flipout.py:370-439) and, run on grad_out, their backward data.
//
// Where the sampled weight lives:
//   stationary   the [BN x K] panel of a (sample, n-block) is sampled into shared memory once per CTA and kept while
//                the CTA walks m-groups: the weight never exists in global memory
//   streamed     long reductions: the S sampled copies of a call are written once by tc2_presample_kernel (same
//                Philox positions, same bf16 arithmetic) and their k-blocks arrive by TMA (a.w_pre); the fused variant
//                (16 producer warps sample every k-block of every item, CTA pairs share the work) stays for tiny calls
//
// Persistent warp-specialised CTA of 768 threads, one per SM (warp ids are issue priorities):
//   warps 0-15   W producers (fused sampling; one thread becomes the W loader when the weights are pre-sampled;
//                FLIP / EPI8 instantiations: warps 12-15 are A flippers / a second epilogue group)
//   warps 16-19  epilogue : tcgen05.ld -> (residual) -> ReLU -> bf16/f32 -> swizzled smem -> TMA store
//   warp  20     A loader : ONE thread issues TMA loads; convolutions use the im2col tensor-map mode
//                (cuTensorMapEncodeIm2col): the 128-pixel x 64-channel tile of one filter tap lands in the
//                128B-swizzled K-major layout tcgen05.mma reads, padding zero-filled by the hardware
//   warps 21-22  MMA      : two issuer warps (tile t of an m-group -> issuer t % 2), each a warp-uniform loop whose
//                elected lane issues tcgen05.mma (M = 128, N = BN in {64, 128, 256}, K = 16) into TMEM accumulators
//   warp  23     residual loader, or the second A loader when the call has no residual input
// mbarrier pipelines: A rings (TMA complete_tx / tcgen05.commit), W ring or stationary panel, accumulator tiles in
// TMEM (512 columns), residual ring.  Items (sample, n-block, m-group) are dealt round-robin to the CTAs when the
// n-block is the fastest index, so the passes over one group of A tiles run on neighbouring SMs at the same time.
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include "bt_common.cuh"
#include "bt_kl.cuh"
#include "bt_tc.cuh"

using namespace btc;

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_SLOT = BM * BK * 2;  // 16 KB
constexpr int BAR_BYTES = 1024;           // mbarriers + TMEM base
constexpr int ONES_BYTES = BM * 32;       // [128 x 16] bf16 tile of the rank-1 "add shift" MMA (32-byte swizzle)
constexpr int SHIFT_BYTES = 256 * 32;     // [BN x 16] bf16 (BN <= 256): per-channel shift split into three bf16 terms
constexpr int CTRL_BYTES = BAR_BYTES + ONES_BYTES + SHIFT_BYTES;
constexpr int STAGE_UNIT = 4 * 4096;  // one staging (or residual) buffer of 32 rows x 128 B for each of the 4 epilogue warps
constexpr int N_PROD_WARPS = 16;  // warps 0..15: W producers; the Flipout instantiation turns warps 12..15 into A flippers
constexpr int N_FLIP_WARPS = 4;
constexpr int N_MMA_WARPS = 2;  // issuer warps: tile t of an m-group belongs to issuer t % 2
// warp groups (setmaxnreg works on aligned groups of 4 warps): 0..3 = W producers, 4 = epilogue,
// 5 = TMA + 2 MMA issuers + one idle warp
// Warp ids are priorities: the SM's issue arbiter prefers the highest warp id among eligible
// warps, so the latency-critical single-lane loops (TMA, MMA issue) sit above the epilogue,
// and the 16 throughput-bound producer warps take whatever issue slots are left.
constexpr int WARP_PROD0 = 0, WARP_EPI0 = N_PROD_WARPS, WARP_TMA = WARP_EPI0 + 4, WARP_MMA = WARP_TMA + 1;
constexpr int NUM_THREADS = (N_PROD_WARPS + 8) * 32;  // 768
// register budget: 768 threads launch with 80 registers each; the control warp group and
// the producers shrink to 72, which lets the epilogue group (64 accumulator columns in flight per thread)
// grow to 120. The pool setmaxnreg draws on is what the CTA was LAUNCHED with
// (768 x 80 = 61440 registers), not the whole register file: 128*120 + 128*72 + 512*72 = 61440.
constexpr int REGS_EPI = 120, REGS_CTRL = 72, REGS_PROD = 72;
constexpr int REGS_EPI8 = 96;  // eight epilogue warps: 256*96 + 128*72 (control) + 384*72 (12 producer warps) = 61440
constexpr int MAX_NA = 8, MAX_RW = 24, MAX_SETS = 8, MAX_MT = 4;
constexpr int SMEM_LIMIT = 232448;
constexpr int AVAIL_BYTES = SMEM_LIMIT - 1024 - CTRL_BYTES;  // staging + residual ring + A ring + W ring / panel

struct Tc2Args {
  const void *mu_p, *sig_p;           // packed [N][K], k = tap*Cin + c, scale folded in; bf16, or fp32 for the tf32 instantiation
  const float* eps_p;                 // caller's eps in the same order (deterministic mode), else null
  const float *mu_b, *rho_b, *eps_b;
  const float *out_scale, *out_shift;  // folded BatchNorm (nullable)
  int out_relu;
  int64_t total_items;
  int N, K, S, NB;   // K: GEMM reduction length = row pitch of the packed parameter arrays
  int Kt;            // true reduction length (pitch of the RNG stream); < K when the patch matrix pads K up to 8
  int Ms;            // GEMM rows per sample: batch * OH * OW
  int x_rows_per_s;  // images (conv) / rows (linear) per sample in x; 0 when all samples share x
  int conv;          // 1: im2col map, 0: tiled 2-d map
  int OH, OW, KW, CB;  // CB = Cin / 64: k-blocks per filter tap
  int sh, sw, ph, pw, dh, dw;
  int num_mt, num_mg, MT, NSETS, nkb, NA, RW, stationary;  // num_mg: m-groups per key as walked by one cluster
  int nb_inner;      // item order: 1 = (sample, m-group, n-block), n-block fastest (streaming only); 0 = m-group fastest;
                     // 2 = (m-group, n-block, sample), sample fastest: layers whose x is shared by all samples
  int spn;           // 1, or 2: x is shared by all samples and N == 64 -> two samples ride side by side on the N axis of one
                     // BN = 128 tile (virtual row v = 64 * (sample within the pair) + channel): every A tile is read once per
                     // PAIR of samples and feeds a full-width MMA.  S then counts sample pairs.
  int a_keep;        // order 2 with an A ring of exactly one item's tiles: the tiles stay in shared memory while
                     // the samples of an m-group go by (the loader re-arms the slots instead of re-loading them)
  int CL;            // CTAs per cluster (1 or 2). CL = 2: the pair walks the same (sample, n-block, k-block) sequence on
                     // neighbouring m-groups, each CTA samples half of every W k-block and stores it into both CTAs' rings
  int NR, NAr;       // A rings (one per active MMA issuer) and slots per ring; NA = NR * NAr
  int has_shift;     // bias and/or folded BatchNorm shift: added by a rank-1 MMA that initialises each accumulator
  int res_bufs;      // residual tiles in flight per epilogue warp (0: no residual input)
  int stage_bufs;    // epilogue staging buffers per warp (2, or 1 when shared memory is tight)
  int pdl;           // launched with programmatic stream serialization: 1 = every role waits for the previous kernel after
                     // its set-up, 2 = the W producers do not (their inputs -- packed parameters of a frozen layer, biases,
                     // BatchNorm terms -- were complete long before) and sample the first k-blocks during the previous kernel's tail
  int epi8;          // eight epilogue warps (EPI8 instantiation): staging holds 2 * stage_bufs buffers per lane quarter
  int w_pre;         // the sampled weights were materialised by tc2_presample_kernel ([S][N][K] bf16 behind tm_a2): one thread of
                     // the producer group streams the W k-blocks in by TMA instead of the 16 warps sampling them per item
  int two_loaders;   // no residual input, two A rings: the residual-loader warp loads ring 1
  int pair_launch;   // host-side only: launch the single-CTA kernel with a cluster dimension of two (placement, not cooperation)
  int rr, rr_dnb, rr_dmg, rr_ds;  // round-robin item assignment (order 1) and the (nb, mg, s) digits of the cluster count
  int tf32;          // fp32 activations: fp32 A tiles, tf32 MMAs (k-blocks of 32 elements), fp32 packed parameters
  int debug;         // BT_TC2_DEBUG bits, honoured by the BT_TC2_PROF=1 instantiation only: 1 skip output stores, 4 skip MMAs,
                     // 8 skip A loads, 16 skip the normal draws, 32 skip the parameter loads, 64 skip the residual loads, 128 skip the
                     // loads of pre-sampled W
  uint64_t seed;
  uint32_t sample_id, layer_id;
  // Flipout (FLIP instantiation): tensor-fed signs (deterministic mode; fp32 +-1 laid out like x / out) or null
  // for Philox bits, and the geometry the A flippers need to find the sign-stream position of a tile row
  const float *sign_in, *sign_out;
  int IH, IW, Cin;                 // dense logical input image (conv) / features (linear)
  int kind_a, kind_o;              // sign streams of the A flippers / of the epilogue (forward: SIGN_IN / SIGN_OUT; dgrad: swapped)
  int a_dil_h, a_dil_w, SH, SW;    // dgrad of a strided conv: the A tensor is grad_out with zeros inserted (factor a_dil); SH x SW is
                                   // the un-inserted image the sign stream (and a tensor-fed sign tensor) is indexed by
  int a2_tma;                      // the sign-flipped A tiles come from a second tensor map (patch-matrix mode) instead of the flippers
  int64_t x_bs, x_hs, x_ws, x_pitch;  // strides of x in elements (tensor-fed sign_in shares them)
};

struct Ctrl {
  uint64_t a_full[MAX_NA], a_empty[MAX_NA];
  uint64_t af_full[MAX_NA];  // Flipout: the sign-flipped copy of the A slot is in place
  uint64_t w_full[MAX_RW + 1], w_empty[MAX_RW + 1];  // slot RW: the shift tile
  uint64_t acc_full[MAX_SETS], acc_empty[MAX_SETS];
  uint64_t res_full[4], res_empty[4];  // residual ring: one 128-row x 128-byte tile per slot, a 32-row slab per epilogue warp
  uint32_t tmem_base;
  uint32_t pad[3];
};
static_assert(sizeof(Ctrl) <= BAR_BYTES, "control block too large");

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_load_im2col_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c, int w, int h,
                                                   int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2], {%7, %8};"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
template <int N>
__device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ bool elect_one() {  // one lane of the (converged) warp
  uint32_t is_leader;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(is_leader));
  return is_leader != 0;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t fma_bf16x2(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ uint32_t pack_bf16_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}

// ---------------------------------------------------------------- parameter packing ----
// mu[n][c][tap] -> mu_p[n][tap*Cin + c] with row pitch Kp (zero beyond K), sigma_p = softplus(rho)
// likewise (sigma does not depend on the sample: hoisted out of the per-sample work), eps likewise.
// With KL = true the same pass also evaluates the layer's KL term, kl(weight) + kl(bias) as forward(return_kl=True)
// returns it (linear_variational.py:183-198): mu and rho are read ONCE for sampling and KL (8 B per weight; the
// separate KL launch read them a second time).  Per-thread fp32 partial sums are flushed into fp64 every 16
// elements, blocks add one fp64 atomic each, the last block to finish writes the two means' sum.
struct PackKl {
  double* acc;                   // [2]: sum over the weight, sum over the bias (zeroed by the caller)
  unsigned int* counter;         // blocks done (zeroed by the caller)
  float* kl_out;
  const float *pmu_t, *psig_t;   // tensor-valued weight priors in the parameter's own layout, or null
  const float *mu_b, *rho_b, *pmu_b_t, *psig_b_t;
  float prior_mu, prior_sigma;
  int64_t w_numel;
  int b_numel;
};

template <typename TP>
__device__ __forceinline__ TP pack_cast(float v);
template <>
__device__ __forceinline__ __nv_bfloat16 pack_cast<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ float pack_cast<float>(float v) { return v; }

template <bool KL, typename TP>
__global__ void __launch_bounds__(256) tc2_pack_kernel(const float* __restrict__ mu, const float* __restrict__ rho,
                                const float* __restrict__ eps, TP* __restrict__ mu_p, TP* __restrict__ sig_p,
                                float* __restrict__ eps_p, const float* __restrict__ scale, int64_t total, int Cin, int T,
                                int Kp, const PackKl kf) {
  const int K = Cin * T;
  const KlPrior prior_c = kl_prior(KL ? kf.prior_sigma : 1.f);
  float part = 0.f;
  double dsum = 0.0;
  int since = 0;
  // (n, k) of the packed index advance incrementally: a 64-bit division per element was most of this kernel's time
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  const int64_t dn = step / Kp;
  const int dk = (int)(step - dn * Kp);
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t n = i / Kp;
  int k = (int)(i - n * Kp);
  for (; i < total; i += step, n += dn, k += dk) {
    if (k >= Kp) { k -= Kp; ++n; }
    float m = 0.f, sg = 0.f, e = 0.f;
    if (k < K) {
      int tap = 0, c = k;
      if (T > 1) { tap = k / Cin; c = k - tap * Cin; }
      const int64_t src = n * K + (int64_t)c * T + tap;
      m = mu[src];
      const float r = rho[src];
      sg = bt_softplus(r);
      if (eps_p) e = eps[src];
      if constexpr (KL) {
        const float pm = kf.pmu_t ? kf.pmu_t[src] : kf.prior_mu;
        part += kl_elem<false>(m, r, pm, kf.psig_t ? kl_prior(kf.psig_t[src]) : prior_c, false).kl;
        if (++since == 16) { dsum += (double)part; part = 0.f; since = 0; }
      }
      if (scale) {  // folded BatchNorm scale of output channel n rides on the sampled weight row
        const float sc = scale[n];
        m *= sc;
        sg *= sc;
      }
    }
    mu_p[i] = pack_cast<TP>(m);
    sig_p[i] = pack_cast<TP>(sg);
    if (eps_p) eps_p[i] = e;
  }
  if constexpr (KL) {
    __shared__ double red[2][8];
    __shared__ bool is_last;
    dsum += (double)part;
    double bsum = 0.0;
    if (blockIdx.x == 0 && kf.mu_b) {   // the bias term: a few hundred elements, one block
      for (int i = threadIdx.x; i < kf.b_numel; i += blockDim.x) {
        const float pm = kf.pmu_b_t ? kf.pmu_b_t[i] : kf.prior_mu;
        bsum += (double)kl_elem<false>(kf.mu_b[i], kf.rho_b[i], pm, kf.psig_b_t ? kl_prior(kf.psig_b_t[i]) : prior_c, false).kl;
      }
    }
    dsum = bt_warp_sum(dsum);
    bsum = bt_warp_sum(bsum);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = dsum; red[1][threadIdx.x >> 5] = bsum; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double a = 0.0, b = 0.0;
      for (int w = 0; w < 8; ++w) { a += red[0][w]; b += red[1][w]; }
      atomicAdd(&kf.acc[0], a);
      if (blockIdx.x == 0) atomicAdd(&kf.acc[1], b);
      __threadfence();
      is_last = atomicAdd(kf.counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (is_last && threadIdx.x == 0) {
      __threadfence();
      double kl = __ldcg(&kf.acc[0]) / (double)kf.w_numel;
      if (kf.mu_b && kf.b_numel > 0) kl += __ldcg(&kf.acc[1]) / (double)kf.b_numel;
      *kf.kl_out = (float)kl;
    }
  }
}

// Backward-data packing: dX = conv(dY [zero-inserted], W') with W'[c][tap'][n] = W[n][c][T-1-tap'] (the
// transposed, tap-reversed filter).  mu_p[c][tap'*Cout + n], sigma_p likewise, and eps_p = the SAME draw
// the forward used for that weight: stream position (n*T + tap)*Cin + c, regenerated here (one Philox call
// per weight; |W| is small next to the activations) or taken from the caller's eps tensor.
__global__ void tc2_pack_dgrad_kernel(const float* __restrict__ mu, const float* __restrict__ rho,
                                      const float* __restrict__ eps, __nv_bfloat16* __restrict__ mu_p, __nv_bfloat16* __restrict__ sig_p,
                                      float* __restrict__ eps_p, int64_t total, int Cin, int Cout, int T, int Kp, BtRngKey key) {
  const int K = Cout * T;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = i / Kp;
    const int k = (int)(i - c * Kp);
    float m = 0.f, sg = 0.f, e = 0.f;
    if (k < K) {
      const int tap2 = k / Cout, n = k - tap2 * Cout;
      const int tap = T - 1 - tap2;
      const int64_t src = ((int64_t)n * Cin + c) * T + tap;
      m = mu[src];
      sg = bt_softplus(rho[src]);
      e = eps ? eps[src] : bt_normal1((uint64_t)(((int64_t)n * T + tap) * Cin + c), key);
    }
    mu_p[i] = __float2bfloat16_rn(m);
    sig_p[i] = __float2bfloat16_rn(sg);
    eps_p[i] = e;
  }
}

// grad_out with (stride - 1) zeros between pixels: dst[b][oh*sh][ow*sw][:] = src[b][oh][ow][:], 16 bytes per thread
// (dst is zero-filled by the caller).
__global__ void tc2_zero_insert_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t chunks, int C8,
                                       int OH, int OW, int sh, int sw, int64_t sbs, int64_t shs, int64_t sws, int HZ, int WZ, int C) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < chunks; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = i / C8;
    const int c8 = (int)(i - r * C8);
    const int ow = (int)(r % OW); r /= OW;
    const int oh = (int)(r % OH);
    const int64_t b = r / OH;
    const uint4 v = *reinterpret_cast<const uint4*>(src + b * sbs + (int64_t)oh * shs + (int64_t)ow * sws + c8 * 8);
    *reinterpret_cast<uint4*>(dst + ((b * HZ + (int64_t)oh * sh) * WZ + (int64_t)ow * sw) * C + c8 * 8) = v;
  }
}

// ---------------------------------------------------------------- patch matrix ----
// Explicit im2col of x into a dense bf16 matrix A[row][k], k = tap*Cin + c, row pitch Kp, for
// convolutions the im2col tensor map cannot describe (C_in not a multiple of 64 -- the RGB stem --
// rank 3, or x not channels-last).  Costs rows*Kp*2 bytes of extra traffic, paid once per call
// and shared by every Monte-Carlo sample of the call when x is shared.
struct PatchGeo {
  int64_t rows, xbs, xcs, xs0, xs1, xs2;
  int Kp, K, Cin, KH, KW;
  int OD, OH, OW, ID, IH, IW;
  int sd, sh, sw, pd, ph, pw, dd, dh, dw;
};

// One thread per 16-byte chunk (8 consecutive k) of the patch matrix: consecutive threads write
// consecutive chunks and read neighbouring input elements.  Index arithmetic is 32-bit whenever
// the chunk count allows (64-bit division is ~10x the cost and was most of this kernel's time).
template <typename TX, typename IDX>
__global__ void __launch_bounds__(256) tc2_patch_kernel(const TX* __restrict__ x, TX* __restrict__ A, const PatchGeo g) {
  const IDX C8 = (IDX)(g.Kp >> 3);
  const IDX total = (IDX)g.rows * C8;
  const IDX stride = (IDX)gridDim.x * blockDim.x;
  for (IDX i = (IDX)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const IDX row = i / C8;
    const int chunk = (int)(i - row * C8);
    IDX r = row;
    const int ow = (int)(r % (IDX)g.OW); r /= (IDX)g.OW;
    const int oh = (int)(r % (IDX)g.OH); r /= (IDX)g.OH;
    const int od = (int)(r % (IDX)g.OD); r /= (IDX)g.OD;
    const TX* xi = x + (int64_t)r * g.xbs;
    const int id0 = od * g.sd - g.pd, ih0 = oh * g.sh - g.ph, iw0 = ow * g.sw - g.pw;
    int k = chunk * 8;
    const int tap = k / g.Cin;
    int c = k - tap * g.Cin;
    int kw = tap % g.KW;
    const int t2 = tap / g.KW;
    int kh = t2 % g.KH, kd = t2 / g.KH;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e, ++k) {
      v[e] = 0.f;
      if (k < g.K) {
        const int id = id0 + kd * g.dd, ih = ih0 + kh * g.dh, iw = iw0 + kw * g.dw;
        if ((unsigned)id < (unsigned)g.ID && (unsigned)ih < (unsigned)g.IH && (unsigned)iw < (unsigned)g.IW)
          v[e] = bt_to_float<TX>(xi[(int64_t)c * g.xcs + (int64_t)id * g.xs0 + (int64_t)ih * g.xs1 + (int64_t)iw * g.xs2]);
      }
      if (++c == g.Cin) { c = 0; if (++kw == g.KW) { kw = 0; if (++kh == g.KH) { kh = 0; ++kd; } } }
    }
    if constexpr (sizeof(TX) == 4) {   // fp32 activations (tf32 path): the patch matrix stays fp32, 32 bytes per thread
      float4* dst = reinterpret_cast<float4*>(A + (int64_t)row * g.Kp + chunk * 8);
      dst[0] = make_float4(v[0], v[1], v[2], v[3]);
      dst[1] = make_float4(v[4], v[5], v[6], v[7]);
    } else {
      uint4 pk;
      pk.x = pack_bf16(v[0], v[1]); pk.y = pack_bf16(v[2], v[3]); pk.z = pack_bf16(v[4], v[5]); pk.w = pack_bf16(v[6], v[7]);
      *reinterpret_cast<uint4*>(A + (int64_t)row * g.Kp + chunk * 8) = pk;
    }
  }
}

// Table-driven variant for the common case (short reduction, small window: the RGB stem).  The k -> (offset
// inside the window, tap position) decode is done ONCE per block into shared memory and the row -> (image,
// od, oh, ow) decode once per 64-row tile, so the inner loop is two shared loads, three compares and one
// global load per element: the generic kernel above spent most of its time in integer division.
constexpr int PT_ROWS = 64;     // rows of the patch matrix per tile
constexpr int PT_MAXK = 2048;   // longest padded reduction the tables hold (16 KB of shared memory)

template <typename TX>
__global__ void __launch_bounds__(256) tc2_patch_tab_kernel(const TX* __restrict__ x, TX* __restrict__ A, const PatchGeo g) {
  __shared__ int tab_off[PT_MAXK];        // element offset of k inside the window; INT_MIN for the K padding
  __shared__ uint32_t tab_pos[PT_MAXK];   // kd*dd << 20 | kh*dh << 10 | kw*dw
  __shared__ int64_t row_base[PT_ROWS];   // element offset of the window's (possibly out-of-image) origin
  __shared__ int row_d[PT_ROWS], row_h[PT_ROWS], row_w[PT_ROWS];
  for (int k = threadIdx.x; k < g.Kp; k += blockDim.x) {
    if (k < g.K) {
      const int tap = k / g.Cin, c = k - tap * g.Cin;
      const int kw = tap % g.KW, t2 = tap / g.KW;
      const int kh = t2 % g.KH, kd = t2 / g.KH;
      tab_off[k] = (int)((int64_t)c * g.xcs + (int64_t)kd * g.dd * g.xs0 + (int64_t)kh * g.dh * g.xs1 + (int64_t)kw * g.dw * g.xs2);
      tab_pos[k] = ((uint32_t)(kd * g.dd) << 20) | ((uint32_t)(kh * g.dh) << 10) | (uint32_t)(kw * g.dw);
    } else {
      tab_off[k] = INT_MIN;
      tab_pos[k] = 0;
    }
  }
  const int C8 = g.Kp >> 3;
  const int64_t tiles = (g.rows + PT_ROWS - 1) / PT_ROWS;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t row0 = tile * PT_ROWS;
    const int nrows = (int)(g.rows - row0 < PT_ROWS ? g.rows - row0 : PT_ROWS);
    __syncthreads();   // tables written / previous tile's rows consumed
    if ((int)threadIdx.x < nrows) {
      int64_t r = row0 + threadIdx.x;
      const int ow = (int)(r % g.OW); r /= g.OW;
      const int oh = (int)(r % g.OH); r /= g.OH;
      const int od = (int)(r % g.OD); r /= g.OD;
      const int id0 = od * g.sd - g.pd, ih0 = oh * g.sh - g.ph, iw0 = ow * g.sw - g.pw;
      row_d[threadIdx.x] = id0; row_h[threadIdx.x] = ih0; row_w[threadIdx.x] = iw0;
      row_base[threadIdx.x] = r * g.xbs + (int64_t)id0 * g.xs0 + (int64_t)ih0 * g.xs1 + (int64_t)iw0 * g.xs2;
    }
    __syncthreads();
    const int chunks = nrows * C8;
    TX* At = A + row0 * g.Kp;   // the tile's rows are contiguous: chunk j sits at At + 8 j
    for (int j = threadIdx.x; j < chunks; j += blockDim.x) {
      const int rl = j / C8;
      const int k0 = (j - rl * C8) * 8;
      const TX* xi = x + row_base[rl];
      const int id0 = row_d[rl], ih0 = row_h[rl], iw0 = row_w[rl];
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int off = tab_off[k0 + e];
        const uint32_t pos = tab_pos[k0 + e];
        const int id = id0 + (int)(pos >> 20), ih = ih0 + (int)((pos >> 10) & 1023u), iw = iw0 + (int)(pos & 1023u);
        const bool ok = off != INT_MIN && (unsigned)id < (unsigned)g.ID && (unsigned)ih < (unsigned)g.IH && (unsigned)iw < (unsigned)g.IW;
        v[e] = ok ? bt_to_float<TX>(xi[off]) : 0.f;
      }
      if constexpr (sizeof(TX) == 4) {
        float4* dst = reinterpret_cast<float4*>(At + (int64_t)j * 8);
        dst[0] = make_float4(v[0], v[1], v[2], v[3]);
        dst[1] = make_float4(v[4], v[5], v[6], v[7]);
      } else {
        uint4 pk;
        pk.x = pack_bf16(v[0], v[1]); pk.y = pack_bf16(v[2], v[3]); pk.z = pack_bf16(v[4], v[5]); pk.w = pack_bf16(v[6], v[7]);
        *reinterpret_cast<uint4*>(At + (int64_t)j * 8) = pk;
      }
    }
  }
}

// ---------------------------------------------------------------- pre-sampled weights ----
// W[s][n][k] = bf16(sigma * eps(s) + mu) for every sample of a call, written once, in the arithmetic and at the
// Philox positions of the in-kernel sampling loop (bit-identical weights).  For the layers whose W k-blocks are
// STREAMED the fused kernel re-samples a [BN x K] panel for every (sample, n-block, m-group) item -- 25 times per
// weight on the 3x3 layers at 14x14 -- and its 16 producer warps, not the tensor core, set the pace (65 % of all
// issue slots, tensor pipe 46 %; profiles/r2_conv3x3_fwd_ncu_metrics.txt).  Materialising S x |W| x 2 bytes costs
// ~3 GB of traffic per ResNet-50 step (0.5 ms) and turns those layers into plain TMA-fed GEMMs.
// One thread per 16-byte chunk (eight weights = one Philox block); blockIdx.y = sample.
__global__ void __launch_bounds__(256) tc2_presample_kernel(const uint4* __restrict__ mu_p, const uint4* __restrict__ sig_p,
                                                            uint4* __restrict__ wq, int64_t chunks, uint64_t seed, uint32_t sample_id,
                                                            uint32_t layer_id, int pdl, const float* __restrict__ eps_p, int flip) {
  // eps_p: the draws come from a packed fp32 tensor in the weights' order (backward data: tc2_pack_dgrad_kernel regenerated
  // them at the FORWARD stream positions); flip: Flipout keeps mu apart, only delta = bf16(sigma * eps) is materialised.
  // Programmatic dependent launch (pdl != 0): this grid may start while the previous kernel of the stream drains.
  // pdl = 1: the packed parameters were written by that kernel -- wait first.  pdl = 2 (frozen layer: the packed
  // parameters are old, nothing here depends on the previous kernel): sample during its tail and wait at the END, so
  // that this grid -- and with it the GEMM that waits on it -- still completes after everything before it.
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (pdl == 1) asm volatile("griddepcontrol.wait;" ::: "memory");
  const int s = blockIdx.y;
  const BtRngKey rk = bt_rng_key(seed, sample_id + (uint32_t)s, layer_id, BT_KIND_EPS_W);
  uint4* __restrict__ dst = wq + (int64_t)s * chunks;
  for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < chunks; c += (int64_t)gridDim.x * blockDim.x) {
    const uint4 sg = __ldg(sig_p + c);
    uint4 mu = make_uint4(0u, 0u, 0u, 0u);
    if (!flip) mu = __ldg(mu_p + c);
    float z[8];
    if (eps_p) {
      const float4 e0 = __ldg(reinterpret_cast<const float4*>(eps_p) + 2 * c), e1 = __ldg(reinterpret_cast<const float4*>(eps_p) + 2 * c + 1);
      z[0] = e0.x; z[1] = e0.y; z[2] = e0.z; z[3] = e0.w; z[4] = e1.x; z[5] = e1.y; z[6] = e1.z; z[7] = e1.w;
    } else {
      bt_normal8((uint64_t)c, rk, z);
    }
    uint4 w;
    w.x = fma_bf16x2(sg.x, pack_bf16(z[0], z[1]), mu.x);
    w.y = fma_bf16x2(sg.y, pack_bf16(z[2], z[3]), mu.y);
    w.z = fma_bf16x2(sg.z, pack_bf16(z[4], z[5]), mu.z);
    w.w = fma_bf16x2(sg.w, pack_bf16(z[6], z[7]), mu.w);
    dst[c] = w;
  }
  if (pdl == 2) asm volatile("griddepcontrol.wait;" ::: "memory");
}

// Padded reduction (patch-matrix mode: K rounded up to 8, the eps stream keeps the TRUE pitch Kt): element by element.
// Only tiny weight tensors take this path (the RGB stem: 64 x 147).
__global__ void __launch_bounds__(256) tc2_presample_padded_kernel(const __nv_bfloat16* __restrict__ mu_p, const __nv_bfloat16* __restrict__ sig_p,
                                                                   __nv_bfloat16* __restrict__ wq, int N, int K, int Kt, uint64_t seed,
                                                                   uint32_t sample_id, uint32_t layer_id, int pdl) {
  if (pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (pdl == 1) asm volatile("griddepcontrol.wait;" ::: "memory");
  const int s = blockIdx.y;
  const BtRngKey rk = bt_rng_key(seed, sample_id + (uint32_t)s, layer_id, BT_KIND_EPS_W);
  const int64_t pairs = (int64_t)N * K / 2;   // K % 8 == 0: a pair never straddles two rows
  const uint32_t* mu2 = reinterpret_cast<const uint32_t*>(mu_p);
  const uint32_t* sg2 = reinterpret_cast<const uint32_t*>(sig_p);
  uint32_t* dst = reinterpret_cast<uint32_t*>(wq) + (int64_t)s * pairs;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < pairs; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(2 * i / K), k = (int)(2 * i - (int64_t)n * K);
    const float z0 = k < Kt ? bt_normal1((uint64_t)n * Kt + k, rk) : 0.f;
    const float z1 = k + 1 < Kt ? bt_normal1((uint64_t)n * Kt + k + 1, rk) : 0.f;
    dst[i] = fma_bf16x2(sg2[i], pack_bf16(z0, z1), mu2[i]);   // the instruction of the in-kernel sampling loop
  }
  if (pdl == 2) asm volatile("griddepcontrol.wait;" ::: "memory");
}

// x * s_in as a bf16 tensor with x's strides (patch-matrix Flipout: the RGB stem, odd C_in): one thread per
// word of the SIGN_IN stream, i.e. 32 consecutive elements of the sample in channels-last order.
struct FlipGeo {
  int64_t words_per_s, elems_per_s, xss, xbs, xcs, xs0, xs1, xs2;
  int C, ID, IH, IW, S, kind;
  uint64_t seed;
  uint32_t sample_id, layer_id;
};
__global__ void __launch_bounds__(256) tc2_flipx_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ xf, const FlipGeo g) {
  const int64_t total = g.words_per_s * g.S;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int s = (int)(i / g.words_per_s);
    const int64_t w = i - (int64_t)s * g.words_per_s;
    const BtRngKey key = bt_rng_key(g.seed, g.sample_id + (uint32_t)s, g.layer_id, g.kind);
    const uint4 r = bt_philox((uint64_t)(w >> 2), key);
    const int q = (int)(w & 3);
    const uint32_t word = q == 0 ? r.x : (q == 1 ? r.y : (q == 2 ? r.z : r.w));
    int64_t li = w << 5;
    int c = (int)(li % g.C);
    int64_t pix = li / g.C;
    int iw = (int)(pix % g.IW); pix /= g.IW;
    int ih = (int)(pix % g.IH); pix /= g.IH;
    int id = (int)(pix % g.ID);
    int64_t b = pix / g.ID;
    for (int j = 0; j < 32 && li < g.elems_per_s; ++j, ++li) {
      const int64_t off = (int64_t)s * g.xss + b * g.xbs + (int64_t)c * g.xcs + (int64_t)id * g.xs0 + (int64_t)ih * g.xs1 + (int64_t)iw * g.xs2;
      const unsigned short v = __bfloat16_as_ushort(x[off]);
      xf[off] = __ushort_as_bfloat16((unsigned short)(v ^ (((word >> bt_sign_bitpos((uint32_t)j)) & 1u) << 15)));
      if (++c == g.C) { c = 0; if (++iw == g.IW) { iw = 0; if (++ih == g.IH) { ih = 0; if (++id == g.ID) { id = 0; ++b; } } } }
    }
  }
}

// Epilogue of one 32-column unit of one accumulator row: ReLU, conversion and the swizzled stores
// into chunks [cb, cb + 32/EPC) of the 128-byte staging row.  Bias and folded BatchNorm never show
// up here: the scale is folded into the sampled weights and the shift is in the accumulator
// already (rank-1 MMA), so for bf16 outputs this is one cvt (with .relu) per two elements.
template <typename TO, bool RELU, bool RES>
__device__ __forceinline__ void epi_unit(const uint32_t (&v)[32], uint32_t rowaddr, uint32_t resaddr, int cb, uint32_t row_sw) {
  constexpr int EPC = 16 / (int)sizeof(TO);  // elements per 16-byte chunk
  constexpr int NCH = 32 / EPC;              // chunks per unit
  uint32_t r[RES ? NCH : 1][4];
  if constexpr (RES) {  // the residual tile sits in shared memory in the staging layout: all loads first
#pragma unroll
    for (int j = 0; j < NCH; ++j)
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(r[j][0]), "=r"(r[j][1]), "=r"(r[j][2]), "=r"(r[j][3])
                   : "r"(resaddr + (((uint32_t)(cb + j) ^ row_sw) << 4)));
  }
#pragma unroll
  for (int j = 0; j < NCH; ++j) {
    float f[EPC];
#pragma unroll
    for (int e = 0; e < EPC; ++e) f[e] = __uint_as_float(v[j * EPC + e]);
    uint32_t w[4];
    if constexpr (sizeof(TO) == 2) {
      if constexpr (RELU && !RES) {  // cvt.rn.relu: the clamp rides on the conversion
        w[0] = pack_bf16_relu(f[0], f[1]); w[1] = pack_bf16_relu(f[2], f[3]);
        w[2] = pack_bf16_relu(f[4], f[5]); w[3] = pack_bf16_relu(f[6], f[7]);
      } else {
        w[0] = pack_bf16(f[0], f[1]); w[1] = pack_bf16(f[2], f[3]); w[2] = pack_bf16(f[4], f[5]); w[3] = pack_bf16(f[6], f[7]);
      }
      if constexpr (RES) {  // bf16(bf16(affine) + residual), then the clamp: what the eager bf16 graph computes
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          __nv_bfloat162 t = __hadd2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]), *reinterpret_cast<const __nv_bfloat162*>(&r[j][i]));
          if constexpr (RELU) t = __hmax2(t, __floats2bfloat162_rn(0.f, 0.f));
          w[i] = *reinterpret_cast<const uint32_t*>(&t);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if constexpr (RES) f[i] += __uint_as_float(r[j][i]);
        if constexpr (RELU) f[i] = fmaxf(f[i], 0.f);
        w[i] = __float_as_uint(f[i]);
      }
    }
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + (((uint32_t)(cb + j) ^ row_sw) << 4)), "r"(w[0]),
                 "r"(w[1]), "r"(w[2]), "r"(w[3])
                 : "memory");
  }
}

// Cycle accounting for BT_TC2_PROF=1 (diagnostics; a separate instantiation, the product kernel
// carries none of it): per CTA and role, cycles spent in each kind of wait.
constexpr int PROF_ROLES = 6, PROF_SLOTS = 8;
__device__ unsigned long long g_tc_prof[160 * PROF_ROLES * PROF_SLOTS];
#define PW(i, ...)                                              \
  do {                                                          \
    if constexpr (PROF) {                                       \
      const long long _t0 = clock64();                          \
      __VA_ARGS__;                                              \
      pc[i] += (unsigned long long)(clock64() - _t0);           \
    } else {                                                    \
      __VA_ARGS__;                                              \
    }                                                           \
  } while (0)

// TF32: fp32 activations and fp32 outputs (the reference's default dtype, linear_variational.py:181 /
// conv_variational.py:384).  The A tiles are fp32 (a 128-byte swizzled row holds 32 elements: a k-block is 32
// deep), the sampled weight is fmaf(sigma, eps, mu) in fp32 rounded to tf32 (cvt.rna), the MMAs are
// tcgen05.mma.kind::tf32 (K = 8 per instruction, fp32 accumulation).  Every byte layout -- slots, swizzle,
// descriptors, staging -- is the bf16 one; only the element counts per 16 / 128 bytes change.
// EPI8: EIGHT epilogue warps.  On the layers whose output (and residual) stream is the bottleneck -- wide 1x1
// expansions with a short reduction -- four epilogue warps, one per TMEM lane quarter and scheduler, are latency bound
// (tcgen05.ld -> convert -> st.shared -> proxy fence -> TMA store, ~1 200-1 700 cycles per 32 x 64 slab: 14 B / cycle /
// SM of output against an HBM share of 25 B / cycle).  Producer warps 12-15 become a second epilogue group: group g
// drains columns [64 g, 64 g + 64) of every 128-column tile, two warps per lane quarter interleave on each scheduler.
template <int BN, typename TO, bool PROF, int CL, bool FLIP, bool TF32 = false, bool EPI8 = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc2_fwd_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_out,
               const __grid_constant__ CUtensorMap tm_res, const __grid_constant__ CUtensorMap tm_a2,
               const __grid_constant__ Tc2Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  Ctrl* ctrl = reinterpret_cast<Ctrl*>(smem_raw + (base - raw));
  // Flipout: every k-block carries TWO weight sub-tiles (mu, then sigma*eps), every A slot a second,
  // sign-flipped copy of the tile, and every output tile two accumulators (mean path, perturbation path)
  static_assert(!(TF32 && (FLIP || PROF)) && (!TF32 || sizeof(TO) == 4), "tf32: Reparameterization, fp32 in / fp32 out");
  static_assert(!EPI8 || (BN == 128 && sizeof(TO) == 2 && !FLIP && !PROF && !TF32), "EPI8: bf16 Reparameterization, BN = 128");
  constexpr int BKE = TF32 ? 32 : 64;   // elements of one k-block (128 bytes of A per row)
  constexpr int EPC16 = TF32 ? 4 : 8;   // weights per 16-byte chunk
  constexpr int NSUBW = FLIP ? 2 : 1;
  constexpr int W_SUB = BN * BK * 2;
  constexpr int W_SLOT = NSUBW * W_SUB;
  constexpr int A_SLOT_T = NSUBW * A_SLOT;
  constexpr int ACC_COLS = NSUBW * BN;
  constexpr int NPW = (FLIP || EPI8) ? N_PROD_WARPS - N_FLIP_WARPS : N_PROD_WARPS;  // producer warps (12 when warps 12-15 have another role)
  constexpr int PT = NPW * 32;
  constexpr int SUBN = 128 / (int)sizeof(TO);  // output columns per 128-byte staging row
  const uint32_t ones_base = base + BAR_BYTES, shift_base = ones_base + ONES_BYTES;
  const uint32_t stage_base = base + CTRL_BYTES + (FLIP ? SHIFT_BYTES : 0);  // Flipout: a second shift tile (perturbation bias)
  const uint32_t res_base = stage_base + (uint32_t)(a.stage_bufs * (EPI8 ? 2 : 1)) * STAGE_UNIT;
  const uint32_t a_base = res_base + (uint32_t)a.res_bufs * STAGE_UNIT;
  const uint32_t w_base = a_base + a.NA * A_SLOT_T;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned long long pc[PROF_SLOTS] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long prof_t0 = PROF ? clock64() : 0;
  auto prof_flush = [&](int role) {
    if constexpr (PROF) {
      pc[PROF_SLOTS - 1] = (unsigned long long)(clock64() - prof_t0);
      if (lane == 0 && blockIdx.x < 160)
        for (int i = 0; i < PROF_SLOTS; ++i) g_tc_prof[(blockIdx.x * PROF_ROLES + role) * PROF_SLOTS + i] = pc[i];
    }
  };

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.NA; ++i) {
      mbar_init(smem_u32(&ctrl->a_full[i]), 1); mbar_init(smem_u32(&ctrl->a_empty[i]), 1);
      mbar_init(smem_u32(&ctrl->af_full[i]), N_FLIP_WARPS);
    }
    for (int i = 0; i < a.RW; ++i) { mbar_init(smem_u32(&ctrl->w_full[i]), a.w_pre ? 1 : NPW); mbar_init(smem_u32(&ctrl->w_empty[i]), N_MMA_WARPS * CL); }
    mbar_init(smem_u32(&ctrl->w_full[a.RW]), NPW);  // the shift tile stays CTA-local
    mbar_init(smem_u32(&ctrl->w_empty[a.RW]), N_MMA_WARPS);
    for (int i = 0; i < a.NSETS; ++i) mbar_init(smem_u32(&ctrl->acc_full[i]), N_MMA_WARPS);
    // acc_empty is per accumulator TILE (set * MT + t): the MMAs of the next item restart on tile t as soon as the
    // epilogue has drained it, instead of waiting for the whole set
    for (int i = 0; i < a.NSETS * a.MT; ++i) mbar_init(smem_u32(&ctrl->acc_empty[i]), EPI8 ? 8 : 4);   // one arrival per epilogue warp (32 serialized arrivals per warp and tile before)
    for (int i = 0; i < 4; ++i) { mbar_init(smem_u32(&ctrl->res_full[i]), 1); mbar_init(smem_u32(&ctrl->res_empty[i]), 4); }
    mbar_fence_init();
  }
  if (warp == WARP_TMA && lane == 0) {
    prefetch_tmap(&tm_a); prefetch_tmap(&tm_out);
    if (a.res_bufs) prefetch_tmap(&tm_res);
    if ((FLIP && a.a2_tma) || a.w_pre) prefetch_tmap(&tm_a2);
    if (FLIP && a.w_pre) prefetch_tmap(&tm_res);
  }
  if (warp == WARP_MMA) tmem_alloc(smem_u32(&ctrl->tmem_base), 512);
  tc_fence_before();
  __syncthreads();
  if constexpr (CL > 1) cluster_sync_all();  // the peer's mbarriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem = ctrl->tmem_base;
  // Programmatic dependent launch: this grid may have been scheduled while the previous kernel of the stream was still
  // draining (its CTAs leave their SMs one by one; ours move in, initialise barriers, allocate TMEM, prefetch the tensor
  // maps above).  Let the NEXT kernel do the same behind us, then wait -- here, before the first byte of global memory
  // that a previous kernel may have written (activations, residual) is touched or any output is stored.
  if (a.pdl) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (!(a.pdl == 2 && warp < NPW)) asm volatile("griddepcontrol.wait;" ::: "memory");
  }

  const int vN = a.N * a.spn;  // width of the (virtual) N axis the tiles are cut from
  const int crank = CL > 1 ? (int)cluster_ctarank() : 0;
  const int cid = blockIdx.x / CL, ncl = gridDim.x / CL;
  // Items are dealt to the clusters in contiguous ranges, or (a.rr, order 1 only) round-robin: cluster c takes items
  // c, c + ncl, ...  With the n-block fastest, the NB passes over one group of A tiles then run on NEIGHBOURING
  // clusters at the same time and all but one are served by the L2, whatever the length of the reduction.  A
  // contiguous range runs them one after the other on one SM, and the 148 A streams in flight between two passes
  // (MT x 128 rows x K each: 150-300 MB at K = 1024-2048) evict the tiles before they are read again.
  const int64_t i0 = a.rr ? (int64_t)cid : a.total_items * cid / ncl;
  const int64_t i1 = a.total_items * (cid + 1) / ncl;
  const int n_items = a.rr ? (int)((a.total_items - cid + ncl - 1) / ncl) : (int)(i1 - i0);
  // first m-tile of this CTA's m-group within cluster step mg (may lie past num_mt for the odd
  // tail of a pair: such an item has no tiles, but its W k-blocks are still produced and consumed)
  auto tile0 = [&](int mg) { return (mg * CL + crank) * a.MT; };
  auto tiles_of = [&](int mg) { return max(0, min(a.MT, a.num_mt - tile0(mg))); };
  // first item of this CTA, decoded once; the role loops then walk (s, nb, mg) incrementally --
  // the single-thread TMA / MMA loops are instruction-latency bound, so no division or modulo
  // by a run-time value may sit inside them.
  // Item order: (sample, n-block, m-group) with the m-group fastest when a sampled panel is kept
  // across m-groups (stationary), (sample, m-group, n-block) with the n-block fastest when W is
  // streamed: then the NB passes over one group of A tiles follow each other on the same SM and
  // all but the first are L2 hits (measured before: every pass re-read A from HBM).
  int s0, nb0, mg0;
  if (a.nb_inner == 2) {
    const int64_t key0 = i0 / a.S;
    s0 = (int)(i0 - key0 * a.S);
    mg0 = (int)(key0 / a.NB);
    nb0 = (int)(key0 - (int64_t)mg0 * a.NB);
  } else if (a.nb_inner) {
    const int64_t key0 = i0 / a.NB;
    nb0 = (int)(i0 - key0 * a.NB);
    s0 = (int)(key0 / a.num_mg);
    mg0 = (int)(key0 - (int64_t)s0 * a.num_mg);
  } else {
    const int64_t key0 = i0 / a.num_mg;
    mg0 = (int)(i0 - key0 * a.num_mg);
    s0 = (int)(key0 / a.NB);
    nb0 = (int)(key0 - (int64_t)s0 * a.NB);
  }
  // advances (s, nb, mg) to the next item; true when a new (sample, n-block) panel starts (stationary order only)
  auto walk = [&](int& ws, int& wnb, int& wmg) -> bool {
    if (a.nb_inner == 2) {
      if (++ws == a.S) { ws = 0; if (++wnb == a.NB) { wnb = 0; ++wmg; } }
      return false;
    }
    if (a.rr) {  // + ncl items in the (s, mg, nb) mixed radix; the digits of ncl come from the host
      wnb += a.rr_dnb;
      int c = 0;
      if (wnb >= a.NB) { wnb -= a.NB; c = 1; }
      wmg += a.rr_dmg + c;
      c = 0;
      if (wmg >= a.num_mg) { wmg -= a.num_mg; c = 1; }
      ws += a.rr_ds + c;
      return false;
    }
    if (a.nb_inner) {
      if (++wnb == a.NB) { wnb = 0; if (++wmg == a.num_mg) { wmg = 0; ++ws; } }
      return false;
    }
    if (++wmg == a.num_mg) { wmg = 0; if (++wnb == a.NB) { wnb = 0; ++ws; } return true; }
    return false;
  };

  const bool epi_a = warp >= WARP_EPI0 && warp < WARP_EPI0 + 4;
  const bool epi_b = EPI8 && warp >= NPW && warp < NPW + 4;
  if (EPI8 && (epi_a || epi_b)) {
    if constexpr (EPI8) {
    // =========================== EPILOGUE, two groups of four warps ===========================
    reg_alloc<REGS_EPI8>();
    const int grp = epi_a ? 0 : 1;   // columns [64 grp, 64 grp + 64) of every tile
    const int q = warp & 3;          // TMEM lane quarter (WARP_EPI0 and NPW are multiples of 4)
    const uint32_t my_stage = stage_base + (uint32_t)((grp * 4 + q) * a.stage_bufs) * 4096u;
    const uint32_t my_res = res_base + (uint32_t)q * 4096u;
    const uint32_t row_sw = (uint32_t)(lane & 7);
    uint32_t nstore = 0, sbuf = 0;
    int s = s0, nb = nb0, mg = mg0, set = 0;
    uint32_t set_phase = 0;
    int rslot = grp;                 // the residual ring alternates sub-tile 0 / 1: group g owns the slots = g (mod 2)
    uint32_t rphase = 0;
    auto epi8_loop = [&](auto relu_c, auto res_c) {
      constexpr bool RELU = decltype(relu_c)::value, RES = decltype(res_c)::value;
      for (int it = 0; it < n_items; ++it) {
        const int mtv = tiles_of(mg);
        if (mtv == 0) { walk(s, nb, mg); continue; }
        mbar_wait(smem_u32(&ctrl->acc_full[set]), set_phase);
        tc_fence_after();
        for (int t = 0; t < mtv; ++t) {
          const int m0 = (tile0(mg) + t) * BM + q * 32;
          const uint32_t ebar = smem_u32(&ctrl->acc_empty[set * a.MT + t]);
          if (m0 >= a.Ms) {   // slab past the sample's last row: only the rings move on
            if constexpr (RES) {
              mbar_wait(smem_u32(&ctrl->res_full[rslot]), rphase);
              if (lane == 0) mbar_arrive(smem_u32(&ctrl->res_empty[rslot]));
              rslot += 2;
              if (rslot >= a.res_bufs) { rslot -= a.res_bufs; rphase ^= 1u; }
            }
            if (lane == 0) mbar_arrive(ebar);
            continue;
          }
          const uint32_t tacc = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)((set * a.MT + t) * ACC_COLS + grp * 64);
          if (nstore >= (uint32_t)a.stage_bufs) {   // the TMA store that last read this staging buffer has drained it
            if (lane == 0) {
              if (a.stage_bufs == 4) tma_store_wait_read<3>();
              else if (a.stage_bufs == 2) tma_store_wait_read<1>();
              else tma_store_wait_read<0>();
            }
            __syncwarp();
          }
          const uint32_t buf = my_stage + sbuf * 4096u;
          const uint32_t rowaddr = buf + (uint32_t)lane * 128u;
          const uint32_t resaddr = my_res + (uint32_t)rslot * (uint32_t)STAGE_UNIT + (uint32_t)lane * 128u;
          uint32_t va[32];
          tmem_ld32(tacc, va);
          if constexpr (RES) mbar_wait(smem_u32(&ctrl->res_full[rslot]), rphase);
          tmem_ld_wait();
          epi_unit<TO, RELU, RES>(va, rowaddr, resaddr, 0, row_sw);
          tmem_ld32(tacc + 32, va);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();             // every lane's reads of its half of tile t are complete: one arrival for the warp
          if (lane == 0) mbar_arrive(ebar);
          epi_unit<TO, RELU, RES>(va, rowaddr, resaddr, 4, row_sw);
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_3d(&tm_out, buf, nb * BN + grp * 64, m0, s);
            tma_store_commit();
          }
          ++nstore;
          if (++sbuf == (uint32_t)a.stage_bufs) sbuf = 0;
          if constexpr (RES) {
            if (lane == 0) mbar_arrive(smem_u32(&ctrl->res_empty[rslot]));
            rslot += 2;
            if (rslot >= a.res_bufs) { rslot -= a.res_bufs; rphase ^= 1u; }
          }
        }
        if (++set == a.NSETS) { set = 0; set_phase ^= 1u; }
        walk(s, nb, mg);
      }
    };
    using T8 = std::true_type;
    using F8 = std::false_type;
    if (a.res_bufs != 0) { if (a.out_relu) epi8_loop(T8{}, T8{}); else epi8_loop(F8{}, T8{}); }
    else { if (a.out_relu) epi8_loop(T8{}, F8{}); else epi8_loop(F8{}, F8{}); }
    if (lane == 0) tma_store_wait_all();
    }
  } else if (epi_a) {
    // =========================== EPILOGUE ===========================
    reg_alloc<REGS_EPI>();
    const int q = warp - WARP_EPI0;  // TMEM lane quarter: WARP_EPI0 is a multiple of 4
    const uint32_t my_stage = stage_base + (uint32_t)(q * a.stage_bufs) * 4096u;
    const uint32_t row_sw = (uint32_t)(lane & 7);
    uint32_t nstore = 0, sbuf = 0;
    const bool relu = a.out_relu != 0;
    constexpr int UPS = SUBN / 32;                    // 32-column units per staged sub-tile (2 bf16, 1 fp32)
    constexpr int CPU = 8 / UPS;                      // 16-byte chunks of the staging row per unit
    int s = s0, nb = nb0, mg = mg0, set = 0;
    uint32_t set_phase = 0;
    const bool has_res = a.res_bufs != 0;
    // Residual tiles arrive through a ring filled by the residual-loader warp (one 128-row TMA box
    // per sub-tile, res_bufs sub-tiles ahead); this warp reads its 32-row slab of each.
    const uint32_t my_res = res_base + (uint32_t)q * 4096u;
    int rslot = 0;
    uint32_t rphase = 0;
    // the whole loop is instantiated per (ReLU, residual) combination: the variant that runs is
    // straight-line code, and the other three never enter the instruction cache
    auto epi_loop = [&](auto relu_c, auto res_c) {
    constexpr bool RELU = decltype(relu_c)::value, RES = decltype(res_c)::value;
    auto emit = [&](const uint32_t (&v)[32], uint32_t rowaddr, uint32_t resaddr, int cb) {
      epi_unit<TO, RELU, RES>(v, rowaddr, resaddr, cb, row_sw);
    };
    for (int it = 0; it < n_items; ++it) {
      const int mtv = tiles_of(mg);
      if (mtv == 0) {  // tail of a CTA pair: nothing was accumulated for this CTA
        walk(s, nb, mg);
        continue;
      }
      PW(0, mbar_wait(smem_u32(&ctrl->acc_full[set]), set_phase));
      tc_fence_after();
      const int ncols = min(BN, vN - nb * BN);
      const int nsub = (ncols + SUBN - 1) / SUBN;  // staged sub-tiles with at least one real column
      for (int t = 0; t < mtv; ++t) {
        const int m0 = (tile0(mg) + t) * BM + q * 32;  // first row of this warp's 32-row slab
        if (m0 >= a.Ms) {  // slab entirely past the sample's last row (warp-uniform): only the residual ring moves on
          if constexpr (RES) {
            for (int sub = 0; sub < nsub; ++sub) {
              mbar_wait(smem_u32(&ctrl->res_full[rslot]), rphase);
              if (lane == 0) mbar_arrive(smem_u32(&ctrl->res_empty[rslot]));
              if (++rslot == a.res_bufs) { rslot = 0; rphase ^= 1u; }
            }
          }
          if (lane == 0) mbar_arrive(smem_u32(&ctrl->acc_empty[set * a.MT + t]));
          continue;
        }
        const uint32_t tacc = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)((set * a.MT + t) * ACC_COLS);
        // TMEM reads are software-pipelined in 32-column units: the load of unit u+1 is in
        // flight while unit u is converted and staged.
        uint32_t va[32], vb[32];
        if constexpr (!FLIP) tmem_ld32(tacc, va);
#pragma unroll 1
        for (int sub = 0; sub < nsub; ++sub) {
          // the TMA store that last read this staging buffer must have drained it
          if (nstore >= (uint32_t)a.stage_bufs) {
            PW(3, {
              if (lane == 0) {
                if (a.stage_bufs == 4) tma_store_wait_read<3>();
                else if (a.stage_bufs == 2) tma_store_wait_read<1>();
                else tma_store_wait_read<0>();
              }
              __syncwarp();
            });
          }
          const uint32_t buf = my_stage + sbuf * 4096u;
          const uint32_t rowaddr = buf + (uint32_t)lane * 128u;
          const uint32_t resaddr = my_res + (uint32_t)rslot * (uint32_t)STAGE_UNIT + (uint32_t)lane * 128u;
          if constexpr (RES) PW(1, mbar_wait(smem_u32(&ctrl->res_full[rslot]), rphase));
          const int c0 = sub * SUBN;
          if constexpr (FLIP) {
            // out = mean + pert * s_out: the two accumulators of the tile are joined here; s_out is
            // one word of the SIGN_OUT stream per row and 32-column unit (position m * N + n,
            // N % 32 == 0), or the caller's +-1 tensor (deterministic mode)
            const int mrow = m0 + lane;
#pragma unroll
            for (int u = 0; u < UPS; ++u) {
              const int c = c0 + 32 * u;
              tmem_ld32(tacc + c, va);
              tmem_ld32(tacc + BN + c, vb);
              uint32_t word = 0;
              const int n0 = nb * BN + c;
              if (a.sign_out) {
                if (mrow < a.Ms) {
                  const float4* sp = reinterpret_cast<const float4*>(a.sign_out + ((int64_t)s * a.Ms + mrow) * a.N + n0);
#pragma unroll
                  for (int j4 = 0; j4 < 8; ++j4) {
                    if (n0 + 4 * j4 + 4 > a.N) break;  // N % 8 == 0: a group of four is all in or all out
                    const float4 f = __ldg(sp + j4);
                    word |= (__float_as_uint(f.x) >> 31) << bt_sign_bitpos(4 * j4 + 0);
                    word |= (__float_as_uint(f.y) >> 31) << bt_sign_bitpos(4 * j4 + 1);
                    word |= (__float_as_uint(f.z) >> 31) << bt_sign_bitpos(4 * j4 + 2);
                    word |= (__float_as_uint(f.w) >> 31) << bt_sign_bitpos(4 * j4 + 3);
                  }
                }
              } else {
                const uint64_t idx = (uint64_t)mrow * (uint64_t)a.N + (uint64_t)n0;
                const BtRngKey key_so = bt_rng_key(a.seed, a.sample_id + s, a.layer_id, a.kind_o);
                const uint4 r = bt_philox(idx >> 7, key_so);
                const int wq = (int)((idx >> 5) & 3);
                word = wq == 0 ? r.x : (wq == 1 ? r.y : (wq == 2 ? r.z : r.w));
                if ((a.N & 31) != 0) {
                  // c_out % 32 != 0: the unit starts inside a word of the stream; gather its 32 bits from
                  // that word and the next one (rare shapes only, e.g. a 1000-class head)
                  const uint32_t sh = (uint32_t)idx & 31u;
                  const uint4 r2 = bt_philox(((idx >> 5) + 1) >> 2, key_so);
                  const int wq2 = (int)(((idx >> 5) + 1) & 3);
                  const uint32_t word2 = wq2 == 0 ? r2.x : (wq2 == 1 ? r2.y : (wq2 == 2 ? r2.z : r2.w));
                  uint32_t g = 0;
#pragma unroll
                  for (int j = 0; j < 32; ++j) {
                    const uint32_t e = sh + (uint32_t)j;
                    const uint32_t src = e < 32u ? word : word2;
                    g |= ((src >> bt_sign_bitpos(e & 31u)) & 1u) << ((j >> 1) + ((j & 1) << 4));
                  }
                  word = g;
                }
              }
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const int bp = (j >> 1) + ((j & 1) << 4);
                va[j] = __float_as_uint(__uint_as_float(va[j]) + __uint_as_float(vb[j] ^ ((word << (31 - bp)) & 0x80000000u)));
              }
              emit(va, rowaddr, resaddr, u * CPU);
            }
          } else if constexpr (UPS == 2) {
            PW(2, tmem_ld_wait());
            tmem_ld32(tacc + c0 + 32, vb);
            PW(4, emit(va, rowaddr, resaddr, 0));
            PW(2, tmem_ld_wait());
            if (sub + 1 < nsub) tmem_ld32(tacc + c0 + 64, va);
            PW(4, emit(vb, rowaddr, resaddr, CPU));
          } else {
            // fp32 output: one unit per sub-tile; ping-pong the two register arrays over pairs
            tmem_ld_wait();
            if (sub & 1) {
              if (sub + 1 < nsub) tmem_ld32(tacc + c0 + 32, va);
              emit(vb, rowaddr, resaddr, 0);
            } else {
              if (sub + 1 < nsub) tmem_ld32(tacc + c0 + 32, vb);
              emit(va, rowaddr, resaddr, 0);
            }
          }
          PW(5, {
            fence_proxy_async();
            __syncwarp();
            if (lane == 0 && !(PROF && (a.debug & 1))) {
              // rows >= Ms and columns >= N are clipped by the map; a stacked pair: column 64 * j + c is channel c of sample 2 s + j
              if (a.spn == 2) tma_store_3d(&tm_out, buf, (nb * BN + c0) & 63, m0, 2 * s + ((nb * BN + c0) >> 6));
              else tma_store_3d(&tm_out, buf, nb * BN + c0, m0, s);
              tma_store_commit();
            }
          });
          ++nstore;
          if (++sbuf == (uint32_t)a.stage_bufs) sbuf = 0;
          if constexpr (RES) {  // every lane is done reading the slab (fence + __syncwarp above): hand it back
            if (lane == 0) mbar_arrive(smem_u32(&ctrl->res_empty[rslot]));
            if (++rslot == a.res_bufs) { rslot = 0; rphase ^= 1u; }
          }
        }
        // every lane's TMEM reads of tile t have completed (tcgen05.wait::ld above, then the warp barrier): hand the tile back
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&ctrl->acc_empty[set * a.MT + t]));
      }
      if (++set == a.NSETS) { set = 0; set_phase ^= 1u; }
      walk(s, nb, mg);
    }
    };
    using T_ = std::true_type;
    using F_ = std::false_type;
    if (has_res) { if (relu) epi_loop(T_{}, T_{}); else epi_loop(F_{}, T_{}); }
    else { if (relu) epi_loop(T_{}, F_{}); else epi_loop(F_{}, F_{}); }
    if (lane == 0) tma_store_wait_all();
    if (warp == WARP_EPI0) prof_flush(0);
  } else if (warp >= WARP_TMA) {
    reg_dealloc<REGS_CTRL>();
    // a.two_loaders: the residual-loader warp (idle without a residual input) loads ring 1, this warp ring 0 -- the single
    // loading thread was busy 70-85 % of the time on the layers bound by the A stream
    constexpr int WARP_RES = WARP_MMA + N_MMA_WARPS;
    if (warp == WARP_TMA || (a.two_loaders && warp == WARP_RES)) {
    const int lid = warp == WARP_TMA ? 0 : 1;
    // =========================== A LOADER (TMA) ===========================
    // warp-uniform loop (every lane waits, the elected lane issues): keeps the loop state in
    // uniform registers and off the divergent-code path.
    const bool leader = elect_one();
    // Every mbarrier has ONE waiter per side that observes each of its phases (parity waits
    // alias when a waiter skips a phase): tile t of an m-group travels through ring t % NR,
    // which only issuer t % NR consumes.
    int s = s0, mg = mg0, nb = nb0;
    int rslot[N_MMA_WARPS] = {0, 0};
    uint32_t rphase[N_MMA_WARPS] = {0, 0};
    const int Cin = a.CB * BKE;
    int prev_mg = -1;
    for (int it = 0; it < n_items; ++it) {
      const int mtv = tiles_of(mg);
      // x shared by the samples, sample-fastest order, ring == one full item: the previous item left
      // exactly these tiles in exactly these slots
      const bool keep = a.a_keep && mg == prev_mg && mtv == a.MT;
      prev_mg = mg;
      int cw[MAX_MT], chh[MAX_MT], cn[MAX_MT];
#pragma unroll
      for (int t = 0; t < MAX_MT; ++t) {
        const int m = (tile0(mg) + t) * BM;
        cw[t] = 0; chh[t] = 0;
        cn[t] = s * a.x_rows_per_s + m;
        if (a.conv) {
          const int r = m / a.OW;
          const int b = r / a.OH;
          cw[t] = (m - r * a.OW) * a.sw - a.pw;
          chh[t] = (r - b * a.OH) * a.sh - a.ph;
          cn[t] = s * a.x_rows_per_s + b;
        }
      }
      int c0 = 0, offw = 0, offh = 0, kw = 0;
      for (int kb = 0; kb < a.nkb; ++kb) {
#pragma unroll
        for (int t = 0; t < MAX_MT; ++t) {
          if (t < mtv && (!a.two_loaders || (t % N_MMA_WARPS) == lid)) {
            const int r = t % N_MMA_WARPS;
            const int ring = a.NR == 2 ? r : 0;
            const int slot = ring * a.NAr + (a.NR == 2 ? rslot[r] : rslot[0]);
            PW(0, mbar_wait(smem_u32(&ctrl->a_empty[slot]), (a.NR == 2 ? rphase[r] : rphase[0]) ^ 1u));
            if (leader) {
              const uint32_t bar = smem_u32(&ctrl->a_full[slot]);
              if (keep || (PROF && (a.debug & 8))) {
                mbar_arrive(bar);
              } else if (FLIP && a.a2_tma) {  // patch-matrix Flipout: the tile and its sign-flipped twin, two loads on one barrier
                mbar_expect_tx(bar, 2 * A_SLOT);
                tma_load_2d(a_base + slot * A_SLOT_T, &tm_a, bar, c0, cn[t]);
                tma_load_2d(a_base + slot * A_SLOT_T + A_SLOT, &tm_a2, bar, c0, cn[t]);
              } else {
                mbar_expect_tx(bar, A_SLOT);
                if (a.conv)
                  tma_load_im2col_4d(a_base + slot * A_SLOT_T, &tm_a, bar, c0, cw[t], chh[t], cn[t], (uint16_t)offw, (uint16_t)offh);
                else
                  tma_load_2d(a_base + slot * A_SLOT_T, &tm_a, bar, c0, cn[t]);
              }
            }
            if (a.NR == 2) { if (++rslot[r] == a.NAr) { rslot[r] = 0; rphase[r] ^= 1u; } }
            else if (++rslot[0] == a.NAr) { rslot[0] = 0; rphase[0] ^= 1u; }
          }
        }
        c0 += BKE;
        if (a.conv && c0 == Cin) {  // next filter tap
          c0 = 0;
          offw += a.dw;
          if (++kw == a.KW) { kw = 0; offw = 0; offh += a.dh; }
        }
      }
      walk(s, nb, mg);
    }
    __syncwarp();
    if (lid == 0) prof_flush(1);
    } else if (warp < WARP_MMA + N_MMA_WARPS) {
    // =========================== MMA ISSUERS ===========================
    const int j = warp - WARP_MMA;  // this issuer owns tiles t with t % N_MMA_WARPS == j
    const bool leader = elect_one();
    constexpr uint32_t idesc = TF32 ? umma_idesc_tf32(BM, BN) : umma_idesc_bf16(BM, BN);
    int mg = mg0, nb_unused = nb0, s_unused = s0, set = 0, aslot = 0, wslot_r = 0;
    uint32_t set_phase = 0, aphase = 0, wphase = 0, sphase = 0;
    uint32_t emp_bits = 0;  // phase of every acc_empty barrier (one per accumulator tile) as seen by this issuer
    const uint32_t acc0 = a.has_shift ? 1u : 0u;  // accumulators start from the shift, not from zero
    const uint64_t ones_desc = umma_desc_sw32(ones_base), shift_desc = umma_desc_sw32(shift_base);
    const uint64_t shift2_desc = umma_desc_sw32(shift_base + SHIFT_BYTES);  // Flipout: perturbation bias
    bool first = true;  // first item of a (sample, n-block) key on this CTA: its W panel is new
    const int ring0 = (a.NR == 2 ? j : 0) * a.NAr;  // first slot of the A ring this issuer consumes
    for (int it = 0; it < n_items; ++it) {
      const bool last = !a.stationary || it + 1 == n_items || mg + 1 == a.num_mg;
      const int mtv = tiles_of(mg);  // 0 for the odd tail of a CTA pair: W slots are still consumed, nothing is issued
      const uint32_t dbase = tmem + (uint32_t)(set * a.MT * ACC_COLS);
      if (a.has_shift && first) {
        PW(1, mbar_wait(smem_u32(&ctrl->w_full[a.RW]), sphase));
        tc_fence_after();
      }
      for (int kb = 0; kb < a.nkb; ++kb) {
        const int wslot = a.stationary ? kb : wslot_r;
        if (first) {
          if constexpr (CL > 1) PW(1, mbar_wait_cluster(smem_u32(&ctrl->w_full[wslot]), wphase));  // half of the rows come from the peer CTA
          else PW(1, mbar_wait(smem_u32(&ctrl->w_full[wslot]), wphase));
          tc_fence_after();
        }
        const uint64_t bdesc = umma_desc_sw128(w_base + wslot * W_SLOT);
#pragma unroll
        for (int t = 0; t < MAX_MT; ++t) {
          if (t < mtv) {
            if ((a.NR == 2 ? (t % N_MMA_WARPS) : 0) == j) {
              if (kb == 0) {
                // first touch of accumulator tile t in this item: the epilogue must have drained it
                const int ti = set * a.MT + t;
                PW(0, mbar_wait(smem_u32(&ctrl->acc_empty[ti]), ((emp_bits >> ti) & 1u) ^ 1u));
                emp_bits ^= 1u << ti;
                tc_fence_after();
                // D = ones[128 x 16] * shift[BN x 16]^T: every row of the accumulator starts as the
                // per-channel shift (bias * scale + BatchNorm shift, split into three bf16 terms)
                if (a.has_shift && leader && !(PROF && (a.debug & 4))) {
                  umma_any<TF32>(dbase + (uint32_t)(t * ACC_COLS), ones_desc, shift_desc, idesc, 0u);
                  if constexpr (FLIP) umma_bf16(dbase + (uint32_t)(t * ACC_COLS + BN), ones_desc, shift2_desc, idesc, 0u);
                }
              }
              const int slot = ring0 + aslot;
              PW(2, mbar_wait(smem_u32(&ctrl->a_full[slot]), aphase));
              if constexpr (FLIP) {  // the flipped copy is in place too
                if (!a.a2_tma) mbar_wait(smem_u32(&ctrl->af_full[slot]), aphase);
              }
              tc_fence_after();
              if (leader) {
                const uint64_t adesc = umma_desc_sw128(a_base + slot * A_SLOT_T);
                const uint32_t d = dbase + (uint32_t)(t * ACC_COLS);
                if (!(PROF && (a.debug & 4))) {
#pragma unroll
                  for (int k = 0; k < BK / 16; ++k)   // 32 bytes of K per MMA: 16 bf16 or 8 tf32
                    umma_any<TF32>(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : acc0);
                  if constexpr (FLIP) {  // perturbation path: (x * s_in) tile against the sigma * eps sub-tile
                    const uint64_t adesc2 = umma_desc_sw128(a_base + slot * A_SLOT_T + A_SLOT);
                    const uint64_t bdesc2 = umma_desc_sw128(w_base + wslot * W_SLOT + W_SUB);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k)
                      umma_bf16(d + (uint32_t)BN, adesc2 + (uint64_t)(2 * k), bdesc2 + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : acc0);
                  }
                }
                umma_commit(smem_u32(&ctrl->a_empty[slot]));
              }
              __syncwarp();
              if (++aslot == a.NAr) { aslot = 0; aphase ^= 1u; }
            }
          }
        }
        if (kb == 0 && a.has_shift) {  // the shift tile has been read by this issuer's rank-1 MMAs of the item
          if (last && leader) umma_commit(smem_u32(&ctrl->w_empty[a.RW]));
          __syncwarp();
          if (last) sphase ^= 1u;
        }
        if (last && leader) {  // the slot is free once BOTH CTAs' tensor cores have read it: its producers sit in both
          if constexpr (CL > 1) umma_commit_multicast(smem_u32(&ctrl->w_empty[wslot]), (uint16_t)3);
          else umma_commit(smem_u32(&ctrl->w_empty[wslot]));
        }
        if (!a.stationary && ++wslot_r == a.RW) { wslot_r = 0; wphase ^= 1u; }
      }
      if (mtv > 0) {
        if (leader) umma_commit(smem_u32(&ctrl->acc_full[set]));
        __syncwarp();
        if (++set == a.NSETS) { set = 0; set_phase ^= 1u; }
      }
      if (a.stationary) {
        if (last) wphase ^= 1u;  // the next key's panel is the next generation of every slot
        first = last;
      }
      walk(s_unused, nb_unused, mg);
    }
    prof_flush(2 + j);
    } else if (a.res_bufs) {
    // =========================== RESIDUAL LOADER ===========================
    // walks the same (item, tile, sub-tile) sequence as the epilogue, res_bufs sub-tiles ahead of it
    const bool leader = elect_one();
    int s = s0, nb = nb0, mg = mg0, slot = 0;
    uint32_t phase = 0;
    for (int it = 0; it < n_items; ++it) {
      const int mtv = tiles_of(mg);
      const int nsub = (min(BN, vN - nb * BN) + SUBN - 1) / SUBN;
      for (int t = 0; t < mtv; ++t) {
        for (int sub = 0; sub < nsub; ++sub) {
          mbar_wait(smem_u32(&ctrl->res_empty[slot]), phase ^ 1u);
          if (leader) {
            const uint32_t bar = smem_u32(&ctrl->res_full[slot]);
            if (PROF && (a.debug & 64)) {
              mbar_arrive(bar);
            } else {
              mbar_expect_tx(bar, (uint32_t)STAGE_UNIT);
              tma_load_3d(res_base + (uint32_t)slot * (uint32_t)STAGE_UNIT, &tm_res, bar, nb * BN + sub * SUBN, (tile0(mg) + t) * BM, s);
            }
          }
          if (++slot == a.res_bufs) { slot = 0; phase ^= 1u; }
        }
      }
      walk(s, nb, mg);
    }
    __syncwarp();
    }
  } else if (FLIP && warp >= NPW) {
    // =========================== A FLIPPERS (Flipout: x * s_in in shared memory) ===========================
    // One thread per tile row: waits for the TMA of an A slot, XORs the sign bits of the 64 bf16 of its
    // row with 64 bits of the SIGN_IN stream and writes the result to the slot's second half, the A
    // operand of the perturbation MMAs.  The stream position of a row is the CHANNELS-LAST index of its
    // first element, (pixel * C_in + c0): 64-aligned, so the row takes words (x, y) or (z, w) of ONE
    // Philox block; the bits of the packed pair (2p, 2p+1) are bits p and p+16 of a word, so the mask of
    // a pair is one shift and the flip rides on a 3-input logic op.
    reg_dealloc<REGS_PROD>();
    const int f = threadIdx.x - NPW * 32;  // tile row
    int s = s0, nb = nb0, mg = mg0;
    int rslot[N_MMA_WARPS] = {0, 0};
    uint32_t rphase[N_MMA_WARPS] = {0, 0};
    const uint32_t row_off = (uint32_t)f * 128u, row_sw = (uint32_t)(f & 7);
    for (int it = 0; it < (a.a2_tma ? 0 : n_items); ++it) {
      const int mtv = tiles_of(mg);
      const BtRngKey key = bt_rng_key(a.seed, a.sample_id + s, a.layer_id, a.kind_a);
      int fb[MAX_MT], fh[MAX_MT], fw[MAX_MT], fm[MAX_MT];
#pragma unroll
      for (int t = 0; t < MAX_MT; ++t) {
        const int m = (tile0(mg) + t) * BM + f;  // row within the sample
        fm[t] = m; fb[t] = 0; fh[t] = 0; fw[t] = 0;
        if (a.conv) {
          const int r = m / a.OW;
          fb[t] = r / a.OH;
          fw[t] = (m - r * a.OW) * a.sw - a.pw;
          fh[t] = (r - fb[t] * a.OH) * a.sh - a.ph;
        }
      }
      int c0 = 0, offw = 0, offh = 0, kw = 0;
      for (int kb = 0; kb < a.nkb; ++kb) {
#pragma unroll
        for (int t = 0; t < MAX_MT; ++t) {
          if (t < mtv) {
            const int r = t % N_MMA_WARPS;
            const int ring = a.NR == 2 ? r : 0;
            const int slot = ring * a.NAr + (a.NR == 2 ? rslot[r] : rslot[0]);
            uint32_t w0 = 0, w1 = 0;
            int ih = fh[t] + offh, iw = fw[t] + offw;
            bool inb = fm[t] < a.Ms && (!a.conv || ((unsigned)ih < (unsigned)a.IH && (unsigned)iw < (unsigned)a.IW));
            if (a.a_dil_h > 1 || a.a_dil_w > 1) {  // zero-inserted grad_out: only every a_dil-th pixel is data
              inb = inb && (ih % a.a_dil_h == 0) && (iw % a.a_dil_w == 0);
              ih /= a.a_dil_h; iw /= a.a_dil_w;
            }
            if (a.sign_in) {  // deterministic mode: the caller's +-1 tensor, laid out like x
              if (inb) {
                const int64_t off = a.conv ? (int64_t)fb[t] * a.x_bs + (int64_t)ih * a.x_hs + (int64_t)iw * a.x_ws + c0
                                           : (int64_t)fm[t] * a.x_pitch + c0;
                const float4* sp = reinterpret_cast<const float4*>(a.sign_in + off);
#pragma unroll
                for (int j4 = 0; j4 < 16; ++j4) {
                  const float4 v = __ldg(sp + j4);
                  const int j = (4 * j4) & 31;
                  const uint32_t bits = ((__float_as_uint(v.x) >> 31) << bt_sign_bitpos(j)) | ((__float_as_uint(v.y) >> 31) << bt_sign_bitpos(j + 1)) |
                                        ((__float_as_uint(v.z) >> 31) << bt_sign_bitpos(j + 2)) | ((__float_as_uint(v.w) >> 31) << bt_sign_bitpos(j + 3));
                  if (j4 < 8) w0 |= bits; else w1 |= bits;
                }
              }
            } else if (inb) {  // zero-filled (padding) and out-of-range rows need no signs
              const uint64_t pix = a.conv ? ((uint64_t)fb[t] * (uint64_t)a.SH + (uint64_t)ih) * (uint64_t)a.SW + (uint64_t)iw : (uint64_t)fm[t];
              const uint64_t idx = pix * (uint64_t)a.Cin + (uint64_t)c0;
              const uint4 rr = bt_philox(idx >> 7, key);
              const bool hi = ((idx >> 6) & 1) != 0;
              w0 = hi ? rr.z : rr.x;
              w1 = hi ? rr.w : rr.y;
            }
            mbar_wait(smem_u32(&ctrl->a_full[slot]), (a.NR == 2 ? rphase[r] : rphase[0]));
            const uint32_t src = a_base + (uint32_t)slot * (uint32_t)A_SLOT_T + row_off;
            uint32_t d[8][4];
#pragma unroll
            for (int c = 0; c < 8; ++c)
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(d[c][0]), "=r"(d[c][1]), "=r"(d[c][2]), "=r"(d[c][3])
                           : "r"(src + (((uint32_t)c ^ row_sw) << 4)));
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const uint32_t wsel = c < 4 ? w0 : w1;
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int pp = (4 * c + i) & 15;  // pair within the word
                d[c][i] ^= (wsel << (15 - pp)) & 0x80008000u;
              }
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(src + (uint32_t)A_SLOT + (((uint32_t)c ^ row_sw) << 4)),
                           "r"(d[c][0]), "r"(d[c][1]), "r"(d[c][2]), "r"(d[c][3])
                           : "memory");
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&ctrl->af_full[slot]));
            if (a.NR == 2) { if (++rslot[r] == a.NAr) { rslot[r] = 0; rphase[r] ^= 1u; } }
            else if (++rslot[0] == a.NAr) { rslot[0] = 0; rphase[0] ^= 1u; }
          }
        }
        c0 += BK;
        if (a.conv && c0 == a.Cin) {  // next filter tap
          c0 = 0;
          offw += a.dw;
          if (++kw == a.KW) { kw = 0; offw = 0; offh += a.dh; }
        }
      }
      walk(s, nb, mg);
    }
  } else {
    // =========================== W PRODUCERS (sample in smem) ===========================
    reg_dealloc<REGS_PROD>();
    const int p = threadIdx.x - WARP_PROD0 * 32;
    const int chunk = p & 7;      // 16-byte chunk (8 consecutive k) of the 128-byte row
    const int row_base = p >> 3;  // rows row_base + 64*i of this CTA's share of the k-block
    constexpr int RPP = PT / 8;                              // rows of a k-block covered by one pass of the producers
    constexpr int ROWS_PER_THREAD = (BN + RPP - 1) / RPP;
    constexpr int rows_cta = BN / CL;        // CL = 2: each CTA of the pair samples half of the rows of a k-block
    const int row_first = crank * rows_cta;
    const uint32_t peer_w_base = CL > 1 ? mapa_shared(w_base, (uint32_t)(crank ^ 1)) : 0u;
    const uint32_t peer_w_full = CL > 1 ? mapa_shared(smem_u32(&ctrl->w_full[0]), (uint32_t)(crank ^ 1)) : 0u;
    int s = s0, nb = nb0, mg = mg0, wslot_r = 0;
    uint32_t wphase = 0, sphase = 0;
    bool first = true;
    if (a.has_shift && p < BM) {  // the constant ones tile (made visible by the first w_full[RW] arrival)
      const uint32_t dst = ones_base + (uint32_t)p * 32u;
      const uint32_t sw = (uint32_t)((p >> 2) & 1) << 4;  // 32-byte swizzle: chunk ^= bit 2 of the row
      if constexpr (TF32)   // [1, 1, 1, 0 | 0 ...] as tf32 words (K = 8 per MMA: the 32-byte row is the whole reduction)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + sw), "r"(0x3F800000u), "r"(0x3F800000u), "r"(0x3F800000u), "r"(0u) : "memory");
      else
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + sw), "r"(0x3F803F80u), "r"(0x00003F80u), "r"(0u), "r"(0u) : "memory");
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (sw ^ 16u)), "r"(0u), "r"(0u), "r"(0u), "r"(0u) : "memory");
    }
    for (int it = 0; it < n_items; ++it) {
      if (first || !a.stationary) {
        if (a.has_shift) {
          PW(0, mbar_wait(smem_u32(&ctrl->w_empty[a.RW]), sphase ^ 1u));
          if (p < NSUBW * BN) {
            // Flipout: threads [BN, 2 BN) fill the second tile, the bias perturbation sigma_b * eps_b of the
            // perturbation accumulator (inside the s_out product, linear_flipout.py:162,172-173)
            const bool pert = FLIP && p >= BN;
            const int vn = nb * BN + (pert ? p - BN : p);
            const int n = a.spn == 2 ? (vn & 63) : vn;                      // channel
            const uint32_t smp = a.spn == 2 ? 2u * s + (vn >> 6) : (uint32_t)s;  // sample of this row
            float sh = 0.f;
            if (vn < vN) {
              if (a.mu_b) {
                const float e = a.eps_b ? a.eps_b[n]
                                        : bt_normal1((uint64_t)n, bt_rng_key(a.seed, a.sample_id + smp, a.layer_id, BT_KIND_EPS_B));
                const float db = bt_softplus(a.rho_b[n]) * e;
                sh = FLIP ? (pert ? db : a.mu_b[n]) : a.mu_b[n] + db;
              }
              if (a.out_scale) sh *= a.out_scale[n];
              if (a.out_shift && !pert) sh += a.out_shift[n];
            }
            // three bf16 terms carry 24 bits of the fp32 shift through the bf16 MMA
            uint32_t w0, w1, w2 = 0u;
            if constexpr (TF32) {   // three tf32 terms (10-bit mantissas, truncation splits exactly): 30 bits of the shift
              w0 = __float_as_uint(sh) & 0xFFFFE000u;
              const float r1 = sh - __uint_as_float(w0);
              w1 = __float_as_uint(r1) & 0xFFFFE000u;
              w2 = __float_as_uint(r1 - __uint_as_float(w1)) & 0xFFFFE000u;
            } else {
            const __nv_bfloat16 h0 = __float2bfloat16_rn(sh);
            const float r1 = sh - __bfloat162float(h0);
            const __nv_bfloat16 h1 = __float2bfloat16_rn(r1);
            const __nv_bfloat16 h2 = __float2bfloat16_rn(r1 - __bfloat162float(h1));
            w0 = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
            w1 = (uint32_t)__bfloat16_as_ushort(h2);
            }
            // (tile rows are 32 bytes and BN = 64 / 128 rows keep the 8-row swizzle phase, so the second tile,
            // SHIFT_BYTES further on, is addressed by its own row index)
            const int prow = pert ? p - BN : p;
            const uint32_t dst = shift_base + (pert ? (uint32_t)SHIFT_BYTES : 0u) + (uint32_t)prow * 32u;
            const uint32_t sw = (uint32_t)((prow >> 2) & 1) << 4;
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + sw), "r"(w0), "r"(w1), "r"(w2), "r"(0u) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (sw ^ 16u)), "r"(0u), "r"(0u), "r"(0u), "r"(0u) : "memory");
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&ctrl->w_full[a.RW]));
          sphase ^= 1u;
        }
        if (a.w_pre) {
          // pre-sampled weights (tc2_presample_kernel): the k-blocks of W[s][nb * BN .. + BN) arrive by TMA with the
          // 128-byte swizzle the sampling loop below writes; rows past N and columns past K are zero-filled by the map
          if (p == 0) {
            for (int kb = 0; kb < a.nkb; ++kb) {
              PW(0, mbar_wait(smem_u32(&ctrl->w_empty[wslot_r]), wphase ^ 1u));
              const uint32_t bar = smem_u32(&ctrl->w_full[wslot_r]);
              if (PROF && (a.debug & 128)) {
                mbar_arrive(bar);
              } else {
                mbar_expect_tx(bar, (uint32_t)W_SLOT);
                if constexpr (FLIP) {   // the mu sub-tile straight from the packed parameters (2-d map behind tm_res), then delta
                  tma_load_2d(w_base + (uint32_t)(wslot_r * W_SLOT), &tm_res, bar, kb * BKE, nb * BN);
                  tma_load_3d(w_base + (uint32_t)(wslot_r * W_SLOT + W_SUB), &tm_a2, bar, kb * BKE, nb * BN, s);
                } else {
                  tma_load_3d(w_base + (uint32_t)(wslot_r * W_SLOT), &tm_a2, bar, kb * BKE, nb * BN, s);
                }
              }
              if (++wslot_r == a.RW) { wslot_r = 0; wphase ^= 1u; }
            }
          }
          first = false;
          walk(s, nb, mg);
          continue;
        }
        const BtRngKey rk0 = bt_rng_key(a.seed, a.sample_id + (uint32_t)(s * a.spn), a.layer_id, BT_KIND_EPS_W);
        // packed bf16 (mu, sigma) of this thread's chunk of k-block kb2, eight weights per row
        uint4 nmu8[ROWS_PER_THREAD], nsg8[ROWS_PER_THREAD];
        auto load_params = [&](int kb2) {
          const int k2 = kb2 * BKE + chunk * EPC16;
#pragma unroll
          for (int i = 0; i < ROWS_PER_THREAD; ++i) {
            const int lrow = row_base + RPP * i;
            const int vn = nb * BN + row_first + lrow;
            const int n = a.spn == 2 ? (vn & 63) : vn;
            nmu8[i] = nsg8[i] = make_uint4(0u, 0u, 0u, 0u);
            if (lrow < rows_cta && k2 < a.K && vn < vN && !(PROF && (a.debug & 32))) {
              const int64_t e = (int64_t)n * a.K + k2;
              // 16 bytes each: eight packed bf16, or four fp32 for the tf32 instantiation
              nmu8[i] = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(a.mu_p) + e * (TF32 ? 4 : 2)));
              nsg8[i] = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(a.sig_p) + e * (TF32 ? 4 : 2)));
            }
          }
        };
        load_params(0);
        for (int kb = 0; kb < a.nkb; ++kb) {
          const int slot = a.stationary ? kb : wslot_r;
          const int k = kb * BKE + chunk * EPC16;
          const bool kin = k < a.K;  // K % 8 == 0: a chunk is all in or all out
          // parameters of this k-block were requested one k-block ago (cur <- nxt); the requests for
          // the next one go out now, a whole k-block of Philox / Box-Muller work ahead of their use
          uint4 mu8[ROWS_PER_THREAD], sg8[ROWS_PER_THREAD];
#pragma unroll
          for (int i = 0; i < ROWS_PER_THREAD; ++i) { mu8[i] = nmu8[i]; sg8[i] = nsg8[i]; }
          if (kb + 1 < a.nkb) load_params(kb + 1);
          if constexpr (CL > 1) PW(0, mbar_wait_cluster(smem_u32(&ctrl->w_empty[slot]), wphase ^ 1u));
          else PW(0, mbar_wait(smem_u32(&ctrl->w_empty[slot]), wphase ^ 1u));
          const long long gen_t0 = PROF ? clock64() : 0;
#pragma unroll
          for (int i = 0; i < ROWS_PER_THREAD; ++i) {
            const int lrow = row_base + RPP * i;
            if (lrow >= rows_cta) break;
            const int row = row_first + lrow;
            const int vn = nb * BN + row;
            const int n = a.spn == 2 ? (vn & 63) : vn;
            BtRngKey rk = rk0;
            if (a.spn == 2) rk.sample_id += (uint32_t)(vn >> 6);  // second sample of the stacked pair
            if constexpr (TF32) {
              // four fp32 weights per 16-byte chunk: w = tf32(fmaf(sigma, eps, mu)), the same stream positions as the
              // bf16 path (half a Philox block per chunk)
              float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
              if (kin && vn < vN) {
                const int64_t e = (int64_t)n * a.K + k;
                if (a.eps_p) {
                  z4 = __ldg(reinterpret_cast<const float4*>(a.eps_p + e));
                } else if (a.Kt == a.K) {
                  z4 = bt_normal4((uint64_t)e >> 2, rk);
                } else {  // padded K (patch matrix): positions of the TRUE stream, element by element (tiny layers only)
                  const uint64_t e0 = (uint64_t)n * a.Kt + k;
                  z4.x = k + 0 < a.Kt ? bt_normal1(e0 + 0, rk) : 0.f;
                  z4.y = k + 1 < a.Kt ? bt_normal1(e0 + 1, rk) : 0.f;
                  z4.z = k + 2 < a.Kt ? bt_normal1(e0 + 2, rk) : 0.f;
                  z4.w = k + 3 < a.Kt ? bt_normal1(e0 + 3, rk) : 0.f;
                }
              }
              const uint32_t t0 = to_tf32(fmaf(__uint_as_float(sg8[i].x), z4.x, __uint_as_float(mu8[i].x)));
              const uint32_t t1 = to_tf32(fmaf(__uint_as_float(sg8[i].y), z4.y, __uint_as_float(mu8[i].y)));
              const uint32_t t2 = to_tf32(fmaf(__uint_as_float(sg8[i].z), z4.z, __uint_as_float(mu8[i].z)));
              const uint32_t t3 = to_tf32(fmaf(__uint_as_float(sg8[i].w), z4.w, __uint_as_float(mu8[i].w)));
              const uint32_t toff = (uint32_t)(slot * W_SLOT) + (uint32_t)row * 128u + (((uint32_t)chunk ^ (uint32_t)(row & 7)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(w_base + toff), "r"(t0), "r"(t1), "r"(t2), "r"(t3) : "memory");
              continue;
            }
            float z[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) z[e] = 0.f;
            if (kin && vn < vN && !(PROF && (a.debug & 16))) {
              const int64_t e = (int64_t)n * a.K + k;
              if (a.eps_p) {
                const float4 e0 = __ldg(reinterpret_cast<const float4*>(a.eps_p + e));
                const float4 e1 = __ldg(reinterpret_cast<const float4*>(a.eps_p + e + 4));
                z[0] = e0.x; z[1] = e0.y; z[2] = e0.z; z[3] = e0.w; z[4] = e1.x; z[5] = e1.y; z[6] = e1.z; z[7] = e1.w;
              } else if (a.Kt == a.K) {
                bt_normal8((uint64_t)e >> 3, rk, z);
              } else {  // padded K (patch matrix): this chunk straddles two Philox blocks of the true stream
                const uint64_t e0 = (uint64_t)n * a.Kt + k;
                float zz[16];
                bt_normal8(e0 >> 3, rk, *reinterpret_cast<float(*)[8]>(&zz[0]));
                bt_normal8((e0 >> 3) + 1, rk, *reinterpret_cast<float(*)[8]>(&zz[8]));
                const int off = (int)(e0 & 7);
#pragma unroll
                for (int o = 0; o < 8; ++o)
                  if (o == off) {
#pragma unroll
                    for (int i2 = 0; i2 < 8; ++i2) z[i2] = (k + i2 < a.Kt) ? zz[o + i2] : 0.f;
                  }
              }
            }
            // w = bf16(sigma * eps + mu) on packed pairs: fp32 fused multiply-add of the bf16 operands, one rounding
            // Flipout keeps the two terms apart: the mu sub-tile as packed, then bf16(sigma * eps)
            const uint32_t w0 = fma_bf16x2(sg8[i].x, pack_bf16(z[0], z[1]), FLIP ? 0u : mu8[i].x);
            const uint32_t w1 = fma_bf16x2(sg8[i].y, pack_bf16(z[2], z[3]), FLIP ? 0u : mu8[i].y);
            const uint32_t w2 = fma_bf16x2(sg8[i].z, pack_bf16(z[4], z[5]), FLIP ? 0u : mu8[i].z);
            const uint32_t w3 = fma_bf16x2(sg8[i].w, pack_bf16(z[6], z[7]), FLIP ? 0u : mu8[i].w);
            const uint32_t off = (uint32_t)(slot * W_SLOT) + (uint32_t)row * 128u + (((uint32_t)chunk ^ (uint32_t)(row & 7)) << 4);
            if constexpr (FLIP) {
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(w_base + off), "r"(mu8[i].x), "r"(mu8[i].y), "r"(mu8[i].z), "r"(mu8[i].w) : "memory");
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(w_base + off + (uint32_t)W_SUB), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
            } else {
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(w_base + off), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
            }
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            const uint32_t bar = smem_u32(&ctrl->w_full[slot]);
            // This warp's four rows of a pass (512 contiguous bytes per sub-tile) travel to the same place in
            // the peer's ring as asynchronous bulk copies that complete (complete_tx) on the PEER's w_full
            // barrier; symmetrically this warp's own arrival expects the bytes its counterpart warp in the
            // peer sends here.  No remote arrive, no remote store, nothing waits on the DSMEM round trip.
            int npieces = 0;  // lane 0: row_base is the first of the warp's four rows
            if constexpr (CL > 1) {
#pragma unroll
              for (int i = 0; i < ROWS_PER_THREAD; ++i) npieces += (row_base + RPP * i < rows_cta) ? 1 : 0;
            }
            if (npieces > 0) {
              mbar_expect_tx(bar, 512u * (uint32_t)(npieces * NSUBW));
#pragma unroll
              for (int i = 0; i < ROWS_PER_THREAD; ++i) {
                if (row_base + RPP * i < rows_cta) {
#pragma unroll
                  for (int sub = 0; sub < NSUBW; ++sub) {
                    const uint32_t off = (uint32_t)(slot * W_SLOT + sub * W_SUB) + (uint32_t)(row_first + row_base + RPP * i) * 128u;
                    bulk_copy_to_cluster(peer_w_base + off, w_base + off, 512u, peer_w_full + (uint32_t)slot * 8u);
                  }
                }
              }
            } else {
              mbar_arrive(bar);
            }
          }
          if constexpr (PROF) pc[1] += (unsigned long long)(clock64() - gen_t0);
          if (!a.stationary && ++wslot_r == a.RW) { wslot_r = 0; wphase ^= 1u; }
        }
        if (a.stationary) wphase ^= 1u;
      }
      first = false;
      if (walk(s, nb, mg)) first = true;
    }
    if (warp == WARP_PROD0) prof_flush(4);
    if (warp == WARP_PROD0 + NPW - 1) prof_flush(5);
  }

  tc_fence_before();
  __syncthreads();
  if constexpr (CL > 1) cluster_sync_all();  // no CTA leaves while its peer may still store into / arrive on its shared memory
  if (warp == WARP_MMA) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------ host side ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct DriverFns {
  EncodeTiledFn tiled = nullptr;
  EncodeIm2colFn im2col = nullptr;
  int driver_version = 0;
  bool ok = false;
};

const DriverFns& driver_fns() {
  static DriverFns f;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      f.tiled = (EncodeTiledFn)p;
    p = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      f.im2col = (EncodeIm2colFn)p;
    cudaDriverGetVersion(&f.driver_version);
    f.ok = f.tiled && f.im2col;
    cudaGetLastError();
  });
  return f;
}

struct Dense {  // the dense channels-last view the tensor maps describe
  int64_t in[2], out[2], k[2], st[2], pd[2], dl[2];
};

// rank 1 is run as rank 2 with a unit H axis.
bool conv_dims(const bt_geom* g, Dense* d, int64_t* xs_h, int64_t* xs_w, int64_t* os_h, int64_t* os_w) {
  if (g->rank == 2) {
    for (int i = 0; i < 2; ++i) {
      d->in[i] = g->in_size[i]; d->out[i] = g->out_size[i]; d->k[i] = g->k_size[i];
      d->st[i] = g->stride[i]; d->pd[i] = g->pad[i]; d->dl[i] = g->dil[i];
    }
    *xs_h = g->x_sp_stride[0]; *xs_w = g->x_sp_stride[1]; *os_h = g->out_sp_stride[0]; *os_w = g->out_sp_stride[1];
    return true;
  }
  if (g->rank == 1) {
    d->in[0] = d->out[0] = d->k[0] = d->st[0] = d->dl[0] = 1; d->pd[0] = 0;
    d->in[1] = g->in_size[0]; d->out[1] = g->out_size[0]; d->k[1] = g->k_size[0];
    d->st[1] = g->stride[0]; d->pd[1] = g->pad[0]; d->dl[1] = g->dil[0];
    *xs_w = g->x_sp_stride[0]; *os_w = g->out_sp_stride[0];
    *xs_h = d->in[1] * *xs_w; *os_h = d->out[1] * *os_w;
    return true;
  }
  return false;
}

}  // namespace

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

// How the A operand (activations) reaches shared memory.
enum AMode { A_NONE = 0, A_GEMM = 1, A_IM2COL = 2, A_PATCH = 3 };

struct Plan {
  AMode mode;
  int64_t rows_per_s;  // GEMM rows per sample
  int64_t x_rows;      // rows (A_GEMM / A_PATCH) or images (A_IM2COL) per sample in the A source
  int64_t pitch;       // A_GEMM: row pitch of x in elements
  int T, K, Kp;        // taps, true reduction length, padded reduction length
  int es;              // bytes per activation element: 2 (bf16 operands) or 4 (fp32 activations, tf32 operands)
  Dense d;
  int64_t xh, xw;
};

static int64_t prod_out(const bt_geom* g) {
  int64_t t = 1;
  for (int i = 0; i < g->rank; ++i) t *= g->out_size[i];
  return t;
}

static bool tc2_plan(const bt_geom* g, const bt_layer_params* p, bool flipout, Plan* pl, const char** why, bool no_gemm = false) {
#define TC_NO(cond, msg) if (cond) { *why = msg; return false; }
  memset(pl, 0, sizeof(*pl));
  // Flipout: the 64 input channels of a tile row must be an aligned run of the sign stream (anything else
  // goes through the patch matrix, built twice: from x and from x * s_in)
  TC_NO(flipout && g->n_samples > 1 && (p->sign_in || p->sign_out), "tensor-fed signs with n_samples > 1");
  TC_NO(flipout && ((p->sign_in == nullptr) != (p->sign_out == nullptr)), "sign_in and sign_out must both be given or both be null");
  TC_NO(flipout && g->rank > 2, "Flipout rank 3");
  TC_NO(g->groups != 1, "groups != 1");
  TC_NO(g->x_dtype != BT_BF16 && g->x_dtype != BT_F32, "activations are neither bf16 nor fp32");
  // fp32 activations run the tf32 instantiation: Reparameterization forward, fp32 in / fp32 out
  TC_NO(g->x_dtype == BT_F32 && (flipout || g->out_dtype != BT_F32), "fp32 activations: Flipout / mixed dtypes run the fp32 kernels");
  TC_NO(g->x_dtype == BT_F32 && env_int("BT_TF32", 1) == 0, "tf32 tensor-core path switched off (BT_TF32=0)");
  const int es = g->x_dtype == BT_F32 ? 4 : 2;
  const int e16 = 16 / es, bke = 128 / es;   // elements per 16 bytes / per k-block
  TC_NO(g->out_chan_stride != 1, "output channels are not the fastest axis (need channels-last)");
  TC_NO(g->c_out % 8 != 0, "c_out is not a multiple of 8");
  TC_NO(g->batch <= 0, "empty batch");
  TC_NO(g->n_samples > 1 && (p->eps_w || p->eps_b), "tensor-fed eps with n_samples > 1");
  TC_NO(!driver_fns().ok, "cuTensorMapEncode* entry points unavailable");
  // the output must be one dense [samples][rows][c_out] matrix
  int64_t dense = g->c_out, taps = 1;
  for (int i = g->rank - 1; i >= 0; --i) {
    TC_NO(g->out_sp_stride[i] != dense, "output is not dense channels-last");
    dense *= g->out_size[i];
    taps *= g->k_size[i];
    TC_NO(g->in_size[i] <= 0 || g->k_size[i] <= 0 || g->stride[i] <= 0 || g->dil[i] <= 0 || g->pad[i] < 0, "bad conv parameters");
    const int64_t expect = (g->in_size[i] + 2 * g->pad[i] - g->dil[i] * (g->k_size[i] - 1) - 1) / g->stride[i] + 1;
    TC_NO(g->out_size[i] != expect || expect <= 0, "out_size inconsistent with the conv arithmetic");
  }
  TC_NO(g->out_batch_stride != dense, "output is not dense channels-last");
  TC_NO(g->n_samples > 1 && g->out_sample_stride != g->batch * dense, "out samples are not stacked densely");
  const bool x_shared = g->n_samples > 1 && g->x_sample_stride == 0;
  TC_NO(g->n_samples > 1 && !x_shared && g->x_sample_stride != g->batch * g->x_batch_stride, "x samples are not stacked densely");
  pl->rows_per_s = g->batch * prod_out(g);
  TC_NO(pl->rows_per_s >= (1ll << 30), "too many output positions");
  pl->T = (int)taps;
  TC_NO(g->c_in * taps >= (1ll << 24), "reduction length too large");
  pl->K = (int)(g->c_in * taps);
  pl->Kp = (pl->K + 7) & ~7;
  pl->es = es;
  const int pin = env_int("BT_TC2_AMODE", 0);
  if (g->rank == 0) {
    if (g->x_chan_stride == 1 && g->x_batch_stride % e16 == 0 && g->c_in % 8 == 0 && (!flipout || g->c_in % 64 == 0) && pin != A_PATCH) {
      pl->mode = A_GEMM; pl->x_rows = g->batch; pl->pitch = g->x_batch_stride;
      return true;
    }
  } else if (g->rank <= 2 && g->x_chan_stride == 1 && g->c_in % bke == 0 && pin != A_PATCH) {
    conv_dims(g, &pl->d, &pl->xh, &pl->xw, &dense, &taps);
    const Dense& d = pl->d;
    bool ok = pl->xw % e16 == 0 && pl->xh % e16 == 0 && g->x_batch_stride % e16 == 0;
    for (int i = 0; i < 2; ++i)
      ok = ok && d.st[i] <= 8 && d.pd[i] <= 127 && (d.k[i] - 1) * d.dl[i] - d.pd[i] <= 128 && (d.k[i] - 1) * d.dl[i] <= 65535;
    if (ok) {
      // A 1x1 / stride 1 / unpadded convolution over a dense tensor IS a GEMM over [pixels, channels]
      if (d.k[0] == 1 && d.k[1] == 1 && d.st[0] == 1 && d.st[1] == 1 && d.pd[0] == 0 && d.pd[1] == 0 &&
          pl->xh == d.in[1] * pl->xw && g->x_batch_stride == d.in[0] * pl->xh && pin != A_IM2COL && !no_gemm) {
        pl->mode = A_GEMM; pl->x_rows = g->batch * d.in[0] * d.in[1]; pl->pitch = pl->xw;
      } else {
        pl->mode = A_IM2COL; pl->x_rows = g->batch;
      }
      return true;
    }
  }
  // anything else with a dense channels-last output: explicit patch matrix, then a GEMM
  TC_NO(flipout && x_shared, "Flipout through the patch matrix with x shared between samples");
  TC_NO(flipout && (p->sign_in != nullptr), "Flipout through the patch matrix with tensor-fed signs");
  TC_NO(taps > 4096, "kernel volume > 4096 taps");
  pl->mode = A_PATCH;
  pl->x_rows = pl->rows_per_s;
  pl->pitch = pl->Kp;
  return true;
#undef TC_NO
}

bool bt_tc_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why) {
  Plan pl;
  return tc2_plan(g, p, flipout, &pl, why);
}

// elements spanned by the x operand of a call (all samples), for the sign-flipped copy of the patch-matrix Flipout
static int64_t x_span_elems(const bt_geom* g) {
  const bool x_shared = g->n_samples > 1 && g->x_sample_stride == 0;
  return g->batch * g->x_batch_stride * (x_shared ? 1 : g->n_samples);
}

// bytes of the caller's workspace the forward needs: packed (mu, sigma, eps) + the patch matrix
// (Flipout: two patch matrices and the sign-flipped copy of x they are built from).  The operand mode is
// decided by the SAME tc2_plan the launch uses (with the most permissive parameters: no tensor-fed draws),
// so the size and the carve-up of tc2_run cannot disagree; a geometry the plan rejects never reaches the
// tensor-core kernels and needs the packed area only.
static int64_t packed_bytes(int64_t n, int64_t kp) { return ((3 * n * kp * 4 + 255) / 256) * 256; }
static int64_t patch_bytes(const bt_geom* g, int64_t rows_per_s, int64_t Kp) {
  const bool x_shared = g->n_samples > 1 && g->x_sample_stride == 0;
  return ((rows_per_s * (x_shared ? 1 : g->n_samples) * Kp * (g->x_dtype == BT_F32 ? 4 : 2) + 255) / 256) * 256;
}
// Pre-sampled weights ([S][N][K] bf16 behind the packed area): geometry-only eligibility, a superset of what the
// launch finally chooses (it also needs a streamed W, Philox-drawn eps, a forward run), so that the size the
// caller allocates always covers the launch's carve-up.
static int64_t wpre_bytes(const bt_geom* g, const Plan& pl, bool flipout) {
  const int64_t K = pl.mode == A_PATCH ? pl.Kp : pl.K;
  if ((flipout && pl.mode == A_PATCH) || g->x_dtype != BT_BF16 || pl.es != 2 || (K & 7)) return 0;
  const int64_t b = (int64_t)g->n_samples * g->c_out * K * 2;
  return b <= (1ll << 31) ? ((b + 255) / 256) * 256 : 0;
}
int64_t bt_tc_workspace_bytes(const bt_geom* g, bool flipout) {
  int64_t taps = 1;
  for (int i = 0; i < g->rank; ++i) taps *= g->k_size[i];
  const int64_t Kp = (g->c_in / (g->groups > 0 ? g->groups : 1) * taps + 7) & ~7ll;
  int64_t bytes = packed_bytes(g->c_out, Kp);
  bt_layer_params none;
  memset(&none, 0, sizeof(none));
  Plan pl;
  const char* why = "";
  if (tc2_plan(g, &none, flipout, &pl, &why)) {
    if (pl.mode == A_PATCH) {
      const int64_t pm = patch_bytes(g, pl.rows_per_s, pl.Kp);
      bytes += pm;
      if (flipout) bytes += pm + ((x_span_elems(g) * 2 + 255) / 256) * 256;
    }
    bytes += wpre_bytes(g, pl, flipout);
  }
  return bytes;
}

template <int BN, typename TO, bool PROF, int CL, bool FLIP, bool TF32 = false, bool EPI8 = false>
static int tc2_launch_one(const CUtensorMap& ta, const CUtensorMap& to, const CUtensorMap& tr, const CUtensorMap& ta2,
                          const Tc2Args& a, int grid, int smem_bytes, cudaStream_t stream) {
  BT_CUDA_OK(cudaFuncSetAttribute(tc2_fwd_kernel<BN, TO, PROF, CL, FLIP, TF32, EPI8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = (size_t)smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  // a.pair_launch: the CL = 1 kernel launched as clusters of two so that the CTAs with consecutive block ids -- which the
  // round-robin item order gives consecutive items, i.e. the same A tiles against neighbouring n-blocks -- sit on one TPC
  attr[0].val.clusterDim.x = (unsigned)(CL == 1 && a.pair_launch && grid % 2 == 0 ? 2 : CL);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = a.pdl ? 2 : 1;
  BT_CUDA_OK(cudaLaunchKernelEx(&cfg, tc2_fwd_kernel<BN, TO, PROF, CL, FLIP, TF32, EPI8>, ta, to, tr, ta2, a));
  BT_LAUNCH_OK(PROF ? "tc2_fwd_kernel(prof)" : (FLIP ? "tc2_fwd_kernel(flipout)" : (TF32 ? "tc2_fwd_kernel(tf32)" : "tc2_fwd_kernel")));
  return BT_OK;
}

template <int BN, typename TO>
static int tc2_launch(const CUtensorMap& ta, const CUtensorMap& to, const CUtensorMap& tr, const CUtensorMap& ta2, const Tc2Args& a,
                      int grid, int smem_bytes, bool flip, cudaStream_t stream) {
  if constexpr (BN == 256) {   // one instantiation: pre-sampled W, single CTA (the plan guarantees both)
    return tc2_launch_one<BN, TO, false, 1, false>(ta, to, tr, ta2, a, grid, smem_bytes, stream);
  } else {
  if constexpr (sizeof(TO) == 4) {
    if (a.tf32)
      return a.CL == 2 ? tc2_launch_one<BN, TO, false, 2, false, true>(ta, to, tr, ta2, a, grid, smem_bytes, stream)
                       : tc2_launch_one<BN, TO, false, 1, false, true>(ta, to, tr, ta2, a, grid, smem_bytes, stream);
  }
  if constexpr (sizeof(TO) == 2 && BN == 128) {
    if (a.epi8)
      return a.CL == 2 ? tc2_launch_one<BN, TO, false, 2, false, false, true>(ta, to, tr, ta2, a, grid, smem_bytes, stream)
                       : tc2_launch_one<BN, TO, false, 1, false, false, true>(ta, to, tr, ta2, a, grid, smem_bytes, stream);
  }
  if (flip)
    return a.CL == 2 ? tc2_launch_one<BN, TO, false, 2, true>(ta, to, tr, ta2, a, grid, smem_bytes, stream)
                     : tc2_launch_one<BN, TO, false, 1, true>(ta, to, tr, ta2, a, grid, smem_bytes, stream);
  if constexpr (sizeof(TO) == 2) {
    if (env_int("BT_TC2_PROF", 0))
      return a.CL == 2 ? tc2_launch_one<BN, TO, true, 2, false>(ta, to, tr, ta2, a, grid, smem_bytes, stream)
                       : tc2_launch_one<BN, TO, true, 1, false>(ta, to, tr, ta2, a, grid, smem_bytes, stream);
  }
  return a.CL == 2 ? tc2_launch_one<BN, TO, false, 2, false>(ta, to, tr, ta2, a, grid, smem_bytes, stream)
                   : tc2_launch_one<BN, TO, false, 1, false>(ta, to, tr, ta2, a, grid, smem_bytes, stream);
  }
}

// Diagnostics: cycle counters of the last BT_TC2_PROF=1 launch, [cta][role][slot] (see bt_b200.h).
extern "C" int bt_debug_tc_profile(unsigned long long* host_out, int n_words) {
  const int total = 160 * PROF_ROLES * PROF_SLOTS;
  if (!host_out || n_words < total) { bt_set_error("bt_debug_tc_profile: need room for %d words", total); return BT_ERR_INVALID; }
  BT_CUDA_OK(cudaDeviceSynchronize());
  BT_CUDA_OK(cudaMemcpyFromSymbol(host_out, g_tc_prof, sizeof(unsigned long long) * total));
  return BT_OK;
}

// Patch-matrix operand mode: the explicit im2col of x at `base` and, for Flipout, a second one built from
// x * s (sign stream `kind`), materialised behind it (x is the SMALL tensor here -- the RGB stem, an odd C_in).
static int build_patch_matrices(const bt_geom* g, const Plan& pl, const bt_layer_params* p, const void* x, bool flipout, int kind,
                                void* base, __nv_bfloat16** pm_out, __nv_bfloat16** pm2_out, cudaStream_t stream) {
  const int S = g->n_samples;
  const bool x_shared = S > 1 && g->x_sample_stride == 0;
  __nv_bfloat16* patches = (__nv_bfloat16*)base;
  *pm_out = patches;
  *pm2_out = nullptr;
  PatchGeo pg;
  memset(&pg, 0, sizeof(pg));
  int64_t in[3] = {1, 1, 1}, o[3] = {1, 1, 1}, k[3] = {1, 1, 1}, st[3] = {1, 1, 1}, pd[3] = {0, 0, 0}, dl[3] = {1, 1, 1}, xs[3] = {0, 0, 0};
  for (int i = 0; i < g->rank; ++i) {
    const int j = 3 - g->rank + i;
    in[j] = g->in_size[i]; o[j] = g->out_size[i]; k[j] = g->k_size[i]; st[j] = g->stride[i]; pd[j] = g->pad[i];
    dl[j] = g->dil[i]; xs[j] = g->x_sp_stride[i];
  }
  pg.rows = pl.rows_per_s * (x_shared ? 1 : S);
  pg.xbs = g->x_batch_stride; pg.xcs = g->x_chan_stride; pg.xs0 = xs[0]; pg.xs1 = xs[1]; pg.xs2 = xs[2];
  pg.Kp = pl.Kp; pg.K = pl.K; pg.Cin = (int)g->c_in; pg.KH = (int)k[1]; pg.KW = (int)k[2];
  pg.OD = (int)o[0]; pg.OH = (int)o[1]; pg.OW = (int)o[2]; pg.ID = (int)in[0]; pg.IH = (int)in[1]; pg.IW = (int)in[2];
  pg.sd = (int)st[0]; pg.sh = (int)st[1]; pg.sw = (int)st[2]; pg.pd = (int)pd[0]; pg.ph = (int)pd[1]; pg.pw = (int)pd[2];
  pg.dd = (int)dl[0]; pg.dh = (int)dl[1]; pg.dw = (int)dl[2];
  const int64_t chunks = pg.rows * (pg.Kp / 8);
  int64_t blocks = (chunks + 255) / 256;
  const int64_t cap = (int64_t)bt_num_sms() * 32;
  if (blocks > cap) blocks = cap;
  // (stacked samples of x are just more images: the batch stride walks across them)
  auto build_patches = [&](const __nv_bfloat16* src, __nv_bfloat16* dst) -> int {
    const bool small = chunks + cap * 256 < 0x7fffffffll;
    // table-driven kernel: the reduction fits its tables, tap positions fit 10 bits, the window fits 31-bit offsets
    const int64_t win = (g->c_in - 1) * pg.xcs + (k[0] - 1) * dl[0] * pg.xs0 + (k[1] - 1) * dl[1] * pg.xs1 + (k[2] - 1) * dl[2] * pg.xs2;
    static const bool tab_on = !(getenv("BT_PATCH_TAB") && atoi(getenv("BT_PATCH_TAB")) == 0);
    const bool tab = tab_on && pg.Kp <= PT_MAXK && pg.Kp >= 8 && win >= 0 && win < (1ll << 30) && pg.xcs >= 0 && pg.xs0 >= 0 && pg.xs1 >= 0 &&
                     pg.xs2 >= 0 && k[0] * dl[0] < 1024 && k[1] * dl[1] < 1024 && k[2] * dl[2] < 1024;
    if (tab) {
      int64_t tb = (pg.rows + PT_ROWS - 1) / PT_ROWS;
      if (tb > (int64_t)bt_num_sms() * 8) tb = (int64_t)bt_num_sms() * 8;
      if (g->x_dtype == BT_F32) tc2_patch_tab_kernel<float><<<(int)tb, 256, 0, stream>>>((const float*)src, (float*)dst, pg);
      else tc2_patch_tab_kernel<__nv_bfloat16><<<(int)tb, 256, 0, stream>>>(src, dst, pg);
      BT_LAUNCH_OK("tc2_patch_tab_kernel");
      return BT_OK;
    }
    if (g->x_dtype == BT_F32) {   // (pointers are typed bf16 for the Flipout bookkeeping below; fp32 never takes that branch)
      if (small) tc2_patch_kernel<float, uint32_t><<<(int)blocks, 256, 0, stream>>>((const float*)src, (float*)dst, pg);
      else tc2_patch_kernel<float, int64_t><<<(int)blocks, 256, 0, stream>>>((const float*)src, (float*)dst, pg);
    } else if (small) {
      tc2_patch_kernel<__nv_bfloat16, uint32_t><<<(int)blocks, 256, 0, stream>>>(src, dst, pg);
    } else {
      tc2_patch_kernel<__nv_bfloat16, int64_t><<<(int)blocks, 256, 0, stream>>>(src, dst, pg);
    }
    BT_LAUNCH_OK("tc2_patch_kernel");
    return BT_OK;
  };
  int rc = build_patches((const __nv_bfloat16*)x, patches);
  if (rc != BT_OK) return rc;
  if (flipout) {
    const int64_t pm_bytes = ((pg.rows * pg.Kp * 2 + 255) / 256) * 256;
    __nv_bfloat16* patches2 = (__nv_bfloat16*)((char*)patches + pm_bytes);
    __nv_bfloat16* xf = (__nv_bfloat16*)((char*)patches2 + pm_bytes);
    FlipGeo fg;
    memset(&fg, 0, sizeof(fg));
    fg.C = (int)g->c_in; fg.ID = (int)in[0]; fg.IH = (int)in[1]; fg.IW = (int)in[2]; fg.S = S;
    fg.elems_per_s = g->batch * g->c_in * in[0] * in[1] * in[2];
    fg.words_per_s = (fg.elems_per_s + 31) / 32;
    fg.xss = g->x_sample_stride; fg.xbs = g->x_batch_stride; fg.xcs = g->x_chan_stride; fg.xs0 = xs[0]; fg.xs1 = xs[1]; fg.xs2 = xs[2];
    fg.seed = p->seed; fg.sample_id = p->sample_id; fg.layer_id = p->layer_id; fg.kind = kind;
    int64_t fb = (fg.words_per_s * S + 255) / 256;
    if (fb > cap) fb = cap;
    tc2_flipx_kernel<<<(int)fb, 256, 0, stream>>>((const __nv_bfloat16*)x, xf, fg);
    BT_LAUNCH_OK("tc2_flipx_kernel");
    rc = build_patches(xf, patches2);
    if (rc != BT_OK) return rc;
    *pm2_out = patches2;
  }
  return BT_OK;
}

// What distinguishes a backward-data run of the kernel from a forward run.
struct RunOpt {
  bool dgrad = false;        // pack W transposed / tap-reversed with regenerated eps (tc2_pack_dgrad_kernel)
  int cin_w = 0, cout_w = 0; // the layer's own (C_in, C_out) when dgrad
  int a_dil_h = 1, a_dil_w = 1, SH = 0, SW = 0;
  int64_t s_bs = 0, s_hs = 0, s_ws = 0;  // strides of the un-inserted grad_out (tensor-fed signs)
  int64_t ws_bytes = 0;      // bytes the caller says `workspace` holds (0: not stated, unchecked)
  float* kl_out = nullptr;   // forward(return_kl=True): kl(weight) + kl(bias) <- evaluated by the packing pass
  void* kl_scratch = nullptr;  // >= 32 bytes of the caller's workspace for the fp64 accumulators
};

static int tc2_run(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, void* workspace,
                   bool flipout, const RunOpt& opt, cudaStream_t stream) {
  const char* why = "";
  Plan pl;
  // (zero-inserted grad_out: the flippers need image coordinates, so a 1x1 filter stays in im2col mode)
  BT_REQUIRE(tc2_plan(g, p, flipout, &pl, &why, opt.a_dil_h > 1 || opt.a_dil_w > 1), BT_ERR_UNSUPPORTED, "tensor-core %s: %s",
             opt.dgrad ? "dgrad" : "forward", why);
  BT_REQUIRE(workspace, BT_ERR_INVALID, "tensor-core forward: workspace required (bt_workspace_bytes)");
  BT_REQUIRE(((uintptr_t)out & 15) == 0 && ((uintptr_t)workspace & 15) == 0 && (pl.mode == A_PATCH || ((uintptr_t)x & 15) == 0),
             BT_ERR_ALIGN, "tensor-core forward: x / out / workspace must be 16-byte aligned");
  const DriverFns& drv = driver_fns();
  Tc2Args a;
  memset(&a, 0, sizeof(a));
  const int S = g->n_samples;
  const bool x_shared = S > 1 && g->x_sample_stride == 0;
  a.N = (int)g->c_out;
  a.S = S;
  a.mu_b = p->mu_b; a.rho_b = p->rho_b; a.eps_b = p->eps_b;
  a.out_scale = p->out_scale; a.out_shift = p->out_shift; a.out_relu = p->out_relu;
  a.seed = p->seed; a.sample_id = p->sample_id; a.layer_id = p->layer_id;
  a.debug = env_int("BT_TC2_DEBUG", 0);
  a.pdl = env_int("BT_PDL", 1) && !env_int("BT_TC2_PROF", 0) ? 1 : 0;
  a.sign_in = flipout ? p->sign_in : nullptr;
  a.sign_out = flipout ? p->sign_out : nullptr;
  a.kind_a = opt.dgrad ? BT_KIND_SIGN_OUT : BT_KIND_SIGN_IN;
  a.kind_o = opt.dgrad ? BT_KIND_SIGN_IN : BT_KIND_SIGN_OUT;
  a.a_dil_h = opt.a_dil_h; a.a_dil_w = opt.a_dil_w;
  a.Cin = (int)g->c_in;
  a.x_pitch = pl.pitch;
  a.Ms = (int)pl.rows_per_s;
  a.Kt = pl.K;
  a.K = pl.mode == A_PATCH ? pl.Kp : pl.K;
  a.tf32 = pl.es == 4 ? 1 : 0;
  const int bke = 128 / pl.es;   // elements per k-block
  const CUtensorMapDataType a_dt = a.tf32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  BT_REQUIRE(!a.tf32 || (!opt.dgrad && !flipout), BT_ERR_UNSUPPORTED, "tensor-core forward: fp32 activations support the Reparameterization forward only");
  a.x_rows_per_s = x_shared ? 0 : (int)pl.x_rows;
  alignas(64) CUtensorMap tm_a, tm_out, tm_res, tm_a2;
  const int es_o = g->out_dtype == BT_F32 ? 4 : 2;

  // ---- workspace carve-up: packed parameters, then (A_PATCH) the patch matrix
  const int64_t wn = (int64_t)a.N * a.K;
  // bf16 packing: mu (2 B), sigma (2 B), eps (4 B) per weight; tf32: three fp32 arrays -- both within 12 B per weight
  __nv_bfloat16* mu_p = (__nv_bfloat16*)workspace;
  __nv_bfloat16* sig_p = a.tf32 ? (__nv_bfloat16*)((float*)workspace + wn) : mu_p + wn;
  float* eps_p = (p->eps_w || opt.dgrad) ? (a.tf32 ? (float*)workspace + 2 * wn : (float*)(sig_p + wn)) : nullptr;
  __nv_bfloat16* patches = (__nv_bfloat16*)((char*)workspace + ((3 * wn * 4 + 255) / 256) * 256);
  bool kl_done = false, packed_frozen = false;
  bool wq_fits = opt.ws_bytes == 0;   // unchecked workspace: sized by bt_workspace_bytes, which counts them
  if (opt.ws_bytes > 0) {  // the carve-up below must fit what the caller allocated (bt_workspace_bytes)
    int64_t need = ((3 * wn * 4 + 255) / 256) * 256;
    if (pl.mode == A_PATCH) {
      const int64_t pm = patch_bytes(g, pl.rows_per_s, pl.Kp);
      need += pm + (flipout ? pm + ((x_span_elems(g) * 2 + 255) / 256) * 256 : 0);
    }
    // room for the pre-sampled weights is optional: a caller that sized its workspace without them gets the fused sampling
    wq_fits = need + wpre_bytes(g, pl, flipout) <= opt.ws_bytes;
    BT_REQUIRE(need <= opt.ws_bytes, BT_ERR_INVALID, "tensor-core %s: workspace of %lld bytes, %lld needed (bt_workspace_bytes)",
               opt.dgrad ? "dgrad" : "forward", (long long)opt.ws_bytes, (long long)need);
  }
  {
    int64_t blocks = (wn + 255) / 256;
    const int64_t cap = (int64_t)bt_num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (opt.dgrad) {
      tc2_pack_dgrad_kernel<<<(int)blocks, 256, 0, stream>>>(p->mu_w, p->rho_w, p->eps_w, mu_p, sig_p, eps_p, wn, opt.cin_w, opt.cout_w, pl.T,
                                                             a.K, bt_rng_key(p->seed, p->sample_id, p->layer_id, BT_KIND_EPS_W));
      BT_LAUNCH_OK("tc2_pack_dgrad_kernel");
    } else if ((p->flags & BT_FLAG_PACKED_VALID) && !p->eps_w) {
      // frozen inference model: the workspace still holds this layer's packed parameters
      packed_frozen = true;
      if (a.pdl && !p->eps_b && pl.mode != A_PATCH) a.pdl = 2;
    } else if (opt.kl_out) {
      BT_REQUIRE(opt.kl_scratch, BT_ERR_INVALID, "tensor-core forward: kl_out needs the workspace's KL scratch");
      PackKl kf;
      memset(&kf, 0, sizeof(kf));
      kf.acc = (double*)opt.kl_scratch;
      kf.counter = (unsigned int*)(kf.acc + 2);
      kf.kl_out = opt.kl_out;
      kf.pmu_t = p->prior_mu_w_t; kf.psig_t = p->prior_sigma_w_t;
      kf.mu_b = p->mu_b; kf.rho_b = p->rho_b; kf.pmu_b_t = p->prior_mu_b_t; kf.psig_b_t = p->prior_sigma_b_t;
      kf.prior_mu = p->prior_mu; kf.prior_sigma = p->prior_sigma;
      kf.w_numel = (int64_t)a.N * pl.K;
      kf.b_numel = p->mu_b ? a.N : 0;
      BT_CUDA_OK(cudaMemsetAsync(opt.kl_scratch, 0, 32, stream));
      if (a.tf32)
        tc2_pack_kernel<true, float><<<(int)blocks, 256, 0, stream>>>(p->mu_w, p->rho_w, p->eps_w, (float*)mu_p, (float*)sig_p, eps_p, p->out_scale,
                                                                      wn, (int)g->c_in, pl.T, a.K, kf);
      else
      tc2_pack_kernel<true, __nv_bfloat16><<<(int)blocks, 256, 0, stream>>>(p->mu_w, p->rho_w, p->eps_w, mu_p, sig_p, eps_p, p->out_scale, wn, (int)g->c_in,
                                                             pl.T, a.K, kf);
      BT_LAUNCH_OK("tc2_pack_kernel(kl)");
      kl_done = true;
    } else {
      PackKl kf;
      memset(&kf, 0, sizeof(kf));
      if (a.tf32)
        tc2_pack_kernel<false, float><<<(int)blocks, 256, 0, stream>>>(p->mu_w, p->rho_w, p->eps_w, (float*)mu_p, (float*)sig_p, eps_p, p->out_scale,
                                                                       wn, (int)g->c_in, pl.T, a.K, kf);
      else
      tc2_pack_kernel<false, __nv_bfloat16><<<(int)blocks, 256, 0, stream>>>(p->mu_w, p->rho_w, p->eps_w, mu_p, sig_p, eps_p, p->out_scale, wn, (int)g->c_in,
                                                              pl.T, a.K, kf);
      BT_LAUNCH_OK("tc2_pack_kernel");
    }
  }
  if (opt.kl_out && !kl_done) {   // packing skipped (frozen layer) or a backward-data run: the stand-alone KL launch
    const int rc = bt_kl_layer(p, (int64_t)a.N * pl.K, p->mu_b ? a.N : 0, opt.kl_out, opt.kl_scratch, stream);
    if (rc != BT_OK) return rc;
  }
  a.mu_p = mu_p; a.sig_p = sig_p; a.eps_p = eps_p;
  a.has_shift = (p->mu_b || p->out_shift) ? 1 : 0;

  const void* a_src = x;
  __nv_bfloat16* patches2 = nullptr;
  if (pl.mode == A_PATCH) {
    __nv_bfloat16* pm = nullptr;
    const int rc = build_patch_matrices(g, pl, p, x, flipout, a.kind_a, patches, &pm, &patches2, stream);
    if (rc != BT_OK) return rc;
    a_src = pm;
    if (patches2) a.a2_tma = 1;
  }

  if (pl.mode == A_GEMM || pl.mode == A_PATCH) {
    a.conv = 0;
    const cuuint64_t rows = (cuuint64_t)pl.x_rows * (x_shared ? 1 : S);
    const cuuint64_t gdim[2] = {(cuuint64_t)a.K, rows};
    const cuuint64_t gstr[1] = {(cuuint64_t)pl.pitch * pl.es};
    const cuuint32_t box[2] = {(cuuint32_t)bke, BM};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = drv.tiled(&tm_a, a_dt, 2, const_cast<void*>(a_src), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(x) failed with %d", (int)r);
    tm_a2 = tm_a;
    if (patches2) {
      const CUresult r2 = drv.tiled(&tm_a2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, patches2, gdim, gstr, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      BT_REQUIRE(r2 == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(x * s_in) failed with %d", (int)r2);
    }
  } else {
    const Dense& d = pl.d;
    a.conv = 1;
    a.CB = (int)g->c_in / bke;
    a.OH = (int)d.out[0]; a.OW = (int)d.out[1]; a.KW = (int)d.k[1];
    a.sh = (int)d.st[0]; a.sw = (int)d.st[1]; a.ph = (int)d.pd[0]; a.pw = (int)d.pd[1]; a.dh = (int)d.dl[0]; a.dw = (int)d.dl[1];
    a.IH = (int)d.in[0]; a.IW = (int)d.in[1];
    a.SH = a.IH; a.SW = a.IW;
    a.x_bs = g->x_batch_stride; a.x_hs = pl.xh; a.x_ws = pl.xw;
    if (opt.a_dil_h > 1 || opt.a_dil_w > 1) { a.SH = opt.SH; a.SW = opt.SW; a.x_bs = opt.s_bs; a.x_hs = opt.s_hs; a.x_ws = opt.s_ws; }
    const cuuint64_t imgs = (cuuint64_t)g->batch * (x_shared ? 1 : S);
    const cuuint64_t gdim[4] = {(cuuint64_t)g->c_in, (cuuint64_t)d.in[1], (cuuint64_t)d.in[0], imgs};
    const cuuint64_t gstr[3] = {(cuuint64_t)pl.xw * pl.es, (cuuint64_t)pl.xh * pl.es, (cuuint64_t)g->x_batch_stride * pl.es};
    // bounding box of the filter's top-left corner, in (W, H) order like the coordinates
    const int lower[2] = {-(int)d.pd[1], -(int)d.pd[0]};
    const int upper[2] = {(int)(d.pd[1] - (d.k[1] - 1) * d.dl[1]), (int)(d.pd[0] - (d.k[0] - 1) * d.dl[0])};
    const cuuint32_t estr[4] = {1, (cuuint32_t)d.st[1], (cuuint32_t)d.st[0], 1};
    const CUresult r = drv.im2col(&tm_a, a_dt, 4, const_cast<void*>(x), gdim, gstr, lower, upper,
                                  (cuuint32_t)bke, BM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed with %d", (int)r);
    // Small-tensor workaround of drivers <= 13.1 (same fix-up CUTLASS applies to im2col descriptors).
    const int64_t span = (int64_t)imgs * g->x_batch_stride * pl.es;
    if (drv.driver_version <= 13010 && span < 131072) reinterpret_cast<uint64_t*>(&tm_a)[1] &= ~(1ull << 21);
    tm_a2 = tm_a;
  }
  {
    const cuuint64_t gdim[3] = {(cuuint64_t)a.N, (cuuint64_t)a.Ms, (cuuint64_t)S};
    const cuuint64_t gstr[2] = {(cuuint64_t)a.N * es_o, (cuuint64_t)a.Ms * a.N * es_o};
    const cuuint32_t box[3] = {(cuuint32_t)(128 / es_o), 32, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = drv.tiled(&tm_out, es_o == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                                 out, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(out) failed with %d", (int)r);
    tm_res = tm_out;  // placeholder when there is no residual (never dereferenced)
    if (p->residual) {  // same shape, dtype and strides as out: the same map over another base
      BT_REQUIRE(((uintptr_t)p->residual & 15) == 0, BT_ERR_ALIGN, "tensor-core forward: residual must be 16-byte aligned");
      const cuuint32_t rbox[3] = {box[0], BM, 1};  // one box = the residual of a whole 128-row tile, 32 rows per epilogue warp
      const CUresult r2 = drv.tiled(&tm_res, es_o == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                                    const_cast<void*>(p->residual), gdim, gstr, rbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      BT_REQUIRE(r2 == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(residual) failed with %d", (int)r2);
    }
  }

  // ---- tiling / pipeline configuration
  const int sms = bt_num_sms();
  a.num_mt = (a.Ms + BM - 1) / BM;
  a.nkb = (a.K + bke - 1) / bke;
  // x shared by all samples and exactly 64 output channels (the ResNet stem under mc_batched): two samples ride side by
  // side on the N axis of a BN = 128 tile, so an A tile is fetched once per pair and feeds a full-width MMA (at N = 64
  // the tensor core reads 128 B of A per cycle at full rate: the layer was bound by that read)
  a.spn = (x_shared && a.N == 64 && S % 2 == 0 && !flipout && !p->residual && !p->eps_w && !p->eps_b && !opt.dgrad &&
           env_int("BT_TC2_SPN", 1) != 0) ? 2 : 1;
  if (a.spn == 2) a.S = S / 2;
  int BN = a.N * a.spn <= 64 ? 64 : 128;
  // shared-memory budget: staging (1, 2 or 4 buffers per epilogue warp) + residual ring (0, 2 or 4)
  // + A ring + W ring / panel
  const int res_min = p->residual ? 2 : 0;
  // Flipout doubles a W slot (mu and sigma*eps sub-tiles), an A slot (the tile and its sign-flipped copy)
  // and the accumulator columns of a tile, and adds a second shift tile
  const int NSUBW = flipout ? 2 : 1;
  const int A_SLOT_T = NSUBW * A_SLOT;
  const int avail = AVAIL_BYTES - (flipout ? SHIFT_BYTES : 0);
  int SG = 1;   // epilogue groups, each with its own staging buffers (2 with eight epilogue warps)
  auto ring_bytes = [&](int stage_bufs, int res_bufs) { return avail - (stage_bufs * SG + res_bufs) * STAGE_UNIT; };
  // a [BN x K] panel that fits next to a 4-deep A ring is generated once per (sample, n-block)
  // (single-buffered epilogue staging is accepted for that)
  auto panel_fits = [&](int bn) { return a.nkb <= MAX_RW && (int64_t)a.nkb * NSUBW * bn * BK * 2 + 4 * A_SLOT_T <= ring_bytes(1, res_min); };
  const int force_bn = env_int("BT_TC2_BN", 0), force_mt = env_int("BT_TC2_MT", 0), force_stat = env_int("BT_TC2_STATIONARY", -1);
  // N > 64 whose [128 x K] panel does not fit streams at BN = 128 rather than keeping a BN = 64
  // panel: the narrow panel re-reads every A tile once per 64 output channels, and the A stream
  // (L2 -> shared memory, then shared memory -> tensor core) is the scarcer resource
  // (1024 -> 256 1x1 at 14x14: 0.34 ms stationary/64, 0.23 ms streaming/128).
  if (force_bn == 64 || force_bn == 128) BN = force_bn;
  // BN = 256 for long reductions into >= 256 channels with pre-sampled W.  The compute-bound layers are bound by the
  // L2 -> SM fabric (~43 B / cycle / SM: a 16 KB A tile takes ~385 cycles to arrive, as long as its four N = 128 MMAs
  // run at the sustained rate; 3x3 256 at 14x14 moved 10 TB/s, profiles/r2_conv3x3_fwd_ncu_metrics.txt): twice the
  // columns per A tile halve the A bytes per FLOP.  Two 256-column accumulators fill the TMEM (MT = 2).  The conditions
  // repeat those of the pre-sampled mode, which the 256-row panel (never fits for >= 4 k-blocks) then always selects.
  // Measured (profiles/r2_exp13_bn256.txt, r2_exp14_bn256_short_k.txt): 3x3 256 @14 0.87 -> 0.74 ms (1 275 TFLOP/s; 1 361 inside the
  // network), 3x3 512 @7 0.94 -> 0.75, 512<-1024 0.90 -> 0.81, 512<-2048 0.47 -> 0.41, 2048<-1024 /2 0.90 -> 0.78.  Short reductions
  // (4-7 k-blocks) gain without a residual input (512<-256 /2 1.28 -> 1.16) and lose with one (1024<-256 0.73 -> 0.90: the single
  // accumulator set no longer hides the residual epilogue), hence the two thresholds.
  if (force_bn != 64 && force_bn != 128 && a.N % 256 == 0 && a.nkb >= env_int("BT_TC2_BN256_NKB", p->residual ? 8 : 4) && g->out_dtype == BT_BF16 && !flipout &&
      !a.tf32 && a.spn == 1 && a.Kt == a.K && !opt.dgrad && !p->eps_w && wq_fits && wpre_bytes(g, pl, flipout) > 0 &&
      !x_shared && a.num_mt >= env_int("BT_TC2_WPRE_MIN", 16) && env_int("BT_TC2_WPRE", 1) != 0 &&
      !env_int("BT_TC2_PROF", 0) && env_int("BT_TC2_BN256", 1) != 0)
    BN = 256;
  // Eight epilogue warps (EPI8 instantiation) where the output stream is the bottleneck: full 128-column tiles,
  // bf16 outputs and ONE k-block (K <= 64: the 256<-64 expansions of layer1, -3 % on the same box).  Longer
  // reductions stream W and need all 16 producer warps: 1024<-256 with 12 producers ran 0.95 -> 1.26 ms
  // (profiles/r2_experiments.txt), so the switch stays where the producers are idle.
  const bool epi8_ok = BN == 128 && !flipout && !a.tf32 && g->out_dtype == BT_BF16 && a.spn == 1 && a.N % 128 == 0 && !env_int("BT_TC2_PROF", 0) &&
                       env_int("BT_TC2_EPI8", 1) != 0;
  a.epi8 = (epi8_ok && a.nkb <= env_int("BT_TC2_EPI8_NKB", 1)) ? 1 : 0;
  if (a.epi8) SG = 2;
  const int WB = NSUBW * BN * BK * 2;
  a.NB = (a.N * a.spn + BN - 1) / BN;
  a.stationary = panel_fits(BN) ? 1 : 0;
  // Wide layers with a short reduction (the 1x1 expansions, N >= 4 n-blocks) stream instead: with
  // the n-block innermost in the item order every group of A tiles is fetched from HBM once rather
  // than once per n-block, and a CTA pair halves the sampling cost that streaming re-pays per item.
  if (a.stationary && a.NB >= 4 && a.nkb >= 2 && a.num_mt >= 4) a.stationary = 0;
  if (force_stat == 0) a.stationary = 0;
  if (force_stat == 1 && panel_fits(BN)) a.stationary = 1;
  // x shared by all samples (the first layer of an MC-batched model), BT_TC2_SINNER=1: samples are walked like
  // n-blocks of one wide GEMM -- (m-group, n-block, sample) order, W streamed -- so an m-group's A tiles are
  // fetched once and, when the reduction is short enough for the ring to hold a whole item, never leave shared
  // memory.  OFF by default: the order re-samples W for every (m-group, sample) item, and on the ResNet stem
  // (3->64 7x7, 32 samples x 128 images) that costs more than the A traffic it saves (7.1 ms against 3.3 ms
  // with one stationary panel per sample; profiles/r1_experiments.txt).
  // With pre-sampled W (below) the order no longer re-samples anything -- the W tiles of an item come from L2 by TMA and the
  // patch matrix of the stem is read ONCE instead of once per sample pair (7.8 GB -> 0.5 GB) -- and it is still slower:
  // stem at 32 samples 2.19 -> 2.56 ms, at 4 samples 0.57 -> 0.62 (profiles/r2_exp17_stem_sample_fastest.txt; the 48 KB of W
  // per two-tile item cost the L2 -> SM fabric what the A tiles did).  BT_TC2_SINNER=1 selects it; bit-identical either way.
  // (backward data draws nothing: its eps come packed from tc2_pack_dgrad_kernel; Flipout materialises delta = sigma * eps only)
  const bool wpre_static = wpre_bytes(g, pl, flipout) > 0 && wq_fits && (opt.dgrad || !p->eps_w) && !a.tf32 && !(flipout && p->residual) &&
                           (!(flipout || opt.dgrad) || env_int("BT_TC2_WPRE_TRAIN", 1) != 0) &&
                           a.num_mt >= env_int("BT_TC2_WPRE_MIN", 16) && env_int("BT_TC2_WPRE", 1) != 0;
  const int sinner_env = env_int("BT_TC2_SINNER", 0);
  const bool s_inner = x_shared && !flipout && a.nkb <= MAX_NA / 2 && sinner_env == 1;
  if (s_inner) a.stationary = 0;
  const int NSLOTS = 512 / (NSUBW * BN);
  // Streamed W on a layer with many m-tiles per sample: the fused kernel would re-sample every [BN x K] panel once per
  // m-group (x25 on the 3x3 layers at 14x14) and its producers, not the tensor core, set the pace.  Sample once into the
  // workspace instead (tc2_presample_kernel) and stream the k-blocks in by TMA; no CTA pairs then (no sampling left to
  // share).  Measured at 32 samples x 128 images, fused -> pre-sampled (profiles/r2_experiments.txt 7): 3x3 128 @28
  // 1.11 -> 0.93 ms, 3x3 256 @14 1.03 -> 0.90, 256<-1024 @14 0.69 -> 0.65, 512<-256 /2 1.42 -> 1.25; at 7x7 (49 m-tiles per
  // sample) the W stream costs more than the sampling it replaces (512<-2048: 0.69 -> 0.80), hence the m-tile floor.
  {
    const bool wpre_ok = wpre_static && !a.epi8 && (a.Kt == a.K || !(opt.dgrad || flipout));   // (the padded-K pass draws from Philox only)
    // a stationary panel walks (sample, n-block, m-group): with two or more n-blocks every A tile comes from HBM once per
    // n-block.  Streaming pre-sampled W with the n-block innermost reads it once (256<-512 at 28x28: 3.3 GB of A).
    if (wpre_ok && a.stationary && a.NB >= 2 && a.nkb >= 2 && force_stat != 1) a.stationary = 0;
    a.w_pre = (wpre_ok && !a.stationary) ? 1 : 0;
    // with the weights pre-sampled the producer warps idle: four of them become the second epilogue group on the layers
    // whose output (+ residual) stream sets the pace (short reductions into wide outputs)
    // (measured: NOT a win -- 512<-128 0.89 -> 0.95 ms, 1024<-256 0.62 -> 0.67, 256<-512 1.05 -> 1.31; with a residual within
    // +-5 %; profiles/r2_exp6_epi8_presampled.txt -- the epilogue of these layers is not short of warps.  Off by default.)
    if (a.w_pre && epi8_ok && a.nkb <= env_int("BT_TC2_EPI8_WPRE_NKB", 0)) { a.epi8 = 1; SG = 2; }
    if (a.w_pre && a.pdl == 2) a.pdl = 1;   // the W loader reads what the kernel just before this one wrote
  }
  if (a.stationary) {
    a.RW = a.nkb;
    a.MT = a.num_mt >= 2 ? 2 : 1;  // one tile per issuer warp
    if (force_mt >= 1 && force_mt <= MAX_MT && force_mt <= NSLOTS) a.MT = force_mt;
  } else {
    a.MT = 1;
    for (int mt = NSLOTS < MAX_MT ? NSLOTS : MAX_MT; mt >= 1; mt >>= 1) {
      const int64_t items = (int64_t)a.S * a.NB * ((a.num_mt + mt - 1) / mt);
      if (items >= sms || mt == 1 || (mt == 2 && a.num_mt >= 2)) { a.MT = mt; break; }
    }
    // short reductions WITH a residual input: two accumulator sets of two tiles, so the (long) epilogue of one item
    // overlaps the MMAs of the next.  Without a residual the epilogue is short and four tiles per sampled k-block win
    // (512<-256 /2 at 56x56: 1.63 -> 1.40 ms, 1024<-256 at 14x14: 0.84 -> 0.73 ms; profiles/r2_experiments.txt)
    if (a.nkb <= 8 && a.MT > 2 && (p->residual || env_int("BT_TC2_SHORTK_MT2", 0))) a.MT = 2;
    // pre-sampled W, 1x1 layers: nothing is saved by sharing a k-block between four tiles, and two sets of two overlap
    // the epilogue with the next item's MMAs (256<-1024 @14: 0.74 ms with four, 0.65 with two; the 3x3 layers and the wide 1x1 layers prefer four)
    if (a.w_pre && pl.T == 1 && a.MT > 2 && a.NB <= 2) a.MT = 2;
    if (s_inner) a.MT = a.num_mt >= 2 ? 2 : 1;
    if (force_mt >= 1 && force_mt <= MAX_MT && force_mt <= NSLOTS) a.MT = force_mt;
    a.RW = (ring_bytes(2, res_min) - 4 * A_SLOT_T) / WB;
    if (a.RW > 4) a.RW = 4;  // deeper W rings measured no faster; the space goes to the A ring
    if (flipout) a.RW = 2;   // 32 KB slots: two W k-blocks, the rest of shared memory goes to the A ring
    if (env_int("BT_TC2_RW", 0) >= 2 && env_int("BT_TC2_RW", 0) < a.RW) a.RW = env_int("BT_TC2_RW", 0);
  }
  a.NSETS = NSLOTS / a.MT;
  if (a.NSETS > MAX_SETS) a.NSETS = MAX_SETS;
  a.num_mg = (a.num_mt + a.MT - 1) / a.MT;
  // Streaming layers are bound by the W producers (~30 issue slots and 2 MUFU per sampled weight):
  // a CTA pair walks the same k-block sequence on neighbouring m-groups and each CTA samples half
  // of every k-block, storing it into both rings over DSMEM.
  a.CL = (!a.stationary && !a.w_pre && a.num_mg >= 2 && sms % 2 == 0) ? 2 : 1;
  if (env_int("BT_TC2_CL", 0) == 1) a.CL = 1;
  if (env_int("BT_TC2_CL", 0) == 2 && a.num_mg >= 2 && sms % 2 == 0 && !a.w_pre) a.CL = 2;
  a.num_mg = (a.num_mg + a.CL - 1) / a.CL;
  a.nb_inner = (!a.stationary && a.NB > 1) ? 1 : 0;
  if (env_int("BT_TC2_NBINNER", -1) == 0) a.nb_inner = 0;
  if (s_inner) a.nb_inner = 2;
  a.stage_bufs = 2;
  a.res_bufs = res_min;
  a.NA = (ring_bytes(2, res_min) - a.RW * WB) / A_SLOT_T;
  if (a.NA < 4 || flipout) { a.stage_bufs = 1; a.NA = (ring_bytes(1, res_min) - a.RW * WB) / A_SLOT_T; }
  if (a.NA > MAX_NA) a.NA = MAX_NA;
  // leftover shared memory first deepens the residual prefetch (its tiles come from L2 / HBM, a
  // microsecond away), then the store staging; a stationary layer with a long A ring trades two
  // of its slots for the former
  {
    int left = ring_bytes(a.stage_bufs, a.res_bufs) - a.RW * WB - a.NA * A_SLOT_T;
    if (a.res_bufs == 2) {
      // Two residual sub-tiles in flight (32 KB per SM) cap the residual stream at ~32 KB per HBM round trip
      // (~2 000 cycles): about half of an SM's share of the HBM bandwidth, and the wide expansions read as many
      // residual bytes as they write outputs (512<-128 at 28x28 ran at 58 % of the HBM bound with two, the
      // 256<-64 layers at 87 % with four).  A streaming layer gives up one of its four W slots for the space.
      if (left < 2 * STAGE_UNIT && !a.stationary && a.RW >= 3 && env_int("BT_TC2_RES4", 1) != 0) { a.RW -= 1; left += WB; }
      if (left >= 2 * STAGE_UNIT) { a.res_bufs = 4; left -= 2 * STAGE_UNIT; }
      else if (a.stationary && a.NA >= 6) { a.res_bufs = 4; a.NA -= 2; left += 2 * A_SLOT_T - 2 * STAGE_UNIT; }
    }
    if (a.stage_bufs == 1 && left >= SG * STAGE_UNIT) { a.stage_bufs = 2; left -= SG * STAGE_UNIT; }
    if (a.stage_bufs == 2 && left >= 2 * SG * STAGE_UNIT) a.stage_bufs = 4;
  }
  if (env_int("BT_TC2_STAGE", 0) == 2 && a.stage_bufs == 4) a.stage_bufs = 2;
  if (env_int("BT_TC2_NA", 0) >= 2 && env_int("BT_TC2_NA", 0) < a.NA) a.NA = env_int("BT_TC2_NA", 0);
  a.NR = a.MT >= 2 ? N_MMA_WARPS : 1;
  a.NAr = a.NA / a.NR;
  a.NA = a.NAr * a.NR;
  if (s_inner && a.NAr >= a.nkb && a.NR * a.nkb >= 2 && (a.MT == 1 || a.MT == a.NR)) {  // one tile per ring: slot kb <-> k-block kb
    a.NAr = a.nkb;
    a.NA = a.NR * a.NAr;
    a.a_keep = 1;
  }
  // second pass over what the roundings above left unused (an odd A slot of a two-ring plan, a W slot given up for a
  // residual ring that still did not fit): 256<-1024-class expansions with a residual ran with 2.5 of 13 units idle
  if (!flipout && env_int("BT_TC2_SMEM2", 1) != 0) {
    auto used = [&]() { return 1024 + CTRL_BYTES + (a.stage_bufs * SG + a.res_bufs) * STAGE_UNIT + a.NA * A_SLOT_T + a.RW * WB; };
    if (a.res_bufs == 2 && SMEM_LIMIT - used() >= 2 * STAGE_UNIT) a.res_bufs = 4;
    if (!a.stationary && a.RW < 4 && SMEM_LIMIT - used() >= WB) a.RW += 1;
    if (a.stage_bufs == 2 && SMEM_LIMIT - used() >= 2 * SG * STAGE_UNIT) a.stage_bufs = 4;
  }
  BT_REQUIRE(a.NA >= 2 && a.RW >= 1 && a.RW <= MAX_RW, BT_ERR_UNSUPPORTED, "tensor-core forward: no room for the pipelines");
  a.total_items = (int64_t)a.S * a.NB * a.num_mg;
  const int ncl = sms / a.CL;
  const int grid = (int)(a.total_items < ncl ? a.total_items : ncl) * a.CL;
  if (a.nb_inner == 1 && grid == ncl * a.CL && env_int("BT_TC2_RR", 1) != 0) {
    a.rr = 1;
    a.rr_dnb = ncl % a.NB;
    a.rr_dmg = (ncl / a.NB) % a.num_mg;
    a.rr_ds = (ncl / a.NB) / a.num_mg;
  }
  a.two_loaders = (a.NR == 2 && a.res_bufs == 0 && !a.a_keep && !flipout && env_int("BT_TC2_LOADERS", 2) == 2) ? 1 : 0;
  a.pair_launch = (a.rr && a.CL == 1 && a.NB % 2 == 0 && env_int("BT_TC2_PAIR", 1) != 0) ? 1 : 0;
  const int smem_bytes = 1024 + CTRL_BYTES + (flipout ? SHIFT_BYTES : 0) + (a.stage_bufs * SG + a.res_bufs) * STAGE_UNIT + a.NA * A_SLOT_T + a.RW * WB;
  BT_REQUIRE(smem_bytes <= SMEM_LIMIT, BT_ERR_UNSUPPORTED, "tensor-core forward: shared memory plan %d exceeds %d", smem_bytes, SMEM_LIMIT);
  if (a.w_pre) {
    // behind the packed area and, in patch-matrix mode, behind the patch matrix
    uint4* wq = reinterpret_cast<uint4*>(reinterpret_cast<char*>(patches) + (pl.mode == A_PATCH ? patch_bytes(g, pl.rows_per_s, pl.Kp) : 0));
    const int64_t chunks = wn / 8;
    int64_t bx = (chunks + 255) / 256;
    const int64_t capx = ((int64_t)sms * 8 + S - 1) / S;
    if (bx > capx) bx = capx < 1 ? 1 : capx;
    {
      // Early start (2) is opt-in: it measured within noise, and it is only safe when the kernel just before this one in the
      // stream does not read this layer's pre-sampled copy (the same frozen layer called back to back would).
      const int ppdl = !a.pdl ? 0 : (packed_frozen && env_int("BT_TC2_WPRE_PDL", 1) == 2 ? 2 : 1);
      cudaLaunchConfig_t cfg;
      memset(&cfg, 0, sizeof(cfg));
      cfg.gridDim = dim3((unsigned)bx, (unsigned)S);
      cfg.blockDim = dim3(256);
      cfg.stream = stream;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = ppdl ? 1 : 0;
      if (a.Kt != a.K)
        BT_CUDA_OK(cudaLaunchKernelEx(&cfg, tc2_presample_padded_kernel, (const __nv_bfloat16*)mu_p, (const __nv_bfloat16*)sig_p, (__nv_bfloat16*)wq,
                                      a.N, a.K, a.Kt, (uint64_t)p->seed, (uint32_t)p->sample_id, (uint32_t)p->layer_id, ppdl));
      else
      BT_CUDA_OK(cudaLaunchKernelEx(&cfg, tc2_presample_kernel, reinterpret_cast<const uint4*>(mu_p), reinterpret_cast<const uint4*>(sig_p), wq, chunks,
                                    (uint64_t)p->seed, (uint32_t)p->sample_id, (uint32_t)p->layer_id, ppdl, (const float*)eps_p, flipout ? 1 : 0));
    }
    BT_LAUNCH_OK("tc2_presample_kernel");
    // a stacked sample pair (spn = 2) reads rows [0, 64) of sample 2 s and of sample 2 s + 1 as ONE 128-row block: the same memory
    const cuuint64_t gdim[3] = {(cuuint64_t)a.K, (cuuint64_t)a.N * a.spn, (cuuint64_t)(S / a.spn)};
    const cuuint64_t gstr[2] = {(cuuint64_t)a.K * 2, (cuuint64_t)a.N * a.spn * a.K * 2};
    const cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)BN, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = drv.tiled(&tm_a2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, wq, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(pre-sampled W) failed with %d", (int)r);
    if (flipout) {   // the mu sub-tiles come from the packed parameters themselves ([N][K] bf16); no residual in this mode: tm_res is free
      const cuuint64_t gdim2[2] = {(cuuint64_t)a.K, (cuuint64_t)a.N};
      const cuuint64_t gstr2[1] = {(cuuint64_t)a.K * 2};
      const cuuint32_t box2[2] = {(cuuint32_t)BK, (cuuint32_t)BN};
      const cuuint32_t estr2[2] = {1, 1};
      const CUresult r2 = drv.tiled(&tm_res, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, mu_p, gdim2, gstr2, box2, estr2, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      BT_REQUIRE(r2 == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(packed mu) failed with %d", (int)r2);
    }
  }
  if (env_int("BT_TC2_LOG", 0))
    fprintf(stderr, "tc2: wpre=%d flip=%d mode=%d S=%d Ms=%d N=%d K=%d(%d) BN=%d stat=%d MT=%d NSETS=%d nkb=%d NA=%dx%d RW=%d NB=%d num_mg=%d items=%lld grid=%d smem=%d stage=%d res=%d CL=%d order=%d keep=%d spn=%d epi8=%d tf32=%d pdl=%d\n",
            a.w_pre, (int)flipout, (int)pl.mode, a.S, a.Ms, a.N, a.K, a.Kt, BN, a.stationary, a.MT, a.NSETS, a.nkb, a.NR, a.NAr, a.RW, a.NB, a.num_mg,
            (long long)a.total_items, grid, smem_bytes, a.stage_bufs, a.res_bufs, a.CL, a.nb_inner, a.a_keep, a.spn, a.epi8, a.tf32, a.pdl);
  const bool o32 = g->out_dtype == BT_F32;
  if (BN == 256) {
    BT_REQUIRE(a.w_pre && a.CL == 1 && !o32 && !flipout && !a.tf32 && !a.epi8, BT_ERR_UNSUPPORTED, "tensor-core forward: BN = 256 plan without pre-sampled W");
    return tc2_launch<256, __nv_bfloat16>(tm_a, tm_out, tm_res, tm_a2, a, grid, smem_bytes, flipout, stream);
  }
  if (BN == 64)
    return o32 ? tc2_launch<64, float>(tm_a, tm_out, tm_res, tm_a2, a, grid, smem_bytes, flipout, stream)
               : tc2_launch<64, __nv_bfloat16>(tm_a, tm_out, tm_res, tm_a2, a, grid, smem_bytes, flipout, stream);
  return o32 ? tc2_launch<128, float>(tm_a, tm_out, tm_res, tm_a2, a, grid, smem_bytes, flipout, stream)
             : tc2_launch<128, __nv_bfloat16>(tm_a, tm_out, tm_res, tm_a2, a, grid, smem_bytes, flipout, stream);
}

int bt_tc_forward(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, void* workspace, int64_t ws_bytes,
                  bool flipout, float* kl_out, void* kl_scratch, cudaStream_t stream) {
  RunOpt opt;
  opt.ws_bytes = ws_bytes;
  opt.kl_out = kl_out;
  opt.kl_scratch = kl_scratch;
  return tc2_run(g, p, x, out, workspace, flipout, opt, stream);
}

// ------------------------------------------------------------------ backward data ----
// dX = conv_transpose(dY, W) [Flipout: conv_transpose(dY, mu) + conv_transpose(dY * s_out, sigma*eps) * s_in] runs on
// the SAME kernel as the forward: a stride-1 convolution of grad_out (zeros inserted between its pixels when the
// layer is strided) with the transposed, tap-reversed filter, whose sampled tile the producers rebuild from the
// regenerated eps.  The geometry below is that convolution's.
static bool dgrad_geom(const bt_geom* g, bt_geom* t, bool* inserted, const char** why) {
#define DG_NO(cond, msg) if (cond) { *why = msg; return false; }
  DG_NO(g->rank > 2, "rank 3");
  DG_NO(g->groups != 1, "groups != 1");
  DG_NO(g->n_samples != 1, "n_samples != 1");
  DG_NO(g->x_dtype != BT_BF16 || g->out_dtype != BT_BF16, "activations / gradients are not bf16");
  memset(t, 0, sizeof(*t));
  t->rank = g->rank; t->groups = 1; t->n_samples = 1; t->x_dtype = BT_BF16; t->out_dtype = BT_BF16;
  t->batch = g->batch; t->c_in = g->c_out; t->c_out = g->c_in;
  *inserted = false;
  for (int i = 0; i < g->rank; ++i) {
    const int64_t pd = g->dil[i] * (g->k_size[i] - 1) - g->pad[i];
    DG_NO(pd < 0, "padding larger than the filter reach");
    t->k_size[i] = g->k_size[i]; t->stride[i] = 1; t->dil[i] = g->dil[i]; t->pad[i] = pd;
    t->out_size[i] = g->in_size[i];
    t->in_size[i] = g->in_size[i] + 2 * g->pad[i] - g->dil[i] * (g->k_size[i] - 1);  // == out_size when stride 1
    if (g->stride[i] != 1) *inserted = true;
    DG_NO((g->out_size[i] - 1) * g->stride[i] + 1 > t->in_size[i], "inconsistent conv arithmetic");
  }
  // grad_x has x's strides (the "out" of this convolution)
  t->out_batch_stride = g->x_batch_stride; t->out_chan_stride = g->x_chan_stride;
  for (int i = 0; i < g->rank; ++i) t->out_sp_stride[i] = g->x_sp_stride[i];
  if (!*inserted) {  // grad_out is read in place
    t->x_batch_stride = g->out_batch_stride; t->x_chan_stride = g->out_chan_stride;
    for (int i = 0; i < g->rank; ++i) t->x_sp_stride[i] = g->out_sp_stride[i];
  } else {           // dense channels-last copy with zeros inserted
    DG_NO(g->out_chan_stride != 1 || g->c_out % 8 != 0, "strided layer: grad_out must be channels-last with c_out % 8 == 0");
    for (int i = 0; i < g->rank; ++i) DG_NO(g->out_sp_stride[i] % 8 != 0, "strided layer: grad_out strides must be 16-byte aligned");
    DG_NO(g->out_batch_stride % 8 != 0, "strided layer: grad_out strides must be 16-byte aligned");
    int64_t st = g->c_out;
    t->x_chan_stride = 1;
    for (int i = g->rank - 1; i >= 0; --i) { t->x_sp_stride[i] = st; st *= t->in_size[i]; }
    t->x_batch_stride = st;
  }
  return true;
#undef DG_NO
}

static bt_layer_params dgrad_params(const bt_layer_params* p) {
  bt_layer_params q = *p;
  // (a ConvTranspose forward keeps its bias: the rank-1 shift MMA adds it per output channel of this run)
  if (!(p->flags & BT_FLAG_TRANSPOSE_BIAS)) q.mu_b = q.rho_b = q.eps_b = nullptr;
  q.out_scale = q.out_shift = nullptr; q.out_relu = 0; q.residual = nullptr;
  q.sign_in = p->sign_out;   // the A operand (grad_out) is flipped by s_out, the result by s_in
  q.sign_out = p->sign_in;
  return q;
}

bool bt_tc_dgrad_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why) {
  bt_geom t;
  bool inserted;
  if (!dgrad_geom(g, &t, &inserted, why)) return false;
  const bt_layer_params q = dgrad_params(p);
  Plan pl;
  if (!tc2_plan(&t, &q, flipout, &pl, why, inserted)) return false;
  if (pl.es != 2) { *why = "fp32 gradients run the fp32 kernels"; return false; }
  if (inserted && pl.mode == A_PATCH && flipout) { *why = "strided Flipout layer through the patch matrix"; return false; }
  return true;
}

static int64_t dgrad_insert_bytes(const bt_geom* t) { return ((t->batch * t->x_batch_stride * 2 + 255) / 256) * 256; }

int64_t bt_tc_dgrad_workspace_bytes(const bt_geom* g, bool flipout) {
  bt_geom t;
  bool inserted;
  const char* why = "";
  if (!dgrad_geom(g, &t, &inserted, &why)) return 0;
  return bt_tc_workspace_bytes(&t, flipout) + (inserted ? dgrad_insert_bytes(&t) : 0);
}

int bt_tc_dgrad(const bt_geom* g, const bt_layer_params* p, const void* grad_out, void* grad_x, void* workspace, int64_t ws_bytes,
                bool flipout, cudaStream_t stream) {
  bt_geom t;
  bool inserted;
  const char* why = "";
  BT_REQUIRE(dgrad_geom(g, &t, &inserted, &why), BT_ERR_UNSUPPORTED, "tensor-core dgrad: %s", why);
  BT_REQUIRE(workspace, BT_ERR_INVALID, "tensor-core dgrad: workspace required (bt_workspace_bytes)");
  const bt_layer_params q = dgrad_params(p);
  RunOpt opt;
  opt.dgrad = true;
  opt.cin_w = (int)g->c_in; opt.cout_w = (int)g->c_out;
  const void* a_src = grad_out;
  void* ws = workspace;
  if (inserted) {
    const int r = g->rank;
    const int64_t OH = r == 2 ? g->out_size[0] : 1, OW = g->out_size[r - 1];
    const int64_t HZ = r == 2 ? t.in_size[0] : 1, WZ = t.in_size[r - 1];
    const int sh = r == 2 ? (int)g->stride[0] : 1, sw = (int)g->stride[r - 1];
    const int64_t ohs = r == 2 ? g->out_sp_stride[0] : 0, ows = g->out_sp_stride[r - 1];
    const int64_t bytes = dgrad_insert_bytes(&t);
    BT_REQUIRE(((uintptr_t)grad_out & 15) == 0, BT_ERR_ALIGN, "tensor-core dgrad: grad_out must be 16-byte aligned");
    BT_CUDA_OK(cudaMemsetAsync(workspace, 0, (size_t)(t.batch * t.x_batch_stride * 2), stream));
    const int C8 = (int)(g->c_out / 8);
    const int64_t chunks = g->batch * OH * OW * C8;
    int64_t blocks = (chunks + 255) / 256;
    const int64_t cap = (int64_t)bt_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    tc2_zero_insert_kernel<<<(int)blocks, 256, 0, stream>>>((const __nv_bfloat16*)grad_out, (__nv_bfloat16*)workspace, chunks, C8, (int)OH,
                                                            (int)OW, sh, sw, g->out_batch_stride, ohs, ows, (int)HZ, (int)WZ, (int)g->c_out);
    BT_LAUNCH_OK("tc2_zero_insert_kernel");
    a_src = workspace;
    ws = (char*)workspace + bytes;
    if (ws_bytes > 0) {
      BT_REQUIRE(ws_bytes > bytes, BT_ERR_INVALID, "tensor-core dgrad: workspace of %lld bytes cannot hold the zero-inserted grad_out (%lld)",
                 (long long)ws_bytes, (long long)bytes);
      ws_bytes -= bytes;
    }
    opt.a_dil_h = sh; opt.a_dil_w = sw; opt.SH = (int)OH; opt.SW = (int)OW;
    opt.s_bs = g->out_batch_stride; opt.s_hs = ohs; opt.s_ws = ows;
  }
  opt.ws_bytes = ws_bytes;
  return tc2_run(&t, &q, a_src, grad_x, ws, flipout, opt, stream);
}

// =====================================================================================================
// Backward weights on tensor cores.
//
//   dW[n][k] = sum_pixels dY[pixel][n] * A[pixel][k],  k = (tap, channel)        (Flipout: a second product
//   dDelta[n][k] = sum (dY * s_out)[pixel][n] * (A(x * s_in))[pixel][k])
//
// The reduction runs over PIXELS, the slow axis of both channels-last operands, so both MMA operands are
// MN-major: a TMA box of 64 pixels x 64 channels lands in shared memory as 64 rows of 128 bytes
// (128-byte swizzle) and is, as it lies, the canonical MN-major SWIZZLE_128B atom of a tcgen05 descriptor
// (8 k-rows of 128 bytes = 1024 bytes = SBO; the next 64 channels one box = 8 KB = LBO further on).  No
// transposition anywhere.  D is [128 k-rows x NT output channels] fp32 in TMEM; each CTA reduces one slice
// of the pixels for one (k-tile, n-tile) and adds it into a packed fp32 accumulator [n][k] (rows of D are
// consecutive k, so a warp's 32 lanes hit 32 consecutive floats).  Flipout flips both operand tiles in
// shared memory exactly like the forward flips its A tile (one thread per 128-byte row, 64 sign bits).
namespace {

constexpr int WG_THREADS = 384;       // warp 0 TMA, warp 1 MMA, warps 4-7 epilogue, warps 8-11 flippers
constexpr int WG_KP = 64;             // pixels per pipeline stage
constexpr int WG_UNIT = WG_KP * 128;  // one box: [64 pixels][64 channels] bf16 = 8 KB

struct WgArgs {
  float *dW, *dD;        // packed accumulators [N][Kp] (dD: Flipout perturbation product)
  int N, Kp, Kvalid;     // output channels; pitch and true length of the k axis
  int64_t P;             // pixels = reduction length
  int64_t steps_total;   // ceil(P / 64)
  int U;                 // 64-wide units of the k axis
  int NT, n_tiles, m_tiles, splits, stages;
  int conv, CB, OH, OW, KW, sh, sw, ph, pw, dh, dw, IH, IW, Cin;
  int a2_tma;            // Flipout: the flipped k-side tiles come from a second matrix by TMA
  uint64_t seed;
  uint32_t sample_id, layer_id;
};

struct WgCtrl {
  uint64_t full[4], flipped[4], empty[4], acc_full;
  uint32_t tmem_base;
};

// MN-major operand, 128-byte swizzle: LBO = bytes between consecutive 64-element groups along M/N,
// SBO = bytes between consecutive groups of 8 k-rows.
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

template <bool FLIP>
__global__ void __launch_bounds__(WG_THREADS, 1)
tc_wgrad_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                const __grid_constant__ CUtensorMap tm_a2, const __grid_constant__ WgArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  WgCtrl* ctrl = reinterpret_cast<WgCtrl*>(smem_raw + (base - raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int NB = a.NT / 64;                       // dY boxes per stage
  const uint32_t A_BYTES = 2u * WG_UNIT, B_BYTES = (uint32_t)NB * WG_UNIT;
  const uint32_t stage_bytes = (A_BYTES + B_BYTES) * (FLIP ? 2u : 1u);
  const uint32_t stage0 = base + 1024u;
  // stage layout: [A: 2 boxes][B: NB boxes] and, for Flipout, the flipped twins [A'][B'] behind them

  int bid = blockIdx.x;
  const int split = bid % a.splits; bid /= a.splits;
  const int nt = bid % a.n_tiles;
  const int mt = bid / a.n_tiles;
  const int64_t step0 = a.steps_total * split / a.splits, step1 = a.steps_total * (split + 1) / a.splits;
  const int n0 = nt * a.NT;
  const int tmem_cols = a.NT * (FLIP ? 2 : 1) < 32 ? 32 : a.NT * (FLIP ? 2 : 1);

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.stages; ++i) {
      mbar_init(smem_u32(&ctrl->full[i]), 1);
      mbar_init(smem_u32(&ctrl->flipped[i]), 4);
      mbar_init(smem_u32(&ctrl->empty[i]), 1);
    }
    mbar_init(smem_u32(&ctrl->acc_full), 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) { prefetch_tmap(&tm_a); prefetch_tmap(&tm_b); if (FLIP && a.a2_tma) prefetch_tmap(&tm_a2); }
  if (warp == 1) tmem_alloc(smem_u32(&ctrl->tmem_base), (uint32_t)tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = ctrl->tmem_base;

  // k-side units of this tile: u = 2*mt + {0, 1}; im2col: unit -> (tap, 64-channel block)
  const int u0 = 2 * mt;
  const int nunits = min(2, a.U - u0);

  if (warp == 0) {
    // =========================== TMA loader ===========================
    const bool leader = elect_one();
    int stage = 0;
    uint32_t phase = 0;
    // image coordinates of the first pixel of the step (im2col mode), advanced incrementally
    int64_t m = step0 * WG_KP;
    int ow = 0, oh = 0, b = 0;
    if (a.conv) {
      const int64_t r = m / a.OW;
      ow = (int)(m - r * a.OW);
      b = (int)(r / a.OH);
      oh = (int)(r - (int64_t)b * a.OH);
    }
    int offw[2] = {0, 0}, offh[2] = {0, 0}, c0[2] = {0, 0};
    for (int uu = 0; uu < 2; ++uu) {
      const int u = u0 + uu;
      if (a.conv) {
        const int tap = u / a.CB, cb = u - tap * a.CB;
        const int kh = tap / a.KW, kw = tap - kh * a.KW;
        offw[uu] = kw * a.dw; offh[uu] = kh * a.dh; c0[uu] = cb * 64;
      } else {
        c0[uu] = u * 64;
      }
    }
    int nb_live = 0;
    for (int j = 0; j < NB; ++j) nb_live += (n0 + 64 * j < a.N) ? 1 : 0;
    const uint32_t tx = (uint32_t)(nunits * (FLIP && a.a2_tma ? 2 : 1) + nb_live) * WG_UNIT;
    for (int64_t step = step0; step < step1; ++step) {
      mbar_wait(smem_u32(&ctrl->empty[stage]), phase ^ 1u);
      if (leader) {
        const uint32_t bar = smem_u32(&ctrl->full[stage]);
        const uint32_t sA = stage0 + (uint32_t)stage * stage_bytes, sB = sA + A_BYTES;
        mbar_expect_tx(bar, tx);
        for (int uu = 0; uu < nunits; ++uu) {
          if (a.conv) {
            tma_load_im2col_4d(sA + uu * WG_UNIT, &tm_a, bar, c0[uu], ow * a.sw - a.pw, oh * a.sh - a.ph, b, (uint16_t)offw[uu], (uint16_t)offh[uu]);
          } else {
            tma_load_2d(sA + uu * WG_UNIT, &tm_a, bar, c0[uu], (int)m);
            if (FLIP && a.a2_tma) tma_load_2d(sA + A_BYTES + B_BYTES + uu * WG_UNIT, &tm_a2, bar, c0[uu], (int)m);
          }
        }
        for (int j = 0; j < NB; ++j)
          if (n0 + 64 * j < a.N) tma_load_2d(sB + j * WG_UNIT, &tm_b, bar, n0 + 64 * j, (int)m);
      }
      m += WG_KP;
      if (a.conv) {
        ow += WG_KP;
        while (ow >= a.OW) { ow -= a.OW; if (++oh == a.OH) { oh = 0; ++b; } }
      }
      if (++stage == a.stages) { stage = 0; phase ^= 1u; }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    const bool leader = elect_one();
    // D = A^T-tile [128 x 16 pixels] * B-tile [16 pixels x NT]: both operands MN-major
    const uint32_t idesc = umma_idesc_bf16(128, (uint32_t)a.NT) | (1u << 15) | (1u << 16);
    int stage = 0;
    uint32_t phase = 0;
    for (int64_t step = step0; step < step1; ++step) {
      mbar_wait(smem_u32(&ctrl->full[stage]), phase);
      if constexpr (FLIP) mbar_wait(smem_u32(&ctrl->flipped[stage]), phase);
      tc_fence_after();
      if (leader) {
        const uint32_t sA = stage0 + (uint32_t)stage * stage_bytes, sB = sA + A_BYTES;
#pragma unroll
        for (int k = 0; k < WG_KP / 16; ++k) {
          const uint64_t ad = umma_desc_mn_sw128(sA + k * 2048, WG_UNIT, 1024);
          const uint64_t bd = umma_desc_mn_sw128(sB + k * 2048, WG_UNIT, 1024);
          umma_bf16(tmem, ad, bd, idesc, (step > step0 || k > 0) ? 1u : 0u);
        }
        if constexpr (FLIP) {
          const uint32_t sA2 = sB + B_BYTES, sB2 = sA2 + A_BYTES;
#pragma unroll
          for (int k = 0; k < WG_KP / 16; ++k) {
            const uint64_t ad = umma_desc_mn_sw128(sA2 + k * 2048, WG_UNIT, 1024);
            const uint64_t bd = umma_desc_mn_sw128(sB2 + k * 2048, WG_UNIT, 1024);
            umma_bf16(tmem + (uint32_t)a.NT, ad, bd, idesc, (step > step0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(smem_u32(&ctrl->empty[stage]));
      }
      __syncwarp();
      if (++stage == a.stages) { stage = 0; phase ^= 1u; }
    }
    if (leader) umma_commit(smem_u32(&ctrl->acc_full));
    __syncwarp();
  } else if (warp >= 4 && warp < 8) {
    // =========================== epilogue: TMEM -> packed fp32 accumulator ===========================
    const int q = warp - 4;
    const int r = q * 32 + lane;                       // row of D = k within the tile
    const int k = (u0 + (r >> 6)) * 64 + (r & 63);
    const bool kvalid = (r >> 6) < nunits && k < a.Kvalid;
    if (step1 > step0) {
      mbar_wait(smem_u32(&ctrl->acc_full), 0);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < a.NT; c += 32) {
        if (n0 + c >= a.N) break;
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
        tmem_ld_wait();
        if (kvalid) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (n0 + c + j < a.N) atomicAdd(a.dW + (int64_t)(n0 + c + j) * a.Kp + k, __uint_as_float(v[j]));
        }
        if constexpr (FLIP) {
          tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(a.NT + c), v);
          tmem_ld_wait();
          if (kvalid) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (n0 + c + j < a.N) atomicAdd(a.dD + (int64_t)(n0 + c + j) * a.Kp + k, __uint_as_float(v[j]));
          }
        }
      }
    }
  } else if (FLIP && warp >= 8) {
    // =========================== flippers ===========================
    const int f = threadIdx.x - 256;
    const BtRngKey key_in = bt_rng_key(a.seed, a.sample_id, a.layer_id, BT_KIND_SIGN_IN);
    const BtRngKey key_out = bt_rng_key(a.seed, a.sample_id, a.layer_id, BT_KIND_SIGN_OUT);
    int stage = 0;
    uint32_t phase = 0;
    const int rows_total = 128 + NB * 64;
    for (int64_t step = step0; step < step1; ++step) {
      const uint32_t sA = stage0 + (uint32_t)stage * stage_bytes;
      const uint32_t twin = A_BYTES + B_BYTES;        // distance from a tile to its flipped copy
      mbar_wait(smem_u32(&ctrl->full[stage]), phase);
      for (int rr = f; rr < rows_total; rr += 128) {
        const int64_t pix = step * WG_KP + (rr & 63);
        uint32_t w0 = 0, w1 = 0;
        bool live;
        uint32_t src;
        if (rr < 128) {  // k-side row: 64 channels of one input pixel of one tap
          const int uu = rr >> 6, u = u0 + uu;
          live = uu < nunits && !a.a2_tma;
          src = sA + (uint32_t)rr * 128u;
          if (live && pix < a.P) {
            uint64_t idx;
            bool inb = true;
            if (a.conv) {
              const int tap = u / a.CB, cb = u - tap * a.CB;
              const int kh = tap / a.KW, kw = tap - kh * a.KW;
              const int64_t r2 = pix / a.OW;
              const int owp = (int)(pix - r2 * a.OW);
              const int bp = (int)(r2 / a.OH);
              const int ohp = (int)(r2 - (int64_t)bp * a.OH);
              const int ih = ohp * a.sh - a.ph + kh * a.dh, iw = owp * a.sw - a.pw + kw * a.dw;
              inb = (unsigned)ih < (unsigned)a.IH && (unsigned)iw < (unsigned)a.IW;
              idx = (((uint64_t)bp * (uint64_t)a.IH + (uint64_t)ih) * (uint64_t)a.IW + (uint64_t)iw) * (uint64_t)a.Cin + (uint64_t)(cb * 64);
            } else {
              idx = (uint64_t)pix * (uint64_t)a.Cin + (uint64_t)(u * 64);
            }
            if (inb) {
              const uint4 rr4 = bt_philox(idx >> 7, key_in);
              const bool hi = ((idx >> 6) & 1) != 0;
              w0 = hi ? rr4.z : rr4.x; w1 = hi ? rr4.w : rr4.y;
            }
          }
        } else {         // dY-side row: 64 output channels of one output pixel
          const int j = (rr - 128) >> 6;
          live = n0 + 64 * j < a.N;
          src = sA + A_BYTES + (uint32_t)(rr - 128) * 128u;
          if (live && pix < a.P) {
            const uint64_t idx = (uint64_t)pix * (uint64_t)a.N + (uint64_t)(n0 + 64 * j);
            const uint4 rr4 = bt_philox(idx >> 7, key_out);
            const bool hi = ((idx >> 6) & 1) != 0;
            w0 = hi ? rr4.z : rr4.x; w1 = hi ? rr4.w : rr4.y;
          }
        }
        if (live) {
          const uint32_t row_sw = (uint32_t)(rr & 7);
          uint32_t d[8][4];
#pragma unroll
          for (int c = 0; c < 8; ++c)
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(d[c][0]), "=r"(d[c][1]), "=r"(d[c][2]), "=r"(d[c][3])
                         : "r"(src + (((uint32_t)c ^ row_sw) << 4)));
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const uint32_t wsel = c < 4 ? w0 : w1;
#pragma unroll
            for (int i = 0; i < 4; ++i) d[c][i] ^= (wsel << (15 - ((4 * c + i) & 15))) & 0x80008000u;
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(src + twin + (((uint32_t)c ^ row_sw) << 4)),
                         "r"(d[c][0]), "r"(d[c][1]), "r"(d[c][2]), "r"(d[c][3])
                         : "memory");
          }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&ctrl->flipped[stage]));
      if (++stage == a.stages) { stage = 0; phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, (uint32_t)tmem_cols);
}

// packed accumulators -> grad_mu / grad_rho in the parameter's own layout [n][c][tap], plus the KL gradient
// (the tensor-core twin of param_grad_finalize_kernel in bt_simt_wgrad.cu)
template <bool FLIP>
__global__ void tc_wgrad_finalize_kernel(const float* __restrict__ dW, const float* __restrict__ dD, float* __restrict__ gmu,
                                         float* __restrict__ grho, const float* __restrict__ mu, const float* __restrict__ rho,
                                         const float* __restrict__ eps, int64_t numel, int T, int Cin, int Kp, BtRngKey key,
                                         const float* __restrict__ grad_kl, const float* __restrict__ pmu_t,
                                         const float* __restrict__ psig_t, float prior_mu, float prior_sigma) {
  const float gk = grad_kl ? __ldg(grad_kl) / (float)numel : 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / ((int64_t)Cin * T);
    const int64_t rem = i - n * (int64_t)Cin * T;
    const int64_t c = rem / T, tap = rem - c * T;
    const int64_t k = tap * Cin + c;
    const float r = rho[i];
    const float sig = bt_softplus(r);
    const float dsig = r > 20.f ? 1.f : bt_sigmoid(r);
    const float e = eps ? eps[i] : bt_normal1((uint64_t)(n * (int64_t)T * Cin + k), key);
    const float dw = dW[n * Kp + k];
    const float dd = FLIP ? dD[n * Kp + k] : dw;
    float om = dw, orr = dd * e * dsig;
    if (grad_kl) {
      const float pm = pmu_t ? pmu_t[i] : prior_mu;
      const float ps = psig_t ? psig_t[i] : prior_sigma;
      const float inv2 = 1.f / (ps * ps);
      om += gk * (mu[i] - pm) * inv2;
      orr += gk * (sig * inv2 - 1.f / sig) * dsig;
    }
    gmu[i] = om;
    grho[i] = orr;
  }
}

}  // namespace

static bool wgrad_plan(const bt_geom* g, const bt_layer_params* p, bool flipout, Plan* pl, const char** why) {
#define WG_NO(cond, msg) if (cond) { *why = msg; return false; }
  WG_NO(g->n_samples != 1, "n_samples != 1");
  WG_NO(g->x_dtype != BT_BF16 || g->out_dtype != BT_BF16, "activations / gradients are not bf16");
  WG_NO(flipout && (p->sign_in || p->sign_out), "tensor-fed signs (deterministic mode runs the fp32 kernel)");
  WG_NO(flipout && g->c_out % 64 != 0, "Flipout needs c_out % 64 == 0");
  // the forward plan checks everything about x and about grad_out's layout (dense channels-last [pixels][c_out])
  WG_NO(!tc2_plan(g, p, flipout, pl, why), *why);
  WG_NO(pl->rows_per_s >= (1ll << 31) - 64, "too many output positions");
  return true;
#undef WG_NO
}

bool bt_tc_wgrad_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why) {
  Plan pl;
  return wgrad_plan(g, p, flipout, &pl, why);
}

// workspace: two packed fp32 accumulators [c_out][Kp] + whatever the patch-matrix mode needs
int64_t bt_tc_wgrad_workspace_bytes(const bt_geom* g, bool flipout) {
  int64_t taps = 1;
  for (int i = 0; i < g->rank; ++i) taps *= g->k_size[i];
  const int64_t Kp = (g->c_in * taps + 7) & ~7ll;
  const int64_t acc = ((g->c_out * Kp * 4 + 255) / 256) * 256;
  return 2 * acc + bt_tc_workspace_bytes(g, flipout);
}

int bt_tc_wgrad(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out, const float* grad_kl,
                float* grad_mu_w, float* grad_rho_w, void* workspace, int64_t ws_bytes, bool flipout, cudaStream_t stream) {
  const char* why = "";
  Plan pl;
  BT_REQUIRE(wgrad_plan(g, p, flipout, &pl, &why), BT_ERR_UNSUPPORTED, "tensor-core wgrad: %s", why);
  BT_REQUIRE(workspace && grad_mu_w && grad_rho_w, BT_ERR_INVALID, "tensor-core wgrad: null buffer");
  BT_REQUIRE(((uintptr_t)grad_out & 15) == 0 && ((uintptr_t)workspace & 15) == 0 && (pl.mode == A_PATCH || ((uintptr_t)x & 15) == 0),
             BT_ERR_ALIGN, "tensor-core wgrad: x / grad_out / workspace must be 16-byte aligned");
  const DriverFns& drv = driver_fns();
  WgArgs a;
  memset(&a, 0, sizeof(a));
  a.N = (int)g->c_out;
  a.Kvalid = pl.K;
  a.Kp = pl.Kp;
  a.P = pl.rows_per_s;
  a.steps_total = (a.P + WG_KP - 1) / WG_KP;
  a.Cin = (int)g->c_in;
  a.seed = p->seed; a.sample_id = p->sample_id; a.layer_id = p->layer_id;
  const int64_t acc_bytes = (((int64_t)a.N * a.Kp * 4 + 255) / 256) * 256;
  a.dW = (float*)workspace;
  a.dD = (float*)((char*)workspace + acc_bytes);
  void* rest = (char*)workspace + 2 * acc_bytes;
  BT_REQUIRE(ws_bytes <= 0 || ws_bytes >= bt_tc_wgrad_workspace_bytes(g, flipout), BT_ERR_INVALID,
             "tensor-core wgrad: workspace of %lld bytes, %lld needed (bt_workspace_bytes)", (long long)ws_bytes,
             (long long)bt_tc_wgrad_workspace_bytes(g, flipout));
  BT_CUDA_OK(cudaMemsetAsync(workspace, 0, (size_t)(flipout ? 2 : 1) * acc_bytes, stream));

  alignas(64) CUtensorMap tm_a, tm_b, tm_a2;
  const void* a_src = x;
  int64_t a_pitch = pl.pitch;
  int a_cols = pl.K;
  if (pl.mode == A_PATCH) {
    // patches sit behind the packed-parameter area the forward would use; wgrad has no use for that area
    __nv_bfloat16 *pm = nullptr, *pm2 = nullptr;
    const int rc = build_patch_matrices(g, pl, p, x, flipout, BT_KIND_SIGN_IN, rest, &pm, &pm2, stream);
    if (rc != BT_OK) return rc;
    a_src = pm;
    a_pitch = pl.Kp;
    a_cols = pl.Kp;
    if (pm2) {
      a.a2_tma = 1;
      const cuuint64_t gdim[2] = {(cuuint64_t)a_cols, (cuuint64_t)a.P};
      const cuuint64_t gstr[1] = {(cuuint64_t)a_pitch * 2};
      const cuuint32_t box[2] = {64, WG_KP};
      const cuuint32_t estr[2] = {1, 1};
      const CUresult r2 = drv.tiled(&tm_a2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, pm2, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      BT_REQUIRE(r2 == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(patches of x * s_in) failed with %d", (int)r2);
    }
  }
  if (pl.mode == A_GEMM || pl.mode == A_PATCH) {
    a.conv = 0;
    const cuuint64_t gdim[2] = {(cuuint64_t)a_cols, (cuuint64_t)a.P};
    const cuuint64_t gstr[1] = {(cuuint64_t)a_pitch * 2};
    const cuuint32_t box[2] = {64, WG_KP};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = drv.tiled(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(a_src), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(x) failed with %d", (int)r);
    a.U = (a_cols + 63) / 64;
  } else {
    const Dense& d = pl.d;
    a.conv = 1;
    a.CB = (int)g->c_in / 64;
    a.OH = (int)d.out[0]; a.OW = (int)d.out[1]; a.KW = (int)d.k[1];
    a.sh = (int)d.st[0]; a.sw = (int)d.st[1]; a.ph = (int)d.pd[0]; a.pw = (int)d.pd[1]; a.dh = (int)d.dl[0]; a.dw = (int)d.dl[1];
    a.IH = (int)d.in[0]; a.IW = (int)d.in[1];
    a.U = pl.T * a.CB;
    const cuuint64_t gdim[4] = {(cuuint64_t)g->c_in, (cuuint64_t)d.in[1], (cuuint64_t)d.in[0], (cuuint64_t)g->batch};
    const cuuint64_t gstr[3] = {(cuuint64_t)pl.xw * 2, (cuuint64_t)pl.xh * 2, (cuuint64_t)g->x_batch_stride * 2};
    const int lower[2] = {-(int)d.pd[1], -(int)d.pd[0]};
    const int upper[2] = {(int)(d.pd[1] - (d.k[1] - 1) * d.dl[1]), (int)(d.pd[0] - (d.k[0] - 1) * d.dl[0])};
    const cuuint32_t estr[4] = {1, (cuuint32_t)d.st[1], (cuuint32_t)d.st[0], 1};
    const CUresult r = drv.im2col(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), gdim, gstr, lower, upper, 64, WG_KP,
                                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed with %d", (int)r);
    const int64_t span = (int64_t)g->batch * g->x_batch_stride * 2;
    if (drv.driver_version <= 13010 && span < 131072) reinterpret_cast<uint64_t*>(&tm_a)[1] &= ~(1ull << 21);
  }
  if (!a.a2_tma) tm_a2 = tm_a;
  {
    const cuuint64_t gdim[2] = {(cuuint64_t)a.N, (cuuint64_t)a.P};
    const cuuint64_t gstr[1] = {(cuuint64_t)a.N * 2};
    const cuuint32_t box[2] = {64, WG_KP};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = drv.tiled(&tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(grad_out), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(grad_out) failed with %d", (int)r);
  }
  a.NT = a.N <= 64 ? 64 : (a.N <= 128 ? 128 : 256);
  if (env_int("BT_WG_NT", 0) == 64 || env_int("BT_WG_NT", 0) == 128) a.NT = env_int("BT_WG_NT", 0);
  a.n_tiles = (a.N + a.NT - 1) / a.NT;
  a.m_tiles = (a.U + 1) / 2;
  const int sms = bt_num_sms();
  const int64_t tiles = (int64_t)a.n_tiles * a.m_tiles;
  // ONE wave of CTAs (one CTA per SM: 133 KB of shared memory), never a CTA more: the first version aimed at two CTAs' worth of
  // work per SM and rounded the split count up -- 9 tiles x 33 = 297 CTAs = two waves and a third with a single CTA (SMs active
  // 63 % of the kernel).  Whole waves: ResNet-18 Flipout training step 19.23 -> 18.31 ms; one wave instead of two (half the
  // fp32 atomics into the packed accumulators): 17.93 (profiles/r2_exp28_wgrad_waves.txt).
  const int waves = env_int("BT_WG_WAVES", 1);
  int64_t splits = env_int("BT_WG_ROUND_UP", 0) ? (waves * sms + tiles - 1) / tiles : (waves * sms) / tiles;
  if (splits > a.steps_total) splits = a.steps_total;
  if (splits < 1) splits = 1;
  if (env_int("BT_WG_SPLITS", 0) > 0) splits = env_int("BT_WG_SPLITS", 0) < a.steps_total ? env_int("BT_WG_SPLITS", 0) : a.steps_total;
  a.splits = (int)splits;
  const int stage_bytes = (2 * WG_UNIT + (a.NT / 64) * WG_UNIT) * (flipout ? 2 : 1);
  a.stages = (SMEM_LIMIT - 2048) / stage_bytes;
  if (a.stages > 4) a.stages = 4;
  if (env_int("BT_WG_STAGES", 0) >= 2 && env_int("BT_WG_STAGES", 0) < a.stages) a.stages = env_int("BT_WG_STAGES", 0);
  BT_REQUIRE(a.stages >= 2, BT_ERR_UNSUPPORTED, "tensor-core wgrad: no room for two pipeline stages");
  const int smem_bytes = 2048 + a.stages * stage_bytes;
  const int64_t grid = tiles * splits;
  BT_REQUIRE(grid < (1ll << 31), BT_ERR_UNSUPPORTED, "tensor-core wgrad: grid too large");
  if (env_int("BT_TC2_LOG", 0))
    fprintf(stderr, "wgrad: flip=%d mode=%d P=%lld N=%d K=%d(%d) U=%d NT=%d tiles=%dx%d splits=%d stages=%d smem=%d\n", (int)flipout,
            (int)pl.mode, (long long)a.P, a.N, a.Kvalid, a.Kp, a.U, a.NT, a.m_tiles, a.n_tiles, a.splits, a.stages, smem_bytes);
  if (a.P > 0) {
    if (flipout) {
      BT_CUDA_OK(cudaFuncSetAttribute(tc_wgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
      tc_wgrad_kernel<true><<<(unsigned)grid, WG_THREADS, smem_bytes, stream>>>(tm_a, tm_b, tm_a2, a);
    } else {
      BT_CUDA_OK(cudaFuncSetAttribute(tc_wgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
      tc_wgrad_kernel<false><<<(unsigned)grid, WG_THREADS, smem_bytes, stream>>>(tm_a, tm_b, tm_a2, a);
    }
    BT_LAUNCH_OK("tc_wgrad_kernel");
  }
  const int64_t wn = (int64_t)a.N * pl.K;
  int64_t fgrid = (wn + 255) / 256;
  if (fgrid > (int64_t)sms * 8) fgrid = (int64_t)sms * 8;
  if (fgrid < 1) fgrid = 1;
  const BtRngKey key = bt_rng_key(p->seed, p->sample_id, p->layer_id, BT_KIND_EPS_W);
  if (flipout)
    tc_wgrad_finalize_kernel<true><<<(int)fgrid, 256, 0, stream>>>(a.dW, a.dD, grad_mu_w, grad_rho_w, p->mu_w, p->rho_w, p->eps_w, wn, pl.T,
                                                                   (int)g->c_in, a.Kp, key, grad_kl, p->prior_mu_w_t, p->prior_sigma_w_t,
                                                                   p->prior_mu, p->prior_sigma);
  else
    tc_wgrad_finalize_kernel<false><<<(int)fgrid, 256, 0, stream>>>(a.dW, a.dD, grad_mu_w, grad_rho_w, p->mu_w, p->rho_w, p->eps_w, wn, pl.T,
                                                                    (int)g->c_in, a.Kp, key, grad_kl, p->prior_mu_w_t, p->prior_sigma_w_t,
                                                                    p->prior_mu, p->prior_sigma);
  BT_LAUNCH_OK("tc_wgrad_finalize_kernel");
  return BT_OK;
}
