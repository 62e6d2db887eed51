// TMA-fed tcgen05 / TMEM implicit GEMM with the weight tile SAMPLED IN SHARED MEMORY (sm_100a).
//
//   out[s][m][n] = sum_k A(x_s)[m][k] * bf16(mu[n][k] + sigma[n][k] * eps_s[n][k]) + (mu_b + sigma_b * eps_b)[n]
//
// replaces, for bf16 channels-last activations, the reference's per-call sequence
// log1p(exp(rho)) -> normal_() -> sigma*eps -> mu+tmp -> F.linear / F.conv2d
// (linear_variational.py:163-181, conv_variational.py:366-385).  The sampled weight never
// exists in global memory.
//
// Persistent warp-specialised CTA, one per SM:
//   warps 0-3   epilogue : tcgen05.ld -> +sampled bias -> bf16/f32 -> swizzled smem -> TMA store
//   warp  4     A loader : ONE thread issues TMA loads; convolutions use the im2col tensor-map
//                          mode (cuTensorMapEncodeIm2col), so the 128-pixel x 64-channel tile of one
//                          filter tap lands in the 128B-swizzled K-major layout tcgen05.mma reads,
//                          padding zero-filled by the hardware; Linear uses a tiled 2-d map
//   warps 5-6   MMA      : two issuer warps (tile t of an m-group -> issuer t % 2), each a
//                          warp-uniform loop whose elected lane issues tcgen05.mma (M=128, N=BN,
//                          K=16) into TMEM accumulators.  Measured on B200 (tools/ubench/pipe2.cu):
//                          one issuing thread costs ~340 cycles per 4-MMA ring slot against 256
//                          cycles of tensor work at N=128, two issuers overlap completely
//   warps 7-14  W producers: mu + sigma * eps -> bf16 -> swizzled smem, eps from in-register
//                          Philox4x32-10 (8 normals per call) or the caller's tensor
// mbarrier pipelines: A ring (TMA complete_tx / tcgen05.commit), W ring or stationary panel,
// accumulator sets in TMEM.  The sampling cost is amortised over the batch either by keeping
// the whole [BN x K] sampled panel in shared memory (stationary) or by keeping MT accumulators
// resident in TMEM so one generated k-block feeds MT m-tiles (streaming).
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "bt_common.cuh"
#include "bt_tc.cuh"

using namespace btc;

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_SLOT = BM * BK * 2;  // 16 KB
constexpr int CTRL_BYTES = 2048;
constexpr int STAGE_BYTES = 4 * 2 * 4096;  // per epilogue warp: 2 buffers of 32 rows x 128 B (1 when smem is tight)
constexpr int N_PROD_WARPS = 8;
constexpr int N_MMA_WARPS = 2;  // issuer warps: tile t of an m-group belongs to issuer t % 2
constexpr int WARP_TMA = 4, WARP_MMA = 5, WARP_PROD0 = WARP_MMA + N_MMA_WARPS;
constexpr int NUM_THREADS = (WARP_PROD0 + N_PROD_WARPS) * 32;  // 480
constexpr int PROD_THREADS = N_PROD_WARPS * 32;
constexpr int MAX_NA = 8, MAX_RW = 24, MAX_SETS = 8, MAX_MT = 4;
constexpr int SMEM_LIMIT = 232448;
constexpr int RING_BYTES = SMEM_LIMIT - 1024 - CTRL_BYTES - STAGE_BYTES;  // A ring + W ring / panel

struct Tc2Args {
  const float *mu_p, *sig_p, *eps_p;  // packed [N][K], k = tap*Cin + c
  const float *mu_b, *rho_b, *eps_b;
  int64_t total_items;
  int N, K, S, NB;
  int Ms;            // GEMM rows per sample: batch * OH * OW
  int x_rows_per_s;  // images (conv) / rows (linear) per sample in x; 0 when all samples share x
  int conv;          // 1: im2col map, 0: tiled 2-d map
  int OH, OW, KW, CB;  // CB = Cin / 64: k-blocks per filter tap
  int sh, sw, ph, pw, dh, dw;
  int num_mt, num_mg, MT, NSETS, nkb, NA, RW, stationary;
  int NR, NAr;       // A rings (one per active MMA issuer) and slots per ring; NA = NR * NAr
  int stage_bufs;    // epilogue staging buffers per warp (2, or 1 when shared memory is tight)
  int debug;         // BT_TC2_DEBUG bits (timing experiments only): 1 = skip the output stores
  uint64_t seed;
  uint32_t sample_id, layer_id;
};

struct Ctrl {
  uint64_t a_full[MAX_NA], a_empty[MAX_NA];
  uint64_t w_full[MAX_RW], w_empty[MAX_RW];
  uint64_t acc_full[MAX_SETS], acc_empty[MAX_SETS];
  uint32_t tmem_base;
  uint32_t pad[3];
  float bias[256];
};
static_assert(sizeof(Ctrl) <= CTRL_BYTES, "control block too large");

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
__device__ __forceinline__ bool elect_one() {  // one lane of the (converged) warp
  uint32_t is_leader;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(is_leader));
  return is_leader != 0;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

template <int BN, typename TO>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc2_fwd_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_out,
               const __grid_constant__ Tc2Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  Ctrl* ctrl = reinterpret_cast<Ctrl*>(smem_raw + (base - raw));
  constexpr int W_SLOT = BN * BK * 2;
  constexpr int SUBN = 128 / (int)sizeof(TO);  // output columns per 128-byte staging row
  const uint32_t stage_base = base + CTRL_BYTES;
  const uint32_t a_base = stage_base + (uint32_t)a.stage_bufs * 16384u;
  const uint32_t w_base = a_base + a.NA * A_SLOT;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < a.NA; ++i) { mbar_init(smem_u32(&ctrl->a_full[i]), 1); mbar_init(smem_u32(&ctrl->a_empty[i]), 1); }
    for (int i = 0; i < a.RW; ++i) { mbar_init(smem_u32(&ctrl->w_full[i]), N_PROD_WARPS); mbar_init(smem_u32(&ctrl->w_empty[i]), N_MMA_WARPS); }
    for (int i = 0; i < a.NSETS; ++i) { mbar_init(smem_u32(&ctrl->acc_full[i]), N_MMA_WARPS); mbar_init(smem_u32(&ctrl->acc_empty[i]), 128); }
    mbar_fence_init();
  }
  if (warp == WARP_TMA && lane == 0) { prefetch_tmap(&tm_a); prefetch_tmap(&tm_out); }
  if (warp == WARP_MMA) tmem_alloc(smem_u32(&ctrl->tmem_base), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = ctrl->tmem_base;

  const int64_t i0 = a.total_items * blockIdx.x / gridDim.x;
  const int64_t i1 = a.total_items * (blockIdx.x + 1) / gridDim.x;
  const int n_items = (int)(i1 - i0);
  // first item of this CTA, decoded once; the role loops then walk (s, nb, mg) incrementally --
  // the single-thread TMA / MMA loops are instruction-latency bound, so no division or modulo
  // by a run-time value may sit inside them.
  const int64_t key0 = i0 / a.num_mg;
  const int mg0 = (int)(i0 - key0 * a.num_mg);
  const int s0 = (int)(key0 / a.NB), nb0 = (int)(key0 - (int64_t)s0 * a.NB);

  if (warp < 4) {
    // =========================== EPILOGUE ===========================
    const int q = warp;
    const uint32_t my_stage = stage_base + (uint32_t)(q * a.stage_bufs) * 4096u;
    const uint32_t row_sw = (uint32_t)(lane & 7);
    uint32_t nstore = 0;
    int s = s0, nb = nb0, mg = mg0, set = 0;
    uint32_t set_phase = 0;
    for (int it = 0; it < n_items; ++it) {
      if (a.mu_b) {
        named_bar_sync(1, 128);  // previous item's readers of ctrl->bias are done
        for (int j = threadIdx.x; j < BN; j += 128) {
          const int n = nb * BN + j;
          float b = 0.f;
          if (n < a.N) {
            const float e = a.eps_b ? a.eps_b[n]
                                    : bt_normal1((uint64_t)n, bt_rng_key(a.seed, a.sample_id + s, a.layer_id, BT_KIND_EPS_B));
            b = a.mu_b[n] + bt_softplus(a.rho_b[n]) * e;
          }
          ctrl->bias[j] = b;
        }
        named_bar_sync(1, 128);
      }
      mbar_wait(smem_u32(&ctrl->acc_full[set]), set_phase);
      tc_fence_after();
      const int mtv = min(a.MT, a.num_mt - mg * a.MT);
      for (int t = 0; t < mtv; ++t) {
        const int m0 = (mg * a.MT + t) * BM + q * 32;  // first row of this warp's 32-row slab
        if (m0 >= a.Ms) break;                          // slab entirely past the sample's last row (warp-uniform)
        const uint32_t tacc = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)((set * a.MT + t) * BN);
#pragma unroll 1
        for (int sub = 0; sub < BN / SUBN; ++sub) {
          const int n0 = nb * BN + sub * SUBN;
          if (n0 >= a.N) break;  // warp-uniform
          uint32_t v[SUBN];
          if constexpr (SUBN == 64) {
            tmem_ld32(tacc + sub * SUBN, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
            tmem_ld32(tacc + sub * SUBN + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
          } else {
            tmem_ld32(tacc + sub * SUBN, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          }
          tmem_ld_wait();
          // the TMA store that last read this staging buffer must have drained it
          uint32_t buf = my_stage;
          if (a.stage_bufs == 2) {
            buf += (nstore & 1u) * 4096u;
            if (nstore >= 2) {
              if (lane == 0) tma_store_wait_read<1>();
              __syncwarp();
            }
          } else if (nstore >= 1) {
            if (lane == 0) tma_store_wait_read<0>();
            __syncwarp();
          }
          const uint32_t rowaddr = buf + (uint32_t)lane * 128u;
#pragma unroll
          for (int j = 0; j < 8; ++j) {  // 16-byte chunks of the 128-byte row
            uint32_t w0, w1, w2, w3;
            if constexpr (SUBN == 64) {
              const float* bs = &ctrl->bias[sub * SUBN + j * 8];
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[j * 8 + e]) + (a.mu_b ? bs[e] : 0.f);
              w0 = pack_bf16(f[0], f[1]); w1 = pack_bf16(f[2], f[3]); w2 = pack_bf16(f[4], f[5]); w3 = pack_bf16(f[6], f[7]);
            } else {
              const float* bs = &ctrl->bias[sub * SUBN + j * 4];
              w0 = __float_as_uint(__uint_as_float(v[j * 4 + 0]) + (a.mu_b ? bs[0] : 0.f));
              w1 = __float_as_uint(__uint_as_float(v[j * 4 + 1]) + (a.mu_b ? bs[1] : 0.f));
              w2 = __float_as_uint(__uint_as_float(v[j * 4 + 2]) + (a.mu_b ? bs[2] : 0.f));
              w3 = __float_as_uint(__uint_as_float(v[j * 4 + 3]) + (a.mu_b ? bs[3] : 0.f));
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + (((uint32_t)j ^ row_sw) << 4)), "r"(w0),
                         "r"(w1), "r"(w2), "r"(w3)
                         : "memory");
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0 && !(a.debug & 1)) {
            tma_store_3d(&tm_out, buf, n0, m0, s);  // rows >= Ms and columns >= N are clipped by the map
            tma_store_commit();
          }
          ++nstore;
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&ctrl->acc_empty[set]));
      if (++set == a.NSETS) { set = 0; set_phase ^= 1u; }
      if (++mg == a.num_mg) { mg = 0; if (++nb == a.NB) { nb = 0; ++s; } }
    }
    if (lane == 0) tma_store_wait_all();
  } else if (warp == WARP_TMA) {
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
    const int Cin = a.CB * BK;
    for (int it = 0; it < n_items; ++it) {
      const int mtv = min(a.MT, a.num_mt - mg * a.MT);
      int cw[MAX_MT], chh[MAX_MT], cn[MAX_MT];
#pragma unroll
      for (int t = 0; t < MAX_MT; ++t) {
        const int m = (mg * a.MT + t) * BM;
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
          if (t < mtv) {
            const int r = t % N_MMA_WARPS;
            const int ring = a.NR == 2 ? r : 0;
            const int slot = ring * a.NAr + (a.NR == 2 ? rslot[r] : rslot[0]);
            mbar_wait(smem_u32(&ctrl->a_empty[slot]), (a.NR == 2 ? rphase[r] : rphase[0]) ^ 1u);
            if (leader) {
              const uint32_t bar = smem_u32(&ctrl->a_full[slot]);
              if (a.debug & 8) {
                mbar_arrive(bar);
              } else {
                mbar_expect_tx(bar, A_SLOT);
                if (a.conv)
                  tma_load_im2col_4d(a_base + slot * A_SLOT, &tm_a, bar, c0, cw[t], chh[t], cn[t], (uint16_t)offw, (uint16_t)offh);
                else
                  tma_load_2d(a_base + slot * A_SLOT, &tm_a, bar, c0, cn[t]);
              }
            }
            if (a.NR == 2) { if (++rslot[r] == a.NAr) { rslot[r] = 0; rphase[r] ^= 1u; } }
            else if (++rslot[0] == a.NAr) { rslot[0] = 0; rphase[0] ^= 1u; }
          }
        }
        c0 += BK;
        if (a.conv && c0 == Cin) {  // next filter tap
          c0 = 0;
          offw += a.dw;
          if (++kw == a.KW) { kw = 0; offw = 0; offh += a.dh; }
        }
      }
      if (++mg == a.num_mg) { mg = 0; if (++nb == a.NB) { nb = 0; ++s; } }
    }
    __syncwarp();
  } else if (warp < WARP_PROD0) {
    // =========================== MMA ISSUERS ===========================
    const int j = warp - WARP_MMA;  // this issuer owns tiles t with t % N_MMA_WARPS == j
    const bool leader = elect_one();
    constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
    int mg = mg0, set = 0, aslot = 0, wslot_r = 0;
    uint32_t set_phase = 0, aphase = 0, wphase = 0;
    bool first = true;  // first item of a (sample, n-block) key on this CTA: its W panel is new
    const int ring0 = (a.NR == 2 ? j : 0) * a.NAr;  // first slot of the A ring this issuer consumes
    for (int it = 0; it < n_items; ++it) {
      const bool last = !a.stationary || it + 1 == n_items || mg + 1 == a.num_mg;
      mbar_wait(smem_u32(&ctrl->acc_empty[set]), set_phase ^ 1u);
      tc_fence_after();
      const int mtv = min(a.MT, a.num_mt - mg * a.MT);
      const uint32_t dbase = tmem + (uint32_t)(set * a.MT * BN);
      for (int kb = 0; kb < a.nkb; ++kb) {
        const int wslot = a.stationary ? kb : wslot_r;
        if (first) {
          mbar_wait(smem_u32(&ctrl->w_full[wslot]), wphase);
          tc_fence_after();
        }
        const uint64_t bdesc = umma_desc_sw128(w_base + wslot * W_SLOT);
#pragma unroll
        for (int t = 0; t < MAX_MT; ++t) {
          if (t < mtv) {
            if ((a.NR == 2 ? (t % N_MMA_WARPS) : 0) == j) {
              const int slot = ring0 + aslot;
              mbar_wait(smem_u32(&ctrl->a_full[slot]), aphase);
              tc_fence_after();
              if (leader) {
                const uint64_t adesc = umma_desc_sw128(a_base + slot * A_SLOT);
                const uint32_t d = dbase + (uint32_t)(t * BN);
                if (!(a.debug & 4)) {
#pragma unroll
                  for (int k = 0; k < BK / 16; ++k)
                    umma_bf16(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
                }
                umma_commit(smem_u32(&ctrl->a_empty[slot]));
              }
              __syncwarp();
              if (++aslot == a.NAr) { aslot = 0; aphase ^= 1u; }
            }
          }
        }
        if (last && leader) umma_commit(smem_u32(&ctrl->w_empty[wslot]));
        if (!a.stationary && ++wslot_r == a.RW) { wslot_r = 0; wphase ^= 1u; }
      }
      if (leader) umma_commit(smem_u32(&ctrl->acc_full[set]));
      __syncwarp();
      if (++set == a.NSETS) { set = 0; set_phase ^= 1u; }
      if (a.stationary) {
        if (last) wphase ^= 1u;  // the next key's panel is the next generation of every slot
        first = last;
      }
      if (++mg == a.num_mg) mg = 0;
    }
  } else {
    // =========================== W PRODUCERS (sample in smem) ===========================
    const int p = threadIdx.x - WARP_PROD0 * 32;
    const int chunk = p & 7;      // 16-byte chunk (8 consecutive k) of the 128-byte row
    const int row_base = p >> 3;  // rows row_base + 32*i
    constexpr int ROWS_PER_THREAD = BN / (PROD_THREADS / 8);
    int s = s0, nb = nb0, mg = mg0, wslot_r = 0;
    uint32_t wphase = 0;
    bool first = true;
    for (int it = 0; it < n_items; ++it) {
      if (first || !a.stationary) {
        const BtRngKey rk = bt_rng_key(a.seed, a.sample_id + s, a.layer_id, BT_KIND_EPS_W);
        for (int kb = 0; kb < a.nkb; ++kb) {
          const int slot = a.stationary ? kb : wslot_r;
          const int k = kb * BK + chunk * 8;
          const bool kin = k < a.K;  // K % 8 == 0: a chunk is all in or all out
          const uint32_t wdst = w_base + slot * W_SLOT;
          float4 mu4[ROWS_PER_THREAD][2], sg4[ROWS_PER_THREAD][2];
#pragma unroll
          for (int i = 0; i < ROWS_PER_THREAD; ++i) {  // parameter loads first (L2 resident), then the slot wait
            const int n = nb * BN + row_base + (PROD_THREADS / 8) * i;
            mu4[i][0] = mu4[i][1] = sg4[i][0] = sg4[i][1] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (kin && n < a.N) {
              const int64_t e = (int64_t)n * a.K + k;
              mu4[i][0] = __ldg(reinterpret_cast<const float4*>(a.mu_p + e));
              mu4[i][1] = __ldg(reinterpret_cast<const float4*>(a.mu_p + e + 4));
              sg4[i][0] = __ldg(reinterpret_cast<const float4*>(a.sig_p + e));
              sg4[i][1] = __ldg(reinterpret_cast<const float4*>(a.sig_p + e + 4));
            }
          }
          mbar_wait(smem_u32(&ctrl->w_empty[slot]), wphase ^ 1u);
#pragma unroll
          for (int i = 0; i < ROWS_PER_THREAD; ++i) {
            const int row = row_base + (PROD_THREADS / 8) * i;
            const int n = nb * BN + row;
            float z[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) z[e] = 0.f;
            if (kin && n < a.N) {
              const int64_t e = (int64_t)n * a.K + k;
              if (a.eps_p) {
                const float4 e0 = __ldg(reinterpret_cast<const float4*>(a.eps_p + e));
                const float4 e1 = __ldg(reinterpret_cast<const float4*>(a.eps_p + e + 4));
                z[0] = e0.x; z[1] = e0.y; z[2] = e0.z; z[3] = e0.w; z[4] = e1.x; z[5] = e1.y; z[6] = e1.z; z[7] = e1.w;
              } else {
                bt_normal8((uint64_t)e >> 3, rk, z);
              }
            }
            const uint32_t w0 = pack_bf16(fmaf(sg4[i][0].x, z[0], mu4[i][0].x), fmaf(sg4[i][0].y, z[1], mu4[i][0].y));
            const uint32_t w1 = pack_bf16(fmaf(sg4[i][0].z, z[2], mu4[i][0].z), fmaf(sg4[i][0].w, z[3], mu4[i][0].w));
            const uint32_t w2 = pack_bf16(fmaf(sg4[i][1].x, z[4], mu4[i][1].x), fmaf(sg4[i][1].y, z[5], mu4[i][1].y));
            const uint32_t w3 = pack_bf16(fmaf(sg4[i][1].z, z[6], mu4[i][1].z), fmaf(sg4[i][1].w, z[7], mu4[i][1].w));
            const uint32_t addr = wdst + (uint32_t)row * 128u + (((uint32_t)chunk ^ (uint32_t)(row & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&ctrl->w_full[slot]));
          if (!a.stationary && ++wslot_r == a.RW) { wslot_r = 0; wphase ^= 1u; }
        }
        if (a.stationary) wphase ^= 1u;
      }
      first = false;
      if (++mg == a.num_mg) { mg = 0; first = true; if (++nb == a.NB) { nb = 0; ++s; } }
    }
  }

  tc_fence_before();
  __syncthreads();
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

bool bt_tc2_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why) {
#define TC_NO(cond, msg) if (cond) { *why = msg; return false; }
  TC_NO(flipout, "Flipout is not on the TMA path");
  TC_NO(g->groups != 1, "groups != 1");
  TC_NO(g->x_dtype != BT_BF16, "activations are not bf16");
  TC_NO(g->rank > 2, "rank 3");
  TC_NO(g->x_chan_stride != 1 || g->out_chan_stride != 1, "channels are not the fastest axis");
  TC_NO(g->c_out % 8 != 0 || g->c_in % 8 != 0, "c_in / c_out not multiples of 8");
  TC_NO(g->batch <= 0, "empty batch");
  TC_NO(g->n_samples > 1 && (p->eps_w || p->eps_b), "tensor-fed eps with n_samples > 1");
  TC_NO(!driver_fns().ok, "cuTensorMapEncode* entry points unavailable");
  const int64_t es_o = g->out_dtype == BT_F32 ? 4 : 2;
  if (g->rank == 0) {
    TC_NO(g->x_batch_stride % 8 != 0, "x row pitch not a multiple of 16 bytes");
    TC_NO(g->out_batch_stride != g->c_out, "output rows are not dense");
    TC_NO(g->n_samples > 1 && g->x_sample_stride != 0 && g->x_sample_stride != g->batch * g->x_batch_stride,
          "x samples are not stacked densely");
    TC_NO(g->n_samples > 1 && g->out_sample_stride != g->batch * g->c_out, "out samples are not stacked densely");
  } else {
    Dense d;
    int64_t xh, xw, oh, ow;
    conv_dims(g, &d, &xh, &xw, &oh, &ow);
    TC_NO(g->c_in % 64 != 0, "c_in is not a multiple of 64 (one 128-byte swizzle row per filter tap)");
    TC_NO(xw % 8 != 0 || xh % 8 != 0 || g->x_batch_stride % 8 != 0, "x strides not multiples of 16 bytes");
    TC_NO(ow != g->c_out || oh != d.out[1] * ow || g->out_batch_stride != d.out[0] * oh, "output is not dense channels-last");
    TC_NO(g->n_samples > 1 && g->x_sample_stride != 0 && g->x_sample_stride != g->batch * g->x_batch_stride,
          "x samples are not stacked densely");
    TC_NO(g->n_samples > 1 && g->out_sample_stride != g->batch * g->out_batch_stride, "out samples are not stacked densely");
    for (int i = 0; i < 2; ++i) {
      TC_NO(d.st[i] > 8, "stride > 8");
      TC_NO(d.pd[i] > 127 || (d.k[i] - 1) * d.dl[i] - d.pd[i] > 128, "padding / dilated extent outside the im2col box range");
      TC_NO((d.k[i] - 1) * d.dl[i] > 65535, "tap offset overflows 16 bits");
    }
    TC_NO(g->batch * d.out[0] * d.out[1] >= (1ll << 31) / 2, "too many output positions");
  }
  TC_NO((g->c_out * es_o) % 16 != 0, "output row pitch not a multiple of 16 bytes");
  (void)p;
#undef TC_NO
  return true;
}

int bt_tc_pack(const bt_layer_params* p, int64_t wn, int Cin, int T, float* mu_p, float* sig_p, float* eps_p,
               cudaStream_t stream);

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

template <int BN, typename TO>
static int tc2_launch(const CUtensorMap& ta, const CUtensorMap& to, const Tc2Args& a, int grid, int smem_bytes,
                      cudaStream_t stream) {
  BT_CUDA_OK(cudaFuncSetAttribute(tc2_fwd_kernel<BN, TO>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  tc2_fwd_kernel<BN, TO><<<grid, NUM_THREADS, smem_bytes, stream>>>(ta, to, a);
  BT_LAUNCH_OK("tc2_fwd_kernel");
  return BT_OK;
}

int bt_tc2_forward(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, void* workspace,
                   bool flipout, cudaStream_t stream) {
  const char* why = "";
  BT_REQUIRE(bt_tc2_supported(g, p, flipout, &why), BT_ERR_UNSUPPORTED, "TMA tensor-core forward: %s", why);
  BT_REQUIRE(workspace, BT_ERR_INVALID, "TMA tensor-core forward: workspace required (bt_workspace_bytes)");
  BT_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0 && ((uintptr_t)workspace & 15) == 0, BT_ERR_ALIGN,
             "TMA tensor-core forward: x / out / workspace must be 16-byte aligned");
  const DriverFns& drv = driver_fns();
  Tc2Args a;
  memset(&a, 0, sizeof(a));
  const int S = g->n_samples;
  const bool x_shared = S > 1 && g->x_sample_stride == 0;
  a.N = (int)g->c_out;
  a.S = S;
  a.mu_b = p->mu_b; a.rho_b = p->rho_b; a.eps_b = p->eps_b;
  a.seed = p->seed; a.sample_id = p->sample_id; a.layer_id = p->layer_id;
  alignas(64) CUtensorMap tm_a, tm_out;
  int T = 1;
  const int es_o = g->out_dtype == BT_F32 ? 4 : 2;
  a.debug = env_int("BT_TC2_DEBUG", 0);
  // A 1x1 / stride 1 / unpadded convolution over a dense channels-last tensor IS a GEMM over
  // [pixels, channels]: the plain tiled map moves it in larger bursts than the im2col walk.
  bool as_gemm = g->rank == 0;
  int64_t gemm_rows = g->batch, gemm_pitch = g->x_batch_stride;
  if (g->rank > 0 && !(a.debug & 2)) {
    Dense d;
    int64_t xh, xw, oh, ow;
    conv_dims(g, &d, &xh, &xw, &oh, &ow);
    if (d.k[0] == 1 && d.k[1] == 1 && d.st[0] == 1 && d.st[1] == 1 && d.pd[0] == 0 && d.pd[1] == 0 &&
        xh == d.in[1] * xw && g->x_batch_stride == d.in[0] * xh) {
      as_gemm = true;
      gemm_rows = g->batch * d.in[0] * d.in[1];
      gemm_pitch = xw;
    }
  }
  if (as_gemm) {
    a.conv = 0;
    a.K = (int)g->c_in;
    a.Ms = (int)gemm_rows;
    a.x_rows_per_s = x_shared ? 0 : (int)gemm_rows;
    const cuuint64_t rows = (cuuint64_t)gemm_rows * (x_shared ? 1 : S);
    const cuuint64_t gdim[2] = {(cuuint64_t)g->c_in, rows};
    const cuuint64_t gstr[1] = {(cuuint64_t)gemm_pitch * 2};
    const cuuint32_t box[2] = {BK, BM};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = drv.tiled(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(x), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeTiled(x) failed with %d", (int)r);
  } else {
    Dense d;
    int64_t xh, xw, oh, ow;
    conv_dims(g, &d, &xh, &xw, &oh, &ow);
    for (int i = 0; i < 2; ++i) {
      const int64_t expect = (d.in[i] + 2 * d.pd[i] - d.dl[i] * (d.k[i] - 1) - 1) / d.st[i] + 1;
      BT_REQUIRE(d.out[i] == expect && d.out[i] > 0, BT_ERR_INVALID, "TMA tensor-core forward: inconsistent out_size");
    }
    a.conv = 1;
    T = (int)(d.k[0] * d.k[1]);
    a.K = (int)g->c_in * T;
    a.CB = (int)g->c_in / BK;
    a.OH = (int)d.out[0]; a.OW = (int)d.out[1]; a.KW = (int)d.k[1];
    a.sh = (int)d.st[0]; a.sw = (int)d.st[1]; a.ph = (int)d.pd[0]; a.pw = (int)d.pd[1]; a.dh = (int)d.dl[0]; a.dw = (int)d.dl[1];
    a.Ms = (int)(g->batch * d.out[0] * d.out[1]);
    a.x_rows_per_s = x_shared ? 0 : (int)g->batch;
    const cuuint64_t imgs = (cuuint64_t)g->batch * (x_shared ? 1 : S);
    const cuuint64_t gdim[4] = {(cuuint64_t)g->c_in, (cuuint64_t)d.in[1], (cuuint64_t)d.in[0], imgs};
    const cuuint64_t gstr[3] = {(cuuint64_t)xw * 2, (cuuint64_t)xh * 2, (cuuint64_t)g->x_batch_stride * 2};
    // bounding box of the filter's top-left corner, in (W, H) order like the coordinates
    const int lower[2] = {-(int)d.pd[1], -(int)d.pd[0]};
    const int upper[2] = {(int)(d.pd[1] - (d.k[1] - 1) * d.dl[1]), (int)(d.pd[0] - (d.k[0] - 1) * d.dl[0])};
    const cuuint32_t estr[4] = {1, (cuuint32_t)d.st[1], (cuuint32_t)d.st[0], 1};
    const CUresult r = drv.im2col(&tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), gdim, gstr, lower, upper,
                                  BK, BM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    BT_REQUIRE(r == CUDA_SUCCESS, BT_ERR_CUDA, "cuTensorMapEncodeIm2col(x) failed with %d", (int)r);
    // Small-tensor workaround of drivers <= 13.1 (same fix-up CUTLASS applies to im2col descriptors).
    const int64_t span = (int64_t)imgs * g->x_batch_stride * 2;
    if (drv.driver_version <= 13010 && span < 131072) reinterpret_cast<uint64_t*>(&tm_a)[1] &= ~(1ull << 21);
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
  }

  // ---- pack (mu, softplus(rho), eps) into the kernel's k order
  const int64_t wn = (int64_t)a.N * a.K;
  float* mu_p = (float*)workspace;
  float* sig_p = mu_p + wn;
  float* eps_p = p->eps_w ? sig_p + wn : nullptr;
  int rc = bt_tc_pack(p, wn, (int)g->c_in, T, mu_p, sig_p, eps_p, stream);
  if (rc != BT_OK) return rc;
  a.mu_p = mu_p; a.sig_p = sig_p; a.eps_p = eps_p;

  // ---- tiling / pipeline configuration
  const int sms = bt_num_sms();
  a.num_mt = (a.Ms + BM - 1) / BM;
  a.nkb = (a.K + BK - 1) / BK;
  int BN = a.N <= 64 ? 64 : 128;
  // a [BN x K] panel that fits next to a 3-deep A ring is generated once per (sample, n-block)
  // (with single-buffered epilogue staging there are 16 KB more)
  auto panel_fits = [&](int bn) { return a.nkb <= MAX_RW && (int64_t)a.nkb * bn * BK * 2 + 4 * A_SLOT <= RING_BYTES + STAGE_BYTES / 2; };
  const int force_bn = env_int("BT_TC2_BN", 0), force_mt = env_int("BT_TC2_MT", 0), force_stat = env_int("BT_TC2_STATIONARY", -1);
  if (BN == 128 && !panel_fits(128) && panel_fits(64)) BN = 64;
  if (force_bn == 64 || force_bn == 128) BN = force_bn;
  const int WB = BN * BK * 2;
  a.NB = (a.N + BN - 1) / BN;
  a.stationary = panel_fits(BN) ? 1 : 0;
  if (force_stat == 0) a.stationary = 0;
  const int NSLOTS = 512 / BN;
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
    if (force_mt >= 1 && force_mt <= MAX_MT && force_mt <= NSLOTS) a.MT = force_mt;
    a.RW = (RING_BYTES - 4 * A_SLOT) / WB;
    if (a.RW > 8) a.RW = 8;
  }
  a.NSETS = NSLOTS / a.MT;
  if (a.NSETS > MAX_SETS) a.NSETS = MAX_SETS;
  a.num_mg = (a.num_mt + a.MT - 1) / a.MT;
  a.stage_bufs = 2;
  a.NA = (RING_BYTES - a.RW * WB) / A_SLOT;
  if (a.NA < 4) { a.stage_bufs = 1; a.NA = (RING_BYTES + STAGE_BYTES / 2 - a.RW * WB) / A_SLOT; }
  if (a.NA > MAX_NA) a.NA = MAX_NA;
  a.NR = a.MT >= 2 ? N_MMA_WARPS : 1;
  a.NAr = a.NA / a.NR;
  a.NA = a.NAr * a.NR;
  if (env_int("BT_TC2_NA", 0) >= 2 && env_int("BT_TC2_NA", 0) < a.NA) a.NA = env_int("BT_TC2_NA", 0);
  BT_REQUIRE(a.NA >= 2 && a.RW >= 1 && a.RW <= MAX_RW, BT_ERR_UNSUPPORTED, "TMA tensor-core forward: no room for the pipelines");
  a.total_items = (int64_t)a.S * a.NB * a.num_mg;
  const int grid = (int)(a.total_items < sms ? a.total_items : sms);
  const int smem_bytes = 1024 + CTRL_BYTES + a.stage_bufs * (STAGE_BYTES / 2) + a.NA * A_SLOT + a.RW * WB;
  if (env_int("BT_TC2_LOG", 0))
    fprintf(stderr, "tc2: conv=%d S=%d Ms=%d N=%d K=%d BN=%d stat=%d MT=%d NSETS=%d nkb=%d NA=%dx%d RW=%d NB=%d num_mg=%d items=%lld grid=%d smem=%d xrows=%d OH=%d OW=%d KW=%d CB=%d s=%d,%d p=%d,%d d=%d,%d x=%p out=%p\n",
            a.conv, a.S, a.Ms, a.N, a.K, BN, a.stationary, a.MT, a.NSETS, a.nkb, a.NR, a.NAr, a.RW, a.NB, a.num_mg,
            (long long)a.total_items, grid, smem_bytes, a.x_rows_per_s, a.OH, a.OW, a.KW, a.CB, a.sh, a.sw, a.ph, a.pw, a.dh, a.dw, x, out);
  const bool o32 = g->out_dtype == BT_F32;
  if (BN == 64)
    return o32 ? tc2_launch<64, float>(tm_a, tm_out, a, grid, smem_bytes, stream)
               : tc2_launch<64, __nv_bfloat16>(tm_a, tm_out, a, grid, smem_bytes, stream);
  return o32 ? tc2_launch<128, float>(tm_a, tm_out, a, grid, smem_bytes, stream)
             : tc2_launch<128, __nv_bfloat16>(tm_a, tm_out, a, grid, smem_bytes, stream);
}
