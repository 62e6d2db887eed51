// Shared host/device helpers for the bt_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/bt_b200.h"

// ---------------------------------------------------------------- errors ----
void bt_set_error(const char* fmt, ...);

#define BT_REQUIRE(cond, code, ...)   \
  do {                                \
    if (!(cond)) {                    \
      bt_set_error(__VA_ARGS__);      \
      return (code);                  \
    }                                 \
  } while (0)

#define BT_CUDA_OK(expr)                                                            \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      bt_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                   __LINE__);                                                       \
      return BT_ERR_CUDA;                                                           \
    }                                                                               \
  } while (0)

void bt_count_launch();  // process-wide tally of this library's kernel launches (bt_launch_count)

#define BT_LAUNCH_OK(name)                                                      \
  do {                                                                          \
    bt_count_launch();                                                          \
    cudaError_t _e = cudaGetLastError();                                        \
    if (_e != cudaSuccess) {                                                    \
      bt_set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));    \
      return BT_ERR_CUDA;                                                       \
    }                                                                           \
  } while (0)

int bt_num_sms();  // cached cudaDevAttrMultiProcessorCount of the current device (148 on B200)

// ------------------------------------------------------------------ math ----
// softplus as the reference writes it, log1p(exp(rho)) (linear_variational.py:163), but
// overflow-safe: for rho > 20, log1p(exp(rho)) == rho to fp32 precision.
__device__ __forceinline__ float bt_softplus(float rho) {
  return rho > 20.f ? rho : log1pf(expf(rho));
}
__device__ __forceinline__ float bt_sigmoid(float x) { return 1.f / (1.f + expf(-x)); }

// ---------------------------------------------------------- Philox4x32-10 ----
// RNG contract (restated in oracle/philox.py): key = (seed_lo, seed_hi),
// counter = (block_lo, block_hi, sample_id, stream_id), stream_id = layer_id<<4 | kind.
enum { BT_KIND_EPS_W = 0, BT_KIND_EPS_B = 1, BT_KIND_SIGN_IN = 2, BT_KIND_SIGN_OUT = 3 };

struct BtRngKey {
  uint32_t k0, k1;     // seed
  uint32_t sample_id;  // counter word 2
  uint32_t stream;     // counter word 3
};

__host__ __device__ __forceinline__ BtRngKey bt_rng_key(uint64_t seed, uint32_t sample_id, uint32_t layer_id,
                                                       int kind) {
  BtRngKey k;
  k.k0 = (uint32_t)seed;
  k.k1 = (uint32_t)(seed >> 32);
  k.sample_id = sample_id;
  k.stream = (layer_id << 4) | (uint32_t)kind;
  return k;
}

__device__ __forceinline__ uint4 bt_philox(uint64_t block, const BtRngKey& key) {
  uint32_t c0 = (uint32_t)block, c1 = (uint32_t)(block >> 32), c2 = key.sample_id, c3 = key.stream;
  uint32_t k0 = key.k0, k1 = key.k1;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
#ifdef BT_PHILOX_WIDE
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2, p0 = (uint64_t)0xD2511F53u * c0;  // IMAD.WIDE: hi and lo in one issue
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1;
    c3 = (uint32_t)p0;
#else
    const uint32_t n0 = __umulhi(0xCD9E8D57u, c2) ^ c1 ^ k0;
    const uint32_t n2 = __umulhi(0xD2511F53u, c0) ^ c3 ^ k1;
    c1 = 0xCD9E8D57u * c2;
    c3 = 0xD2511F53u * c0;
#endif
    c0 = n0;
    c2 = n2;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

// Box-Muller on ONE 32-bit word: u0 from the high 16 bits sets the radius, u1 from the low
// 16 bits the angle; u = (h + 0.5) * 2^-16 in (0,1), built without an int->float conversion
// as 1.m - 1.  Radius tops out at sqrt(2 ln 2^17) = 4.85 sigma, ample for bf16 weights, and
// one Philox4x32 call now yields EIGHT normals (the sampling cost bounds the fused GEMM).
//   z0 = r cos(2 pi (u1 - 1/2)),  z1 = r sin(2 pi (u1 - 1/2)),  r = sqrt(-2 ln u0)
__device__ __forceinline__ float bt_sqrt_approx(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float2 bt_box_muller16(uint32_t w) {
  const float x0 = __uint_as_float(0x3F800040u | ((w >> 9) & 0x007FFF80u));  // 1 + u0
  const float x1 = __uint_as_float(0x3F800040u | ((w << 7) & 0x007FFF80u));  // 1 + u1
  const float rad = bt_sqrt_approx(-1.3862943611198906f * __log2f(x0 - 1.0f));
  float s, c;
  __sincosf(fmaf(x1, 6.2831853071795864f, -9.4247779607693797f), &s, &c);  // 2 pi (u1 - 1/2)
  return make_float2(rad * c, rad * s);
}

// Eight N(0,1) values for elements 8*block .. 8*block+7 of the stream.
__device__ __forceinline__ void bt_normal8(uint64_t block, const BtRngKey& key, float (&z)[8]) {
  const uint4 r = bt_philox(block, key);
  const float2 a = bt_box_muller16(r.x), b = bt_box_muller16(r.y), c = bt_box_muller16(r.z), d = bt_box_muller16(r.w);
  z[0] = a.x; z[1] = a.y; z[2] = b.x; z[3] = b.y; z[4] = c.x; z[5] = c.y; z[6] = d.x; z[7] = d.y;
}

// Four N(0,1) values for elements 4*quad .. 4*quad+3 (half a Philox block; the generic
// tensor-core kernel stages float4 groups).
__device__ __forceinline__ float4 bt_normal4(uint64_t quad, const BtRngKey& key) {
  const uint4 r = bt_philox(quad >> 1, key);
  const float2 a = bt_box_muller16((quad & 1) ? r.z : r.x), b = bt_box_muller16((quad & 1) ? r.w : r.y);
  return make_float4(a.x, a.y, b.x, b.y);
}

// One N(0,1) value for stream element idx (wasteful: 1 of 8 lanes used; SIMT path only).
__device__ __forceinline__ float bt_normal1(uint64_t idx, const BtRngKey& key) {
  const uint4 r = bt_philox(idx >> 3, key);
  const int wsel = (int)((idx >> 1) & 3);
  const uint32_t w = wsel == 0 ? r.x : (wsel == 1 ? r.y : (wsel == 2 ? r.z : r.w));
  const float2 z = bt_box_muller16(w);
  return (idx & 1) ? z.y : z.x;
}

// Rademacher sign of stream element idx: block idx / 128 -> four 32-bit words; element j = idx % 32 of
// word (idx / 32) % 4 reads bit (j / 2) + 16 * (j % 2); 1 -> -1.  With this layout the sign bits of the
// bf16 PAIR (2p, 2p+1) are bits p and p+16 of the word: (word << (15 - p)) & 0x80008000 is the XOR mask
// that flips both halves of the packed pair (the tensor-core Flipout path flips activations that way).
__device__ __forceinline__ int bt_sign_bitpos(uint32_t j) { return (int)((j >> 1) + ((j & 1u) << 4)); }
__device__ __forceinline__ float bt_sign1(uint64_t idx, const BtRngKey& key) {
  const uint4 r = bt_philox(idx >> 7, key);
  const int w = (int)((idx >> 5) & 3);
  const uint32_t word = w == 0 ? r.x : (w == 1 ? r.y : (w == 2 ? r.z : r.w));
  return ((word >> bt_sign_bitpos((uint32_t)idx & 31u)) & 1u) ? -1.f : 1.f;
}
// Position of element (b, c, d, h, w) of a logical [B, C, D, H, W] activation tensor in the sign
// streams: CHANNELS-LAST order, whatever the tensor's memory layout -- the 64 channels of one pixel
// that make a row of a tensor-core A tile are then 64 consecutive stream elements (two words of one
// Philox block), and so are 32 consecutive output channels of one accumulator row.
__device__ __forceinline__ uint64_t bt_sign_index(uint64_t b, int64_t C, int64_t c, int D, int d, int H, int h, int W,
                                                  int w) {
  return ((((b * (uint64_t)D + (uint64_t)d) * (uint64_t)H + (uint64_t)h) * (uint64_t)W + (uint64_t)w) * (uint64_t)C) + (uint64_t)c;
}

// --------------------------------------------------------------- dtypes ----
template <typename T>
__device__ __forceinline__ float bt_to_float(T v);
template <>
__device__ __forceinline__ float bt_to_float<float>(float v) { return v; }
template <>
__device__ __forceinline__ float bt_to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
__device__ __forceinline__ T bt_from_float(float v);
template <>
__device__ __forceinline__ float bt_from_float<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 bt_from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float bt_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double bt_warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Internal launchers shared between translation units (defined in bt_kl.cu / bt_simt.cu / bt_tc.cu).
int bt_kl_layer(const bt_layer_params* p, int64_t w_numel, int64_t b_numel, float* kl_out, void* workspace,
                cudaStream_t stream);
int bt_simt_forward(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, bool flipout,
                    cudaStream_t stream);
int bt_simt_backward(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                     const float* grad_kl, const bt_layer_grads* grads, void* workspace, bool flipout,
                     cudaStream_t stream);
bool bt_tc_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why);

// kl_out non-null: forward(return_kl=True) -- the KL term is evaluated by the parameter-packing pass (kl_scratch: >= 32
// bytes of the caller's workspace)
int bt_tc_forward(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, void* workspace, int64_t ws_bytes,
                  bool flipout, float* kl_out, void* kl_scratch, cudaStream_t stream);
int64_t bt_tc_workspace_bytes(const bt_geom* g, bool flipout);
// backward data on the tensor-core kernel (grad_out convolved with the transposed, regenerated weights)
bool bt_tc_dgrad_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why);
int64_t bt_tc_dgrad_workspace_bytes(const bt_geom* g, bool flipout);
// backward weights on tensor cores (MN-major operands, split over pixels, packed fp32 accumulators in the workspace)
bool bt_tc_wgrad_supported(const bt_geom* g, const bt_layer_params* p, bool flipout, const char** why);
int64_t bt_tc_wgrad_workspace_bytes(const bt_geom* g, bool flipout);
int bt_tc_wgrad(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out, const float* grad_kl,
                float* grad_mu_w, float* grad_rho_w, void* workspace, int64_t ws_bytes, bool flipout, cudaStream_t stream);
int bt_tc_dgrad(const bt_geom* g, const bt_layer_params* p, const void* grad_out, void* grad_x, void* workspace, int64_t ws_bytes,
                bool flipout, cudaStream_t stream);
