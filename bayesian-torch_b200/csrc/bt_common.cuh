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

#define BT_LAUNCH_OK(name)                                                      \
  do {                                                                          \
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
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    c1 = (uint32_t)p1;
    c3 = (uint32_t)p0;
    c0 = n0;
    c2 = n2;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

// Box-Muller on two 32-bit words: u = ((r >> 8) + 0.5) * 2^-24 in (0,1).
__device__ __forceinline__ float2 bt_box_muller(uint32_t r0, uint32_t r1) {
  const float u0 = ((float)(r0 >> 8) + 0.5f) * 5.9604644775390625e-08f;
  const float u1 = ((float)(r1 >> 8) + 0.5f) * 5.9604644775390625e-08f;
  const float rad = sqrtf(-2.f * __logf(u0));
  float s, c;
  __sincosf(6.2831853071795864f * u1, &s, &c);
  return make_float2(rad * c, rad * s);
}

// Four N(0,1) values for elements 4*block .. 4*block+3 of the stream.
__device__ __forceinline__ float4 bt_normal4(uint64_t block, const BtRngKey& key) {
  const uint4 r = bt_philox(block, key);
  const float2 a = bt_box_muller(r.x, r.y);
  const float2 b = bt_box_muller(r.z, r.w);
  return make_float4(a.x, a.y, b.x, b.y);
}

// One N(0,1) value for stream element idx (wasteful: 1 of 4 lanes used; SIMT path only).
__device__ __forceinline__ float bt_normal1(uint64_t idx, const BtRngKey& key) {
  const float4 z = bt_normal4(idx >> 2, key);
  const int l = (int)(idx & 3);
  return l == 0 ? z.x : (l == 1 ? z.y : (l == 2 ? z.z : z.w));
}

// Rademacher sign of stream element idx: bit (idx % 128) of the 128-bit block idx / 128; 1 -> -1.
__device__ __forceinline__ float bt_sign1(uint64_t idx, const BtRngKey& key) {
  const uint4 r = bt_philox(idx >> 7, key);
  const int w = (int)((idx >> 5) & 3);
  const uint32_t word = w == 0 ? r.x : (w == 1 ? r.y : (w == 2 ? r.z : r.w));
  return ((word >> (idx & 31)) & 1u) ? -1.f : 1.f;
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
int bt_tc_forward(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, void* workspace,
                  bool flipout, cudaStream_t stream);
int64_t bt_tc_workspace_bytes(const bt_geom* g);
