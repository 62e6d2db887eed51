// Small element-wise / reduction kernels around the layer path: RNG stream export,
// MOPED rho init, Monte-Carlo predictive reduction.
#include "bt_common.cuh"

namespace {

__global__ void philox_normal_fill_kernel(float* __restrict__ out, int64_t n, BtRngKey key) {
  const int64_t nblk = (n + 3) >> 2;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblk; b += (int64_t)gridDim.x * blockDim.x) {
    const float4 z = bt_normal4((uint64_t)b, key);
    const int64_t i = b << 2;
    if (i + 3 < n && ((uintptr_t)out & 15) == 0) {
      *reinterpret_cast<float4*>(out + i) = z;
    } else {
      const float v[4] = {z.x, z.y, z.z, z.w};
      for (int j = 0; j < 4 && i + j < n; ++j) out[i + j] = v[j];
    }
  }
}

__global__ void philox_sign_fill_kernel(float* __restrict__ out, int64_t n, BtRngKey key) {
  // one Philox call covers 128 elements; a thread writes the 32 elements of one word
  const int64_t nwords = (n + 31) >> 5;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += (int64_t)gridDim.x * blockDim.x) {
    const uint4 r = bt_philox((uint64_t)(w >> 2), key);
    const int q = (int)(w & 3);
    const uint32_t word = q == 0 ? r.x : (q == 1 ? r.y : (q == 2 ? r.z : r.w));
    const int64_t i0 = w << 5;
    for (int j = 0; j < 32 && i0 + j < n; ++j) out[i0 + j] = ((word >> bt_sign_bitpos((uint32_t)j)) & 1u) ? -1.f : 1.f;
  }
}

// rho = log(expm1(delta*|w|) + 1e-20)   (utils/util.py:63-69)
__global__ void moped_rho_kernel(const float* __restrict__ w, float* __restrict__ rho, int64_t n, float delta) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    rho[i] = logf(expm1f(delta * fabsf(w[i])) + 1e-20f);
}

// One block per batch row b of logits[S,B,C].  Phase 1: the warps split the samples and reduce each
// row's softmax max / denominator into shared memory.  Phase 2: every thread owns columns
// c = tid, tid + blockDim, ... and adds p, p^2 over the samples IN SAMPLE ORDER (no atomics:
// the [B,C] accumulators and the result are deterministic), plus its share of sum_c p log p.
constexpr int MC_THREADS = 256, MC_SCHUNK = 64;
template <typename T>
__global__ void __launch_bounds__(MC_THREADS) mc_accumulate_kernel(const T* __restrict__ logits, int64_t S, int64_t B, int64_t C, int softmax,
                                     float* __restrict__ sum_p, float* __restrict__ sum_p2,
                                     float* __restrict__ sum_plogp) {
  __shared__ float s_mx[MC_SCHUNK], s_inv[MC_SCHUNK], s_red[MC_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = MC_THREADS / 32;
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    float plogp = 0.f;
    for (int64_t sc = 0; sc < S; sc += MC_SCHUNK) {
      const int ns = (int)(S - sc < MC_SCHUNK ? S - sc : MC_SCHUNK);
      if (softmax) {
        for (int s = warp; s < ns; s += nw) {
          const T* row = logits + ((sc + s) * B + b) * C;
          float mx = -INFINITY;
          for (int64_t c = lane; c < C; c += 32) mx = fmaxf(mx, bt_to_float<T>(row[c]));
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
          float d = 0.f;
          for (int64_t c = lane; c < C; c += 32) d += expf(bt_to_float<T>(row[c]) - mx);
          d = bt_warp_sum(d);
          if (lane == 0) { s_mx[s] = mx; s_inv[s] = 1.f / d; }
        }
      }
      __syncthreads();
      for (int64_t c = threadIdx.x; c < C; c += MC_THREADS) {
        float ap = 0.f, ap2 = 0.f;
        for (int s = 0; s < ns; ++s) {
          const float v = bt_to_float<T>(logits[((sc + s) * B + b) * C + c]);
          const float pr = softmax ? expf(v - s_mx[s]) * s_inv[s] : v;
          ap += pr;
          ap2 += pr * pr;
          plogp += pr * logf(pr + 1e-15f);
        }
        if (sum_p) sum_p[b * C + c] += ap;
        if (sum_p2) sum_p2[b * C + c] += ap2;
      }
      __syncthreads();
    }
    if (sum_plogp) {  // fixed-order block reduction: warp shuffles, then warp 0 over the per-warp partials
      plogp = bt_warp_sum(plogp);
      if (lane == 0) s_red[warp] = plogp;
      __syncthreads();
      if (warp == 0) {
        float t = lane < nw ? s_red[lane] : 0.f;
        t = bt_warp_sum(t);
        if (lane == 0) sum_plogp[b] += t;
      }
      __syncthreads();
    }
  }
}

__global__ void mc_finalize_kernel(const float* __restrict__ sum_p, const float* __restrict__ sum_p2,
                                   const float* __restrict__ sum_plogp, float inv_s, int64_t B, int64_t C,
                                   float* __restrict__ mean, float* __restrict__ var, float* __restrict__ entropy,
                                   float* __restrict__ mi) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t b = warp; b < B; b += nwarps) {
    float h = 0.f;
    for (int64_t c = lane; c < C; c += 32) {
      const float m = sum_p[b * C + c] * inv_s;
      if (mean) mean[b * C + c] = m;
      if (var && sum_p2) var[b * C + c] = fmaxf(sum_p2[b * C + c] * inv_s - m * m, 0.f);
      h -= m * logf(m + 1e-15f);
    }
    h = bt_warp_sum(h);
    if (lane == 0) {
      if (entropy) entropy[b] = h;
      if (mi && sum_plogp) mi[b] = h + sum_plogp[b] * inv_s;
    }
  }
}

// y = relu(y + r), 16 bytes per thread per step.
template <typename T>
__global__ void add_relu_kernel(T* __restrict__ y, const T* __restrict__ r, int64_t n) {
  constexpr int V = 16 / (int)sizeof(T);
  const int64_t nv = n / V;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    uint4 a = reinterpret_cast<const uint4*>(y)[i];
    const uint4 b = __ldcs(reinterpret_cast<const uint4*>(r) + i);
    T* av = reinterpret_cast<T*>(&a);
    const T* bv = reinterpret_cast<const T*>(&b);
#pragma unroll
    for (int e = 0; e < V; ++e) av[e] = bt_from_float<T>(fmaxf(bt_to_float<T>(av[e]) + bt_to_float<T>(bv[e]), 0.f));
    reinterpret_cast<uint4*>(y)[i] = a;
  }
  for (int64_t i = nv * V + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    y[i] = bt_from_float<T>(fmaxf(bt_to_float<T>(y[i]) + bt_to_float<T>(r[i]), 0.f));
}

// Channels-last max pooling, 8 channels (16 bytes) per thread: the stem's MaxPool2d(3, 2, 1)
// sits between two Bayesian layers of every ResNet and torch's NHWC kernel runs it at ~0.7 TB/s.
// K = 3 (every ResNet): the nine taps are loaded first, all in flight at once, then reduced -- the generic loop below issues
// load -> max -> load and ran latency-bound at 48 % of the DRAM rate (profiles/r2_maxpool_pool_ncu_metrics.txt).
template <typename IDX>
__global__ void __launch_bounds__(256) maxpool2d_k3_nhwc_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                                     int64_t total64, int C8, int IH, int IW, int OH, int OW, int st, int pd) {
  const IDX total = (IDX)total64, stride = (IDX)gridDim.x * blockDim.x;
  const uint32_t ninf = 0xFF80FF80u;   // (-inf, -inf) in bf16
  for (IDX i = (IDX)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c8 = (int)(i % (IDX)C8);
    IDX r = i / (IDX)C8;
    const int ow = (int)(r % (IDX)OW); r /= (IDX)OW;
    const int oh = (int)(r % (IDX)OH); r /= (IDX)OH;
    const __nv_bfloat16* xb = x + (int64_t)r * IH * IW * C8 * 8 + c8 * 8;
    uint4 v[9];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const int ih = oh * st - pd + kh;
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int iw = ow * st - pd + kw;
        const bool ok = (unsigned)ih < (unsigned)IH && (unsigned)iw < (unsigned)IW;
        v[kh * 3 + kw] = ok ? __ldg(reinterpret_cast<const uint4*>(xb + ((int64_t)ih * IW + iw) * C8 * 8)) : make_uint4(ninf, ninf, ninf, ninf);
      }
    }
    __nv_bfloat162 m[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) m[e] = reinterpret_cast<const __nv_bfloat162*>(&v[0])[e];
#pragma unroll
    for (int t = 1; t < 9; ++t) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[t]);
#pragma unroll
      for (int e = 0; e < 4; ++e) m[e] = __hmax2_nan(m[e], h[e]);  // NaN propagates like torch's max_pool2d
    }
    reinterpret_cast<uint4*>(y)[i] = *reinterpret_cast<const uint4*>(m);
  }
}

template <typename IDX>
__global__ void __launch_bounds__(256) maxpool2d_nhwc_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                                  int64_t total64, int C8, int IH, int IW, int OH, int OW, int k,
                                                                  int st, int pd) {
  const IDX total = (IDX)total64, stride = (IDX)gridDim.x * blockDim.x;
  for (IDX i = (IDX)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c8 = (int)(i % (IDX)C8);
    IDX r = i / (IDX)C8;
    const int ow = (int)(r % (IDX)OW); r /= (IDX)OW;
    const int oh = (int)(r % (IDX)OH); r /= (IDX)OH;
    const __nv_bfloat16* xb = x + (int64_t)r * IH * IW * C8 * 8 + c8 * 8;
    const __nv_bfloat162 ninf = __floats2bfloat162_rn(-INFINITY, -INFINITY);
    __nv_bfloat162 m[4] = {ninf, ninf, ninf, ninf};  // max is exact in bf16: no widening needed
    for (int kh = 0; kh < k; ++kh) {
      const int ih = oh * st - pd + kh;
      if ((unsigned)ih >= (unsigned)IH) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int iw = ow * st - pd + kw;
        if ((unsigned)iw >= (unsigned)IW) continue;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(xb + ((int64_t)ih * IW + iw) * C8 * 8));
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
        for (int e = 0; e < 4; ++e) m[e] = __hmax2_nan(m[e], h[e]);  // NaN propagates like torch's max_pool2d
      }
    }
    reinterpret_cast<uint4*>(y)[i] = *reinterpret_cast<const uint4*>(m);
  }
}

int grid_for(int64_t work_items, int threads) {
  int64_t blocks = (work_items + threads - 1) / threads;
  const int64_t cap = (int64_t)bt_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

}  // namespace

extern "C" int bt_philox_normal_fill(float* out, int64_t n, uint64_t seed, uint32_t sample_id, uint32_t layer_id,
                                     int kind, bt_stream_t stream) {
  BT_REQUIRE(out && n >= 0 && kind >= 0 && kind < 16, BT_ERR_INVALID, "bt_philox_normal_fill: bad argument");
  if (n == 0) return BT_OK;
  philox_normal_fill_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, (cudaStream_t)stream>>>(
      out, n, bt_rng_key(seed, sample_id, layer_id, kind));
  BT_LAUNCH_OK("philox_normal_fill_kernel");
  return BT_OK;
}

extern "C" int bt_philox_sign_fill(float* out, int64_t n, uint64_t seed, uint32_t sample_id, uint32_t layer_id,
                                   int kind, bt_stream_t stream) {
  BT_REQUIRE(out && n >= 0 && kind >= 0 && kind < 16, BT_ERR_INVALID, "bt_philox_sign_fill: bad argument");
  if (n == 0) return BT_OK;
  philox_sign_fill_kernel<<<grid_for((n + 31) / 32, 256), 256, 0, (cudaStream_t)stream>>>(
      out, n, bt_rng_key(seed, sample_id, layer_id, kind));
  BT_LAUNCH_OK("philox_sign_fill_kernel");
  return BT_OK;
}

extern "C" int bt_moped_rho(const float* w, float* rho, int64_t n, float delta, bt_stream_t stream) {
  BT_REQUIRE(w && rho && n >= 0, BT_ERR_INVALID, "bt_moped_rho: bad argument");
  if (n == 0) return BT_OK;
  moped_rho_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(w, rho, n, delta);
  BT_LAUNCH_OK("moped_rho_kernel");
  return BT_OK;
}

extern "C" int bt_maxpool2d_nhwc(const void* x, void* y, int64_t batch, int64_t channels, int64_t ih, int64_t iw,
                                 int kernel, int stride, int pad, int dtype, bt_stream_t stream) {
  BT_REQUIRE(x && y && batch >= 0 && channels > 0 && ih > 0 && iw > 0 && kernel > 0 && stride > 0 && pad >= 0 && 2 * pad <= kernel,
             BT_ERR_INVALID, "bt_maxpool2d_nhwc: bad argument");
  BT_REQUIRE(dtype == BT_BF16 && channels % 8 == 0, BT_ERR_UNSUPPORTED, "bt_maxpool2d_nhwc: bf16 with channels %% 8 == 0 only");
  BT_REQUIRE((((uintptr_t)x | (uintptr_t)y) & 15) == 0, BT_ERR_ALIGN, "bt_maxpool2d_nhwc: pointers must be 16-byte aligned");
  const int64_t oh = (ih + 2 * pad - kernel) / stride + 1, ow = (iw + 2 * pad - kernel) / stride + 1;
  BT_REQUIRE(oh > 0 && ow > 0, BT_ERR_INVALID, "bt_maxpool2d_nhwc: empty output");
  const int64_t total = batch * oh * ow * (channels / 8);
  if (total == 0) return BT_OK;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)bt_num_sms() * 64;
  if (blocks > cap) blocks = cap;
  if (kernel == 3 && total + cap * 256 < 0x7fffffffll)
    maxpool2d_k3_nhwc_bf16_kernel<uint32_t><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)y, total, (int)(channels / 8), (int)ih, (int)iw, (int)oh, (int)ow, stride, pad);
  else if (total + cap * 256 < 0x7fffffffll)
    maxpool2d_nhwc_bf16_kernel<uint32_t><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)y, total, (int)(channels / 8), (int)ih, (int)iw, (int)oh, (int)ow, kernel, stride, pad);
  else
    maxpool2d_nhwc_bf16_kernel<int64_t><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)y, total, (int)(channels / 8), (int)ih, (int)iw, (int)oh, (int)ow, kernel, stride, pad);
  BT_LAUNCH_OK("maxpool2d_nhwc_bf16_kernel");
  return BT_OK;
}

extern "C" int bt_add_relu(void* y, const void* r, int64_t n, int dtype, bt_stream_t stream) {
  BT_REQUIRE(y && r && n >= 0, BT_ERR_INVALID, "bt_add_relu: bad argument");
  BT_REQUIRE((((uintptr_t)y | (uintptr_t)r) & 15) == 0, BT_ERR_ALIGN, "bt_add_relu: pointers must be 16-byte aligned");
  if (n == 0) return BT_OK;
  if (dtype == BT_BF16)
    add_relu_kernel<__nv_bfloat16><<<grid_for(n / 8 + 1, 256), 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)y, (const __nv_bfloat16*)r, n);
  else if (dtype == BT_F32)
    add_relu_kernel<float><<<grid_for(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>((float*)y, (const float*)r, n);
  else
    BT_REQUIRE(false, BT_ERR_UNSUPPORTED, "bt_add_relu: dtype %d", dtype);
  BT_LAUNCH_OK("add_relu_kernel");
  return BT_OK;
}

extern "C" int bt_mc_accumulate(const void* logits, int dtype, int64_t S, int64_t B, int64_t C, int apply_softmax,
                                float* sum_p, float* sum_p2, float* sum_plogp, bt_stream_t stream) {
  BT_REQUIRE(logits && S >= 0 && B >= 0 && C > 0, BT_ERR_INVALID, "bt_mc_accumulate: bad argument");
  if (S == 0 || B == 0) return BT_OK;
  const int grid = (int)(B < 65535 ? B : 65535);
  if (dtype == BT_F32)
    mc_accumulate_kernel<float><<<grid, MC_THREADS, 0, (cudaStream_t)stream>>>((const float*)logits, S, B, C, apply_softmax,
                                                                         sum_p, sum_p2, sum_plogp);
  else if (dtype == BT_BF16)
    mc_accumulate_kernel<__nv_bfloat16><<<grid, MC_THREADS, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, S, B, C, apply_softmax, sum_p, sum_p2, sum_plogp);
  else
    BT_REQUIRE(false, BT_ERR_UNSUPPORTED, "bt_mc_accumulate: dtype %d", dtype);
  BT_LAUNCH_OK("mc_accumulate_kernel");
  return BT_OK;
}

extern "C" int bt_mc_finalize(const float* sum_p, const float* sum_p2, const float* sum_plogp, int64_t S_total,
                              int64_t B, int64_t C, float* mean, float* var, float* entropy, float* mutual_info,
                              bt_stream_t stream) {
  BT_REQUIRE(sum_p && S_total > 0 && B >= 0 && C > 0, BT_ERR_INVALID, "bt_mc_finalize: bad argument");
  if (B == 0) return BT_OK;
  mc_finalize_kernel<<<grid_for(B * 32, 256), 256, 0, (cudaStream_t)stream>>>(
      sum_p, sum_p2, sum_plogp, 1.f / (float)S_total, B, C, mean, var, entropy, mutual_info);
  BT_LAUNCH_OK("mc_finalize_kernel");
  return BT_OK;
}
