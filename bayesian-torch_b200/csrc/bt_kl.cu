// Fused multi-tensor KL(N(mu, softplus(rho)) || N(mu_p, sigma_p)) reduction.
//
// Replaces the ~10 element-wise ATen launches + .mean() per tensor of
// BaseVariationalLayer_.kl_div (base_variational_layer.py:70-72) and the Python loop
// of get_kl_loss (models/dnn_to_bnn.py:157-165) with one HBM-bound pass:
// algorithmic traffic 8 B / element (mu + rho read once), +8 B with tensor priors,
// +8 B when gradients are written.
//
// Work decomposition: every tensor is cut into chunks of KL_CHUNK elements; a
// grid-stride loop over the global chunk list keeps all SMs busy whatever the mix of
// tensor sizes (ResNet-50: 54 tensors from 4 K to 2.4 M elements).  Each thread
// accumulates <= 32 elements in fp32, the block reduces with warp shuffles and issues
// ONE fp64 atomicAdd per chunk into the tensor's accumulator; the last block to finish
// turns the fp64 sums into the fp32 means kl_div returns and their total.
#include "bt_common.cuh"
#include "bt_kl.cuh"

#define KL_MAX_TENSORS 192
#define KL_THREADS 256
#define KL_VEC_ITERS 8
#define KL_CHUNK (KL_THREADS * 4 * KL_VEC_ITERS) /* 8192 elements = 64 KB of mu+rho */

struct KlTable {
  const float* mu[KL_MAX_TENSORS];
  const float* rho[KL_MAX_TENSORS];
  const float* pmu[KL_MAX_TENSORS];  // tensor priors or NULL
  const float* psig[KL_MAX_TENSORS];
  float* dmu[KL_MAX_TENSORS];  // gradient outputs or NULL
  float* drho[KL_MAX_TENSORS];
  int64_t numel[KL_MAX_TENSORS];
  int32_t chunk_end[KL_MAX_TENSORS];  // exclusive prefix end, in chunks
  int32_t n;
  int32_t total_chunks;
  int32_t q_is_sigma;
  int32_t laplace;  // the fork's 'laplace' prior: KL(N(mu, sigma) || Laplace(0, 1)), analytic (base_variational_layer.py:74-97)
  float prior_mu, prior_sigma;
};

// KL(N(mu, sigma) || Laplace(mu_p = 0, b = 1)) as the fork's kl_div('laplace') writes it
// (base_variational_layer.py:74-97; location and scale are hard-coded there):
//   log(2b) - 0.5 log(2 pi sigma^2) - 0.5 + E|w| / b,   E|w| = sigma sqrt(2/pi) exp(-mu^2 / 2 sigma^2) + mu erf(mu / (sigma sqrt 2))
// (1 - 2 Phi(-z) = erf(z / sqrt 2)).  d/dmu = erf(mu / (sigma sqrt 2)),  d/dsigma = -1/sigma + sqrt(2/pi) exp(-mu^2 / 2 sigma^2).
template <bool kGrad>
__device__ __forceinline__ KlElem kl_elem_laplace(float mu, float rho, bool q_is_sigma) {
  float sig, dsig_drho;
  if (q_is_sigma) {
    sig = rho;
    dsig_drho = 1.f;
  } else {
    const float t = expf(fminf(rho, 30.f));
    sig = rho > 20.f ? rho : log1pf(t);
    dsig_drho = rho > 20.f ? 1.f : t / (1.f + t);
  }
  const float inv = 1.f / sig;
  const float z = mu * inv;
  const float g = 0.7978845608028654f * expf(-0.5f * z * z);  // sqrt(2/pi) exp(-mu^2 / 2 sigma^2)
  const float er = erff(z * 0.7071067811865476f);
  KlElem e;
  e.kl = (0.6931471805599453f - 0.9189385332046727f - logf(sig) - 0.5f) + (sig * g + mu * er);
  e.dmu = kGrad ? er : 0.f;
  e.drho = kGrad ? (g - inv) * dsig_drho : 0.f;
  return e;
}

template <bool kGrad, bool kTensorPrior, bool kLaplace = false>
__global__ void __launch_bounds__(KL_THREADS) kl_normal_multi_kernel(const __grid_constant__ KlTable tab,
                                                                      double* __restrict__ acc,
                                                                      unsigned int* __restrict__ done_counter,
                                                                      float* __restrict__ kl_each,
                                                                      float* __restrict__ kl_total,
                                                                      const float* __restrict__ grad_scale) {
  __shared__ float warp_part[KL_THREADS / 32];
  __shared__ bool is_last;
  const int tid = threadIdx.x;
  const float gscale = (kGrad && grad_scale != nullptr) ? __ldg(grad_scale) : 1.f;
  const KlPrior prior_c = kl_prior(tab.prior_sigma);

  for (int chunk = blockIdx.x; chunk < tab.total_chunks; chunk += gridDim.x) {
    // which tensor owns this chunk (binary search over the chunk prefix)
    int lo = 0, hi = tab.n - 1;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (chunk < tab.chunk_end[mid]) hi = mid; else lo = mid + 1;
    }
    const int t = lo;
    const int first_chunk = t == 0 ? 0 : tab.chunk_end[t - 1];
    const int64_t numel = tab.numel[t];
    const int64_t base = (int64_t)(chunk - first_chunk) * KL_CHUNK;
    const int64_t end = min(numel, base + (int64_t)KL_CHUNK);
    const float* __restrict__ mu = tab.mu[t];
    const float* __restrict__ rho = tab.rho[t];
    const float* __restrict__ pmu = kTensorPrior ? tab.pmu[t] : nullptr;
    const float* __restrict__ psig = kTensorPrior ? tab.psig[t] : nullptr;
    float* __restrict__ dmu = kGrad ? tab.dmu[t] : nullptr;
    float* __restrict__ drho = kGrad ? tab.drho[t] : nullptr;
    const float gs = kGrad ? gscale / (float)numel : 0.f;
    const bool qsig = tab.q_is_sigma != 0;
    // 128-bit path needs 16 B aligned bases (torch allocations are; sliced views may not be)
    const bool vec_ok = ((((uintptr_t)mu | (uintptr_t)rho | (uintptr_t)pmu | (uintptr_t)psig | (uintptr_t)dmu |
                           (uintptr_t)drho) & 15) == 0);
    float part = 0.f;
    if (vec_ok && end - base == KL_CHUNK) {
      float4 m4[KL_VEC_ITERS], r4[KL_VEC_ITERS];
#pragma unroll
      for (int i = 0; i < KL_VEC_ITERS; ++i) {  // all loads first: 16 x 16 B in flight per thread
        const int64_t off = base + ((int64_t)i * KL_THREADS + tid) * 4;
        m4[i] = __ldcs(reinterpret_cast<const float4*>(mu + off));
        r4[i] = __ldcs(reinterpret_cast<const float4*>(rho + off));
      }
#pragma unroll
      for (int i = 0; i < KL_VEC_ITERS; ++i) {
        const int64_t off = base + ((int64_t)i * KL_THREADS + tid) * 4;
        float4 pm4 = make_float4(tab.prior_mu, tab.prior_mu, tab.prior_mu, tab.prior_mu);
        float4 ps4 = make_float4(tab.prior_sigma, tab.prior_sigma, tab.prior_sigma, tab.prior_sigma);
        if (kTensorPrior) {
          if (pmu) pm4 = __ldcs(reinterpret_cast<const float4*>(pmu + off));
          if (psig) ps4 = __ldcs(reinterpret_cast<const float4*>(psig + off));
        }
        const bool tp = kTensorPrior && psig != nullptr;
        KlElem e0, e1, e2, e3;
        if constexpr (kLaplace) {
          e0 = kl_elem_laplace<kGrad>(m4[i].x, r4[i].x, qsig);
          e1 = kl_elem_laplace<kGrad>(m4[i].y, r4[i].y, qsig);
          e2 = kl_elem_laplace<kGrad>(m4[i].z, r4[i].z, qsig);
          e3 = kl_elem_laplace<kGrad>(m4[i].w, r4[i].w, qsig);
        } else {
          e0 = kl_elem<kGrad>(m4[i].x, r4[i].x, pm4.x, tp ? kl_prior(ps4.x) : prior_c, qsig);
          e1 = kl_elem<kGrad>(m4[i].y, r4[i].y, pm4.y, tp ? kl_prior(ps4.y) : prior_c, qsig);
          e2 = kl_elem<kGrad>(m4[i].z, r4[i].z, pm4.z, tp ? kl_prior(ps4.z) : prior_c, qsig);
          e3 = kl_elem<kGrad>(m4[i].w, r4[i].w, pm4.w, tp ? kl_prior(ps4.w) : prior_c, qsig);
        }
        part += (e0.kl + e1.kl) + (e2.kl + e3.kl);
        if (kGrad) {
          if (dmu) __stcs(reinterpret_cast<float4*>(dmu + off), make_float4(e0.dmu * gs, e1.dmu * gs, e2.dmu * gs, e3.dmu * gs));
          if (drho) __stcs(reinterpret_cast<float4*>(drho + off), make_float4(e0.drho * gs, e1.drho * gs, e2.drho * gs, e3.drho * gs));
        }
      }
    } else {  // ragged tail chunk or unaligned tensor: scalar, still coalesced
      for (int64_t i = base + tid; i < end; i += KL_THREADS) {
        const float pm = (kTensorPrior && pmu) ? pmu[i] : tab.prior_mu;
        const float ps = (kTensorPrior && psig) ? psig[i] : tab.prior_sigma;
        const KlElem e = kLaplace ? kl_elem_laplace<kGrad>(mu[i], rho[i], qsig)
                                  : kl_elem<kGrad>(mu[i], rho[i], pm, (kTensorPrior && psig) ? kl_prior(ps) : prior_c, qsig);
        part += e.kl;
        if (kGrad) {
          if (dmu) dmu[i] = e.dmu * gs;
          if (drho) drho[i] = e.drho * gs;
        }
      }
    }
    part = bt_warp_sum(part);
    if ((tid & 31) == 0) warp_part[tid >> 5] = part;
    __syncthreads();
    if (tid < 32) {
      float v = tid < KL_THREADS / 32 ? warp_part[tid] : 0.f;
      v = bt_warp_sum(v);
      if (tid == 0) atomicAdd(&acc[t], (double)v);
    }
    __syncthreads();
  }

  // last-block-done: convert the fp64 sums into the fp32 means kl_div returns
  if (tid == 0) {
    __threadfence();
    const unsigned int prev = atomicAdd(done_counter, 1u);
    is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    double tot = 0.0;
    for (int t = tid; t < tab.n; t += KL_THREADS) {
      const double mean = __ldcg(&acc[t]) / (double)tab.numel[t];
      if (kl_each) kl_each[t] = (float)mean;
      tot += mean;
    }
    // deterministic block sum of <= KL_MAX_TENSORS per-tensor means
    __shared__ double dpart[KL_THREADS / 32];
    tot = bt_warp_sum(tot);
    if ((tid & 31) == 0) dpart[tid >> 5] = tot;
    __syncthreads();
    if (tid == 0 && kl_total) {
      double s = 0.0;
      for (int w = 0; w < KL_THREADS / 32; ++w) s += dpart[w];
      *kl_total = (float)s;
    }
  }
}

static int64_t kl_workspace_bytes(int n) { return ((int64_t)n * 8 + 8 + 255) / 256 * 256; }

int64_t bt_kl_workspace_bytes_internal(int n) { return kl_workspace_bytes(n < 1 ? 1 : n); }

static int kl_launch(int n, const float* const* mu, const float* const* rho, const int64_t* numel, int q_is_sigma,
                     float prior_mu, float prior_sigma, const float* const* pmu_t, const float* const* psig_t,
                     float* kl_each, float* kl_total, float* const* d_mu, float* const* d_rho,
                     const float* grad_scale, void* workspace, cudaStream_t stream, bool laplace = false) {
  BT_REQUIRE(n >= 1 && n <= KL_MAX_TENSORS, BT_ERR_INVALID, "bt_kl_normal_multi: n=%d outside [1,%d]", n,
             KL_MAX_TENSORS);
  BT_REQUIRE(mu && rho && numel && workspace, BT_ERR_INVALID, "bt_kl_normal_multi: null table or workspace");
  BT_REQUIRE(laplace || prior_sigma > 0.f || psig_t, BT_ERR_INVALID, "bt_kl_normal_multi: prior_sigma must be > 0");
  KlTable tab;
  bool any_tensor_prior = false;
  const bool grad = (d_mu != nullptr) || (d_rho != nullptr);
  int64_t chunks = 0;
  for (int i = 0; i < n; ++i) {
    BT_REQUIRE(mu[i] && rho[i] && numel[i] > 0, BT_ERR_INVALID, "bt_kl_normal_multi: tensor %d is null or empty", i);
    tab.mu[i] = mu[i];
    tab.rho[i] = rho[i];
    tab.pmu[i] = pmu_t ? pmu_t[i] : nullptr;
    tab.psig[i] = psig_t ? psig_t[i] : nullptr;
    any_tensor_prior |= (tab.pmu[i] != nullptr) || (tab.psig[i] != nullptr);
    tab.dmu[i] = d_mu ? d_mu[i] : nullptr;
    tab.drho[i] = d_rho ? d_rho[i] : nullptr;
    tab.numel[i] = numel[i];
    chunks += (numel[i] + KL_CHUNK - 1) / KL_CHUNK;
    BT_REQUIRE(chunks < (1ll << 31), BT_ERR_UNSUPPORTED, "bt_kl_normal_multi: too many elements");
    tab.chunk_end[i] = (int32_t)chunks;
  }
  tab.n = n;
  tab.total_chunks = (int32_t)chunks;
  tab.q_is_sigma = q_is_sigma;
  tab.laplace = laplace ? 1 : 0;
  tab.prior_mu = prior_mu;
  tab.prior_sigma = prior_sigma;

  double* acc = reinterpret_cast<double*>(workspace);
  unsigned int* counter = reinterpret_cast<unsigned int*>(acc + n);
  BT_CUDA_OK(cudaMemsetAsync(workspace, 0, (size_t)n * 8 + 8, stream));
  // grid: a multiple of the SM count, 8 resident CTAs/SM of 256 threads hide HBM latency
  const int sms = bt_num_sms();
  int grid = (int)(chunks < (int64_t)sms * 8 ? chunks : (int64_t)sms * 8);
  if (grid < 1) grid = 1;
#define KL_GO(G, T)                                                                                         \
  kl_normal_multi_kernel<G, T><<<grid, KL_THREADS, 0, stream>>>(tab, acc, counter, kl_each, kl_total, grad_scale)
  if (laplace) {
    if (grad) kl_normal_multi_kernel<true, false, true><<<grid, KL_THREADS, 0, stream>>>(tab, acc, counter, kl_each, kl_total, grad_scale);
    else kl_normal_multi_kernel<false, false, true><<<grid, KL_THREADS, 0, stream>>>(tab, acc, counter, kl_each, kl_total, grad_scale);
  } else if (grad) {
    if (any_tensor_prior) KL_GO(true, true); else KL_GO(true, false);
  } else {
    if (any_tensor_prior) KL_GO(false, true); else KL_GO(false, false);
  }
#undef KL_GO
  BT_LAUNCH_OK("kl_normal_multi_kernel");
  return BT_OK;
}

extern "C" int bt_kl_normal_multi(int n, const float* const* mu, const float* const* rho, const int64_t* numel,
                                  int q_is_sigma, float prior_mu, float prior_sigma,
                                  const float* const* prior_mu_t, const float* const* prior_sigma_t, float* kl_each,
                                  float* kl_total, float* const* d_mu, float* const* d_rho, const float* grad_scale,
                                  void* workspace, bt_stream_t stream) {
  return kl_launch(n, mu, rho, numel, q_is_sigma, prior_mu, prior_sigma, prior_mu_t, prior_sigma_t, kl_each, kl_total,
                   d_mu, d_rho, grad_scale, workspace, (cudaStream_t)stream);
}

extern "C" int bt_kl_laplace_multi(int n, const float* const* mu, const float* const* rho, const int64_t* numel,
                                   int q_is_sigma, float* kl_each, float* kl_total, float* const* d_mu, float* const* d_rho,
                                   const float* grad_scale, void* workspace, bt_stream_t stream) {
  return kl_launch(n, mu, rho, numel, q_is_sigma, 0.f, 1.f, nullptr, nullptr, kl_each, kl_total, d_mu, d_rho, grad_scale,
                   workspace, (cudaStream_t)stream, true);
}

// kl(weight) + kl(bias) of one layer, as forward(return_kl=True) returns it.
int bt_kl_layer(const bt_layer_params* p, int64_t w_numel, int64_t b_numel, float* kl_out, void* workspace,
                cudaStream_t stream) {
  const float* mu[2] = {p->mu_w, p->mu_b};
  const float* rho[2] = {p->rho_w, p->rho_b};
  const float* pm[2] = {p->prior_mu_w_t, p->prior_mu_b_t};
  const float* ps[2] = {p->prior_sigma_w_t, p->prior_sigma_b_t};
  const int64_t numel[2] = {w_numel, b_numel};
  const int n = p->mu_b ? 2 : 1;
  return kl_launch(n, mu, rho, numel, 0, p->prior_mu, p->prior_sigma, pm, ps, nullptr, kl_out, nullptr, nullptr,
                   nullptr, workspace, stream);
}

// ---------------------------------------------------------------- allocation-free KL ----
// bt_kl_normal_multi_static: bt_kl_normal_multi for callers that cannot allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  The fp64 accumulators + done counter the kernel needs live
// in a file-scope device array instead of a caller workspace; every GPU that loads the library holds these
// KL_STATIC_BYTES bytes until the process exits.  All calls on one device SHARE the array: it is cleared on the
// caller's stream before every launch (it still holds the previous call's sums), and two calls that overlap on
// one GPU (different streams) race on it.
#define KL_STATIC_MAX_TENSORS 192
#define KL_STATIC_BYTES (((KL_STATIC_MAX_TENSORS * 8 + 8 + 255) / 256) * 256) /* 1792 bytes */
#define KL_STATIC_MAX_DEVICES 64
static_assert(KL_STATIC_MAX_TENSORS <= KL_MAX_TENSORS, "the static variant cannot exceed the kernel's table");

static __device__ __align__(256) unsigned char g_kl_static_scratch[KL_STATIC_BYTES];

// The array's name is only a host-side stand-in in host code: its device address is looked up once per device,
// with that device current, and cached.
static int kl_static_scratch_ptr(void** out) {
  static void* cached[KL_STATIC_MAX_DEVICES] = {nullptr};
  int dev = -1;
  BT_CUDA_OK(cudaGetDevice(&dev));
  BT_REQUIRE(dev >= 0 && dev < KL_STATIC_MAX_DEVICES, BT_ERR_UNSUPPORTED, "bt_kl_normal_multi_static: device %d outside [0,%d)", dev,
             KL_STATIC_MAX_DEVICES);
  void* ptr = __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE);
  if (ptr == nullptr) {
    BT_CUDA_OK(cudaGetSymbolAddress(&ptr, g_kl_static_scratch));
    BT_REQUIRE(ptr != nullptr, BT_ERR_CUDA, "bt_kl_normal_multi_static: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&cached[dev], ptr, __ATOMIC_RELEASE);   // every thread looks up the same address: a benign race
  }
  *out = ptr;
  return BT_OK;
}

extern "C" int bt_kl_normal_multi_static(int n, const float* const* mu, const float* const* rho, const int64_t* numel,
                                         int q_is_sigma, float prior_mu, float prior_sigma,
                                         const float* const* prior_mu_t, const float* const* prior_sigma_t, float* kl_each,
                                         float* kl_total, float* const* d_mu, float* const* d_rho, const float* grad_scale,
                                         bt_stream_t stream) {
  BT_REQUIRE(n >= 1 && n <= KL_STATIC_MAX_TENSORS, BT_ERR_INVALID,
             "bt_kl_normal_multi_static: n=%d outside [1,%d] (the built-in scratch array holds %d accumulators; "
             "use bt_kl_normal_multi with a caller workspace beyond that)", n, KL_STATIC_MAX_TENSORS, KL_STATIC_MAX_TENSORS);
  BT_REQUIRE(kl_workspace_bytes(n) <= (int64_t)KL_STATIC_BYTES, BT_ERR_INVALID,
             "bt_kl_normal_multi_static: %lld scratch bytes needed, %d built in", (long long)kl_workspace_bytes(n), (int)KL_STATIC_BYTES);
  void* scratch = nullptr;
  const int rc = kl_static_scratch_ptr(&scratch);
  if (rc != BT_OK) return rc;
  // kl_launch clears the n accumulators + counter with a memset on `stream`, then launches: no allocation anywhere
  return kl_launch(n, mu, rho, numel, q_is_sigma, prior_mu, prior_sigma, prior_mu_t, prior_sigma_t, kl_each, kl_total,
                   d_mu, d_rho, grad_scale, scratch, (cudaStream_t)stream);
}
