// Fused LSTM time step of LSTMReparameterization / LSTMFlipout (fp32).
//
// The reference unrolls the cell in Python (variational_layers/rnn_variational.py:112-153,
// flipout_layers/rnn_flipout.py:102-155): per time step two Bayesian linear layers --
//   gates = ih(x_t) + hh(h_{t-1}),  each drawing FRESH weights (and signs) on the call --
// followed by ~12 element-wise launches (four slices, three sigmoids, two tanhs, the cell update).
// Here ONE launch per step does all of it: both contractions accumulate into the same gate tile,
// the weights are sampled while they are staged (never written to global memory), and the
// epilogue applies the gate non-linearities, updates the cell and writes h_t / c_t straight into
// their [batch, seq, hidden] slots.
//
// Tiling: a CTA owns 64 batch rows x 16 hidden units and computes all FOUR gates (i, f, g, o) of
// those units -- columns j, H + j, 2H + j, 3H + j of the [4H, K] weight -- so the cell update needs no
// exchange.  256 threads, 4 rows x 1 unit x 4 gates per thread, fp32 FFMA in k order (the parity
// path: same arithmetic order as two F.linear calls summed, rtol 1e-5 against the reference).
//
// lstm_gates_bwd_kernel is the element-wise half of the step's backward: from d h_t, d c_t and the
// saved gate activations it forms the pre-activation gate gradients [batch, 4H] and d c_{t-1}; the
// two contractions of the backward are the linear layers' own bt_linear_*_bwd calls on that tensor
// (they regenerate the step's weights from the same keys).
#include "bt_common.cuh"

namespace {

constexpr int LB = 64, LJ = 16, LK = 16, LNT = 256;

struct LstmLayerArgs {
  const float *mu_w, *rho_w, *mu_b, *rho_b, *eps_w, *eps_b, *sign_in, *sign_out;
  BtRngKey key_w, key_b, key_si, key_so;
  int K;  // reduction length of this layer (in_features of ih, hidden of hh)
};

struct LstmArgs {
  LstmLayerArgs l[2];  // ih, hh
  const float* a[2];   // x_t [batch, K_ih], h_{t-1} [batch, H]
  int64_t a_rs[2];     // their row strides in elements
  const float* c_prev;
  int64_t c_rs;
  float *h_out, *c_out, *gates_out;  // gates_out nullable: [batch, 4H] activations i, f, g, o
  int64_t ho_rs, co_rs;
  int64_t batch;
  int H;
};

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + expf(-v)); }

template <bool kFlip>
__global__ void __launch_bounds__(LNT) lstm_step_kernel(const LstmArgs p) {
  constexpr int NA = kFlip ? 2 : 1;           // Flipout: the operand and its sign-flipped copy / mu and sigma*eps
  __shared__ __align__(16) float As[NA][LK][LB];
  __shared__ __align__(16) float Bs[NA][LK][4 * LJ + 4];
  const int tid = threadIdx.x;
  const int64_t b0 = (int64_t)blockIdx.x * LB;
  const int j0 = blockIdx.y * LJ;
  const int tx = tid % LJ, ty = tid / LJ;     // hidden unit j0 + tx; batch rows b0 + 4 * ty + {0..3}
  const int H = p.H;

  float acc[3][4][4];                         // [mean | ih perturbation | hh perturbation][row][gate]
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int g = 0; g < 4; ++g) acc[a][r][g] = 0.f;

  // loaders: A tile [64 rows x 16 k]: thread -> row tid % 64, k lanes tid / 64 + 4 i;
  //          W tile [64 columns (gate-major: 4 x 16 units) x 16 k]: thread -> k lane tid % 16, columns tid / 16 + 16 i
  const int a_row = tid % LB, a_k0 = tid / LB;
  const int w_k = tid % LK, w_c0 = tid / LK;
#pragma unroll 1
  for (int layer = 0; layer < 2; ++layer) {
    const LstmLayerArgs& L = p.l[layer];
    const float* __restrict__ A = p.a[layer];
    const int64_t rs = p.a_rs[layer];
    const int K = L.K;
    const int64_t row = b0 + a_row;
    for (int k0 = 0; k0 < K; k0 += LK) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = k0 + a_k0 + 4 * i;
        float v = 0.f, v2 = 0.f;
        if (row < p.batch && k < K) {
          v = A[row * rs + k];
          if (kFlip) {
            const float sg = L.sign_in ? L.sign_in[row * K + k] : bt_sign1((uint64_t)row * (uint64_t)K + (uint64_t)k, L.key_si);
            v2 = v * sg;
          }
        }
        As[0][a_k0 + 4 * i][a_row] = v;
        if (kFlip) As[NA - 1][a_k0 + 4 * i][a_row] = v2;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int col = w_c0 + 16 * i;        // gate col / 16, unit col % 16
        const int j = j0 + (col % LJ);
        const int n = (col / LJ) * H + j;     // row of the [4H, K] weight
        const int k = k0 + w_k;
        float w = 0.f, w2 = 0.f;
        if (j < H && k < K) {
          const int64_t idx = (int64_t)n * K + k;
          const float mu = __ldg(L.mu_w + idx);
          const float sig = bt_softplus(__ldg(L.rho_w + idx));
          const float eps = L.eps_w ? __ldg(L.eps_w + idx) : bt_normal1((uint64_t)idx, L.key_w);
          if (kFlip) { w = mu; w2 = sig * eps; } else { w = fmaf(sig, eps, mu); }
        }
        Bs[0][w_k][col] = w;
        if (kFlip) Bs[NA - 1][w_k][col] = w2;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < LK; ++kk) {
        const float4 a4 = *reinterpret_cast<const float4*>(&As[0][kk][ty * 4]);
        const float av[4] = {a4.x, a4.y, a4.z, a4.w};
        float bv[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) bv[g] = Bs[0][kk][g * LJ + tx];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int g = 0; g < 4; ++g) acc[0][r][g] = fmaf(av[r], bv[g], acc[0][r][g]);
        if (kFlip) {
          const float4 f4 = *reinterpret_cast<const float4*>(&As[NA - 1][kk][ty * 4]);
          const float fv[4] = {f4.x, f4.y, f4.z, f4.w};
          float dv[4];
#pragma unroll
          for (int g = 0; g < 4; ++g) dv[g] = Bs[NA - 1][kk][g * LJ + tx];
          if (layer == 0) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
              for (int g = 0; g < 4; ++g) acc[1][r][g] = fmaf(fv[r], dv[g], acc[1][r][g]);
          } else {
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
              for (int g = 0; g < 4; ++g) acc[2][r][g] = fmaf(fv[r], dv[g], acc[2][r][g]);
          }
        }
      }
      __syncthreads();
    }
  }

  // ---- epilogue: sampled biases, Flipout recombination, gates, cell update
  const int j = j0 + tx;
  if (j >= H) return;
  float b_mean[2][4], b_pert[2][4];
#pragma unroll
  for (int layer = 0; layer < 2; ++layer) {
    const LstmLayerArgs& L = p.l[layer];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      b_mean[layer][g] = 0.f; b_pert[layer][g] = 0.f;
      if (L.mu_b) {
        const int n = g * H + j;
        const float e = L.eps_b ? L.eps_b[n] : bt_normal1((uint64_t)n, L.key_b);
        b_mean[layer][g] = L.mu_b[n];
        b_pert[layer][g] = bt_softplus(L.rho_b[n]) * e;
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int64_t b = b0 + ty * 4 + r;
    if (b >= p.batch) continue;
    float pre[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const int n = g * H + j;
      if (kFlip) {
        // each layer: x.mu + mu_b + ((x*s_in).(sigma*eps) + sigma_b*eps_b) * s_out   (linear_flipout.py:168-174)
        float v = acc[0][r][g] + b_mean[0][g] + b_mean[1][g];
#pragma unroll
        for (int layer = 0; layer < 2; ++layer) {
          const LstmLayerArgs& L = p.l[layer];
          const uint64_t so = (uint64_t)b * (uint64_t)(4 * H) + (uint64_t)n;
          const float sg = L.sign_out ? L.sign_out[so] : bt_sign1(so, L.key_so);
          v += (acc[1 + layer][r][g] + b_pert[layer][g]) * sg;
        }
        pre[g] = v;
      } else {
        // (ih(x) + b_ih) + (hh(h) + b_hh): the reference adds the two layer outputs
        pre[g] = (acc[0][r][g]) + ((b_mean[0][g] + b_pert[0][g]) + (b_mean[1][g] + b_pert[1][g]));
      }
    }
    const float i_t = sigmoidf_(pre[0]), f_t = sigmoidf_(pre[1]), g_t = tanhf(pre[2]), o_t = sigmoidf_(pre[3]);
    const float c_prev = p.c_prev ? p.c_prev[b * p.c_rs + j] : 0.f;
    const float c_t = f_t * c_prev + i_t * g_t;
    const float h_t = o_t * tanhf(c_t);
    p.c_out[b * p.co_rs + j] = c_t;
    p.h_out[b * p.ho_rs + j] = h_t;
    if (p.gates_out) {
      float* gq = p.gates_out + b * (int64_t)(4 * H) + j;
      gq[0] = i_t; gq[H] = f_t; gq[2 * H] = g_t; gq[3 * H] = o_t;
    }
  }
}

// d pre-activation gates [batch, 4H] and d c_{t-1} from d h_t, d c_t and the saved activations.
__global__ void lstm_gates_bwd_kernel(const float* __restrict__ gates, const float* __restrict__ c_prev, int64_t c_rs,
                                      const float* __restrict__ c_t, int64_t ct_rs, const float* __restrict__ dh, int64_t dh_rs,
                                      const float* __restrict__ dc, int64_t dc_rs, float* __restrict__ dpre,
                                      float* __restrict__ dc_prev, int64_t batch, int H) {
  const int64_t total = batch * H;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / H;
    const int j = (int)(i - b * H);
    const float* gq = gates + b * (int64_t)(4 * H) + j;
    const float it = gq[0], ft = gq[H], gt = gq[2 * H], ot = gq[3 * H];
    const float th = tanhf(c_t[b * ct_rs + j]);
    const float dht = dh ? dh[b * dh_rs + j] : 0.f;
    const float dct = (dc ? dc[b * dc_rs + j] : 0.f) + dht * ot * (1.f - th * th);
    const float cp = c_prev ? c_prev[b * c_rs + j] : 0.f;
    float* dq = dpre + b * (int64_t)(4 * H) + j;
    dq[0] = dct * gt * it * (1.f - it);
    dq[H] = dct * cp * ft * (1.f - ft);
    dq[2 * H] = dct * it * (1.f - gt * gt);
    dq[3 * H] = dht * th * ot * (1.f - ot);
    dc_prev[i] = dct * ft;
  }
}

LstmLayerArgs layer_args(const bt_lstm_layer* l, int K) {
  LstmLayerArgs a;
  a.mu_w = l->mu_w; a.rho_w = l->rho_w; a.mu_b = l->mu_b; a.rho_b = l->rho_b;
  a.eps_w = l->eps_w; a.eps_b = l->eps_b; a.sign_in = l->sign_in; a.sign_out = l->sign_out;
  a.key_w = bt_rng_key(l->seed, l->sample_id, l->layer_id, BT_KIND_EPS_W);
  a.key_b = bt_rng_key(l->seed, l->sample_id, l->layer_id, BT_KIND_EPS_B);
  a.key_si = bt_rng_key(l->seed, l->sample_id, l->layer_id, BT_KIND_SIGN_IN);
  a.key_so = bt_rng_key(l->seed, l->sample_id, l->layer_id, BT_KIND_SIGN_OUT);
  a.K = K;
  return a;
}

}  // namespace

extern "C" int bt_lstm_step_fwd(const bt_lstm_step* s, bt_stream_t stream) {
  BT_REQUIRE(s, BT_ERR_INVALID, "bt_lstm_step_fwd: null argument");
  BT_REQUIRE(s->batch >= 0 && s->hidden > 0 && s->in_features > 0, BT_ERR_INVALID, "bt_lstm_step_fwd: bad sizes");
  BT_REQUIRE(s->ih.mu_w && s->ih.rho_w && s->hh.mu_w && s->hh.rho_w, BT_ERR_INVALID, "bt_lstm_step_fwd: mu_w / rho_w are required");
  BT_REQUIRE((s->ih.mu_b == nullptr) == (s->ih.rho_b == nullptr) && (s->hh.mu_b == nullptr) == (s->hh.rho_b == nullptr),
             BT_ERR_INVALID, "bt_lstm_step_fwd: mu_b and rho_b must both be given or both be null");
  BT_REQUIRE(s->x && s->h_prev && s->h_out && s->c_out, BT_ERR_INVALID, "bt_lstm_step_fwd: null tensor pointer");
  if (s->flipout)
    BT_REQUIRE((s->ih.sign_in == nullptr) == (s->ih.sign_out == nullptr) && (s->hh.sign_in == nullptr) == (s->hh.sign_out == nullptr),
               BT_ERR_INVALID, "bt_lstm_step_fwd: sign_in and sign_out must both be given or both be null");
  if (s->batch == 0) return BT_OK;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    bt_set_error("bt_lstm_step_fwd: no CUDA device (there is no CPU fallback)");
    return BT_ERR_CUDA;
  }
  LstmArgs a;
  a.l[0] = layer_args(&s->ih, s->in_features);
  a.l[1] = layer_args(&s->hh, s->hidden);
  a.a[0] = s->x; a.a_rs[0] = s->x_row_stride;
  a.a[1] = s->h_prev; a.a_rs[1] = s->h_row_stride;
  a.c_prev = s->c_prev; a.c_rs = s->c_row_stride;
  a.h_out = s->h_out; a.c_out = s->c_out; a.gates_out = s->gates_out;
  a.ho_rs = s->h_out_row_stride; a.co_rs = s->c_out_row_stride;
  a.batch = s->batch; a.H = s->hidden;
  const int64_t gx = (s->batch + LB - 1) / LB;
  const int gy = (s->hidden + LJ - 1) / LJ;
  BT_REQUIRE(gx <= 0x7fffffffll && gy <= 65535, BT_ERR_UNSUPPORTED, "bt_lstm_step_fwd: grid too large");
  dim3 grid((unsigned)gx, (unsigned)gy);
  if (s->flipout) lstm_step_kernel<true><<<grid, LNT, 0, (cudaStream_t)stream>>>(a);
  else lstm_step_kernel<false><<<grid, LNT, 0, (cudaStream_t)stream>>>(a);
  BT_LAUNCH_OK("lstm_step_kernel");
  return BT_OK;
}

extern "C" int bt_lstm_step_bwd_gates(int64_t batch, int hidden, const float* gates, const float* c_prev, int64_t c_prev_row_stride,
                                      const float* c_t, int64_t c_t_row_stride, const float* grad_h, int64_t grad_h_row_stride,
                                      const float* grad_c, int64_t grad_c_row_stride, float* grad_pre, float* grad_c_prev,
                                      bt_stream_t stream) {
  BT_REQUIRE(batch >= 0 && hidden > 0, BT_ERR_INVALID, "bt_lstm_step_bwd_gates: bad sizes");
  BT_REQUIRE(gates && c_t && grad_pre && grad_c_prev, BT_ERR_INVALID, "bt_lstm_step_bwd_gates: null tensor pointer");
  if (batch == 0) return BT_OK;
  const int64_t total = batch * hidden;
  int64_t blocks = (total + 255) / 256;
  const int64_t cap = (int64_t)bt_num_sms() * 8;
  if (blocks > cap) blocks = cap;
  lstm_gates_bwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(gates, c_prev, c_prev_row_stride, c_t, c_t_row_stride, grad_h,
                                                                      grad_h_row_stride, grad_c, grad_c_row_stride, grad_pre, grad_c_prev,
                                                                      batch, hidden);
  BT_LAUNCH_OK("lstm_gates_bwd_kernel");
  return BT_OK;
}
