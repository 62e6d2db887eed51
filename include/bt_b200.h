/* bt_b200.h -- C ABI of the B200-native Monte-Carlo variational-layer hot path.
 *
 * The reference (godhj93/bayesian-torch) is pure Python over ATen and has no FFI
 * of its own; each entry point below names the reference Python interface it
 * replaces (paths relative to /root/reference/bayesian_torch/).  A reference
 * maintainer binds these with ctypes -- see INTEGRATION.md.
 *
 * Conventions
 *  - plain pointers and sizes only; every data pointer is a DEVICE pointer owned
 *    by the caller (PyTorch's caching allocator in the shipped host code);
 *  - everything is enqueued on `stream` (a cudaStream_t passed as void*); no entry
 *    point synchronises, allocates device memory or keeps global mutable state,
 *    so calls are re-entrant and CUDA-graph capturable;
 *  - return value: BT_OK or a negative bt_status; bt_last_error() returns a
 *    thread-local message for the last failure on the calling thread;
 *  - there is NO CPU fallback: a machine without a CUDA device gets BT_ERR_CUDA.
 */
#ifndef BT_B200_H_
#define BT_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BT_ABI_VERSION 5
#define BT_FLAG_PACKED_VALID 1
/* bt_conv_*_bwd run as the FORWARD of a ConvTranspose layer (grad_x only, see the table in INTEGRATION.md): the layer's
 * sampled bias -- mu_b + softplus(rho_b) eps_b per output channel; Flipout: mu_b + softplus(rho_b) eps_b s_out -- is
 * added to grad_x in the kernel epilogue (conv_variational.py:941-989, conv_flipout.py:757-831). */
#define BT_FLAG_TRANSPOSE_BIAS 2
/* bt_conv_*_bwd run as the BACKWARD of such a layer: grad_mu_b / grad_rho_b reduce the tensor passed as `x` (the
 * gradient of the layer's output) instead of `grad_out`. */
#define BT_FLAG_BIAS_GRAD_X_SIDE 4

typedef void* bt_stream_t; /* cudaStream_t */

typedef enum bt_status {
  BT_OK = 0,
  BT_ERR_INVALID = -1,     /* bad argument (null pointer, negative size, ...)     */
  BT_ERR_UNSUPPORTED = -2, /* valid request the selected path cannot execute      */
  BT_ERR_CUDA = -3,        /* CUDA runtime / driver error (message has the string) */
  BT_ERR_ALIGN = -4        /* pointer or row pitch not aligned as the path needs   */
} bt_status;

typedef enum bt_dtype { BT_F32 = 0, BT_BF16 = 1 } bt_dtype;

/* Which kernel family executes a layer op.
 *  BT_PATH_SIMT_F32: fp32 FFMA implicit GEMM, any rank/stride/groups/layout --
 *                    the deterministic parity mode (rtol 1e-5 vs the reference).
 *  BT_PATH_TC_BF16 : tcgen05/TMEM implicit GEMM, bf16 operands, fp32 accumulate,
 *                    weights sampled tile-by-tile in shared memory.
 *  BT_PATH_AUTO    : TC when the problem qualifies (bf16 activations, channels-last,
 *                    C_in % 64 == 0, groups == 1), else SIMT. */
typedef enum bt_path { BT_PATH_AUTO = 0, BT_PATH_SIMT_F32 = 1, BT_PATH_TC_BF16 = 2 } bt_path;

/* Geometry of one layer call.  rank 0 = Linear (x is [batch, c_in]); rank 1..3 =
 * Conv{1,2,3}d with spatial dims in (D,H,W) order, the first `rank` entries used.
 * Strides are in ELEMENTS of the logical tensor [sample][batch][channel][spatial...],
 * so both contiguous (NCHW) and channels-last (NHWC) tensors are described without
 * a copy.  n_samples > 1 runs S Monte-Carlo weight samples in one launch: sample s
 * uses RNG sample id (sample_id + s), reads x + s*x_sample_stride (0 = all samples
 * share x) and writes out + s*out_sample_stride. */
typedef struct bt_geom {
  int32_t rank;
  int32_t groups;
  int32_t n_samples;
  int32_t x_dtype;   /* bt_dtype */
  int32_t out_dtype; /* bt_dtype */
  int32_t reserved;
  int64_t batch;
  int64_t c_in, c_out;
  int64_t in_size[3], out_size[3], k_size[3], stride[3], pad[3], dil[3];
  int64_t x_sample_stride, x_batch_stride, x_chan_stride, x_sp_stride[3];
  int64_t out_sample_stride, out_batch_stride, out_chan_stride, out_sp_stride[3];
} bt_geom;

/* The variational parameters of one layer (fp32, contiguous, reference layout
 * [c_out, c_in/groups, k...]), its noise source and its prior.
 *  eps_w/eps_b  : non-NULL -> deterministic mode, the caller's N(0,1) draws are used
 *                 (the reference's eps_weight/eps_kernel/eps_bias buffers,
 *                 linear_variational.py:92-100); NULL -> in-register Philox4x32-10
 *                 keyed (seed, sample_id + s, layer_id<<4 | kind, element index).
 *  sign_in/out  : Flipout only, same convention (+-1 floats shaped like x / out,
 *                 same strides as x / out); NULL -> Philox Rademacher bits.
 *  prior_*_t    : optional tensor-valued priors (after utils/util.py:MOPED());
 *                 NULL -> the scalars prior_mu / prior_sigma are used. */
typedef struct bt_layer_params {
  const float* mu_w;
  const float* rho_w;
  const float* mu_b;  /* NULL: no bias */
  const float* rho_b;
  const float* eps_w;
  const float* eps_b;
  const float* sign_in;
  const float* sign_out;
  const float* prior_mu_w_t;
  const float* prior_sigma_w_t;
  const float* prior_mu_b_t;
  const float* prior_sigma_b_t;
  float prior_mu;
  float prior_sigma;
  uint64_t seed;
  uint32_t sample_id;
  uint32_t layer_id;
  /* Inference-time output epilogue (forward only; NULL / 0 = off):
   *   out = act((conv(x, W) + b) * out_scale[n] + out_shift[n] + residual),  act = ReLU when out_relu != 0.
   * Folds the eval-mode BatchNorm (scale = gamma / sqrt(var + eps), shift = beta - mean * scale)
   * and ReLU that follow every Bayesian conv of the reference's ResNets
   * (models/bayesian/resnet_variational_large.py:97-119) into the producing kernel, so the
   * activation tensor is written once instead of three times.  residual (NULL = none) is the
   * identity branch of a residual block (resnet_variational_large.py:113-118, out += identity;
   * relu): a tensor with out's dtype, shape and strides; for bf16 the sum is rounded the way
   * the eager graph rounds it, bf16(bf16(affine) + residual). */
  const float* out_scale;
  const float* out_shift;
  int32_t out_relu;
  /* bit 0 (BT_FLAG_PACKED_VALID): the caller's workspace still holds the packed (mu, sigma) copy the
   * tensor-core forward made on an earlier call with the same parameters, out_scale and geometry
   * (frozen inference models): the packing launch is skipped.  Ignored with tensor-fed eps. */
  int32_t flags;
  const void* residual;
  /* Size in bytes of the `workspace` passed with this call (what the caller allocated from
   * bt_workspace_bytes).  0 = not stated.  When stated, every entry point checks its carve-up of the
   * workspace against it and returns BT_ERR_INVALID instead of writing past the end. */
  int64_t workspace_bytes;
} bt_layer_params;

/* Gradient outputs of a backward call (all fp32 device buffers shaped like the
 * matching parameter; NULL = not wanted).  grad_x has x's dtype and strides. */
typedef struct bt_layer_grads {
  void* grad_x;
  float* grad_mu_w;
  float* grad_rho_w;
  float* grad_mu_b;
  float* grad_rho_b;
} bt_layer_grads;

/* ---- library ------------------------------------------------------------- */
int bt_abi_version(void);
const char* bt_last_error(void);
/* Releases the lazily created per-device contexts (cached TMA descriptors). */
int bt_shutdown(void);
/* Number of kernels this library has launched in this process so far (all threads). */
int64_t bt_launch_count(void);
/* Bytes of scratch the caller must pass as `workspace` to the call named `op`
 * ("kl", "reparam_fwd", "reparam_bwd", "flipout_fwd", "flipout_bwd") for this
 * geometry; 0 when none is needed. */
int64_t bt_workspace_bytes(const char* op, const bt_geom* g, int n_tensors);

/* Which kernel family a forward call with these arguments will run on (BT_PATH_SIMT_F32 or
 * BT_PATH_TC_BF16) given the requested `path`; negative bt_status if it cannot run at all. */
int bt_select_path(const bt_geom* g, const bt_layer_params* p, int flipout, int path);

/* ---- KL: replaces BaseVariationalLayer_.kl_div(..., 'normal') + .mean()
 *      (layers/base_variational_layer.py:55-72), every layer's kl_loss()
 *      (linear_variational.py:146-158, conv_variational.py:353-360) and the Python
 *      loop of get_kl_loss() (models/dnn_to_bnn.py:157-165) with ONE launch over a
 *      table of tensors.
 *  mu[i], rho[i]   : device pointers to tensor i (numel[i] fp32 elements); host arrays.
 *  q_is_sigma      : rho[i] already holds sigma (kl_div called directly) instead of rho.
 *  prior_mu_t/prior_sigma_t : NULL, or host arrays whose entries are NULL or device
 *                    pointers to tensor-valued priors of tensor i.
 *  kl_each         : [n] device floats <- mean over tensor i (what kl_div returns).
 *  kl_total        : NULL or 1 device float <- sum_i kl_each[i] (what get_kl_loss returns).
 *  d_mu[i]/d_rho[i]: NULL, or host arrays of device buffers <- d kl_total / d mu, d rho
 *                    multiplied by *grad_scale (device scalar; NULL = 1).
 *  workspace       : bt_workspace_bytes("kl", NULL, n) bytes. */
int bt_kl_normal_multi(int n, const float* const* mu, const float* const* rho, const int64_t* numel,
                       int q_is_sigma, float prior_mu, float prior_sigma,
                       const float* const* prior_mu_t, const float* const* prior_sigma_t,
                       float* kl_each, float* kl_total, float* const* d_mu, float* const* d_rho,
                       const float* grad_scale, void* workspace, bt_stream_t stream);

/* ---- bt_kl_normal_multi without a caller workspace and without any allocation.
 *      NO COUNTERPART IN THE REFERENCE: added for callers
 *      that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 *      that forbid it).  The fp64 accumulators + done counter live in a file-scope device array
 *      (bt_kl.cu) that every GPU loading the library holds until exit: 1792 bytes, hence
 *      n <= BT_KL_STATIC_MAX_TENSORS; larger n is refused with BT_ERR_INVALID (use
 *      bt_kl_normal_multi).  The entry point calls no cudaMalloc / cudaFree / allocating library
 *      routine: one memset on `stream` that clears the array (it keeps the last call's sums),
 *      then the launch.  Every call on a device shares the array: calls that may overlap on one
 *      GPU (different streams without an ordering between them) race on it -- serialise them.
 *      Other arguments and results exactly as bt_kl_normal_multi. */
#define BT_KL_STATIC_MAX_TENSORS 192
int bt_kl_normal_multi_static(int n, const float* const* mu, const float* const* rho, const int64_t* numel,
                              int q_is_sigma, float prior_mu, float prior_sigma,
                              const float* const* prior_mu_t, const float* const* prior_sigma_t,
                              float* kl_each, float* kl_total, float* const* d_mu, float* const* d_rho,
                              const float* grad_scale, bt_stream_t stream);

/* ---- KL(N(mu, sigma) || Laplace(0, 1)): replaces the 'laplace' branch the fork added to
 *      BaseVariationalLayer_.kl_div (base_variational_layer.py:74-97; location 0 and scale 1
 *      are hard-coded there, whatever mu_p / sigma_p the caller passes), analytic form
 *      log 2 - 0.5 log(2 pi sigma^2) - 0.5 + E|w|, mean over elements.  Arguments as
 *      bt_kl_normal_multi without the prior. */
int bt_kl_laplace_multi(int n, const float* const* mu, const float* const* rho, const int64_t* numel,
                        int q_is_sigma, float* kl_each, float* kl_total, float* const* d_mu, float* const* d_rho,
                        const float* grad_scale, void* workspace, bt_stream_t stream);

/* ---- Reparameterization forward: replaces LinearReparameterization.forward
 *      (linear_variational.py:160-204) and Conv{1,2,3}dReparameterization.forward
 *      (conv_variational.py:197-231, 362-407, 785-819):
 *      out = conv(x, mu + softplus(rho)*eps) + (mu_b + softplus(rho_b)*eps_b).
 *      kl_out: NULL, or 1 device float <- kl(weight) + kl(bias) as forward returns. */
int bt_linear_reparam_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                          float* kl_out, void* workspace, int path, bt_stream_t stream);
int bt_conv_reparam_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                        float* kl_out, void* workspace, int path, bt_stream_t stream);

/* ---- Reparameterization backward: what autograd derives from the reference's op
 *      sequence: grad_x = conv_transpose(grad_out, W);
 *      grad_mu_w = wgrad(x, grad_out); grad_rho_w = grad_mu_w * eps * sigmoid(rho);
 *      bias likewise.  grad_kl: NULL or device scalar -- when given, d(kl)/d(mu,rho)
 *      * grad_kl is added.  W and eps are REGENERATED from the same (seed, sample,
 *      layer) key or re-read from eps_w; nothing is saved by forward. */
int bt_linear_reparam_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                          const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                          bt_stream_t stream);
int bt_conv_reparam_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                        const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                        bt_stream_t stream);

/* ---- Flipout: replaces LinearFlipout.forward (flipout_layers/linear_flipout.py:145-197)
 *      and Conv2dFlipout.forward (flipout_layers/conv_flipout.py:370-439):
 *      out = conv(x, mu) + mu_b + (conv(x*s_in, softplus(rho)*eps) + softplus(rho_b)*eps_b) * s_out,
 *      both contractions in one kernel. */
int bt_linear_flipout_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                          float* kl_out, void* workspace, int path, bt_stream_t stream);
int bt_conv2d_flipout_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                          float* kl_out, void* workspace, int path, bt_stream_t stream);
int bt_linear_flipout_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                          const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                          bt_stream_t stream);
int bt_conv2d_flipout_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                          const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                          bt_stream_t stream);

/* ---- Fused LSTM time step: replaces one iteration of the unrolled cell of LSTMReparameterization.forward
 *      (variational_layers/rnn_variational.py:112-153) / LSTMFlipout.forward (flipout_layers/rnn_flipout.py:102-155):
 *      gates = ih(x_t) + hh(h_prev) with both layers' weights (and Flipout signs) freshly sampled in the kernel,
 *      i, f, o = sigmoid, g = tanh, c_t = f c_prev + i g, h_t = o tanh(c_t) -- ONE launch instead of two linear
 *      layers and ~12 element-wise launches.  fp32 tensors; rows may be strided (x_t is a slice of [batch, seq, in],
 *      h_out / c_out are slots of [batch, seq, hidden]).  Per layer the noise source follows bt_layer_params:
 *      eps / sign pointers non-NULL -> the caller's draws, NULL -> Philox keyed (seed, sample_id, layer_id). */
typedef struct bt_lstm_layer {
  const float* mu_w;  /* [4 * hidden, K] */
  const float* rho_w;
  const float* mu_b;  /* [4 * hidden] or NULL */
  const float* rho_b;
  const float* eps_w;
  const float* eps_b;
  const float* sign_in;  /* Flipout: [batch, K] contiguous */
  const float* sign_out; /* Flipout: [batch, 4 * hidden] contiguous */
  uint64_t seed;
  uint32_t sample_id;
  uint32_t layer_id;
} bt_lstm_layer;

typedef struct bt_lstm_step {
  int64_t batch;
  int32_t in_features;
  int32_t hidden;
  int32_t flipout;
  int32_t reserved;
  bt_lstm_layer ih, hh;
  const float* x;       /* x_t [batch, in_features] */
  int64_t x_row_stride; /* elements */
  const float* h_prev;  /* [batch, hidden] */
  int64_t h_row_stride;
  const float* c_prev;  /* [batch, hidden]; NULL = zeros */
  int64_t c_row_stride;
  float* h_out;
  int64_t h_out_row_stride;
  float* c_out;
  int64_t c_out_row_stride;
  float* gates_out; /* NULL, or [batch, 4 * hidden] contiguous <- the activations i, f, g, o (saved for backward) */
} bt_lstm_step;

int bt_lstm_step_fwd(const bt_lstm_step* step, bt_stream_t stream);
/* Element-wise half of the step's backward: grad_pre [batch, 4 * hidden] (gradient w.r.t. the pre-activation gates)
 * and grad_c_prev [batch, hidden] from grad_h / grad_c (either may be NULL = zero), the saved activations and c_t.
 * The two contractions of the backward are bt_linear_*_bwd calls with grad_out := grad_pre (same keys). */
int bt_lstm_step_bwd_gates(int64_t batch, int hidden, const float* gates, const float* c_prev, int64_t c_prev_row_stride,
                           const float* c_t, int64_t c_t_row_stride, const float* grad_h, int64_t grad_h_row_stride,
                           const float* grad_c, int64_t grad_c_row_stride, float* grad_pre, float* grad_c_prev,
                           bt_stream_t stream);

/* ---- Monte-Carlo reduction: replaces torch.stack(output_mc) -> softmax -> mean(dim=0)
 *      (examples/main_bayesian_cifar_dnn2bnn.py:545,554-556) and the numpy
 *      predictive_entropy / mutual_information (utils/util.py:41-60).
 *      logits [S,B,C] (bt_dtype) -> ACCUMULATES into fp32 sum_p[B,C], sum_p2[B,C] and
 *      sum_plogp[B] (each nullable; caller zeroes them once, all-reduces them across
 *      ranks, then calls bt_mc_finalize).  apply_softmax = 0 reduces the raw values. */
int bt_mc_accumulate(const void* logits, int dtype, int64_t S, int64_t B, int64_t C, int apply_softmax,
                     float* sum_p, float* sum_p2, float* sum_plogp, bt_stream_t stream);
/* mean = sum_p/S; var = sum_p2/S - mean^2; entropy = -sum_c mean log(mean+1e-15);
 * mutual_info = entropy + sum_plogp/S.  Outputs nullable. */
int bt_mc_finalize(const float* sum_p, const float* sum_p2, const float* sum_plogp, int64_t S_total,
                   int64_t B, int64_t C, float* mean, float* var, float* entropy, float* mutual_info,
                   bt_stream_t stream);

/* ---- Residual join of a ResNet block: y = relu(y + r) in place, one pass (replaces the
 *      out += identity; out = relu(out) pair of torchvision / resnet_variational_large.py:113-118). */
int bt_add_relu(void* y, const void* r, int64_t n, int dtype, bt_stream_t stream);
/* ---- Channels-last max pooling between two Bayesian layers (the ResNet stem's MaxPool2d(3, 2, 1)):
 *      x [batch, ih, iw, channels] -> y [batch, oh, ow, channels], floor mode, bf16, channels % 8 == 0. */
int bt_maxpool2d_nhwc(const void* x, void* y, int64_t batch, int64_t channels, int64_t ih, int64_t iw,
                      int kernel, int stride, int pad, int dtype, bt_stream_t stream);

/* ---- RNG stream export (test / interop hook): writes the exact N(0,1) or +-1
 *      values the fused kernels generate in registers for elements [0,n) of stream
 *      (seed, sample_id, layer_id<<4|kind), so they can be fed to the reference's eps
 *      buffers.  kind: 0 eps_weight, 1 eps_bias, 2 sign_in, 3 sign_out. */
int bt_philox_normal_fill(float* out, int64_t n, uint64_t seed, uint32_t sample_id, uint32_t layer_id,
                          int kind, bt_stream_t stream);
int bt_philox_sign_fill(float* out, int64_t n, uint64_t seed, uint32_t sample_id, uint32_t layer_id,
                        int kind, bt_stream_t stream);

/* ---- MOPED init: replaces get_rho (utils/util.py:63-69): rho = log(expm1(delta*|w|) + 1e-20). */
int bt_moped_rho(const float* w, float* rho, int64_t n, float delta, bt_stream_t stream);

/* Diagnostics (not on any product path): with BT_TC2_PROF=1 in the environment the tensor-core
 * forward runs an instrumented instantiation that records, per CTA and warp role, the cycles
 * spent in each kind of pipeline wait.  Copies the counters of the last such launch to
 * host_out[160 ctas][6 roles][8 slots]; roles: 0 epilogue, 1 TMA loader, 2/3 MMA issuers,
 * 4/5 first/last W-producer warp; the last slot of each role is its total cycle count. */
int bt_debug_tc_profile(unsigned long long* host_out, int n_words);

#ifdef __cplusplus
}
#endif
#endif /* BT_B200_H_ */
