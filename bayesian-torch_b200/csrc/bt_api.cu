// extern "C" surface of libbt_b200.so: argument validation, path selection, error text.
#include <stdarg.h>
#include <string.h>

#include "bt_common.cuh"

int64_t bt_kl_workspace_bytes_internal(int n);

static thread_local char g_err[512] = "";

void bt_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int bt_num_sms() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

#include <atomic>
static std::atomic<long long> g_launches{0};
void bt_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" int64_t bt_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }

extern "C" int bt_abi_version(void) { return BT_ABI_VERSION; }
extern "C" const char* bt_last_error(void) { return g_err; }
extern "C" int bt_shutdown(void) { return BT_OK; }

static int64_t weight_numel(const bt_geom* g) {
  int64_t t = 1;
  for (int i = 0; i < g->rank; ++i) t *= g->k_size[i];
  return g->c_out * (g->c_in / g->groups) * t;
}

#define KL_SCRATCH 256 /* bytes reserved at the head of every layer workspace for the fused KL */

extern "C" int64_t bt_workspace_bytes(const char* op, const bt_geom* g, int n_tensors) {
  if (!op) return 0;
  if (!strcmp(op, "kl")) return bt_kl_workspace_bytes_internal(n_tensors);
  if (!strcmp(op, "reparam_fwd") || !strcmp(op, "flipout_fwd"))
    return KL_SCRATCH + (g ? bt_tc_workspace_bytes(g, !strcmp(op, "flipout_fwd")) : 0);
  if (!strcmp(op, "reparam_bwd") || !strcmp(op, "flipout_bwd"))
  {  // dgrad and wgrad run one after the other on the stream and share the area
    if (!g) return KL_SCRATCH;
    const bool flip = !strcmp(op, "flipout_bwd");
    const int64_t d = bt_tc_dgrad_workspace_bytes(g, flip), w = bt_tc_wgrad_workspace_bytes(g, flip);
    return KL_SCRATCH + (d > w ? d : w);
  }
  return 0;
}

static int check_common(const bt_geom* g, const bt_layer_params* p, const void* x, const void* out, const char* who) {
  BT_REQUIRE(g && p, BT_ERR_INVALID, "%s: null argument", who);
  BT_REQUIRE((x && out) || g->batch == 0, BT_ERR_INVALID, "%s: null tensor pointer", who);
  BT_REQUIRE(p->mu_w && p->rho_w, BT_ERR_INVALID, "%s: mu_w / rho_w are required", who);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    bt_set_error("%s: no CUDA device (there is no CPU fallback)", who);
    return BT_ERR_CUDA;
  }
  return BT_OK;
}

extern "C" int bt_select_path(const bt_geom* g, const bt_layer_params* p, int flipout, int path) {
  BT_REQUIRE(g && p, BT_ERR_INVALID, "bt_select_path: null argument");
  const char* why = "";
  const bool tc = bt_tc_supported(g, p, flipout != 0, &why);
  if (path == BT_PATH_TC_BF16) {
    BT_REQUIRE(tc, BT_ERR_UNSUPPORTED, "tensor-core path unavailable: %s", why);
    return BT_PATH_TC_BF16;
  }
  if (path == BT_PATH_AUTO) return tc ? BT_PATH_TC_BF16 : BT_PATH_SIMT_F32;
  BT_REQUIRE(path == BT_PATH_SIMT_F32, BT_ERR_INVALID, "bt_select_path: unknown path %d", path);
  return BT_PATH_SIMT_F32;
}

static int forward_common(const bt_geom* g, const bt_layer_params* p, const void* x, void* out, float* kl_out,
                          void* workspace, int path, bool flipout, const char* who, bt_stream_t stream) {
  int rc = check_common(g, p, x, out, who);
  if (rc != BT_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  BT_REQUIRE(!kl_out || workspace, BT_ERR_INVALID, "%s: kl_out needs a workspace (bt_workspace_bytes)", who);
  // the tensor-core forward evaluates the KL term inside its parameter-packing pass (mu / rho read once); the
  // fp32 kernels read the parameters tile by tile, many times: there the KL is its own HBM-bound launch
  const char* why0 = "";
  const bool on_tc = workspace && (path == BT_PATH_TC_BF16 || path == BT_PATH_AUTO) && bt_tc_supported(g, p, flipout, &why0);
  if (kl_out && !on_tc) {
    rc = bt_kl_layer(p, weight_numel(g), p->mu_b ? g->c_out : 0, kl_out, workspace, st);
    if (rc != BT_OK) return rc;
  }
  void* tc_ws = workspace ? (char*)workspace + KL_SCRATCH : nullptr;
  BT_REQUIRE(p->workspace_bytes == 0 || p->workspace_bytes >= KL_SCRATCH, BT_ERR_INVALID, "%s: workspace_bytes %lld is smaller than the KL scratch",
             who, (long long)p->workspace_bytes);
  const int64_t tc_ws_bytes = p->workspace_bytes > 0 ? p->workspace_bytes - KL_SCRATCH : 0;
  const char* why = "";
  if (path == BT_PATH_TC_BF16) {
    BT_REQUIRE(bt_tc_supported(g, p, flipout, &why), BT_ERR_UNSUPPORTED, "%s: tensor-core path unavailable: %s", who,
               why);
    return bt_tc_forward(g, p, x, out, tc_ws, tc_ws_bytes, flipout, kl_out, workspace, st);
  }
  if (path == BT_PATH_AUTO && bt_tc_supported(g, p, flipout, &why) && tc_ws)
    return bt_tc_forward(g, p, x, out, tc_ws, tc_ws_bytes, flipout, kl_out, workspace, st);
  BT_REQUIRE(path == BT_PATH_AUTO || path == BT_PATH_SIMT_F32, BT_ERR_INVALID, "%s: unknown path %d", who, path);
  return bt_simt_forward(g, p, x, out, flipout, st);
}

static int backward_common(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                           const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path, bool flipout,
                           const char* who, bt_stream_t stream) {
  int rc = check_common(g, p, x, grad_out, who);
  if (rc != BT_OK) return rc;
  BT_REQUIRE(grads, BT_ERR_INVALID, "%s: null grads", who);
  BT_REQUIRE(!p->out_scale && !p->out_shift && !p->out_relu && !p->residual, BT_ERR_UNSUPPORTED,
             "%s: the folded BatchNorm / ReLU / residual epilogue is inference-only", who);
  cudaStream_t st = (cudaStream_t)stream;
  BT_REQUIRE(p->workspace_bytes == 0 || p->workspace_bytes >= KL_SCRATCH, BT_ERR_INVALID, "%s: workspace_bytes %lld is smaller than the KL scratch",
             who, (long long)p->workspace_bytes);
  const int64_t tc_ws_bytes = p->workspace_bytes > 0 ? p->workspace_bytes - KL_SCRATCH : 0;
  bt_layer_grads rest = *grads;
  // grad_x on the tensor-core kernel when the problem qualifies (bf16 channels-last activations): the forward
  // kernel run on grad_out with the transposed, regenerated weights
  if (grads->grad_x && path != BT_PATH_SIMT_F32 && workspace) {
    const char* why = "";
    if (bt_tc_dgrad_supported(g, p, flipout, &why)) {
      rc = bt_tc_dgrad(g, p, grad_out, grads->grad_x, (char*)workspace + KL_SCRATCH, tc_ws_bytes, flipout, st);
      if (rc != BT_OK) return rc;
      rest.grad_x = nullptr;
    } else {
      BT_REQUIRE(path != BT_PATH_TC_BF16, BT_ERR_UNSUPPORTED, "%s: tensor-core backward unavailable: %s", who, why);
    }
  }
  // weight gradients likewise (bias gradients stay on the small SIMT reduction)
  if (grads->grad_mu_w && grads->grad_rho_w && path != BT_PATH_SIMT_F32 && workspace) {
    const char* why = "";
    if (bt_tc_wgrad_supported(g, p, flipout, &why)) {
      rc = bt_tc_wgrad(g, p, x, grad_out, grad_kl, grads->grad_mu_w, grads->grad_rho_w, (char*)workspace + KL_SCRATCH, tc_ws_bytes, flipout, st);
      if (rc != BT_OK) return rc;
      rest.grad_mu_w = rest.grad_rho_w = nullptr;
    } else {
      BT_REQUIRE(path != BT_PATH_TC_BF16, BT_ERR_UNSUPPORTED, "%s: tensor-core backward unavailable: %s", who, why);
    }
  }
  if (!rest.grad_x && !rest.grad_mu_w && !rest.grad_rho_w && !rest.grad_mu_b && !rest.grad_rho_b) return BT_OK;
  return bt_simt_backward(g, p, x, grad_out, grad_kl, &rest, workspace, flipout, st);
}

extern "C" int bt_linear_reparam_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                                     float* kl_out, void* workspace, int path, bt_stream_t stream) {
  BT_REQUIRE(g && g->rank == 0, BT_ERR_INVALID, "bt_linear_reparam_fwd: geometry rank must be 0");
  return forward_common(g, p, x, out, kl_out, workspace, path, false, "bt_linear_reparam_fwd", stream);
}
extern "C" int bt_conv_reparam_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                                   float* kl_out, void* workspace, int path, bt_stream_t stream) {
  BT_REQUIRE(g && g->rank >= 1 && g->rank <= 3, BT_ERR_INVALID, "bt_conv_reparam_fwd: geometry rank must be 1..3");
  return forward_common(g, p, x, out, kl_out, workspace, path, false, "bt_conv_reparam_fwd", stream);
}
extern "C" int bt_linear_flipout_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                                     float* kl_out, void* workspace, int path, bt_stream_t stream) {
  BT_REQUIRE(g && g->rank == 0, BT_ERR_INVALID, "bt_linear_flipout_fwd: geometry rank must be 0");
  return forward_common(g, p, x, out, kl_out, workspace, path, true, "bt_linear_flipout_fwd", stream);
}
extern "C" int bt_conv2d_flipout_fwd(const bt_geom* g, const bt_layer_params* p, const void* x, void* out,
                                     float* kl_out, void* workspace, int path, bt_stream_t stream) {
  BT_REQUIRE(g && g->rank >= 1 && g->rank <= 3, BT_ERR_INVALID, "bt_conv2d_flipout_fwd: geometry rank must be 1..3");
  return forward_common(g, p, x, out, kl_out, workspace, path, true, "bt_conv2d_flipout_fwd", stream);
}

extern "C" int bt_linear_reparam_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                                     const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                                     bt_stream_t stream) {
  BT_REQUIRE(g && g->rank == 0, BT_ERR_INVALID, "bt_linear_reparam_bwd: geometry rank must be 0");
  return backward_common(g, p, x, grad_out, grad_kl, grads, workspace, path, false, "bt_linear_reparam_bwd", stream);
}
extern "C" int bt_conv_reparam_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                                   const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                                   bt_stream_t stream) {
  BT_REQUIRE(g && g->rank >= 1 && g->rank <= 3, BT_ERR_INVALID, "bt_conv_reparam_bwd: geometry rank must be 1..3");
  return backward_common(g, p, x, grad_out, grad_kl, grads, workspace, path, false, "bt_conv_reparam_bwd", stream);
}
extern "C" int bt_linear_flipout_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                                     const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                                     bt_stream_t stream) {
  BT_REQUIRE(g && g->rank == 0, BT_ERR_INVALID, "bt_linear_flipout_bwd: geometry rank must be 0");
  return backward_common(g, p, x, grad_out, grad_kl, grads, workspace, path, true, "bt_linear_flipout_bwd", stream);
}
extern "C" int bt_conv2d_flipout_bwd(const bt_geom* g, const bt_layer_params* p, const void* x, const void* grad_out,
                                     const float* grad_kl, const bt_layer_grads* grads, void* workspace, int path,
                                     bt_stream_t stream) {
  BT_REQUIRE(g && g->rank >= 1 && g->rank <= 3, BT_ERR_INVALID, "bt_conv2d_flipout_bwd: geometry rank must be 1..3");
  return backward_common(g, p, x, grad_out, grad_kl, grads, workspace, path, true, "bt_conv2d_flipout_bwd", stream);
}
