// Geometry normalisation shared by the SIMT fp32 kernels (forward / dgrad / wgrad).
#pragma once
#include "bt_common.cuh"

#define BT_MAX_TAPS 512

// Every layer is normalised to a 3-d convolution: Linear = 1x1x1 conv over a
// [batch, c_in, 1, 1, 1] tensor, Conv1d uses the W slot, Conv2d the H,W slots.
struct SGeo {
  int64_t batch, M;  // M = batch * OD*OH*OW (output positions)
  int64_t K;         // Cg * T  (reduction length of the forward contraction)
  int32_t Cg, Ng, groups, T, n_samples;
  int32_t ID, IH, IW, OD, OH, OW, KD, KH, KW;
  int32_t sd, sh, sw, pd, ph, pw, dd, dh, dw;
  int64_t xss, xbs, xcs, xs0, xs1, xs2;
  int64_t oss, obs, ocs, os0, os1, os2;
};

struct SParams {
  const float *mu_w, *rho_w, *mu_b, *rho_b, *eps_w, *eps_b, *sign_in, *sign_out;
  const float *out_scale, *out_shift;  // folded BatchNorm epilogue (forward only)
  int out_relu;
  const void* residual;  // out-shaped, out dtype (forward only)
  int dgrad_bias;        // backward-data run as the FORWARD of a ConvTranspose layer: add the sampled bias to grad_x
  uint64_t seed;
  uint32_t sample_id, layer_id;
};

static inline int bt_sgeo_from(const bt_geom* g, SGeo* o, const char* who) {
  BT_REQUIRE(g != nullptr, BT_ERR_INVALID, "%s: null geometry", who);
  BT_REQUIRE(g->rank >= 0 && g->rank <= 3, BT_ERR_INVALID, "%s: rank %d outside [0,3]", who, g->rank);
  BT_REQUIRE(g->batch >= 0 && g->c_in > 0 && g->c_out > 0 && g->groups > 0, BT_ERR_INVALID,
             "%s: bad sizes batch=%lld c_in=%lld c_out=%lld groups=%d", who, (long long)g->batch,
             (long long)g->c_in, (long long)g->c_out, g->groups);
  // same message as the reference constructors (conv_variational.py:270-273)
  BT_REQUIRE(g->c_in % g->groups == 0 && g->c_out % g->groups == 0, BT_ERR_INVALID, "%s: invalid in_channels size",
             who);
  BT_REQUIRE(g->n_samples >= 1, BT_ERR_INVALID, "%s: n_samples must be >= 1", who);
  int64_t in[3] = {1, 1, 1}, out[3] = {1, 1, 1}, k[3] = {1, 1, 1}, s[3] = {1, 1, 1}, p[3] = {0, 0, 0},
          d[3] = {1, 1, 1}, xs[3] = {0, 0, 0}, os[3] = {0, 0, 0};
  for (int i = 0; i < g->rank; ++i) {
    const int j = 3 - g->rank + i;
    in[j] = g->in_size[i]; out[j] = g->out_size[i]; k[j] = g->k_size[i];
    s[j] = g->stride[i]; p[j] = g->pad[i]; d[j] = g->dil[i];
    xs[j] = g->x_sp_stride[i]; os[j] = g->out_sp_stride[i];
    BT_REQUIRE(in[j] > 0 && k[j] > 0 && s[j] > 0 && d[j] > 0 && p[j] >= 0, BT_ERR_INVALID,
               "%s: bad conv parameters in spatial dim %d", who, i);
    const int64_t expect = (in[j] + 2 * p[j] - d[j] * (k[j] - 1) - 1) / s[j] + 1;
    BT_REQUIRE(out[j] == expect && out[j] > 0, BT_ERR_INVALID, "%s: out_size[%d]=%lld but conv arithmetic gives %lld",
               who, i, (long long)out[j], (long long)expect);
  }
  const int64_t T = k[0] * k[1] * k[2];
  BT_REQUIRE(T <= BT_MAX_TAPS, BT_ERR_UNSUPPORTED, "%s: kernel volume %lld > %d taps", who, (long long)T, BT_MAX_TAPS);
  o->batch = g->batch;
  o->M = g->batch * out[0] * out[1] * out[2];
  o->groups = g->groups;
  o->Cg = (int32_t)(g->c_in / g->groups);
  o->Ng = (int32_t)(g->c_out / g->groups);
  o->T = (int32_t)T;
  o->K = (int64_t)o->Cg * T;
  o->n_samples = g->n_samples;
  o->ID = (int32_t)in[0]; o->IH = (int32_t)in[1]; o->IW = (int32_t)in[2];
  o->OD = (int32_t)out[0]; o->OH = (int32_t)out[1]; o->OW = (int32_t)out[2];
  o->KD = (int32_t)k[0]; o->KH = (int32_t)k[1]; o->KW = (int32_t)k[2];
  o->sd = (int32_t)s[0]; o->sh = (int32_t)s[1]; o->sw = (int32_t)s[2];
  o->pd = (int32_t)p[0]; o->ph = (int32_t)p[1]; o->pw = (int32_t)p[2];
  o->dd = (int32_t)d[0]; o->dh = (int32_t)d[1]; o->dw = (int32_t)d[2];
  o->xss = g->x_sample_stride; o->xbs = g->x_batch_stride; o->xcs = g->x_chan_stride;
  o->xs0 = xs[0]; o->xs1 = xs[1]; o->xs2 = xs[2];
  o->oss = g->out_sample_stride; o->obs = g->out_batch_stride; o->ocs = g->out_chan_stride;
  o->os0 = os[0]; o->os1 = os[1]; o->os2 = os[2];
  return BT_OK;
}

static inline SParams bt_sparams_from(const bt_layer_params* p) {
  SParams s;
  s.mu_w = p->mu_w; s.rho_w = p->rho_w; s.mu_b = p->mu_b; s.rho_b = p->rho_b;
  s.eps_w = p->eps_w; s.eps_b = p->eps_b; s.sign_in = p->sign_in; s.sign_out = p->sign_out;
  s.out_scale = p->out_scale; s.out_shift = p->out_shift; s.out_relu = p->out_relu; s.residual = p->residual;
  s.dgrad_bias = (p->flags & BT_FLAG_TRANSPOSE_BIAS) && p->mu_b ? 1 : 0;
  s.seed = p->seed; s.sample_id = p->sample_id; s.layer_id = p->layer_id;
  return s;
}
