// KL(N(mu, softplus(rho)) || N(mu_p, sigma_p)) per element: shared by the multi-tensor KL kernel (bt_kl.cu) and the
// parameter-packing pass of the tensor-core forward (bt_tc.cu), which evaluates the layer's KL while it reads
// mu / rho anyway (one pass over the parameters: 8 B per weight for sampling AND KL).
#pragma once
#include "bt_common.cuh"

struct KlElem {
  float kl, dmu, drho;
};

// Per-prior constants of the KL term: log(sigma_p) and 1 / sigma_p^2.  With the reference's
// constant-filled priors they are loop invariants (computed once per thread, not per element).
struct KlPrior {
  float log_ps, inv2;
};
__device__ __forceinline__ KlPrior kl_prior(float ps) {
  KlPrior p;
  p.log_ps = logf(ps);
  p.inv2 = __frcp_rn(ps * ps);
  return p;
}

// The pass is HBM-bound only if an element costs ~20 issue slots: softplus and the two logs use
// the MUFU approximations (ex2 / lg2, relative error ~2^-22) with a series for small exp(rho),
// where log(1 + t) computed as such would lose the low bits of t.  Checked against the fp64
// oracle at rtol 1e-5 on the mean (tests/test_parity_gpu.py).
template <bool kGrad>
__device__ __forceinline__ KlElem kl_elem(float mu, float rho, float pm, const KlPrior pr, bool q_is_sigma) {
  // sigma = log1p(exp(rho)) (linear_variational.py:163); sigmoid(rho) = t/(1+t) with t = exp(rho)
  float sig, dsig_drho;
  if (q_is_sigma) {
    sig = rho;
    dsig_drho = 1.f;
  } else {
    const float t = __expf(fminf(rho, 30.f));
    const float series = t * (1.f - t * (0.5f - t * (0.33333334f - 0.25f * t)));  // log1p(t), t < 2^-5: error < t^5/5
    const float big = __logf(1.f + t);
    sig = rho > 20.f ? rho : (t < 0.03125f ? series : big);
    dsig_drho = kGrad ? (rho > 20.f ? 1.f : __fdividef(t, 1.f + t)) : 0.f;
  }
  const float d = mu - pm;
  KlElem e;
  // log(sigma_p) - log(sigma_q) + (sigma_q^2 + (mu_q-mu_p)^2) / (2 sigma_p^2) - 0.5
  e.kl = (pr.log_ps - __logf(sig)) + 0.5f * (sig * sig + d * d) * pr.inv2 - 0.5f;
  if (kGrad) {
    e.dmu = d * pr.inv2;
    e.drho = (sig * pr.inv2 - __frcp_rn(sig)) * dsig_drho;
  } else {
    e.dmu = 0.f;
    e.drho = 0.f;
  }
  return e;
}

