// bb_alpha.cuh -- random-walk Metropolis step on alpha when it is unknown (`true_alpha <= 0`),
// device form of BridgeRegression::sample_alpha_marg / llh_alpha_marg,
// /root/reference/Code/C/BridgeRegression.cpp:468-503 (step width ep = 0.1, BridgeRegression.h:155).
//
//   s_j     = log|beta_j / tau|
//   llh(a)  = P log a - P lgamma(1/a) - sum_j exp(a s_j)
//   a_new   ~ U(max(0, a_old - ep), min(1, a_old + ep))                      (block 0 of stage ST_ALPHA)
//   log acc = llh(a_new) - llh(a_old) + log dbeta(a_new) - log dbeta(a_old) + log d_old - log d_new
//   keep a_old when unif() > exp(log acc)                                   (block 1)
//
// The two sums over j are the only O(P) work; every sweep kernel accumulates them with the threads that
// hold beta (alpha_terms) and reduces them the way it reduces the tau statistic; alpha_propose and
// alpha_decide are O(1) and evaluated by one lane (or redundantly by all lanes of a warp).
#pragma once
#include "bb_rng.cuh"

namespace bb {

constexpr double ALPHA_EP = 0.1;

struct AlphaProposal {
  double a_old, a_new;
  double log_d;   // log(d_old) - log(d_new): ratio of the two proposal interval widths
};

__device__ __forceinline__ AlphaProposal alpha_propose(const Stream& s, double a_old) {
  AlphaProposal p;
  p.a_old = a_old;
  const double l_new = fmax(0.0, a_old - ALPHA_EP), r_new = fmin(1.0, a_old + ALPHA_EP);
  const double d_new = r_new - l_new;
  p.a_new = l_new + d_new * s.unif(0);                     // r.flat(l_new, r_new)
  const double l_old = fmax(0.0, p.a_new - ALPHA_EP), r_old = fmin(1.0, p.a_new + ALPHA_EP);
  p.log_d = log(r_old - l_old) - log(d_new);
  return p;
}

// adds coefficient j's terms exp(a s_j) of the two log-likelihoods
__device__ __forceinline__ void alpha_terms(double beta, double tau, const AlphaProposal& p, double& sum_new, double& sum_old) {
  const double s = log(fabs(beta / tau));
  sum_new += exp(p.a_new * s);
  sum_old += exp(p.a_old * s);
}

__device__ __forceinline__ double alpha_log_dbeta(double x, double a, double b) {   // log(r.d_beta(x, a, b))
  return log(exp(lgamma(a + b) - lgamma(a) - lgamma(b) + (a - 1.0) * log(x) + (b - 1.0) * log1p(-x)));
}

static __device__ __noinline__ double alpha_decide(const Stream& s, const AlphaProposal& p, double sum_new, double sum_old, int P,
                                            double pr_a, double pr_b) {
  const double pp = (double)P;
  const double llh_new = pp * log(p.a_new) - pp * lgamma(1.0 / p.a_new) - sum_new;
  const double llh_old = pp * log(p.a_old) - pp * lgamma(1.0 / p.a_old) - sum_old;
  const double log_accept = llh_new - llh_old + alpha_log_dbeta(p.a_new, pr_a, pr_b) - alpha_log_dbeta(p.a_old, pr_a, pr_b) + p.log_d;
  return (s.unif(1) > exp(log_accept)) ? p.a_old : p.a_new;
}

}  // namespace bb
