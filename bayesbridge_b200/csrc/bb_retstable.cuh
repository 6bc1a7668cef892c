// bb_retstable.cuh -- exponentially tilted positive-stable draw (Devroye 2009 double rejection),
// device form of /root/reference/Code/C/retstable.cpp:94-271 (helpers sinc_MM :18-29, A_ :40-47,
// BdB0 :73-77) with V0 = 1 as BridgeRegression.cpp:509 calls it.
//
// The draw is organised as ROUNDS so that a kernel can compact the unfinished items of a CTA
// between rounds (the "retry queue"): one round = one inner trial (3 Philox blocks) and, if the
// inner trial accepts, one outer trial (2 more blocks).  The block index k advances exactly as the
// sequential reference consumes its base draws, so the accepted value does not depend on how the
// rounds were scheduled.
#pragma once
#include "bb_rng.cuh"

namespace bb {

__device__ __forceinline__ double sinc_mm(double x) {
  const double ax = fabs(x);
  if (ax < 0.006) {
    if (x == 0.) return 1.;
    const double x2 = x * x;
    if (ax < 2e-4) return 1. - x2 / 6.;
    return 1. - x2 / 6. * (1. - x2 / 20.);
  }
  return sin(x) / x;
}

constexpr double RS_PI = 3.14159265358979323846;
constexpr double RS_SQRT_PI = 1.7724538509055159;   // sqrt(M_PI)
constexpr double RS_C1 = 1.2533141373155001;        // sqrt(M_PI_2)
constexpr double RS_SQRT2 = 1.41421356237309504880;

// per-item constants (retstable.cpp:121-147); h = beta_j^2/tau^2, alpha = alpha_bridge/2
struct RetConst {
  double h, alpha, b, lambda_alpha, gamma, sgamma, xi, psi, w1, w2, w3;
  __device__ __forceinline__ void init(double h_, double alpha_) {
    h = h_; alpha = alpha_;
    const double c2 = 2. + RS_C1;
    b = (1. - alpha) / alpha;
    lambda_alpha = pow(h, alpha);
    gamma = lambda_alpha * alpha * (1. - alpha);
    sgamma = sqrt(gamma);
    const double c3 = c2 * sgamma;
    xi = (1. + RS_SQRT2 * c3) / RS_PI;
    psi = c3 * exp(-gamma * RS_PI * RS_PI / 8.) / RS_SQRT_PI;
    w1 = RS_C1 * xi / sgamma;
    w2 = 2. * RS_SQRT_PI * psi;
    w3 = xi * RS_PI;
  }
};

// One round starting at block k.  Returns 2 = accepted (value in `out`), otherwise the number of
// blocks consumed is added to k and the caller retries: 0 = inner rejected, 1 = outer rejected.
__device__ inline int retstable_round(const RetConst& c, const Stream& s, uint32_t& k, double& out) {
  const double alpha = c.alpha, Ialpha = 1. - alpha;
  // ---- inner trial (retstable.cpp:162-207) ----
  const double V = s.unif(k);
  double U;
  if (c.gamma >= 1) {
    if (V < c.w1 / (c.w1 + c.w2)) U = fabs(s.norm(k + 1)) / c.sgamma;
    else { const double W_ = s.unif(k + 1); U = RS_PI * (1. - W_ * W_); }
  } else {
    const double W_ = s.unif(k + 1);
    if (V < c.w3 / (c.w2 + c.w3)) U = RS_PI * W_;
    else U = RS_PI * (1. - W_ * W_);
  }
  const double W = s.unif(k + 2);
  const double sA = sinc_mm(alpha * U), sI = sinc_mm(Ialpha * U), sU = sinc_mm(U);
  const double den = pow(sA, alpha) * pow(sI, Ialpha);      // BdB0 (retstable.cpp:73-77)
  const double zeta = sqrt(sU / den);
  const double z = 1 / (1 - pow(1 + alpha * zeta / c.sgamma, -1 / alpha));
  double rho = RS_PI * exp(-c.lambda_alpha * (1. - 1. / (zeta * zeta))) / ((1. + RS_C1) * c.sgamma / zeta + z);
  double d = 0.;
  if (U >= 0 && c.gamma >= 1) d += c.xi * exp(-c.gamma * U * U / 2.);
  if (U > 0 && U < RS_PI) d += c.psi / sqrt(RS_PI - U);
  if (U >= 0 && U <= RS_PI && c.gamma < 1) d += c.xi;
  rho *= d;
  const double Z = W * rho;
  if (!(U < RS_PI && Z <= 1.)) { k += 3; return 0; }
  // ---- outer trial (retstable.cpp:212-256) ----
  const double Aval = pow(Ialpha * sI, Ialpha) * pow(alpha * sA, alpha) / sU;   // A_ (retstable.cpp:40-47)
  const double a = pow(Aval, 1. / (1. - alpha));
  const double m = pow(c.b / a, alpha) * c.lambda_alpha;
  const double delta = sqrt(m * alpha / a);
  const double a1 = delta * RS_C1, a2 = delta, a3 = z / a;
  const double ssum = a1 + a2 + a3;
  const double V_ = s.unif(k + 3);
  double N_ = 0., E_ = 0., X;
  if (V_ < a1 / ssum) { N_ = s.norm(k + 4); X = m - delta * fabs(N_); }
  else if (V_ < (a1 + a2) / ssum) X = m + delta * s.unif(k + 4);
  else { E_ = -log(s.unif(k + 4)); X = m + delta + E_ * a3; }
  const double E = -log(Z);
  double cc = a * (X - m);
  cc += (m != 0) ? c.h * (pow(X, -1. * c.b) - pow(m, -1. * c.b)) : 0.0;
  if (X < m) cc -= N_ * N_ / 2.;
  else if (X > m + delta) cc -= E_;
  k += 5;
  if (!(X >= 0 && cc <= E)) return 1;
  out = exp(-c.b * log(X));   // exp(1/alpha*log(V0) - b*log(X)) with V0 = 1 (retstable.cpp:270)
  return 2;
}

// sequential form, one thread per item
__device__ inline double retstable_scalar(const Stream& s, double h, double alpha, uint32_t& status) {
  if (alpha == 1.) return 1.0;  // point mass at V0 (retstable.cpp:104-110)
  RetConst c; c.init(h, alpha);
  uint32_t k = 0; double out = 0.;
  while (k < K_CAP - 8) {
    if (retstable_round(c, s, k, out) == 2) return out;
  }
  status |= ST_TRIAL_CAP;
  return nan("");
}

}  // namespace bb
