// bb_rng.cuh -- counter-based Philox4x32-10 and the derived draws used by every kernel.
//
// Replaces the role of the reference's RNG library (RNG.hpp, included at
// /root/reference/Code/C/BridgeRegression.h:68; call sites BridgeRegression.cpp:103,120-124,
// 137-143,273,280,392,449,462,519,568 and retstable.cpp:172-235).  The draw a kernel consumes for
// (chain, iteration, stage, item, trial k) is one Philox block
//     ctr = { (stage<<24)|k , item , iter , chain } ,  key = { seed_lo, seed_hi }
// so results do not depend on how chains are scheduled or sharded over GPUs.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "bb_math.cuh"

namespace bb {

enum Stage : uint32_t {
  ST_TAU = 1, ST_SIG2 = 2, ST_OMEGA = 3, ST_U = 4, ST_BETA = 5, ST_LAMBDA = 6,
  ST_BETA_STB = 7, ST_TAU_INIT = 8, ST_ALPHA = 9, ST_SCALAR = 10
};

constexpr uint32_t K_CAP = 1u << 24;  // a rejection loop that reaches this many blocks is flagged
constexpr double TWO_PI = 6.283185307179586;
constexpr double SQRT_2PI = 2.5066282746310002;

struct RngKey { uint32_t seed_lo, seed_hi; };

// status bits reported per chain
enum Status : uint32_t { OK = 0, ST_NAN = 1, ST_EMPTY_INTERVAL = 2, ST_TRIAL_CAP = 4, ST_CHOL_FAIL = 8 };

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__device__ __forceinline__ double u53(uint32_t lo, uint32_t hi) {
  const uint64_t x = (static_cast<uint64_t>(hi) << 32) | lo;
  return (static_cast<double>(x >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}

// One item's stream: block(k) for k = 0,1,2,...
struct Stream {
  uint32_t k0, k1;     // key
  uint32_t stage_hi;   // stage << 24
  uint32_t item, iter, chain;
  __device__ __forceinline__ Stream(RngKey key, uint32_t chain_, uint32_t iter_, uint32_t stage, uint32_t item_)
      : k0(key.seed_lo), k1(key.seed_hi), stage_hi(stage << 24), item(item_), iter(iter_), chain(chain_) {}
  __device__ __forceinline__ void block(uint32_t k, double& ua, double& ub) const {
    uint32_t o[4];
    philox4x32_10(stage_hi | (k & (K_CAP - 1)), item, iter, chain, k0, k1, o);
    ua = u53(o[0], o[1]);
    ub = u53(o[2], o[3]);
  }
  __device__ __forceinline__ double unif(uint32_t k) const { double a, b; block(k, a, b); return a; }
  __device__ __forceinline__ double norm(uint32_t k) const {
    double a, b; block(k, a, b);
    return sqrt(-2.0 * log(a)) * cos(TWO_PI * b);
  }
};

__device__ __forceinline__ double box_muller(double ua, double ub) { return sqrt(-2.0 * log(ua)) * cos(TWO_PI * ub); }

// ---------------------------------------------------------------------------------------------
// Truncated normal, x ~ N(0,1) on [a,b]  (Robert 1995; the reference documents that choice at
// /root/reference/Code/BBPackage/BayesBridge/man/rtnorm.Rd:13,46-47).  One trial = one block.
// `tn_setup` classifies the interval once; `tn_trial` evaluates trial k and says whether it accepts.
// The same pair serves the scalar loop (one thread walks k) and the warp-parallel form
// (lane l evaluates trial base+l; the lowest accepting lane wins), which yields the identical
// draw because acceptance of trial k does not depend on earlier trials.
// ---------------------------------------------------------------------------------------------
struct TnPlan {
  int mode;        // 0 normal proposal, 1 uniform around 0, 2 exponential tail, 3 uniform tail, 4 degenerate, 5 nan
  double a, b;     // bounds in the working orientation (tail modes: 0 < a < b)
  double lam;      // exponential rate
  double sgn;      // +1 / -1 (left tail is mirrored)
};

__device__ __forceinline__ TnPlan tn_setup(double a, double b) {
  TnPlan p; p.sgn = 1.0; p.lam = 0.0; p.a = a; p.b = b;
  if (a != a || b != b) { p.mode = 5; return p; }
  if (!(a < b)) { p.mode = 4; return p; }
  if (a > 0.0 || b < 0.0) {
    if (b < 0.0) { p.sgn = -1.0; p.a = -b; p.b = -a; }
    const double aa = p.a;
    const double sq = sqrt(aa * aa + 4.0);
    p.lam = 0.5 * (aa + sq);
    const double thresh = aa + exp(0.5 + 0.25 * (aa * aa - aa * sq)) / p.lam;
    p.mode = (p.b > thresh) ? 2 : 3;
    return p;
  }
  p.mode = (b - a >= SQRT_2PI) ? 0 : 1;
  return p;
}

// evaluate one trial from its two uniforms; returns acceptance, proposal in x (working orientation)
__device__ __forceinline__ bool tn_eval(const TnPlan& p, double ua, double ub, double& x) {
  switch (p.mode) {
    case 0: x = box_muller(ua, ub); return x >= p.a && x <= p.b;
    case 1: x = p.a + (p.b - p.a) * ua; return ub <= exp(-0.5 * x * x);
    case 2: { x = p.a - log(ua) / p.lam; const double t = x - p.lam; return x <= p.b && ub <= exp(-0.5 * t * t); }
    default: x = p.a + (p.b - p.a) * ua; return ub <= exp(0.5 * (p.a * p.a - x * x));
  }
}
__device__ __forceinline__ bool tn_trial(const TnPlan& p, const Stream& s, uint32_t k, double& x) {
  double ua, ub;
  s.block(k, ua, ub);
  return tn_eval(p, ua, ub, x);
}

// ---- latency-trimmed forms for the sweep (bb_tri.cuh): same plan, same trial, shorter dependent chain ----
// tn_setup_fast decides exponential-vs-uniform tail proposals without sqrt/exp/div whenever the interval is
// clearly on one side of Robert's threshold: thresh - a = exp(E)/lam with E in (0, 1/2] and
// lam in [max(1,a), a+1], so  w max(1,a) > e^(1/2)  => mode 2  and  w (a+1) < 1  => mode 3  (w = b - a);
// only in between is the threshold evaluated.  Identical decisions to tn_setup.
__device__ __forceinline__ TnPlan tn_setup_fast(double a, double b) {
  TnPlan p; p.sgn = 1.0; p.lam = 0.0; p.a = a; p.b = b;
  if (a != a || b != b) { p.mode = 5; return p; }
  if (!(a < b)) { p.mode = 4; return p; }
  if (a > 0.0 || b < 0.0) {
    if (b < 0.0) { p.sgn = -1.0; p.a = -b; p.b = -a; }
    const double aa = p.a, w = p.b - aa;
    if (w * (aa + 1.0) < 1.0) { p.mode = 3; return p; }
    const double sq = x_sqrt(aa * aa + 4.0);
    p.lam = 0.5 * (aa + sq);
    if (w * fmax(1.0, aa) > 1.648721271) { p.mode = 2; return p; }
    const double thresh = aa + x_exp(0.5 + 0.25 * (aa * aa - aa * sq)) * x_rcp(p.lam);
    p.mode = (p.b > thresh) ? 2 : 3;
    return p;
  }
  p.mode = (b - a >= SQRT_2PI) ? 0 : 1;
  return p;
}
// What a trial needs from its two uniforms, computed as soon as the Philox block is there (before the bounds of
// the step are known): nlu = -log(ua), lub = log(ub) and the Box-Muller normal xn (straight-line math, bb_math.cuh).
struct TnDraw {
  double ua, ub, nlu, lub, xn;
  __device__ __forceinline__ void set(double ua_, double ub_) {
    ua = ua_; ub = ub_;
    nlu = -x_log(ua_); lub = x_log(ub_);
    xn = x_sqrt(2.0 * nlu) * x_cos2pi(ub_);
  }
};
// one trial from a prepared TnDraw; the acceptance tests `ub <= exp(t)` of tn_eval are taken in log form,
// `log(ub) <= t` (same event up to the last ulp).
__device__ __forceinline__ bool tn_eval_log(const TnPlan& p, const TnDraw& d, double& x) {
  const double ua = d.ua, nlu = d.nlu, lub = d.lub;
  if (p.mode & 1) {   // 1: uniform proposal around 0, 3: uniform proposal in the tail -- log density (c0 - x^2)/2, c0 = 0 | a^2
    x = p.a + (p.b - p.a) * ua;
    const double c0 = p.mode == 3 ? p.a * p.a : 0.0;
    return lub <= 0.5 * (c0 - x * x);
  }
  if (p.mode == 0) { x = d.xn; return x >= p.a && x <= p.b; }
  x = p.a + nlu * x_rcp(p.lam);
  const double t = x - p.lam;
  return x <= p.b && lub <= -0.5 * t * t;
}

// scalar form: one thread
__device__ inline double tnorm_std_scalar(const Stream& s, double a, double b, uint32_t& status) {
  const TnPlan p = tn_setup(a, b);
  if (p.mode == 5) { status |= ST_NAN; return a + b; }
  if (p.mode == 4) { if (a > b) status |= ST_EMPTY_INTERVAL; return a; }
  double x = 0.0;
  for (uint32_t k = 0; k < K_CAP - 1; ++k)
    if (tn_trial(p, s, k, x)) return p.sgn * x;
  status |= ST_TRIAL_CAP;
  return nan("");
}

// warp form: all 32 lanes of the calling warp must call it with identical arguments
__device__ inline double tnorm_std_warp(const Stream& s, double a, double b, uint32_t& status) {
  const TnPlan p = tn_setup(a, b);
  if (p.mode == 5) { status |= ST_NAN; return a + b; }
  if (p.mode == 4) { if (a > b) status |= ST_EMPTY_INTERVAL; return a; }
  const uint32_t lane = threadIdx.x & 31u;
  for (uint32_t base = 0; base < K_CAP - 32; base += 32) {
    double x;
    const bool acc = tn_trial(p, s, base + lane, x);
    const uint32_t m = __ballot_sync(0xffffffffu, acc);
    if (m) {
      const int src = __ffs(m) - 1;
      x = __shfl_sync(0xffffffffu, x, src);
      return p.sgn * x;
    }
  }
  status |= ST_TRIAL_CAP;
  return nan("");
}

// tnorm(left, right, mu, sd) as the reference calls it (BridgeRegression.cpp:273,392)
template <bool WARP>
__device__ __forceinline__ double tnorm(const Stream& s, double left, double right, double mu, double sd, uint32_t& status) {
  const double a = (left - mu) / sd, b = (right - mu) / sd;
  const double x = WARP ? tnorm_std_warp(s, a, b, status) : tnorm_std_scalar(s, a, b, status);
  double out = mu + sd * x;
  if (out < left) out = left;
  if (out > right) out = right;
  return out;
}

// ---------------------------------------------------------------------------------------------
// Gamma(shape, rate): exponential forms for shape 1 and 2, otherwise Marsaglia-Tsang (2000)
// without the squeeze; trial t consumes blocks first+2t (normal) and first+2t+1 (uniform).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool mt_trial(const Stream& s, uint32_t first, uint32_t t, double d, double c, double& dv) {
  double na, nb;
  s.block(first + 2 * t, na, nb);
  const double x = x_box_muller(na, nb);     // straight-line math (bb_math.cuh): the trial is one basic block
  const double u = s.unif(first + 2 * t + 1);
  double v = 1.0 + c * x;
  const bool pos = v > 0.0;
  v = v * v * v;
  dv = d * v;
  return pos && x_log(u) < 0.5 * x * x + d - dv + d * x_log(pos ? v : 1.0);
}

template <bool WARP>
__device__ inline double gamma_rate(const Stream& s, double shape, double rate, uint32_t& status) {
  if (shape == 1.0) return -log(s.unif(0)) / rate;
  if (shape == 2.0) { double a, b; s.block(0, a, b); return -log(a * b) / rate; }
  double boost = 1.0, sh = shape;
  uint32_t first = 0;
  if (shape < 1.0) { boost = pow(s.unif(0), 1.0 / shape); sh = shape + 1.0; first = 1; }
  const double d = sh - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
  double dv = 0.0;
  if (WARP) {
    const uint32_t lane = threadIdx.x & 31u;
    for (uint32_t base = 0; base < (K_CAP >> 1) - 32; base += 32) {
      const bool acc = mt_trial(s, first, base + lane, d, c, dv);
      const uint32_t m = __ballot_sync(0xffffffffu, acc);
      if (m) { dv = __shfl_sync(0xffffffffu, dv, __ffs(m) - 1); return boost * dv / rate; }
    }
  } else {
    for (uint32_t t = 0; t < (K_CAP >> 1) - 1; ++t)
      if (mt_trial(s, first, t, d, c, dv)) return boost * dv / rate;
  }
  status |= ST_TRIAL_CAP;
  return nan("");
}

}  // namespace bb
