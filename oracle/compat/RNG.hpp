// RNG.hpp -- ORACLE compat shim (test infrastructure; never shipped, never on the product path).
//
// Stand-in for the reference author's un-vendored `RNG` library (#include "RNG.hpp" at
// /root/reference/Code/C/BridgeRegression.h:68, BridgeWrapper.h:66, retstable.h:7; built from
// $(HOME)/Code/RNG, /root/reference/Code/C/Makefile:4,66-73).  The method names and argument
// orders are the ones the reference calls (BridgeRegression.cpp:103,120-124,137-143,160,273,280,
// 392,449,462,472,490,498,502,519,568; retstable.cpp:172-235; BridgeWrapper.cpp:775-957).
// The arithmetic behind them is NOT available (parity unpinned) and is defined by
// oracle/port/bbo_draws.h, which this class forwards to.
//
// Two modes:
//   sequential (default-constructed, as the reference drivers do with `RNG r;`): one running
//       Philox counter seeded from bb_compat_seed()/bb_compat_chain();
//   keyed (test harness): the harness positions (iter, stage) and an item policy so that the
//       block consumed for coordinate j, trial k is Philox(seed; stage|k, j, iter, chain),
//       the same block the CUDA sampler consumes.
#ifndef BB_COMPAT_RNG_HPP
#define BB_COMPAT_RNG_HPP

#include "Matrix.h"
#include "../port/bbo_draws.h"

extern "C" {
void bb_compat_set_seed(unsigned long long seed, unsigned int chain);
unsigned long long bb_compat_seed();
unsigned int bb_compat_chain();
}

class RNG {
 public:
  enum Policy { ITEM_PER_CALL = 0, ITEM_ON_UNIF = 1, ITEM_MANUAL = 2 };

 private:
  bbo_rng s_;
  Policy pol_;
  uint next_item_;
  void top() { if (!s_.sequential && pol_ == ITEM_PER_CALL) next_item(); }

 public:
  RNG() : pol_(ITEM_PER_CALL), next_item_(0) {
    bbo_rng_key(&s_, bb_compat_seed(), bb_compat_chain());
    s_.sequential = 1;
  }
  RNG(unsigned long long seed, unsigned int chain) : pol_(ITEM_PER_CALL), next_item_(0) { bbo_rng_key(&s_, seed, chain); }
  void position(uint iter, uint stage, Policy pol) { bbo_rng_at(&s_, iter, stage, 0); pol_ = pol; next_item_ = 0; }
  void next_item() { s_.item = next_item_++; s_.k = 0; }
  bbo_rng* raw() { return &s_; }

  // base draws
  double unif() { if (!s_.sequential && pol_ == ITEM_ON_UNIF) next_item(); else top(); return bbo_unif(&s_); }
  double norm(double m, double sd) { top(); return m + sd * bbo_norm(&s_); }
  void norm(MatrixFrame M, double m, double sd) { // one item per element (BridgeRegression.cpp:568)
    for (uint i = 0; i < M.vol(); ++i) { if (!s_.sequential) next_item(); M(i) = m + sd * bbo_norm(&s_); }
  }
  // derived draws
  double flat(double a = 0.0, double b = 1.0) { top(); return bbo_flat(&s_, a, b); }
  double expon_rate(double rate) { top(); return bbo_expon_rate(&s_, rate); }
  double gamma_rate(double shape, double rate) { if (pol_ != ITEM_ON_UNIF) top(); return bbo_gamma_rate(&s_, shape, rate); }
  double igamma(double shape, double scale) { top(); return bbo_igamma(&s_, shape, scale); }
  double tnorm(double left, double right, double mu, double sd) {
    top();
    return bbo_tnorm(&s_, left, right, mu, sd);
  }
  double tnorm(double left, double mu, double sd) { top(); return bbo_tnorm_left(&s_, left, mu, sd); }
  double texpon_rate(double left, double rate) { top(); return bbo_texpon_rate_left(&s_, left, rate); }
  double texpon_rate(double left, double right, double rate) { top(); return bbo_texpon_rate_both(&s_, left, right, rate); }
  double rtgamma_rate(double shape, double rate, double right) { top(); return bbo_rtgamma_rate(&s_, shape, rate, right); }
  // special functions the reference reaches through RNG (BridgeRegression.cpp:472,498)
  static double Gamma(double x, bool use_log = false) { return use_log ? lgamma(x) : tgamma(x); }
  static double d_beta(double x, double a, double b) {
    return exp(lgamma(a + b) - lgamma(a) - lgamma(b) + (a - 1) * log(x) + (b - 1) * log1p(-x));
  }
};

#endif
