/*
 * bbo_draws.h -- ORACLE (test infrastructure, never shipped, never on the product path).
 *
 * CPU statement, in plain C, of the random-draw SPECIFICATION shared by the oracle and the
 * CUDA sampler:  Philox4x32-10 counter layout, base draws (uniform / normal) and the derived
 * draws (flat, expon, gamma, igamma, truncated normal, ...).
 *
 * Why this is a specification and not a port: the reference draws through the author's RNG
 * library (`RNG.hpp`, included at /root/reference/Code/C/BridgeRegression.h:68 and
 * BridgeWrapper.h:66), which is NOT in the reference tree (built from $(HOME)/Code/RNG,
 * /root/reference/Code/C/Makefile:4,66-73).  No version pin, no golden vectors:
 * **parity unpinned** at this level.  The reference documents only that `rtnorm` follows
 * Robert (1995) (/root/reference/Code/BBPackage/BayesBridge/man/rtnorm.Rd:13,46-47) and that
 * `rtexp.left == left + rexp(rate)` (/root/reference/Code/R/TestSamplers.R:24-39).
 * The call sites that consume these draws are cited in bbo_port.c.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this file.
 *
 * Counter layout (one Philox block = 128 bits = two 53-bit uniforms `ua`,`ub`):
 *     key  = { seed_lo, seed_hi }
 *     ctr  = { (stage << 24) | k ,  item ,  iter ,  chain }
 * k = index of the block inside one item (rejection trial counter), item = coordinate j,
 * iter = Gibbs iteration, chain = global chain id.  In the triangle beta stage the item of
 * sweep s (0..betaburn), coordinate i is s*P + i.
 * Every base draw consumes exactly one block:  unif() -> ua ;  norm() -> Box-Muller(ua, ub).
 */
#ifndef BBO_DRAWS_H
#define BBO_DRAWS_H

#include <math.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  BBO_ST_TAU = 1,      /* sample_tau_marg           */
  BBO_ST_SIG2 = 2,     /* sample_sig2               */
  BBO_ST_OMEGA = 3,    /* sample_omega              */
  BBO_ST_U = 4,        /* sample_u                  */
  BBO_ST_BETA = 5,     /* rtnorm_gibbs / sample_beta_ortho (truncated normals) */
  BBO_ST_LAMBDA = 6,   /* sample_lambda (retstable) */
  BBO_ST_BETA_STB = 7, /* sample_beta_stable / _ortho (normals) */
  BBO_ST_TAU_INIT = 8, /* the tau pre-draw before burn-in */
  BBO_ST_ALPHA = 9,    /* sample_alpha_marg         */
  BBO_ST_SCALAR = 10   /* stand-alone scalar samplers (rtnorm, rtexpon, ...) */
};

#define BBO_K_CAP (1u << 24) /* trial cap: k must stay below 2^24 */
#define BBO_RET_CAP (1u << 18) /* trial cap of the tilted-stable double rejection (same as the device's RET_CAP) */
#define BBO_RTG_CAP (1u << 16) /* trial cap of the right-truncated gamma's naive rejection */

typedef struct {
  uint32_t seed_lo, seed_hi; /* key */
  uint32_t chain, iter, stage, item;
  uint32_t k;                /* next block inside the item */
  /* sequential mode (end-to-end reference runs): stage==0, a 64-bit running counter */
  uint64_t seq;
  int sequential;
} bbo_rng;

static inline void bbo_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline double bbo_u53(uint32_t lo, uint32_t hi) {
  uint64_t x = ((uint64_t)hi << 32) | lo;
  return ((double)(x >> 11) + 0.5) * (1.0 / 9007199254740992.0); /* (0,1), never 0 or 1 */
}

/* fetch the next block of the current item */
static inline void bbo_block(bbo_rng *r, double *ua, double *ub) {
  uint32_t ctr[4], key[2], out[4];
  key[0] = r->seed_lo; key[1] = r->seed_hi;
  if (r->sequential) {
    ctr[0] = (uint32_t)r->seq; ctr[1] = (uint32_t)(r->seq >> 32); ctr[2] = 0xFFFFFFFFu; ctr[3] = r->chain;
    r->seq++;
  } else {
    ctr[0] = (r->stage << 24) | (r->k & (BBO_K_CAP - 1)); ctr[1] = r->item; ctr[2] = r->iter; ctr[3] = r->chain;
    r->k++;
  }
  bbo_philox4x32_10(ctr, key, out);
  *ua = bbo_u53(out[0], out[1]);
  *ub = bbo_u53(out[2], out[3]);
}

static inline void bbo_rng_key(bbo_rng *r, uint64_t seed, uint32_t chain) {
  r->seed_lo = (uint32_t)seed; r->seed_hi = (uint32_t)(seed >> 32);
  r->chain = chain; r->iter = 0; r->stage = 0; r->item = 0; r->k = 0; r->seq = 0; r->sequential = 0;
}
static inline void bbo_rng_at(bbo_rng *r, uint32_t iter, uint32_t stage, uint32_t item) {
  r->iter = iter; r->stage = stage; r->item = item; r->k = 0;
}

#define BBO_TWO_PI 6.283185307179586
#define BBO_SQRT_2PI 2.5066282746310002

static inline double bbo_bm(double ua, double ub) { return sqrt(-2.0 * log(ua)) * cos(BBO_TWO_PI * ub); }

/* ---- base draws ---- */
static inline double bbo_unif(bbo_rng *r) { double a, b; bbo_block(r, &a, &b); return a; }
static inline double bbo_norm(bbo_rng *r) { double a, b; bbo_block(r, &a, &b); return bbo_bm(a, b); }

/* ---- derived draws ---- */
static inline double bbo_flat(bbo_rng *r, double a, double b) { return a + (b - a) * bbo_unif(r); }
static inline double bbo_expon_rate(bbo_rng *r, double rate) { return -log(bbo_unif(r)) / rate; }

/* Marsaglia & Tsang (2000) without the squeeze step; one trial = norm block + unif block. */
static inline double bbo_gamma_rate(bbo_rng *r, double shape, double rate) {
  if (shape == 1.0) return -log(bbo_unif(r)) / rate;
  if (shape == 2.0) { double a, b; bbo_block(r, &a, &b); return -log(a * b) / rate; }
  double boost = 1.0, sh = shape;
  if (shape < 1.0) { boost = pow(bbo_unif(r), 1.0 / shape); sh = shape + 1.0; }
  double d = sh - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
  for (;;) {
    double x = bbo_norm(r);
    double u = bbo_unif(r);
    double v = 1.0 + c * x;
    if (v > 0.0) {
      v = v * v * v;
      if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) return boost * d * v / rate;
    }
    if (r->k >= BBO_K_CAP - 2) return NAN;
  }
}
static inline double bbo_igamma(bbo_rng *r, double shape, double scale) {
  return scale / bbo_gamma_rate(r, shape, 1.0);
}

/* right tail: x ~ N(0,1) on [a,b], 0 < a < b (b may be +inf).  Robert (1995) Prop. 2.3. */
static inline double bbo_tnorm_tail(bbo_rng *r, double a, double b) {
  double sq = sqrt(a * a + 4.0);
  double lam = 0.5 * (a + sq);
  double thresh = a + exp(0.5 + 0.25 * (a * a - a * sq)) / lam;
  double ua, ub;
  if (b > thresh) { /* translated-exponential proposal */
    for (;;) {
      bbo_block(r, &ua, &ub);
      double x = a - log(ua) / lam;
      double t = x - lam;
      if (x <= b && ub <= exp(-0.5 * t * t)) return x;
      if (r->k >= BBO_K_CAP - 1) return NAN;
    }
  } else { /* uniform proposal */
    for (;;) {
      bbo_block(r, &ua, &ub);
      double x = a + (b - a) * ua;
      if (ub <= exp(0.5 * (a * a - x * x))) return x;
      if (r->k >= BBO_K_CAP - 1) return NAN;
    }
  }
}

/* x ~ N(0,1) truncated to [a,b] (standardised bounds). */
static inline double bbo_tnorm_std(bbo_rng *r, double a, double b) {
  if (a != a || b != b) return NAN;
  if (!(a < b)) return a; /* empty or degenerate interval: documented choice, see DESIGN.md */
  if (a > 0.0) return bbo_tnorm_tail(r, a, b);
  if (b < 0.0) return -bbo_tnorm_tail(r, -b, -a);
  double ua, ub;
  if (b - a >= BBO_SQRT_2PI) { /* normal proposal */
    for (;;) {
      bbo_block(r, &ua, &ub);
      double x = bbo_bm(ua, ub);
      if (x >= a && x <= b) return x;
      if (r->k >= BBO_K_CAP - 1) return NAN;
    }
  } else { /* uniform proposal, density maximal at 0 */
    for (;;) {
      bbo_block(r, &ua, &ub);
      double x = a + (b - a) * ua;
      if (ub <= exp(-0.5 * x * x)) return x;
      if (r->k >= BBO_K_CAP - 1) return NAN;
    }
  }
}

static inline double bbo_tnorm(bbo_rng *r, double left, double right, double mu, double sd) {
  double a = (left - mu) / sd, b = (right - mu) / sd;
  double x = bbo_tnorm_std(r, a, b);
  double out = mu + sd * x;
  /* keep the draw inside the box after the affine map's rounding */
  if (out < left) out = left;
  if (out > right) out = right;
  return out;
}
static inline double bbo_tnorm_left(bbo_rng *r, double left, double mu, double sd) {
  double out = mu + sd * bbo_tnorm_std(r, (left - mu) / sd, INFINITY);
  return out < left ? left : out;
}

static inline double bbo_texpon_rate_left(bbo_rng *r, double left, double rate) {
  return left + bbo_expon_rate(r, rate); /* TestSamplers.R:24-39 identity */
}
static inline double bbo_texpon_rate_both(bbo_rng *r, double left, double right, double rate) {
  double b = -expm1(rate * (left - right)); /* 1 - exp(-rate (right-left)) */
  return left - log1p(-bbo_unif(r) * b) / rate;
}
/* Ga(shape, rate) truncated to (0, right]: plain rejection (Notes/bbnotes.tex:792-793). */
static inline double bbo_rtgamma_rate(bbo_rng *r, double shape, double rate, double right) {
  for (;;) {
    double x = bbo_gamma_rate(r, shape, rate);
    if (x <= right || x != x) return x;
    if (r->k >= BBO_RTG_CAP - 64) return NAN;   /* bounded cost, same cap as the device (bb_scalar.cu RTG_CAP) */
  }
}

#ifdef __cplusplus
}
#endif
#endif
