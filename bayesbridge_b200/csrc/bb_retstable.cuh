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
#include "bb_math.cuh"

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

// Powers are taken in the log domain: pow(x, y) = exp(y log x) costs ~80 instructions against ~250 for
// pow(), and the round needs the logarithms of sinc(alpha U), sinc((1-alpha) U), sinc(U) only once for BdB0
// (retstable.cpp:73-77), A_ (:40-47), a, m and m^-b together.  Relative deviation from the reference's
// pow() calls <= |y log x| 2^-52 (~1e-14 here), far inside the 1e-10 parity tolerance.

// constants of alpha alone (alpha = alpha_bridge/2), shared by every item of a stage
struct RetAlpha {
  double alpha, Ialpha, b, inv_alpha, inv_Ialpha, lA, lI, log_b;
  __device__ __forceinline__ void init(double alpha_) {
    alpha = alpha_; Ialpha = 1. - alpha_;
    b = (1. - alpha_) / alpha_;                // retstable.cpp:123
    inv_alpha = 1. / alpha_; inv_Ialpha = 1. / (1. - alpha_);
    lA = log(alpha_); lI = log(1. - alpha_); log_b = log(b);
  }
};

// per-item constants (retstable.cpp:121-147); h = beta_j^2/tau^2
struct RetConst {
  double h, lambda_alpha, log_la, gamma, sgamma, inv_sgamma, xi, psi;
  double r1, r3;   // branch thresholds w1/(w1+w2) (gamma >= 1) and w3/(w2+w3) (gamma < 1) of the proposal for U (:162-175)
  // 1: every constant is an ordinary finite number, the straight-line round (bb_math.cuh) applies; 0 (h = 0, overflow): libm round
  int simple;
  // log_v0 = log(V0): 0 for the sweep (V0 = 1, BridgeRegression.cpp:509); general V0 only for the exported scalar sampler
  __device__ __forceinline__ void init(double h_, const RetAlpha& A, double log_v0 = 0.0) {
    h = h_;
    const double c2 = 2. + RS_C1;
    log_la = A.alpha * log(h) + log_v0;        // log(lambda^alpha V0); -inf when h = 0
    lambda_alpha = exp(log_la);                // pow(h, alpha) * V0   (:131)
    gamma = lambda_alpha * A.alpha * A.Ialpha; // :140
    sgamma = sqrt(gamma);
    const double c3 = c2 * sgamma;
    xi = (1. + RS_SQRT2 * c3) / RS_PI;
    psi = c3 * exp(-gamma * RS_PI * RS_PI / 8.) / RS_SQRT_PI;
    const double w1 = RS_C1 * xi / sgamma;
    const double w2 = 2. * RS_SQRT_PI * psi;
    const double w3 = xi * RS_PI;
    r1 = w1 / (w1 + w2);
    r3 = w3 / (w2 + w3);
    inv_sgamma = 1. / sgamma;
    simple = h > 1e-250 && h < 1e250 && sgamma > 1e-120 && lambda_alpha < 1e100 && r1 == r1 && r3 == r3;
  }
};

// One round starting at block k, as straight-line code: the inner trial (3 Philox blocks,
// retstable.cpp:162-207) and the outer trial (2 more blocks, :212-256) are BOTH evaluated, the
// outer one speculatively, so that its exp/log chain overlaps the inner one instead of waiting
// behind a branch.  Returns 0 = inner trial rejected (3 blocks consumed), 1 = outer trial rejected
// (5 blocks), 2 = accepted (value in `out`).  The blocks are the ones the sequential reference would
// consume at that position of its stream, so the accepted value does not depend on scheduling.
static __device__ __noinline__ int retstable_round_libm(const RetAlpha& A, const RetConst& c, const Stream& s, uint32_t k, double& out) {
  const double alpha = A.alpha, Ialpha = A.Ialpha;
  double u0, u0b, u1a, u1b, u2, u2b, u3, u3b, u4a, u4b;
  s.block(k, u0, u0b); s.block(k + 1, u1a, u1b); s.block(k + 2, u2, u2b); s.block(k + 3, u3, u3b); s.block(k + 4, u4a, u4b);
  // ---- inner trial ----
  const double V = u0;
  double U;
  if (c.gamma >= 1) {
    if (V < c.r1) U = fabs(box_muller(u1a, u1b)) / c.sgamma;
    else U = RS_PI * (1. - u1a * u1a);
  } else {
    if (V < c.r3) U = RS_PI * u1a;
    else U = RS_PI * (1. - u1a * u1a);
  }
  const double W = u2;
  const double sA = sinc_mm(alpha * U), sI = sinc_mm(Ialpha * U), sU = sinc_mm(U);
  const double lsA = log(sA), lsI = log(sI), lsU = log(sU);
  const double lden = alpha * lsA + Ialpha * lsI;            // log BdB0 denominator (retstable.cpp:73-77)
  const double izeta2 = exp(lden) / sU;                      // 1 / zeta^2
  const double zeta = sqrt(1. / izeta2);
  // z = 1 / (1 - (1 + alpha zeta / sqrt(gamma))^(-1/alpha))
  const double z = -1. / expm1(-log1p(alpha * zeta / c.sgamma) * A.inv_alpha);
  double rho = RS_PI * exp(-c.lambda_alpha * (1. - izeta2)) / ((1. + RS_C1) * c.sgamma / zeta + z);
  double d = 0.;
  if (U >= 0 && c.gamma >= 1) d += c.xi * exp(-c.gamma * U * U / 2.);
  if (U > 0 && U < RS_PI) d += c.psi / sqrt(RS_PI - U);
  if (U >= 0 && U <= RS_PI && c.gamma < 1) d += c.xi;
  rho *= d;
  const double Z = W * rho;
  const bool inner_ok = (U < RS_PI && Z <= 1.);
  // ---- outer trial (speculative) ----
  // A_ = ((1-alpha) sI)^(1-alpha) (alpha sA)^alpha / sU (:40-47);  a = A_^(1/(1-alpha));  m = (b/a)^alpha lambda^alpha
  const double lAval = Ialpha * (A.lI + lsI) + alpha * (A.lA + lsA) - lsU;
  const double la = lAval * A.inv_Ialpha;
  const double a = exp(la);
  const double lm = alpha * (A.log_b - la) + c.log_la;
  const double m = exp(lm);
  const double delta = sqrt(m * alpha / a);
  const double a1 = delta * RS_C1, a2 = delta, a3 = z / a;
  const double ssum = a1 + a2 + a3;
  const double V_ = u3;
  double N_ = 0., E_ = 0., X;
  if (V_ < a1 / ssum) { N_ = box_muller(u4a, u4b); X = m - delta * fabs(N_); }
  else if (V_ < (a1 + a2) / ssum) X = m + delta * u4a;
  else { E_ = -log(u4a); X = m + delta + E_ * a3; }
  const double E = -log(Z);
  out = exp(-A.b * log(X));   // X^-b = exp(1/alpha*log(V0) - b*log(X)) with V0 = 1 (retstable.cpp:270)
  double cc = a * (X - m);
  cc += (m != 0) ? c.h * (out - exp(-A.b * lm)) : 0.0;       // h (X^-b - m^-b)  (:248)
  if (X < m) cc -= N_ * N_ / 2.;
  else if (X > m + delta) cc -= E_;
  if (!inner_ok) return 0;
  return (X >= 0 && cc <= E) ? 2 : 1;
}

// The same round as ONE basic block (bb_math.cuh): every alternative of the reference's if/else chains is
// computed and selected, reciprocals replace divisions where the operand is ready early (exp(-la) for 1/a,
// exp(lm - la) for m/a), and the only branch left is the rarely taken normal proposal for U.  The independent
// chains (five Philox blocks, three sinc/log pairs, the outer trial's constants) overlap; measured latency of
// a round drops from ~8000 to the length of its critical path.  Decisions and draws agree with the libm round
// to rounding error (each elementary function within a few ulp).  Requires c.simple.
__device__ __forceinline__ int retstable_round_fast(const RetAlpha& A, const RetConst& c, const Stream& s, uint32_t k, double& out) {
  const double alpha = A.alpha, Ialpha = A.Ialpha;
  double u0, u0b, u1a, u1b, u2, u2b, u3, u3b, u4a, u4b;
  s.block(k, u0, u0b); s.block(k + 1, u1a, u1b); s.block(k + 2, u2, u2b); s.block(k + 3, u3, u3b); s.block(k + 4, u4a, u4b);
  // ---- inner trial (retstable.cpp:162-207) ----
  const bool ge1 = c.gamma >= 1;
  double U;
  if (ge1 && u0 < c.r1) U = fabs(x_box_muller(u1a, u1b)) * c.inv_sgamma;
  else U = (!ge1 && u0 < c.r3) ? RS_PI * u1a : RS_PI * (1. - u1a * u1a);
  const double sA = x_sinc(alpha * U), sI = x_sinc(Ialpha * U), sU = x_sinc(U);
  const double lsA = x_log(sA), lsI = x_log(sI), lsU = x_log(sU);
  const double lden = alpha * lsA + Ialpha * lsI;            // log BdB0 denominator (:73-77)
  const double izeta2 = x_exp(lden) * x_rcp(sU);             // 1 / zeta^2
  double izeta, zeta;                                        // 1/zeta = sqrt(izeta2), zeta
  x_sqrt_pair(izeta2, izeta, zeta);
  const double z = -x_rcp(x_expm1(-x_log1p(alpha * zeta * c.inv_sgamma) * A.inv_alpha));
  const double rho = RS_PI * x_exp(-c.lambda_alpha * (1. - izeta2)) * x_rcp((1. + RS_C1) * c.sgamma * izeta + z);
  const double dg = c.xi * (ge1 ? x_exp(-0.5 * c.gamma * U * U) : 1.0);
  const double dp = U < RS_PI ? c.psi * x_rsqrt(RS_PI - U) : 0.0;
  const double Z = u2 * rho * (dg + dp);
  const bool inner_ok = (U < RS_PI && Z <= 1.);
  // ---- outer trial, speculative (:212-256) ----
  const double lAval = Ialpha * (A.lI + lsI) + alpha * (A.lA + lsA) - lsU;
  const double la = lAval * A.inv_Ialpha;
  const double lm = alpha * (A.log_b - la) + c.log_la;
  const double a = x_exp(la), ia = x_exp(-la), m = x_exp(lm);
  const double delta = x_sqrt(alpha * x_exp(lm - la));       // sqrt(m alpha / a)
  const double a1 = delta * RS_C1, a2 = delta, a3 = z * ia;
  const double rs = x_rcp(a1 + a2 + a3);
  const double l4 = x_log(u4a);
  const bool b1 = u3 < a1 * rs, b2 = !b1 && u3 < (a1 + a2) * rs;
  const double N_ = b1 ? x_sqrt(-2.0 * l4) * x_cos2pi(u4b) : 0.0;
  const double E_ = (b1 || b2) ? 0.0 : -l4;
  const double X = b1 ? m - delta * fabs(N_) : (b2 ? m + delta * u4a : m + delta + E_ * a3);
  const double E = -x_log(Z);
  out = x_exp(-A.b * x_log(X));                              // X^-b (retstable.cpp:270 with V0 = 1)
  double cc = a * (X - m);
  cc += (m != 0) ? c.h * (out - x_exp(-A.b * lm)) : 0.0;     // h (X^-b - m^-b)  (:248)
  cc -= X < m ? 0.5 * N_ * N_ : (X > m + delta ? E_ : 0.0);
  if (!inner_ok) return 0;
  return (X >= 0 && cc <= E) ? 2 : 1;
}
__device__ __forceinline__ int retstable_round_at(const RetAlpha& A, const RetConst& c, const Stream& s, uint32_t k, double& out) {
  if (__builtin_expect(c.simple, 1)) return retstable_round_fast(A, c, s, k, out);
  return retstable_round_libm(A, c, s, k, out);
}

// Trial cap of the double rejection: 2^18 Philox blocks = at least 52000 rounds.  A round is accepted with probability >= 0.15
// for every (h, alpha) the algorithm is meant for, so the cap is only reached when the constants have degenerated (NaN / inf
// from a collapsed chain); the draw then comes back NaN with ST_TRIAL_CAP after about 0.1 s instead of spinning through the
// 2^24 blocks of the counter space for minutes (the reference loops for ever there, polling the interrupt, retstable.cpp:158-166).
constexpr uint32_t RET_CAP = 1u << 18;

// sequential form, one thread per item
__device__ inline double retstable_scalar(const Stream& s, double h, double alpha, uint32_t& status, double V0 = 1.0) {
  if (alpha == 1.) return V0;  // point mass at V0 (retstable.cpp:104-110)
  RetAlpha A; A.init(alpha);
  const double log_v0 = V0 == 1.0 ? 0.0 : log(V0);
  RetConst c; c.init(h, A, log_v0);
  uint32_t k = 0; double out = 0.;
  while (k < RET_CAP - 8) {
    const int r = retstable_round_at(A, c, s, k, out);
    if (r == 2) return V0 == 1.0 ? out : exp(log_v0 * A.inv_alpha) * out;   // V0^(1/alpha) X^-b (retstable.cpp:270)
    k += r == 0 ? 3u : 5u;
  }
  status |= ST_TRIAL_CAP;
  return nan("");
}

// ---------------------------------------------------------------------------------------------
// CTA-wide lambda stage with a compacted retry queue and speculative look-ahead.
//
// The stream position of an item after some rejected rounds is k0 + 3a + 5b.  Every such offset
// {0,3,5,6,8,9,10,...} can be evaluated independently because block indices are counters, so when
// fewer items than threads are unfinished, G = 2..32 lanes take ONE item and evaluate the rounds
// at G reachable offsets at once; a shuffle walk then follows the path the sequential
// sampler would have taken (0 -> +3 on an inner rejection, +5 on an outer one) and either finds
// the accepting round or leaves the item at the first offset that was not covered.  The result is
// bit-identical to the one-round-at-a-time schedule; only the number of passes (each costing the
// latency of one round, the bottleneck: the rejection tail of max_j trials_j) shrinks.
// ---------------------------------------------------------------------------------------------
// The G offsets evaluated for one item: the G positions the sequential sampler is MOST LIKELY to visit (an inner
// rejection, +3, has probability ~0.72 per round, an outer one, +5, ~0.09): 0,3,6,9,12,15,18,11,14,8,17,21,20,5,...
// Any set gives the same draws; this one needs about one pass in twenty fewer than the first G reachable positions.
// Tables packed 6 bits per entry.
__device__ __forceinline__ int rs_unpack(unsigned long long w0, unsigned long long w1, unsigned long long w2, unsigned long long w3, int i) {
  const int q = i / 10;
  const unsigned long long w = q == 0 ? w0 : q == 1 ? w1 : q == 2 ? w2 : w3;
  return (int)((w >> (6 * (i - 10 * q))) & 63ull);
}
__device__ __forceinline__ int rs_reach(int i) {        // i-th offset, i < 32
  return rs_unpack(148105469521846464ull, 584744086747301201ull, 657330275283769566ull, 1058ull, i);
}
__device__ __forceinline__ int rs_slot_of(int pos) {    // index of the lane group member that evaluates `pos`, 63: none
  return pos >= 40 ? 63 : rs_unpack(56853699704127424ull, 488123725904626175ull, 331371938842247884ull, 1142504548344608532ull, pos);
}
__device__ __forceinline__ int rs_levels(int G) {       // longest walk through the first G offsets: max offset / 3 + 1
  return G >= 32 ? 13 : G >= 16 ? 9 : G >= 8 ? 7 : G >= 4 ? 4 : G >= 2 ? 2 : 1;
}

#ifdef BB_PHASES
__device__ unsigned long long bb_phase_cyc[16];   // development: [0..7] stb64 phases, [8..11] lambda stage: round, walk, queue, passes
#define BB_LAM_MARK(k) do { if (blockIdx.x == 0 && threadIdx.x == 0) { const long long now_ = clock64(); bb_phase_cyc[8 + k] += (unsigned long long)(now_ - lam_t); lam_t = now_; } } while (0)
#else
#define BB_LAM_MARK(k) do { } while (0)
#endif

// NT: threads serving one chain (multiple of 32); every thread of the CTA must call.  beta: per-item
// coordinate values (shared or global), lam_out: where 2*draw goes (shared or global).
// q_s: [2][qld] ints, qn_s: [2] ints (both zero on entry and on exit), kk_s: [P] block positions.
// rc_s: optional [P] cache of the per-item constants (nullptr: recompute every pass).
#ifdef BB_LAMBDA_STATS
__device__ unsigned long long bb_lambda_stats[8];
#endif
template <int NT>
__device__ inline void lambda_stage_cta(const int tid, int P, const double* beta, double* lam_out, double tau, double alpha_half,
                                        const RngKey key, uint32_t gch, uint32_t t, RetConst* rc_s,
                                        int* q_s, int qld, int* qn_s, uint32_t* kk_s, uint32_t& status, int j0 = 0, int gmax = 32) {
  // j0: index of beta[0] / lam_out[0] among the chain's coefficients (a CTA may serve a slice of them): the Philox item of
  // coefficient j is j0 + j, so the draws do not depend on how the coefficients are split over CTAs
  // `tid` = index of the thread among the NT threads that serve THIS chain (a CTA may hold several chains,
  // each with its own queues; the barriers are CTA-wide and the loop runs until every chain is done)
  int cur_q = 0, n_act = P;
  bool first = true;
  if (alpha_half == 1.) {   // point mass at V0 = 1 (retstable.cpp:104-110)
    for (int j = tid; j < P; j += NT) lam_out[j] = 2.0;
    __syncthreads();
    return;
  }
  if (!(tau > 0.0) || !(tau < 1e300)) {   // a chain that has already broken down (NaN / zero / infinite tau): no trials, flag it
    for (int j = tid; j < P; j += NT) lam_out[j] = nan("");
    status |= ST_NAN;
    n_act = 0;                            // (no early return: other chains of the CTA still need this chain at the barriers)
  }
  RetAlpha A; A.init(alpha_half);
  // the two offset tables as shared-memory bytes: the walk below is a chain of dependent look-ups (13 levels at G = 32), and
  // unpacking a 6-bit field of a 64-bit constant costs ~300 cycles per level against ~80 for a shared-memory load
  __shared__ unsigned char lut_slot[64], lut_reach[32];
  if (threadIdx.x < 64) lut_slot[threadIdx.x] = (unsigned char)rs_slot_of((int)threadIdx.x);
  if (threadIdx.x < 32) lut_reach[threadIdx.x] = (unsigned char)rs_reach((int)threadIdx.x);
  __syncthreads();
  bool any = true;
#ifdef BB_PHASES
  long long lam_t = clock64();
#endif
  while (any) {
    BB_LAM_MARK(3);
#ifdef BB_PHASES
    if (blockIdx.x == 0 && threadIdx.x == 0) bb_phase_cyc[12] += 1;
#endif
    int G = 1;
    while (G < gmax && n_act * (G * 2) <= NT) G *= 2;   // gmax < 32: less speculative look-ahead (the draws do not depend on it)
    const int per_pass = NT / G;
    const int sub = tid & (G - 1);
    const int nlvl = rs_levels(G);
    for (int base = 0; base < n_act; base += per_pass) {   // (no pass for a chain that is already done: n_act = 0)
      const int slot = base + tid / G;
      const bool valid = slot < n_act;
      int j = 0; uint32_t k0 = 0; int code = 2; double out = 0.0;
      if (valid) {
        j = first ? slot : q_s[cur_q * qld + slot];
        k0 = first ? 0u : kk_s[j];
        RetConst rc;
        if (rc_s && !first) rc = rc_s[j];
        else {
          const double b = beta[j];
          rc.init(b * b / (tau * tau), A);   // h = beta_j^2 / tau^2 (BridgeRegression.cpp:509)
          if (rc_s && sub == 0) rc_s[j] = rc;
        }
        Stream s(key, gch, t, ST_LAMBDA, (uint32_t)(j0 + j));
        code = retstable_round_at(A, rc, s, k0 + (uint32_t)lut_reach[sub], out);
      }
      BB_LAM_MARK(0);
      // walk the sequential path through the G evaluated offsets (uniform trip count: shuffles)
      int pos = 0, res = -1;
#ifdef BB_LAMBDA_STATS
      int st_in = 0, st_out = 0;
#endif
      for (int lvl = 0; lvl < nlvl; ++lvl) {
        const int idx = lut_slot[pos < 63 ? pos : 63];
        const bool covered = idx < G;
        const int c = __shfl_sync(0xffffffffu, code, covered ? idx : 0, G);
        if (res == -1) {
          if (!covered) res = -2;
          else if (c == 2) res = idx;
          else {
            pos += (c == 0) ? 3 : 5;
#ifdef BB_LAMBDA_STATS
            if (c == 0) ++st_in; else ++st_out;
#endif
          }
        }
      }
#ifdef BB_LAMBDA_STATS
      // development counters (trials per draw of the tilted-stable stage): rounds of the SEQUENTIAL sampler's
      // path rejected at the inner / outer trial, accepted draws, rounds evaluated by the lanes (speculation included), passes
      if (valid) atomicAdd(&bb_lambda_stats[3], 1ull);
      if (valid && sub == 0) {
        atomicAdd(&bb_lambda_stats[0], (unsigned long long)st_in); atomicAdd(&bb_lambda_stats[1], (unsigned long long)st_out);
        if (res >= 0) atomicAdd(&bb_lambda_stats[2], 1ull);
      }
      if (tid == 0 && base == 0) atomicAdd(&bb_lambda_stats[4], 1ull);
#endif
      BB_LAM_MARK(1);
      const double val = __shfl_sync(0xffffffffu, out, res >= 0 ? res : 0, G);
      if (valid && sub == 0) {
        if (res >= 0) lam_out[j] = 2 * val;
        else if (k0 + (uint32_t)pos >= RET_CAP - 64) { lam_out[j] = nan(""); status |= ST_TRIAL_CAP; }
        else {
          kk_s[j] = k0 + (uint32_t)pos;
          q_s[(cur_q ^ 1) * qld + atomicAdd(&qn_s[cur_q ^ 1], 1)] = j;
        }
      }
    }
    __syncthreads();
    cur_q ^= 1;
    n_act = qn_s[cur_q];
    first = false;
    if (tid == 0) qn_s[cur_q ^ 1] = 0;   // the queue the next pass fills
    any = __syncthreads_or(n_act > 0);
    BB_LAM_MARK(2);
  }
}

}  // namespace bb
