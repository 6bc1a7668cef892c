// bb_stb.cuh -- the stable (normal scale mixture) Gibbs sweep for P <= 16*TB, one chain per CTA,
// everything on chip.
//
// One launch advances `nchains` chains by `niter` iterations of
//   tau -> sig2 -> lambda -> beta                 (/root/reference/Code/C/BridgeWrapper.cpp:287-298)
// with   sample_tau_marg  BridgeRegression.cpp:453-465     sample_sig2  :436-450
//        sample_lambda    :506-510 -> retstable_LD retstable.cpp:94-271
//        sample_beta_stable :552-575  (Cholesky of X'X + diag(lambda sig2/tau^2), triangular solves).
//
// CTA layout: the lower triangle of the (P_pad+1) x P_pad augmented matrix [A ; Xy'] is cut into
// TB x TB register tiles, one per thread (136 tiles + 16 for the Xy row = 152 threads, 5 warps).
// Right-looking Cholesky: step k publishes column k through a double-buffered smem vector (one
// __syncthreads per step) and every tile applies the rank-1 update with 1/pivot, so no sqrt sits
// on the critical path; the Xy row turns into w = L^-1 Xy for free.  Since the reference's two
// back-substitutions use the same factor, beta = U^-1 (w + sig * eps) is ONE back-substitution.
// lambda: round-based tilted-stable draws; unfinished coordinates are compacted into a smem retry
// queue between rounds so that finished lanes do not idle the warp (bb_retstable.cuh).
#pragma once
#include "bb_launch.h"
#include "bb_retstable.cuh"
#include "bb_alpha.cuh"

namespace bb {


constexpr int STB_NB = 16;                                   // tiles per dimension
constexpr int STB_TILES = STB_NB * (STB_NB + 1) / 2;         // 136 lower-triangular tiles
constexpr int STB_THREADS = 160;                             // 136 + 16 (Xy row) rounded up to warps

template <int TB>
__global__ void __launch_bounds__(STB_THREADS, (TB == 4 ? 4 : 1)) stb_small_kernel(const StbLaunch L) {
  constexpr int PP = STB_NB * TB;   // padded dimension
  extern __shared__ double smem[];
  const int P = L.m.P;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = STB_THREADS / 32;

  double* Ls = smem;                       // [PP][PP] row-major: Ls[k*PP + i] = L(k,i), i <= k
  double* colbuf = Ls + PP * PP;           // [2][PP + TB]
  double* beta_s = colbuf + 2 * (PP + TB); // [PP]
  double* lam_s = beta_s + PP;             // [PP]
  double* rhs_s = lam_s + PP;              // [PP]   w + sig*eps
  double* idiag_s = rhs_s + PP;            // [PP]   1 / L(k,k)
  double* red_s = idiag_s + PP;            // [3][8]
  double* sc_s = red_s + 24;               // tau, sig2
  int* q_s = reinterpret_cast<int*>(sc_s + 2);   // [2][PP] retry queues
  int* qn_s = q_s + 2 * PP;                      // [2] counts
  uint32_t* kk_s = reinterpret_cast<uint32_t*>(qn_s + 2);   // [PP] next Philox block per coordinate
  RetConst* rc_s = reinterpret_cast<RetConst*>(kk_s + PP + 2);   // [PP] per-coordinate constants of the tilted-stable draw (8-byte aligned: 3 PP + 4 ints)

  const int ch = blockIdx.x;
  if (ch >= L.nchains) return;
  const uint32_t gch = L.chain0 + (uint32_t)ch;

  // tile coordinates
  int ti, tj;
  if (tid < STB_TILES) {
    ti = (int)((sqrt(8.0 * tid + 1.0) - 1.0) * 0.5);
    while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
    while (ti * (ti + 1) / 2 > tid) --ti;
    tj = tid - ti * (ti + 1) / 2;
  } else { ti = STB_NB; tj = tid - STB_TILES; }
  const bool tile_on = tid < STB_TILES + STB_NB;
  const bool xy_row = tile_on && ti == STB_NB;

  const bool know_tau = L.h.true_tau > 0, know_sig2 = L.h.true_sig2 > 0, know_alpha = L.h.true_alpha > 0;
  const bool do_tau = (L.stages & 1) && !know_tau, do_sig2 = (L.stages & 2) && !know_sig2;
  const bool do_lambda = L.stages & 32, do_beta = L.stages & 16, do_alpha = (L.stages & 64) && !know_alpha;
  double alpha = know_alpha ? L.h.true_alpha : L.alpha[ch];   // unknown alpha: chain state, moved by the Metropolis step below
  uint32_t status = 0;

  for (int j = tid; j < PP; j += STB_THREADS) {
    beta_s[j] = j < P ? L.beta[(size_t)ch * P + j] : 0.0;
    lam_s[j] = j < P ? L.lambda[(size_t)ch * P + j] : 1.0;
  }
  if (tid == 0) { sc_s[0] = know_tau ? L.h.true_tau : L.tau[ch]; sc_s[1] = know_sig2 ? L.h.true_sig2 : L.sig2[ch]; }
  __syncthreads();

  for (int it = -L.pre_tau; it < L.niter; ++it) {
    const bool pre = it < 0;
    const uint32_t t = L.iter0 + (pre ? 0u : (uint32_t)it);

    // ---- load the X'X tile; partial sums for tau and the residual sum of squares ----
    double A[TB][TB];
    double p_tau = 0.0, p_q = 0.0, p_g = 0.0;
    if (tile_on && !xy_row) {
#pragma unroll
      for (int s = 0; s < TB; ++s)
#pragma unroll
        for (int r = 0; r < TB; ++r) {
          const int gi = ti * TB + r, gj = tj * TB + s;
          A[r][s] = (gi < P && gj < P) ? __ldg(L.m.XX + gi + (size_t)gj * P) : (gi == gj ? 1.0 : 0.0);
        }
      if (!pre && do_sig2) {   // delta' XX delta over this tile (delta = beta - bc)
        double q = 0.0;
#pragma unroll
        for (int r = 0; r < TB; ++r) {
          const int gi = ti * TB + r;
          const double di = gi < P ? beta_s[gi] - __ldg(L.m.bc + gi) : 0.0;
          double rowsum = 0.0;
#pragma unroll
          for (int s = 0; s < TB; ++s) {
            const int gj = tj * TB + s;
            const double dj = gj < P ? beta_s[gj] - __ldg(L.m.bc + gj) : 0.0;
            rowsum = fma(A[r][s], dj, rowsum);
          }
          q = fma(rowsum, di, q);
        }
        p_q = (ti == tj) ? q : 2.0 * q;
      }
    } else if (xy_row) {
#pragma unroll
      for (int s = 0; s < TB; ++s)
#pragma unroll
        for (int r = 0; r < TB; ++r) {
          const int gj = tj * TB + s;
          A[r][s] = (r == 0 && gj < P) ? __ldg(L.m.Xy + gj) : 0.0;
        }
    } else {
#pragma unroll
      for (int s = 0; s < TB; ++s)
#pragma unroll
        for (int r = 0; r < TB; ++r) A[r][s] = 0.0;
    }
    for (int j = tid; j < P; j += STB_THREADS) {
      const double b = beta_s[j];
      p_tau += exp(alpha * log(fabs(b)));                       // BridgeRegression.cpp:459
      p_g = fma(b - __ldg(L.m.bc + j), __ldg(L.m.gc + j), p_g);
    }
    p_tau = warp_sum(p_tau); p_q = warp_sum(p_q); p_g = warp_sum(p_g);
    if (lane == 0) { red_s[warp] = p_tau; red_s[8 + warp] = p_q; red_s[16 + warp] = p_g; }
    __syncthreads();
    if (warp == 0 && (do_tau || (pre && !know_tau))) {          // BridgeRegression.cpp:453-465
      double st = lane < NW ? red_s[lane] : 0.0;
      st = warp_sum(st);
      const double shape = L.h.nu_shape + ((double)P) / alpha;
      const double rate = L.h.nu_rate + st;
      Stream s(L.key, gch, t, pre ? ST_TAU_INIT : ST_TAU, 0);
      const double nu = gamma_rate<true>(s, shape, rate, status);
      if (lane == 0) sc_s[0] = exp(-1.0 * log(nu) / alpha);
    }
    if (warp == 1 && !pre && do_sig2) { // BridgeRegression.cpp:436-450, rss = r0 - 2 d'g + d'XXd
      double sq = lane < NW ? red_s[8 + lane] : 0.0, sg = lane < NW ? red_s[16 + lane] : 0.0;
      sq = warp_sum(sq); sg = warp_sum(sg);
      const double rss = L.m.r0b - 2.0 * sg + sq;
      const double shape = L.h.sig2_shape + 0.5 * L.m.N;
      const double scale = L.h.sig2_scale + 0.5 * rss;
      Stream s(L.key, gch, t, ST_SIG2, 0);
      const double g = gamma_rate<true>(s, shape, 1.0, status);
      if (lane == 0) sc_s[1] = scale / g;
    }
    if (tid == 0) { qn_s[0] = 0; qn_s[1] = 0; }
    __syncthreads();
    if (pre) continue;
    const double tau = sc_s[0], sig2 = sc_s[1];

    // ---- lambda_j = 2 * retstable(beta_j^2 / tau^2, alpha/2): retry queue + speculative look-ahead ----
    if (do_lambda) lambda_stage_cta<STB_THREADS>(tid, P, beta_s, lam_s, tau, 0.5 * alpha, L.key, gch, t, rc_s, q_s, PP, qn_s, kk_s, status);

    if (do_beta) {
    // ---- A = X'X + diag(lambda sig2 / tau^2)   (BridgeRegression.cpp:554-556) ----
    if (tile_on && !xy_row && ti == tj) {
#pragma unroll
      for (int r = 0; r < TB; ++r) {
        const int g = ti * TB + r;
        if (g < P) A[r][r] += lam_s[g] * sig2 / (tau * tau);
      }
    }
    // normals for the noise term (BridgeRegression.cpp:567-568), one item per coordinate
    double eps = 0.0;
    if (tid < P) { Stream s(L.key, gch, t, ST_BETA_STB, (uint32_t)tid); eps = s.norm(0); }

    // ---- right-looking Cholesky on register tiles, one barrier per column ----
    const double sig = sqrt(sig2);
    for (int k = 0; k < PP; ++k) {
      const int kb = k / TB, kl = k - kb * TB;
      double* cb = colbuf + (k & 1) * (PP + TB);
      if (tile_on && tj == kb) {
#pragma unroll
        for (int r = 0; r < TB; ++r) {
          double v = 0.0;
#pragma unroll
          for (int s = 0; s < TB; ++s) if (s == kl) v = A[r][s];
          cb[ti * TB + r] = v;
        }
      }
      __syncthreads();
      const double piv = cb[k];
      if (tile_on && tj >= kb) {
        const double ip = 1.0 / piv;
        double li[TB], lj[TB];
#pragma unroll
        for (int r = 0; r < TB; ++r) { const int g = ti * TB + r; li[r] = g > k ? cb[g] : 0.0; }
#pragma unroll
        for (int s = 0; s < TB; ++s) { const int g = tj * TB + s; lj[s] = g > k ? cb[g] * ip : 0.0; }
#pragma unroll
        for (int r = 0; r < TB; ++r)
#pragma unroll
          for (int s = 0; s < TB; ++s) A[r][s] = fma(-li[r], lj[s], A[r][s]);
      }
      if (tile_on && tj == kb) {   // finalise column k of L (off the critical path)
        if (!(piv > 0.0)) status |= ST_CHOL_FAIL;
        const double rs = rsqrt(piv);
#pragma unroll
        for (int r = 0; r < TB; ++r) {
          const int g = ti * TB + r;
          if (g >= k) {
            const double l = g == k ? piv * rs : cb[g] * rs;
            if (g < PP) Ls[g * PP + k] = l;
            else if (g == PP) rhs_s[k] = l;   // w_k = (L^-1 Xy)_k ; sig*eps is added below
            if (g == k) idiag_s[k] = rs;
          }
        }
      }
    }
    __syncthreads();
    if (tid < PP) rhs_s[tid] = tid < P ? fma(sig, eps, rhs_s[tid]) : 0.0;   // w + sig * eps
    __syncthreads();

    // ---- back-substitution  L' x = rhs   (the reference's trsm(U, ., 'U','L'), :565,570) ----
    if (warp == 0) {
      constexpr int Q = PP / 32;
      double r[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) r[q] = rhs_s[lane + 32 * q];
      for (int k = PP - 1; k >= 0; --k) {
        const int q0 = k >> 5;
        double rk = 0.0;
#pragma unroll
        for (int q = 0; q < Q; ++q) if (q == q0) rk = r[q];
        const double xk = __shfl_sync(0xffffffffu, rk, k & 31) * idiag_s[k];
        if (lane == 0) beta_s[k] = xk;
#pragma unroll
        for (int q = 0; q < Q; ++q) {
          const int i = lane + 32 * q;
          if (i < k) r[q] = fma(-Ls[k * PP + i], xk, r[q]);
        }
      }
    }
    __syncthreads();

    }  // do_beta

    // ---- alpha (only when it is unknown): BridgeRegression.cpp:477-503 ----
    if (do_alpha) {
      if (warp == 0) {
        Stream sa(L.key, gch, t, ST_ALPHA, 0);
        const AlphaProposal pr = alpha_propose(sa, alpha);
        double sn = 0.0, so = 0.0;
        for (int j = lane; j < P; j += 32) alpha_terms(beta_s[j], tau, pr, sn, so);
        sn = warp_sum(sn); so = warp_sum(so);
        if (lane == 0) red_s[0] = alpha_decide(sa, pr, sn, so, P, L.h.alpha_a, L.h.alpha_b);
      }
      __syncthreads();
      alpha = red_s[0];
      __syncthreads();
    }

    // ---- store ----
    const bool store = L.sink.thin > 0 && ((it + 1) % L.sink.thin) == 0;
    if (store) {
      const long long slot = L.sink.s0 + (it + 1) / L.sink.thin - 1;
      for (int j = tid; j < P; j += STB_THREADS) {
        const size_t o = ((size_t)ch * L.sink.chain_stride + slot) * P + j;
        if (L.sink.beta) L.sink.beta[o] = beta_s[j];
        if (L.sink.lambda) L.sink.lambda[o] = lam_s[j];
        if (L.sink.sum) { L.sink.sum[(size_t)ch * P + j] += beta_s[j]; L.sink.sumsq[(size_t)ch * P + j] += beta_s[j] * beta_s[j]; }
      }
      if (tid == 0) {
        const size_t o = (size_t)ch * L.sink.chain_stride + slot;
        if (L.sink.sig2) L.sink.sig2[o] = sig2;
        if (L.sink.tau) L.sink.tau[o] = tau;
        if (L.sink.alpha) L.sink.alpha[o] = alpha;
      }
    }
  }

  __syncthreads();
  for (int j = tid; j < P; j += STB_THREADS) {
    L.beta[(size_t)ch * P + j] = beta_s[j];
    L.lambda[(size_t)ch * P + j] = lam_s[j];
  }
  if (tid == 0) { L.tau[ch] = sc_s[0]; L.sig2[ch] = sc_s[1]; L.alpha[ch] = alpha; }
  status = __reduce_or_sync(0xffffffffu, status);
  if (lane == 0 && status) atomicOr(L.status + ch, status);
}

}  // namespace bb
