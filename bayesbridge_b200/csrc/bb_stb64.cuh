// bb_stb64.cuh -- the stable (normal scale mixture) Gibbs sweep for P <= 64: one chain per CTA of
// 128 threads, everything on chip.
//
// One launch advances `nchains` chains by `niter` iterations of
//   tau -> sig2 -> lambda -> beta                 (/root/reference/Code/C/BridgeWrapper.cpp:287-298)
// with   sample_tau_marg  BridgeRegression.cpp:453-465     sample_sig2  :436-450
//        sample_lambda    :506-510 -> retstable_LD retstable.cpp:94-271   (bb_retstable.cuh)
//        sample_beta_stable :552-575  (Cholesky of X'X + diag(lambda sig2/tau^2), triangular solves).
//
// The kernel is bound by instruction issue and by the length of its dependent chains, not by
// arithmetic (P^3/3 = 87k flops per chain-iteration), so the Cholesky is organised for FEW
// instructions and ONE barrier per column:
//   * the 64 x 64 system sits in shared memory with row stride 65 (conflict-free by rows and by columns);
//   * thread (t, h) = (tid & 63, tid >> 6) owns row t and the columns j = h (mod 2) of it: at step k it
//     applies  S(t,j) -= S(t,k)/S(k,k) * S(j,k)  for k < j <= t  -- S(j,k) is one broadcast LDS;
//   * columns are left UNSCALED during the factorisation (L(t,k) = S(t,k) rs_k with rs_k = S(k,k)^-1/2
//     published per step), so a step never rewrites what other threads still read and one
//     __syncthreads per column suffices; the thread that completes the next pivot publishes its
//     reciprocal and rsqrt before the barrier, off everyone else's critical path;
//   * w = L^-1 Xy (the reference's first trsm, :564) rides along as one extra FMA per row and step, and
//     since both of the reference's back-substitutions use the same factor, beta = U^-1 (w + sig eps) is
//     ONE back-substitution (:565,:570-574), done by one warp with the row of L broadcast from smem.
#pragma once
#include "bb_launch.h"
#include "bb_retstable.cuh"
#include "bb_alpha.cuh"

namespace bb {

constexpr int S64_T = 128;     // threads per CTA
constexpr int S64_P = 64;      // padded dimension
constexpr int S64_LD = 68;     // row stride of the system in shared memory (68 = 4 mod 16: DMMA fragment loads are 2-way at worst)
constexpr int S64_PB = 12;     // row stride of a published 8 x 8 tile / inverse diagonal block

__host__ __device__ inline size_t stb64_smem_bytes() {   // per chain; a multiple of 16
  return 15 + (size_t)(S64_P * S64_LD + 7 * S64_P + 2 * 8 * 8 * S64_PB + 32) * sizeof(double) + (size_t)(3 * S64_P + 4) * sizeof(int) + (size_t)S64_P * sizeof(RetConst) & ~(size_t)15;
}


// Cholesky factor and inverse factor of one 8 x 8 diagonal block, in registers, every lane redundantly
// (all indices static; no shuffles on the 8-pivot dependent chain).  W = L^-1 (lower) goes to Wout[8][S64_PB].
__device__ __forceinline__ void chol8_inverse(const double* __restrict__ tile, double* __restrict__ Wout, uint32_t& status) {
  double a[8][8], rs[8];
#pragma unroll
  for (int r = 0; r < 8; ++r)
#pragma unroll
    for (int c = 0; c <= r; ++c) a[r][c] = tile[r * S64_PB + c];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const double piv = a[k][k];
    if (!(piv > 0.0)) status |= ST_CHOL_FAIL;
    rs[k] = x_rsqrt(piv);
#pragma unroll
    for (int r = k + 1; r < 8; ++r) a[r][k] *= rs[k];
#pragma unroll
    for (int r = k + 1; r < 8; ++r)
#pragma unroll
      for (int c = k + 1; c <= r; ++c) a[r][c] = fma(-a[r][k], a[c][k], a[r][c]);
  }
  // W = L^-1 column by column: W(c,c) = 1/L(c,c), W(r,c) = -1/L(r,r) sum_{m=c}^{r-1} L(r,m) W(m,c)
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    double w[8];
    w[c] = rs[c];
#pragma unroll
    for (int r = c + 1; r < 8; ++r) {
      double sum = 0.0;
#pragma unroll
      for (int m = c; m < r; ++m) sum = fma(a[r][m], w[m], sum);
      w[r] = -rs[r] * sum;
    }
#pragma unroll
    for (int r = c; r < 8; ++r) Wout[r * S64_PB + c] = w[r];
  }
}

// Phase timing (development, -DBB_PHASES): thread 0 of CTA 0 accumulates the cycles between the phase marks of its chain.
#ifdef BB_PHASES
#define BB_PHASE_MARK(k) do { if (blockIdx.x == 0 && threadIdx.x == 0) { const long long now_ = clock64(); bb_phase_cyc[k] += (unsigned long long)(now_ - ph_t); ph_t = now_; } } while (0)
#else
#define BB_PHASE_MARK(k) do { } while (0)
#endif

// NCH chains share one CTA (128 threads each).  All barriers are CTA-wide, so the chains of an SM walk
// through the kernel's phases together and the instruction cache holds ONE phase at a time; with four
// independent CTAs per SM the same code ran at four different places and `stall_no_instruction` was the
// largest stall reason after the barriers (profiles/r1_c2_*.md).
template <bool TC, int NCH>
__global__ void __launch_bounds__(S64_T * NCH, 4 / NCH) stb64_kernel(const StbLaunch L) {
  extern __shared__ double smem_all[];
  const int P = L.m.P;
  const int cslot = threadIdx.x / S64_T;
  double* smem = smem_all + (size_t)cslot * (stb64_smem_bytes() / sizeof(double));
  const int tid = threadIdx.x - cslot * S64_T, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = S64_T / 32;
  constexpr int PP = S64_P, LD = TC ? S64_LD : 65;   // the scalar path reads columns: stride 65 is conflict-free there

  double* S = smem;                        // [PP][LD]  lower triangle: unscaled columns of the factor
  double* beta_s = S + PP * LD;            // [PP]
  double* lam_s = beta_s + PP;             // [PP]
  double* dl_s = lam_s + PP;               // [PP]   beta - bc
  double* rhs_s = dl_s + PP;               // [PP]   w + sig*eps
  double* rs_s = rhs_s + PP;               // [PP]   1 / sqrt(pivot_k) = 1 / L(k,k)
  double* ip_s = rs_s + PP;                // [PP]   1 / pivot_k
  double* rb_s = ip_s + PP;                // [PP]   forward-substitution residual r_k when row k becomes the pivot row
  double* pb_s = rb_s + PP;                // [8][8][PB] published tiles of the current block column (TC path)
  double* wi_s = pb_s + 8 * 8 * S64_PB;    // [8][8][PB] inverses of the diagonal blocks of the factor (TC path)
  double* red_s = wi_s + 8 * 8 * S64_PB;   // [3][8]
  double* sc_s = red_s + 24;               // tau, sig2 (+ pad to 32)
  int* q_s = reinterpret_cast<int*>(red_s + 32);            // [2][PP] retry queues
  int* qn_s = q_s + 2 * PP;                                 // [2] counts
  uint32_t* kk_s = reinterpret_cast<uint32_t*>(qn_s + 2);   // [PP] next Philox block per coordinate
  RetConst* rc_s = reinterpret_cast<RetConst*>(kk_s + PP + 2);   // [PP]

  const bool valid = blockIdx.x * NCH + cslot < L.nchains;       // a padding slot repeats the last chain, stores nothing
  const int ch = valid ? blockIdx.x * NCH + cslot : L.nchains - 1;
  const uint32_t gch = L.chain0 + (uint32_t)ch;
  const int t = tid & 63, h = tid >> 6;    // row, column parity

  const bool know_tau = L.h.true_tau > 0, know_sig2 = L.h.true_sig2 > 0, know_alpha = L.h.true_alpha > 0;
  const bool do_tau = (L.stages & 1) && !know_tau, do_sig2 = (L.stages & 2) && !know_sig2;
  const bool do_lambda = L.stages & 32, do_beta = L.stages & 16, do_alpha = (L.stages & 64) && !know_alpha;
  double alpha = know_alpha ? L.h.true_alpha : L.alpha[ch];   // unknown alpha: chain state, moved by the Metropolis step below
  uint32_t status = 0;

  for (int j = tid; j < PP; j += S64_T) {
    beta_s[j] = j < P ? L.beta[(size_t)ch * P + j] : 0.0;
    lam_s[j] = j < P ? L.lambda[(size_t)ch * P + j] : 1.0;
  }
  if (tid == 0) { sc_s[0] = know_tau ? L.h.true_tau : L.tau[ch]; sc_s[1] = know_sig2 ? L.h.true_sig2 : L.sig2[ch]; qn_s[0] = 0; qn_s[1] = 0; }
  for (int e = tid; e < 8 * 8 * S64_PB; e += S64_T) wi_s[e] = 0.0;   // the upper triangles stay zero
  __syncthreads();

  const int bar_id = 1 + cslot;   // named barrier of this chain's 128 threads (inside the factorisation only)

#ifdef BB_PHASES
  long long ph_t = clock64();
#endif
  for (int it = -L.pre_tau; it < L.niter; ++it) {
    const bool pre = it < 0;
    const uint32_t tt = L.iter0 + (pre ? 0u : (uint32_t)it);
    BB_PHASE_MARK(7);

    // ---- partial sums for tau and the residual sum of squares ----
    if (tid < PP) dl_s[tid] = tid < P ? beta_s[tid] - __ldg(L.m.bc + tid) : 0.0;
    // Warps 2 and 3 have nothing to do until the lambda stage (the partial sums and the gamma draws below belong to
    // warps 0 and 1): in those two windows they stage X'X (lower 8 x 8 tiles, diagonal tiles in full) into S -- free
    // until the panel step of the factorisation writes it -- and draw the normals of the noise term (:567-568), so
    // that after the lambda stage the tile fragments come from shared memory instead of 18 global loads per thread.
    const bool stager = TC && do_beta && !pre && warp >= 2;
    auto stage_rows = [&](int i0, int i1) {
      const int u = tid - 64, er = u & 7, ec = u >> 3;
#pragma unroll 1
      for (int i = i0; i < i1; ++i) {
        const int row = 8 * i + er;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (j <= i) {
            const int col = 8 * j + ec;
            S[row * LD + col] = (row < P && col < P) ? __ldg(L.m.XX + row + (size_t)col * P) : (row == col ? 1.0 : 0.0);
          }
        }
      }
    };
    __syncthreads();
    if (stager) stage_rows(0, 5);
    double p_tau = 0.0, p_q = 0.0, p_g = 0.0;
    if (tid < P) {
      const double b = beta_s[tid];
      p_tau = x_exp(alpha * x_log(fabs(b)));                       // BridgeRegression.cpp:459
      if (!pre && do_sig2) {                                   // rss = r0 - 2 d'g + d'XXd, d = beta - bc
        double s = 0.0;
        for (int j = 0; j < P; ++j) s = fma(__ldg(L.m.XX + tid + (size_t)j * P), dl_s[j], s);
        p_q = dl_s[tid] * s;
        p_g = dl_s[tid] * __ldg(L.m.gc + tid);
      }
    }
    p_tau = warp_sum(p_tau); p_q = warp_sum(p_q); p_g = warp_sum(p_g);
    if (lane == 0) { red_s[warp] = p_tau; red_s[8 + warp] = p_q; red_s[16 + warp] = p_g; }
    __syncthreads();
    if (stager) {
      stage_rows(5, 8);
      const int u = tid - 64;
      double e = 0.0;
      if (u < P) { Stream s(L.key, gch, tt, ST_BETA_STB, (uint32_t)u); e = s.norm(0); }
      ip_s[u] = e;
    }
    if (warp == 0 && (do_tau || (pre && !know_tau))) {          // BridgeRegression.cpp:453-465
      double st = lane < NW ? red_s[lane] : 0.0;
      st = warp_sum(st);
      const double shape = L.h.nu_shape + ((double)P) / alpha;
      const double rate = L.h.nu_rate + st;
      Stream s(L.key, gch, tt, pre ? ST_TAU_INIT : ST_TAU, 0);
      const double nu = gamma_rate<true>(s, shape, rate, status);
      if (lane == 0) sc_s[0] = exp(-1.0 * log(nu) / alpha);
    }
    if (warp == 1 && !pre && do_sig2) {                         // BridgeRegression.cpp:436-450
      double sq = lane < NW ? red_s[8 + lane] : 0.0, sg = lane < NW ? red_s[16 + lane] : 0.0;
      sq = warp_sum(sq); sg = warp_sum(sg);
      const double rss = L.m.r0b - 2.0 * sg + sq;
      const double shape = L.h.sig2_shape + 0.5 * L.m.N;
      const double scale = L.h.sig2_scale + 0.5 * rss;
      Stream s(L.key, gch, tt, ST_SIG2, 0);
      const double g = gamma_rate<true>(s, shape, 1.0, status);
      if (lane == 0) sc_s[1] = scale / g;
    }
    __syncthreads();
    if (pre) continue;
    const double tau = sc_s[0], sig2 = sc_s[1];
    BB_PHASE_MARK(0);

    // ---- lambda_j = 2 * retstable(beta_j^2 / tau^2, alpha/2): retry queue + speculative look-ahead ----
    if (do_lambda) lambda_stage_cta<S64_T>(tid, P, beta_s, lam_s, tau, 0.5 * alpha, L.key, gch, tt, rc_s, q_s, PP, qn_s, kk_s, status);
    else __syncthreads();
    BB_PHASE_MARK(1);

    if (do_beta && !TC) {
      // ---- S = X'X + diag(lambda sig2 / tau^2)   (BridgeRegression.cpp:554-556), lower triangle ----
      const double f = sig2 / (tau * tau);
      for (int j = h; j <= t; j += 2) {
        double v = (t < P && j < P) ? __ldg(L.m.XX + t + (size_t)j * P) : (t == j ? 1.0 : 0.0);
        if (t == j && t < P) v += lam_s[t] * f;
        S[t * LD + j] = v;
        if (t == 0) {   // first pivot (thread (0,0))
          if (!(v > 0.0)) status |= ST_CHOL_FAIL;
          ip_s[0] = 1.0 / v; rs_s[0] = rsqrt(v);
        }
      }
      double r = (h == 0 && t < P) ? __ldg(L.m.Xy + t) : 0.0;   // forward-substitution residual of row t
      if (tid == 0) rb_s[0] = r;
      // normals for the noise term (BridgeRegression.cpp:567-568), one item per coordinate
      double eps = 0.0;
      if (tid < P) { Stream s(L.key, gch, tt, ST_BETA_STB, (uint32_t)tid); eps = s.norm(0); }
      __syncthreads();

      // ---- right-looking Cholesky with unscaled columns, one barrier per column ----
      for (int k = 0; k < PP - 1; ++k) {
        if (t > k) {
          const double lip = S[t * LD + k] * ip_s[k];          // S(t,k) / pivot_k  =  L(t,k) rs_k
          int j = k + 1 + ((k + 1 + h) & 1);                   // first column > k of my parity
          double* row = S + t * LD;
          const double* colk = S + k;
#pragma unroll 4
          for (; j <= t; j += 2) row[j] = fma(-lip, colk[j * LD], row[j]);
          if (h == 0) r = fma(-lip, rb_s[k], r);               // r_t -= L(t,k) w_k
          if (t == k + 1) {                                    // row k+1 is the next pivot row
            if (h == ((k + 1) & 1)) {
              const double piv = row[t];
              if (!(piv > 0.0)) status |= ST_CHOL_FAIL;
              ip_s[k + 1] = 1.0 / piv; rs_s[k + 1] = rsqrt(piv);
            }
            if (h == 0) rb_s[k + 1] = r;
          }
        }
        __syncthreads();
      }
      // w = L^-1 Xy is rb_s[k] rs_k;  rhs = w + sig eps
      const double sig = sqrt(sig2);
      if (tid < PP) rhs_s[tid] = tid < P ? fma(sig, eps, rb_s[tid] * rs_s[tid]) : 0.0;
      __syncthreads();

      // ---- back-substitution  L' x = rhs  with L(k,i) = S(k,i) rs_i :  x_i = (rhs_i - rs_i acc_i) rs_i,
      //      acc_i = sum_{k>i} S(k,i) x_k   (the reference's trsm(U, ., 'U','L'), :565,570) ----
      if (warp == 0) {
        const double rs0 = rs_s[lane], rs1 = rs_s[lane + 32];
        const double rh0 = rhs_s[lane], rh1 = rhs_s[lane + 32];
        double acc0 = 0.0, acc1 = 0.0;
        for (int k = PP - 1; k >= 0; --k) {
          const double sk0 = S[k * LD + lane], sk1 = S[k * LD + lane + 32];   // row k (independent of x_k: issued early)
          const double cand = (k >= 32) ? (rh1 - rs1 * acc1) * rs1 : (rh0 - rs0 * acc0) * rs0;
          const double xk = __shfl_sync(0xffffffffu, cand, k & 31);
          if (lane == 0) beta_s[k] = xk;
          if (lane < k) acc0 = fma(sk0, xk, acc0);
          if (lane + 32 < k) acc1 = fma(sk1, xk, acc1);
        }
      }
      __syncthreads();
    }  // do_beta, scalar path

    if (do_beta && TC) {
      // ================= blocked Cholesky on the FP64 tensor cores (DMMA m8n8k4) =================
      // The 36 lower 8 x 8 tiles of  X'X + diag(lambda sig2 / tau^2)  (BridgeRegression.cpp:554-556) live in
      // registers as DMMA accumulator fragments: tile n = i(i+1)/2 + j belongs to warp n & 3, slot n >> 2.
      // Block column s:  publish tiles (i,s) -> warp 0 factors the diagonal block and inverts its factor
      // (registers, static indices) -> panel L(i,s) = T(i,s) W_s' by DMMA -> trailing T(i,j) -= L(i,s) L(j,s)'
      // by DMMA.  3 barriers per block column, among the chain's own 4 warps (named barrier).
      const double f = sig2 / (tau * tau);
      const int fr = lane >> 2, fc = 2 * (lane & 3);      // accumulator fragment: row fr, columns fc, fc+1
      const int ar = lane >> 2, ak = lane & 3;            // A/B fragment: row ar, k = ak
      double c0[9], c1[9];
      int ti[9], tj[9];
#pragma unroll
      for (int sl = 0; sl < 9; ++sl) {
        const int n = sl * 4 + warp;   // recomputed every iteration: keeping 18 indices (even packed in two words)
        int i = 0;                     // alive across the lambda stage costs more in spills than these few integer ops
        while ((i + 1) * (i + 2) / 2 <= n) ++i;
        ti[sl] = i; tj[sl] = n - i * (i + 1) / 2;
        const int row = 8 * i + fr, col = 8 * tj[sl] + fc;
        const double2 v = *reinterpret_cast<const double2*>(S + row * LD + col);   // staged during the scalar stage
        double v0 = v.x, v1 = v.y;
        if (row == col && row < P) v0 += lam_s[row] * f;
        if (row == col + 1 && row < P) v1 += lam_s[row] * f;
        c0[sl] = v0; c1[sl] = v1;
      }
      if (tid < PP) rhs_s[tid] = tid < P ? __ldg(L.m.Xy + tid) : 0.0;
      BB_PHASE_MARK(2);

      for (int s = 0; s < 8; ++s) {
        // (1) publish the tiles of block column s
#pragma unroll
        for (int sl = 0; sl < 9; ++sl)
          if (tj[sl] == s) *reinterpret_cast<double2*>(pb_s + (ti[sl] * 8 + fr) * S64_PB + fc) = make_double2(c0[sl], c1[sl]);
        bar_sync(bar_id, S64_T);
        // (2) diagonal block: factor and inverse factor (warp 0; four warps doing it redundantly to save the barrier
        //     was slower: the FP64 pipe of the SM is shared by the 16 warps of the CTA)
        if (warp == 0) chol8_inverse(pb_s + s * 8 * S64_PB, wi_s + s * 8 * S64_PB, status);
        bar_sync(bar_id, S64_T);
        // (3) panel: L(i,s) = T(i,s) W_s'  for i > s, tile i by warp i & 3
        for (int i = s + 1 + ((warp - (s + 1)) & 3); i < 8; i += 4) {
          double d0 = 0.0, d1 = 0.0;
#pragma unroll
          for (int k0 = 0; k0 < 8; k0 += 4)
            dmma_m8n8k4(d0, d1, pb_s[(i * 8 + ar) * S64_PB + k0 + ak], wi_s[(s * 8 + ar) * S64_PB + k0 + ak]);
          *reinterpret_cast<double2*>(S + (8 * i + fr) * LD + 8 * s + fc) = make_double2(d0, d1);
        }
        bar_sync(bar_id, S64_T);
        // (4) trailing update of my tiles (i,j), i >= j > s
#pragma unroll
        for (int sl = 0; sl < 9; ++sl)
          if (tj[sl] > s) {
#pragma unroll
            for (int k0 = 0; k0 < 8; k0 += 4)
              dmma_m8n8k4(c0[sl], c1[sl], -S[(8 * ti[sl] + ar) * LD + 8 * s + k0 + ak], S[(8 * tj[sl] + ar) * LD + 8 * s + k0 + ak]);
          }
      }
      __syncthreads();
      BB_PHASE_MARK(3);
      // ---- w = L^-1 Xy (the reference's first trsm, :564), blocked with the inverse diagonal blocks ----
      const double sig = sqrt(sig2);
      if (warp == 0) {
        for (int s = 0; s < 8; ++s) {
          if (lane < 8) {
            double w = 0.0;
#pragma unroll
            for (int c = 0; c < 8; ++c) w = fma(wi_s[(s * 8 + lane) * S64_PB + c], rhs_s[8 * s + c], w);   // W is lower: zeros above
            rb_s[8 * s + lane] = w;
          }
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const int row = lane + 32 * q;
            if (row >= 8 * (s + 1)) {
              double r = rhs_s[row];
#pragma unroll
              for (int c = 0; c < 8; ++c) r = fma(-S[row * LD + 8 * s + c], rb_s[8 * s + c], r);
              rhs_s[row] = r;
            }
          }
          __syncwarp();
        }
        // rhs = w + sig eps;  then  L' x = rhs  backwards (:565, :570-574)
        rhs_s[lane] = lane < P ? fma(sig, ip_s[lane], rb_s[lane]) : 0.0;      // ip_s: the normals drawn in the scalar stage
      }
      __syncthreads();
      if (warp == 1) rhs_s[32 + lane] = (32 + lane) < P ? fma(sig, ip_s[32 + lane], rb_s[32 + lane]) : 0.0;
      __syncthreads();
      if (warp == 0) {
        for (int s = 7; s >= 0; --s) {
          if (lane < 8) {
            double x = 0.0;
#pragma unroll
            for (int r = 0; r < 8; ++r) x = fma(wi_s[(s * 8 + r) * S64_PB + lane], rhs_s[8 * s + r], x);   // W' y
            beta_s[8 * s + lane] = x;
          }
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const int col = lane + 32 * q;
            if (col < 8 * s) {
              double y = rhs_s[col];
#pragma unroll
              for (int r = 0; r < 8; ++r) y = fma(-S[(8 * s + r) * LD + col], beta_s[8 * s + r], y);
              rhs_s[col] = y;
            }
          }
          __syncwarp();
        }
      }
      __syncthreads();
      BB_PHASE_MARK(4);
    }  // do_beta, tensor-core path

    // ---- alpha (only when it is unknown): BridgeRegression.cpp:477-503 ----
    if (do_alpha) {
      if (warp == 0) {
        Stream sa(L.key, gch, tt, ST_ALPHA, 0);
        const AlphaProposal pr = alpha_propose(sa, alpha);
        double sn = 0.0, so = 0.0;
        for (int j = lane; j < P; j += 32) alpha_terms(beta_s[j], tau, pr, sn, so);
        sn = warp_sum(sn); so = warp_sum(so);
        if (lane == 0) sc_s[2] = alpha_decide(sa, pr, sn, so, P, L.h.alpha_a, L.h.alpha_b);
      }
      __syncthreads();
      alpha = sc_s[2];
    }

    // ---- store ----
    const bool store = valid && L.sink.thin > 0 && ((it + 1) % L.sink.thin) == 0;
    if (store) {
      const long long slot = L.sink.s0 + (it + 1) / L.sink.thin - 1;
      for (int j = tid; j < P; j += S64_T) {
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
    BB_PHASE_MARK(5);
  }

  __syncthreads();
  if (!valid) return;
  for (int j = tid; j < P; j += S64_T) {
    L.beta[(size_t)ch * P + j] = beta_s[j];
    L.lambda[(size_t)ch * P + j] = lam_s[j];
  }
  if (tid == 0) { L.tau[ch] = sc_s[0]; L.sig2[ch] = sc_s[1]; L.alpha[ch] = alpha; }
  status = __reduce_or_sync(0xffffffffu, status);
  if (lane == 0 && status) atomicOr(L.status + ch, status);
}

}  // namespace bb
