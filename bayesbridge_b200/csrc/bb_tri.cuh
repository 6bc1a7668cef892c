// bb_tri.cuh -- the triangle-mixture Gibbs sweep as one persistent kernel.
//
// One launch advances `nchains` independent chains by `niter` iterations of
//   tau -> sig2 -> omega -> u -> beta            (/root/reference/Code/C/BridgeWrapper.cpp:168-189)
// with   sample_tau_marg  BridgeRegression.cpp:453-465      sample_sig2  :436-450
//        sample_omega     :130-147     sample_u  :97-110     sample_beta  :405-433
//        rtnorm_gibbs (SVD form, the innermost loop)  :235-286.
//
// The rotated sweep, restated.  Step i of :250-283 needs, over all j,
//     left_j  = (v>0 ? -(b_j + r_j) : -(b_j - r_j)) / |v|,   right_j = (v>0 ? b_j - r_j : b_j + r_j) / |v|,
//     v = V(j,i),  r_j = beta_j - v z_i   (beta_j = current V z, kept incrementally: Notes/bbnotes.tex:603-612).
// With q = 1/v both signs of v collapse to
//     left_j = z_i - (beta_j q + b_j |q|),      right_j = z_i + (b_j |q| - beta_j q),
// so  lmax = z_i - min_j fma(beta_j, q, b_j|q|),  rmin = z_i + min_j fma(-beta_j, q, b_j|q|):
// one multiply, two FMAs and two minima per (i,j), plus the FMA beta_j += v (z_i' - z_i) after the
// draw: 6 FP64 operations per (i,j).  Where v = 0 (or j is padding) q = +-inf and both candidates
// come out +inf or NaN, which fmin ignores -- the reference's division by zero (:261-262) likewise
// produces bounds that never bind.
//
// Layout.  A CTA holds NG groups; a group is W warps that own C chains: lane l of warp w keeps
// coordinates j = jj*32W + 32w + l (jj < J) of all C chains in registers (beta, b), so each column
// of {V, 1/V} is read once per group-step and the C independent chains interleave in one
// instruction stream (latency hiding).  Columns are streamed from L2 into a shared-memory ring by
// one producer thread with TMA bulk copies (cp.async.bulk + mbarrier complete_tx); a column is
// released as soon as every warp has copied its slice into registers.  Per step: local minima ->
// 5 shuffle levels -> (W > 1: partials through smem, one named barrier) -> ONE warp of the group
// draws all C truncated normals at once (32/C lanes = parallel rejection trials per chain; the
// lowest accepting trial wins, which is what the sequential sampler would have returned) ->
// delta_c through smem -> axpy.  When the whole {V,1/V} fits (P <= 64) the ring is filled once.
#pragma once
#include <cfloat>
#include <math_constants.h>
#include "bb_launch.h"
#include "bb_alpha.cuh"

namespace bb {

// min that keeps `m` when `x` is NaN: one DSETP + two SEL (fmin() carries a NaN fix-up and costs twice that)
__device__ __forceinline__ double dmin_keep(double m, double x) { return (x < m) ? x : m; }
__device__ __forceinline__ double shfl_xor_d(double v, int o) { return __shfl_xor_sync(0xffffffffu, v, o); }

// Warp-wide minimum of one double on the integer REDUX unit: the doubles are mapped to their order-preserving 64-bit
// integer image, the high words are reduced (redux.sync.min.u32), then the low words among the lanes that tie on the high
// word.  Two dependent REDUX instead of five shuffle rounds of (2 SHFL + DSETP + 2 SEL); every lane receives the result.
// Inputs must not be NaN (the per-lane minima never are: dmin_keep drops NaN candidates, the start value is +inf).
// -0.0 orders below +0.0 here; both bound the same interval.  Measured (scripts/mb/mb_reduce.cu): 4 values 126 cycles
// against 230 for the transposed butterfly.
__device__ __forceinline__ uint32_t redux_min_u32(uint32_t v) {
  uint32_t r; asm volatile("redux.sync.min.u32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v)); return r;
}
__device__ __forceinline__ double warp_min_redux(double x) {
  const long long b = __double_as_longlong(x);
  const unsigned long long k = (unsigned long long)(b ^ ((b >> 63) | (long long)0x8000000000000000ULL));
  const uint32_t hi = (uint32_t)(k >> 32), lo = (uint32_t)k;
  const uint32_t mh = redux_min_u32(hi);
  const uint32_t ml = redux_min_u32(hi == mh ? lo : 0xffffffffu);
  const unsigned long long m = ((unsigned long long)mh << 32) | ml;
  const long long r = (long long)(m ^ ((m >> 63) ? 0x8000000000000000ULL : 0xffffffffffffffffULL));
  return __longlong_as_double(r);
}

// Reduce NV = 2C per-lane minima over the 32 lanes with a TRANSPOSED butterfly: at the first log2(NV)
// levels a lane keeps one half of its values and hands the other half to its partner, so the number of
// live values halves together with the distance; plain butterflies finish the last value.  On return
// lane l holds the complete minimum of value index  (l >> (5 - log2 NV))  -- for C = 2: lanes 0-7 L0,
// 8-15 R0, 16-23 L1, 24-31 R1 -- i.e. the {L,R} pair of chain c sits in the lanes of draw slice c.
// 2C = 4: 7 shuffle rounds of one double instead of 20.
template <int C>
__device__ __forceinline__ double reduce_minima(const double (&lmn)[C], const double (&rmn)[C], int lane) {
#ifndef BB_TRI_BUTTERFLY
  if (C <= 2) {   // same lane layout as the butterflies below: value v = (L_0, R_0[, L_1, R_1]) in lanes [v 32/(2C), (v+1) 32/(2C))
    double out = warp_min_redux(lmn[0]);
    const double r0 = warp_min_redux(rmn[0]);
    if (C == 1) return (lane & 16) ? r0 : out;
    const double l1 = warp_min_redux(lmn[C - 1]), r1 = warp_min_redux(rmn[C - 1]);
    return (lane & 16) ? ((lane & 8) ? r1 : l1) : ((lane & 8) ? r0 : out);
  }
#endif
  if (C == 1) {
    const bool hi = lane & 16;
    double keep = hi ? rmn[0] : lmn[0], send = hi ? lmn[0] : rmn[0];
    keep = dmin_keep(keep, shfl_xor_d(send, 16));
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) keep = dmin_keep(keep, shfl_xor_d(keep, o));
    return keep;                                   // lanes 0-15: L, lanes 16-31: R
  } else if (C == 2) {
    const bool h16 = lane & 16, h8 = lane & 8;
    double k0 = h16 ? lmn[1] : lmn[0], k1 = h16 ? rmn[1] : rmn[0];
    const double s0 = h16 ? lmn[0] : lmn[1], s1 = h16 ? rmn[0] : rmn[1];
    k0 = dmin_keep(k0, shfl_xor_d(s0, 16)); k1 = dmin_keep(k1, shfl_xor_d(s1, 16));
    double keep = h8 ? k1 : k0;
    const double send = h8 ? k0 : k1;
    keep = dmin_keep(keep, shfl_xor_d(send, 8));
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) keep = dmin_keep(keep, shfl_xor_d(keep, o));
    return keep;                                   // lanes 0-7 L0, 8-15 R0, 16-23 L1, 24-31 R1
  } else {
    // generic fallback (C = 4): full butterflies, then pick
    double out = 0.0;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      double a = lmn[c], b = rmn[c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { a = dmin_keep(a, shfl_xor_d(a, o)); b = dmin_keep(b, shfl_xor_d(b, o)); }
      if ((lane >> 3) == c) out = (lane & 4) ? b : a;
    }
    return out;                                    // lanes 8c..8c+3: L_c, 8c+4..8c+7: R_c
  }
}

// Phase timing (development, -DBB_TRI_PHASES): one thread of CTA 0 (BB_TRI_PHASE_TID, default 0) accumulates, in shared memory,
// the cycles between the marks of the coordinate step; [0..7] steps in which its warp draws, [8..15] the other steps:
// 7 = loop top + TMA issue, 0 = column wait + load, 1 = candidates + butterfly, 2 = barrier 1, 3 = draw (or RNG look-ahead),
// 4 = barrier 2, 5 = axpy; [6] / [14] count the steps.  Each mark costs about 40 cycles itself.
#ifdef BB_TRI_PHASES
#ifndef BB_TRI_PHASE_TID
#define BB_TRI_PHASE_TID 0
#endif
__device__ unsigned long long tri_phase_cyc[32];
#define BB_TRI_MARK(k) do { if (ph_on) { const long long now_ = clock64(); ph_acc[ph_half + (k)] += (unsigned long long)(now_ - ph_t); ph_t = now_; } } while (0)
#else
#define BB_TRI_MARK(k) do { } while (0)
#endif

// shared-memory footprint of one CTA (bytes); mirrored by the host launcher
__host__ __device__ inline size_t tri_smem_bytes(int P, int LD, int S, int NG, int W, int C) {
  const size_t CH = (size_t)NG * C;
  return (size_t)S * (LD + 2) * 16 + 2 * CH * P * 8 + (size_t)(2 * NG * W * C * 2 + NG * C + NG * C * 3) * 8 + (size_t)2 * S * 8;
}

// registers live per SM sub-partition (16384 each, warps dealt round-robin over the 4 of them)
constexpr int tri_maxnreg(int threads) {
  const int wps = (threads + 127) / 128;
  return 16384 / (wps * 32) >= 255 ? 255 : (16384 / (wps * 32) / 8) * 8;
}

template <int J, int W, int C, int NGMAX>
__global__ void __maxnreg__(tri_maxnreg(32 * W * NGMAX)) tri_sweep_kernel(const TriLaunch L, const int S) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int P = L.m.P, LD = L.m.ldv;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int nwarp = blockDim.x >> 5;
  const int NG = nwarp / W;
  const int g = wid / W, w = wid - g * W;
  const int CH = NG * C;

  const int LQ = LD + 2;                                        // column record: LD pairs {V(j,i), 1/V(j,i)} + {d_i, mean_i} + {1/d_i, 0}
  double2* ring = reinterpret_cast<double2*>(smem_raw);        // [S][LQ]
  double* z_s = reinterpret_cast<double*>(ring + (size_t)S * LQ);   // [CH][P]
  double* u_s = z_s + (size_t)CH * P;                          // [CH][P]
  double* red = u_s + (size_t)CH * P;                          // [2][NG][W][C][2]
  double* bc = red + 2 * NG * W * C * 2;                       // [NG][C]    delta of the current step
  double* sc = bc + NG * C;                                    // [NG][C][2] tau, sig2
  double* al_s = sc + NG * C * 2;                              // [NG][C]    alpha of each chain
  uint64_t* full = reinterpret_cast<uint64_t*>(al_s + NG * C);      // [S]
  uint64_t* empty = full + S;                                  // [S]

  const bool know_tau = L.h.true_tau > 0, know_sig2 = L.h.true_sig2 > 0, know_alpha = L.h.true_alpha > 0;
  const bool do_alpha = (L.stages & 64) && !know_alpha;   // unknown alpha: chain state moved by a Metropolis step per iteration
  const bool do_tau = (L.stages & 1) && !know_tau, do_sig2 = (L.stages & 2) && !know_sig2;
  const bool do_omega = L.stages & 4, do_u = L.stages & 8, do_beta = L.stages & 16;
  const int nsweep = (L.h.betaburn > 0 ? L.h.betaburn : 0) + 1;
  const bool resident = S >= P;
  const long long sweep_steps = do_beta ? (long long)L.niter * nsweep * P : 0;
  const long long total_steps = (L.init_z ? P : 0) + sweep_steps;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], (uint32_t)(NG * W)); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (w == 0 && lane < C) {
    const int ch = blockIdx.x * CH + g * C + lane;
    al_s[g * C + lane] = know_alpha ? L.h.true_alpha : L.alpha[ch < L.nchains ? ch : L.nchains - 1];
  }
  __syncthreads();

  // ======================= column stream: thread 0 issues the TMA copies =======================
  // Step m reuses the stage of step m - S.  Thread 0 issues step n + S - D when it enters step n, after
  // every warp has released step n - D: S - D columns of prefetch, D steps of slack between the groups.
  const bool issuer = tid == 0;
  const uint32_t col_bytes = (uint32_t)LQ * 16u;
  const long long nload = resident ? (total_steps < P ? total_steps : (long long)P) : total_steps;
  const int D = S > 1 ? S / 2 : 1;
  int is_ = 0, icol = 0; uint32_t iuse = 0; long long inext = 0;   // issuer cursor: stage, column, use count, step
  auto issue_one = [&]() {
    if (iuse > 0) mbar_wait(&empty[is_], (iuse - 1) & 1);
    mbar_arrive_expect_tx(&full[is_], col_bytes);
    tma_load_1d(ring + (size_t)is_ * LQ, L.m.VQ + (size_t)icol * LQ * 2, col_bytes, &full[is_]);
    if (++is_ == S) { is_ = 0; ++iuse; }
    if (++icol == P) icol = 0;
    ++inext;
  };
  if (issuer) while (inext < nload && inext < S) issue_one();

  // ======================= consumers =======================
  const int bar_id = 1 + g, bar_n = 32 * W;
  auto group_sync = [&]() { if (W > 1) bar_sync(bar_id, bar_n); else __syncwarp(); };
  constexpr int SL = 32 / C;                 // lanes (= parallel rejection trials) per chain in the draw
  const int dc = lane / SL, sl = lane - dc * SL;
  const uint32_t slmask = (SL == 32) ? 0xffffffffu : (((1u << SL) - 1u) << (dc * SL));

  const int cbase = blockIdx.x * CH + g * C;
  int lch[C]; bool valid[C];
#pragma unroll
  for (int c = 0; c < C; ++c) { valid[c] = cbase + c < L.nchains; lch[c] = valid[c] ? cbase + c : (L.nchains - 1); }
  const int lch_d = (cbase + dc < L.nchains) ? cbase + dc : (L.nchains - 1);
  const uint32_t gch_d = L.chain0 + (uint32_t)lch_d;
  double* zg = z_s + (size_t)(g * C) * P;    // this group's z rows
  double* ug = u_s + (size_t)(g * C) * P;
  double* redg0 = red + (size_t)(g * W) * C * 2;                       // buffer 0, [W][C][2]
  double* redg1 = red + (size_t)(NG * W + g * W) * C * 2;              // buffer 1

  // ---------------- load state ----------------
  double cur[C][J], bb_[C][J];
#pragma unroll
  for (int c = 0; c < C; ++c)
#pragma unroll
    for (int jj = 0; jj < J; ++jj) {
      const int j = jj * 32 * W + 32 * w + lane;
      cur[c][jj] = j < P ? L.beta[(size_t)lch[c] * P + j] : 0.0;
      bb_[c][jj] = 1.0;
      if (j < P) ug[c * P + j] = L.u[(size_t)lch[c] * P + j];
    }
  double tau_c[C], sig2_c[C];
#pragma unroll
  for (int c = 0; c < C; ++c) {
    tau_c[c] = know_tau ? L.h.true_tau : L.tau[lch[c]];
    sig2_c[c] = know_sig2 ? L.h.true_sig2 : L.sig2[lch[c]];
  }
  uint32_t status = 0;        // bits for chain dc (the chain this lane serves in the draw)

  // ring cursor
  int rs = 0; uint32_t ruse = 0; long long nstep = 0;
  double2 vq[J], dm, dm2;   // dm = {d_i, mean_i}, dm2 = {1/d_i, -}
#ifdef BB_TRI_PHASES
  bool ph_on = false; int ph_half = 0; long long ph_t = 0;
  __shared__ unsigned long long ph_acc[32];
  if (tid < 32) ph_acc[tid] = 0;
  __syncthreads();
#endif
  // Thread 0 keeps the ring S - D columns ahead; called where warp 0 is idle (see the sweep), at most two copies per call.
  auto maybe_issue = [&]() {
    if (!issuer) return;
    for (int k = 0; k < 2 && inext < nload && inext <= nstep + S - D; ++k) issue_one();
  };
  // `waited`: a warp of this group has already observed the column's barrier and a group barrier has passed since.
  auto next_column = [&](bool waited) {   // wait for the column, copy this lane's slice into registers, release the stage
    BB_TRI_MARK(7);
    if (!waited && (!resident || nstep < P)) mbar_wait(&full[rs], ruse & 1);
    const double2* col = ring + (size_t)rs * LQ;
    dm = col[LD]; dm2 = col[LD + 1];
#pragma unroll
    for (int jj = 0; jj < J; ++jj) {
      const int j = jj * 32 * W + 32 * w + lane;
      vq[jj] = j < LD ? col[j] : make_double2(0.0, CUDART_INF);
    }
    if (!resident) { __syncwarp(); if (lane == 0) mbar_arrive(&empty[rs]); }
    if (++rs == S) { rs = 0; ++ruse; }
    ++nstep;
  };

  // ---------------- z = V' beta (BridgeRegression.cpp:246) from the registers, or from the state ----------------
  if (L.init_z) {
    for (int i = 0; i < P; ++i) {
      maybe_issue();
      next_column(false);
      double pz[C];
#pragma unroll
      for (int c = 0; c < C; ++c) {
        double s = 0.0;
#pragma unroll
        for (int jj = 0; jj < J; ++jj) s = fma(vq[jj].x, cur[c][jj], s);
        pz[c] = warp_sum(s);
      }
      if (W > 1) {
        double* rb = (i & 1) ? redg1 : redg0;
        if (lane == 0) {
#pragma unroll
          for (int c = 0; c < C; ++c) rb[(w * C + c) * 2] = pz[c];
        }
        bar_sync(bar_id, bar_n);
        if (w == 0 && lane < C) { double s = 0.0; for (int ww = 0; ww < W; ++ww) s += rb[(ww * C + lane) * 2]; zg[lane * P + i] = s; }
      } else if (lane == 0) {
#pragma unroll
        for (int c = 0; c < C; ++c) zg[c * P + i] = pz[c];
      }
    }
  } else {
    for (int c = 0; c < C; ++c)
      for (int i = 32 * w + lane; i < P; i += 32 * W) zg[c * P + i] = L.z[(size_t)lch[c] * P + i];
  }
  group_sync();

  for (int it = -L.pre_tau; it < L.niter; ++it) {
    const bool pre = it < 0;  // the tau pre-draw before burn-in (BridgeWrapper.cpp:142)
    const uint32_t t = L.iter0 + (pre ? 0u : (uint32_t)it);

    // ================= tau and sig2 (scalars per chain) =================
    {
      double st[C], sr[C];
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const double alpha = al_s[g * C + c];
        double a = 0.0, r = 0.0;
#pragma unroll
        for (int jj = 0; jj < J; ++jj)
          if (jj * 32 * W + 32 * w + lane < P) a += x_exp(alpha * x_log(fabs(cur[c][jj])));   // BridgeRegression.cpp:459
        if (!pre && do_sig2)
          for (int i = 32 * w + lane; i < P; i += 32 * W) {   // ||y - X beta||^2 = r0 + sum d_i^2 (z_i - zc_i)^2
            const double dz = zg[c * P + i] - __ldg(L.m.zc + i);
            r = fma(__ldg(L.m.d2 + i) * dz, dz, r);
          }
        st[c] = warp_sum(a); sr[c] = warp_sum(r);
      }
      if (W > 1) {
        if (lane == 0) {
#pragma unroll
          for (int c = 0; c < C; ++c) { redg0[(w * C + c) * 2] = st[c]; redg0[(w * C + c) * 2 + 1] = sr[c]; }
        }
        bar_sync(bar_id, bar_n);
      }
      if (w == 0) {
#pragma unroll
        for (int c = 0; c < C; ++c) {
          double a = st[c], r = sr[c];
          if (W > 1) { a = 0.0; r = 0.0; for (int ww = 0; ww < W; ++ww) { a += redg0[(ww * C + c) * 2]; r += redg0[(ww * C + c) * 2 + 1]; } }
          const uint32_t gch = L.chain0 + (uint32_t)lch[c];
          const double alpha = al_s[g * C + c];
          double tau = tau_c[c], sig2 = sig2_c[c];
          uint32_t stt = 0;
          if (do_tau || (pre && !know_tau)) {  // BridgeRegression.cpp:453-465
            const double shape = L.h.nu_shape + ((double)P) / alpha;
            const double rate = L.h.nu_rate + a;
            Stream s(L.key, gch, t, pre ? ST_TAU_INIT : ST_TAU, 0);
            const double nu = gamma_rate<true>(s, shape, rate, stt);
            tau = exp(-1.0 * log(nu) / alpha);
          }
          if (!pre && do_sig2) {  // BridgeRegression.cpp:436-450
            const double shape = L.h.sig2_shape + 0.5 * L.m.N;
            const double scale = L.h.sig2_scale + 0.5 * (L.m.r0z + r);
            Stream s(L.key, gch, t, ST_SIG2, 0);
            sig2 = scale / gamma_rate<true>(s, shape, 1.0, stt);
          }
          if (lane == 0) { sc[(g * C + c) * 2] = tau; sc[(g * C + c) * 2 + 1] = sig2; if (stt) atomicOr(L.status + lch[c], stt); }
        }
      }
      group_sync();
#pragma unroll
      for (int c = 0; c < C; ++c) { tau_c[c] = sc[(g * C + c) * 2]; sig2_c[c] = sc[(g * C + c) * 2 + 1]; }
    }
    if (pre) continue;

    const bool store = L.sink.thin > 0 && ((it + 1) % L.sink.thin) == 0;
    const long long slot = L.sink.s0 + (it + 1) / (L.sink.thin > 0 ? L.sink.thin : 1) - 1;

    // ================= omega, u, box half-widths (elementwise) =================
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const uint32_t gch = L.chain0 + (uint32_t)lch[c];
      const double tau = tau_c[c], alpha = al_s[g * C + c];
#pragma unroll
      for (int jj = 0; jj < J; ++jj) {
        const int j = jj * 32 * W + 32 * w + lane;
        if (j < P) {
          const double ab = fabs(cur[c][jj]);
          const double uo = ug[c * P + j];
          // sample_omega, BridgeRegression.cpp:130-147
          double om; bool one = true;
          if (do_omega) {
            const double a_j = x_exp(alpha * x_log(ab / ((1 - uo) * tau)));
            const double prob = alpha / (1 + alpha * a_j);
            Stream so(L.key, gch, t, ST_OMEGA, (uint32_t)j);
            double e1, e2; so.block(1, e1, e2);
            one = so.unif(0) > prob;
            om = -x_log(one ? e1 : e1 * e2) + a_j;
          } else om = L.omega[(size_t)lch[c] * P + j];
          // sample_u, BridgeRegression.cpp:97-110
          double un = uo;
          if (do_u) {
            Stream su(L.key, gch, t, ST_U, (uint32_t)j);
            const double lom = x_log(om) / alpha;
            const double right = 1 - ab / tau * x_exp(-lom);
            un = right * su.unif(0);
            ug[c * P + j] = un;
          }
          // b_j, BridgeRegression.cpp:409-411
          bb_[c][jj] = (1.0 - un) * x_exp(x_log(om) / alpha) * tau;
          if (valid[c]) {
            if (store) {
              const size_t o = ((size_t)lch[c] * L.sink.chain_stride + slot) * P + j;
              if (L.sink.omega) L.sink.omega[o] = om;
              if (L.sink.shape) L.sink.shape[o] = do_omega ? (one ? 1.0 : 2.0) : L.shape[(size_t)lch[c] * P + j];
              if (L.sink.u) L.sink.u[o] = un;
            }
            if (it == L.niter - 1 && do_omega) {
              L.omega[(size_t)lch[c] * P + j] = om;
              L.shape[(size_t)lch[c] * P + j] = one ? 1.0 : 2.0;
            }
          }
        }
      }
    }

    // ================= beta: rotated coordinate sweeps (BridgeRegression.cpp:235-286) =================
    double sig_d = sig2_c[0];
#pragma unroll
    for (int c = 1; c < C; ++c) if (dc == c) sig_d = sig2_c[c];
    sig_d = sqrt(sig_d);
    const double isig_d = 1.0 / sig_d;
    // RNG look-ahead: the Philox block of a lane's first trial and its two logarithms do not depend on the
    // bounds.  With W > 1 the warp that draws step i+1 prepares them while the drawer of step i is busy.
    TnDraw td; td.ua = 0.5; td.ub = 0.5; td.nlu = 0.0; td.lub = 0.0; td.xn = 0.0;
    bool have_rng = false;
    bool col_waited = false;   // the next column's barrier was observed by the idle warps of this group during the last draw
    for (int sw = 0; do_beta && sw < nsweep; ++sw) {
      for (int i = 0; i < P; ++i) {
        const bool drawer = (W == 1) || (w == i % W);
#ifdef BB_TRI_PHASES
        ph_on = blockIdx.x == 0 && tid == BB_TRI_PHASE_TID;
        ph_half = drawer ? 0 : 16;
        ph_t = clock64();
        if (ph_on) ph_acc[ph_half + 6] += 1;
#endif
        Stream s(L.key, gch_d, t, ST_BETA, (uint32_t)(sw * P + i));
        if (drawer && !have_rng) { double ua, ub; s.block((uint32_t)sl, ua, ub); td.set(ua, ub); }
        have_rng = false;
        if (W == 1) maybe_issue();
        next_column(col_waited);
        col_waited = false;
        BB_TRI_MARK(0);
        double di = 0.0, mean = 0.0, sd = 1.0, isd = 1.0;
        if (drawer) { di = dm.x; mean = dm.y; sd = sig_d * dm2.x; isd = di * isig_d; }   // sd = sig/d_i (:269) by reciprocals
        // ---- candidates: min_j fma(beta_j, q, b_j|q|) and min_j fma(-beta_j, q, b_j|q|) ----
        double lmn[C], rmn[C];
#pragma unroll
        for (int c = 0; c < C; ++c) { lmn[c] = CUDART_INF; rmn[c] = CUDART_INF; }
#pragma unroll
        for (int jj = 0; jj < J; ++jj) {
          const double q = vq[jj].y, aq = fabs(q);
#pragma unroll
          for (int c = 0; c < C; ++c) {
            const double gq = bb_[c][jj] * aq;
            lmn[c] = dmin_keep(lmn[c], fma(cur[c][jj], q, gq));
            rmn[c] = dmin_keep(rmn[c], fma(-cur[c][jj], q, gq));
          }
        }
        BB_TRI_MARK(8);
        // after the transposed butterfly the first half of draw slice c holds L_c, the second half R_c
        double keep = reduce_minima<C>(lmn, rmn, lane);
        BB_TRI_MARK(1);
        if (W > 1) {
          if ((sl & (SL / 2 - 1)) == 0) redg0[(w * C + dc) * 2 + (sl >= SL / 2 ? 1 : 0)] = keep;
          bar_sync(bar_id, bar_n);
          BB_TRI_MARK(2);
          // Idle until the draw is done.  The idle warps service the column stream: thread 0 issues the TMA copies, and every
          // idle warp waits for the NEXT column's barrier here (90 cycles even when the data has long landed), so that after
          // the group barrier below nobody -- the drawer included -- has to wait for it at the top of the next step.
          if (!drawer) {
            maybe_issue();
            if (nstep < total_steps && (!resident || nstep < P)) mbar_wait(&full[rs], ruse & 1);   // rs / nstep already point at the next step
          }
          col_waited = true;
          if (!drawer && i + 1 < P && w == (i + 1) % W) {   // ... and the next drawer prepares its first trial
            Stream sn(L.key, gch_d, t, ST_BETA, (uint32_t)(sw * P + i + 1));
            double ua, ub; sn.block((uint32_t)sl, ua, ub); td.set(ua, ub);
            have_rng = true;
          }
        }
        if (drawer) {
          // this lane serves chain dc: combine the partial minima
          double lq, rq;
          if (W > 1) {
            lq = CUDART_INF; rq = CUDART_INF;
#pragma unroll
            for (int ww = 0; ww < W; ++ww) { lq = dmin_keep(lq, redg0[(ww * C + dc) * 2]); rq = dmin_keep(rq, redg0[(ww * C + dc) * 2 + 1]); }
          } else {
            const double other = shfl_xor_d(keep, SL / 2);
            lq = sl < SL / 2 ? keep : other; rq = sl < SL / 2 ? other : keep;
          }
          const double zold = zg[dc * P + i];
          const double lm = zold - lq, rm = zold + rq;
          BB_TRI_MARK(9);
          double znew;
          if (di > 1e-16) {   // BridgeRegression.cpp:268-273
            const double a = (lm - mean) * isd, b = (rm - mean) * isd;
            double x = 0.0;
            bool fin = false;
            uint32_t base = 0;
            // Fast path, branch-free: the plans with a UNIFORM proposal -- mode 3 (interval in a tail and narrow: 91 % of the steps
            // at c3) and mode 1 (interval around 0 and narrower than sqrt(2 pi): 9 %) -- classified and evaluated without a branch,
            // so that the two chains of a group never diverge over the plan (tn_setup_fast is five dependent compare-and-branch
            // steps, taken by the two slices of the warp one after the other when they disagree).  Same decisions, same
            // arithmetic and same trial order as tn_setup_fast + tn_eval_log; anything else falls through to them.
            const bool neg = b < 0.0;
            const bool tail = (a > 0.0) || neg;
            const double qa = neg ? -b : a, qb = neg ? -a : b;       // working orientation (tn_setup_fast: p.a, p.b)
            const double qw = qb - qa;
            const bool m3 = tail && (qw * (qa + 1.0) < 1.0);
            const bool m1 = !tail && !(qw >= SQRT_2PI);
            const bool quick = (a < b) && (m1 || m3);
            bool general = true;
            if (__all_sync(0xffffffffu, quick)) {
              const double xt = qa + (qb - qa) * td.ua;
              const double c0 = m3 ? qa * qa : 0.0;
              const bool acc_ = td.lub <= 0.5 * (c0 - xt * xt);
              const uint32_t all = __ballot_sync(0xffffffffu, acc_);
              const uint32_t mine = all & slmask;
              const int src = mine ? (__ffs(mine) - 1) : lane;
              const double xs = __shfl_sync(0xffffffffu, xt, src);
              if (mine) { x = (neg ? -1.0 : 1.0) * xs; fin = true; }
              bool done = true;
#pragma unroll
              for (int c = 0; c < C; ++c) done = done && ((all >> (c * SL)) & ((SL == 32) ? 0xffffffffu : ((1u << SL) - 1u))) != 0u;
              general = !done;
              if (general) {                                         // a slice without an accepting trial among its SL lanes: next block
                base = SL;
                if (!fin) { double ua, ub; s.block(base + (uint32_t)sl, ua, ub); td.set(ua, ub); }
              }
            }
            BB_TRI_MARK(10);
            if (general) {
              const TnPlan p = tn_setup_fast(a, b);
              if (p.mode == 5) { status |= ST_NAN; x = a + b; fin = true; }
              else if (p.mode == 4) { if (a > b) status |= ST_EMPTY_INTERVAL; x = a; fin = true; }
              while (true) {
                double xt = 0.0;
                const bool acc_ = !fin && tn_eval_log(p, td, xt);
                const uint32_t all = __ballot_sync(0xffffffffu, acc_ || fin);   // a finished slice reports all its lanes
                const uint32_t mine = all & slmask;
                const int src = mine ? (__ffs(mine) - 1) : lane;
                const double xs = __shfl_sync(0xffffffffu, xt, src);
                if (!fin && mine) { x = p.sgn * xs; fin = true; }
                bool done = true;                                               // every slice has an accepting (or finished) lane
#pragma unroll
                for (int c = 0; c < C; ++c) done = done && ((all >> (c * SL)) & ((SL == 32) ? 0xffffffffu : ((1u << SL) - 1u))) != 0u;
                if (done) break;
                base += SL;
                if (base >= K_CAP - 32) { if (!fin) { status |= ST_TRIAL_CAP; x = nan(""); fin = true; } break; }
                if (!fin) { double ua, ub; s.block(base + (uint32_t)sl, ua, ub); td.set(ua, ub); }
              }
            }
            BB_TRI_MARK(11);
            znew = mean + sd * x;
            if (znew < lm) znew = lm;
            if (znew > rm) znew = rm;
          } else {            // flat(lmax, rmin), BridgeRegression.cpp:280
            const double u0 = __shfl_sync(0xffffffffu, td.ua, dc * SL);
            znew = lm + (rm - lm) * u0;
          }
          if (sl == 0) { zg[dc * P + i] = znew; bc[g * C + dc] = znew - zold; }
        }
        BB_TRI_MARK(3);
        group_sync();
        BB_TRI_MARK(4);
        // ---- beta += V[:,i] (z_i' - z_i) ----
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const double dl = bc[g * C + c];
#pragma unroll
          for (int jj = 0; jj < J; ++jj) cur[c][jj] = fma(vq[jj].x, dl, cur[c][jj]);
        }
        if (W == 1) __syncwarp();   // bc is rewritten by the next step's draw
        BB_TRI_MARK(5);
      }
    }

    // ================= alpha, only when it is unknown (BridgeRegression.cpp:477-503) =================
    if (do_alpha) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        Stream sa(L.key, L.chain0 + (uint32_t)lch[c], t, ST_ALPHA, 0);
        const AlphaProposal pr = alpha_propose(sa, al_s[g * C + c]);
        double sn = 0.0, so = 0.0;
#pragma unroll
        for (int jj = 0; jj < J; ++jj)
          if (jj * 32 * W + 32 * w + lane < P) alpha_terms(cur[c][jj], tau_c[c], pr, sn, so);
        sn = warp_sum(sn); so = warp_sum(so);
        if (W > 1) {
          if (lane == 0) { redg0[(w * C + c) * 2] = sn; redg0[(w * C + c) * 2 + 1] = so; }
          bar_sync(bar_id, bar_n);
          if (w == 0) { sn = 0.0; so = 0.0; for (int ww = 0; ww < W; ++ww) { sn += redg0[(ww * C + c) * 2]; so += redg0[(ww * C + c) * 2 + 1]; } }
        }
        if (w == 0 && lane == 0) al_s[g * C + c] = alpha_decide(sa, pr, sn, so, P, L.h.alpha_a, L.h.alpha_b);
      }
      group_sync();
    }

    // ================= store =================
    if (store) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        if (!valid[c]) continue;
#pragma unroll
        for (int jj = 0; jj < J; ++jj) {
          const int j = jj * 32 * W + 32 * w + lane;
          if (j < P) {
            const size_t o = ((size_t)lch[c] * L.sink.chain_stride + slot) * P + j;
            if (L.sink.beta) L.sink.beta[o] = cur[c][jj];
            if (L.sink.sum) { L.sink.sum[(size_t)lch[c] * P + j] += cur[c][jj]; L.sink.sumsq[(size_t)lch[c] * P + j] += cur[c][jj] * cur[c][jj]; }
          }
        }
        if (w == 0 && lane == 0) {
          const size_t o = (size_t)lch[c] * L.sink.chain_stride + slot;
          if (L.sink.sig2) L.sink.sig2[o] = sig2_c[c];
          if (L.sink.tau) L.sink.tau[o] = tau_c[c];
          if (L.sink.alpha) L.sink.alpha[o] = al_s[g * C + c];
        }
      }
    }
  }

#ifdef BB_TRI_PHASES
  if (blockIdx.x == 0 && tid == BB_TRI_PHASE_TID) for (int k = 0; k < 32; ++k) tri_phase_cyc[k] += ph_acc[k];
#endif
  // ---------------- write state back ----------------
  group_sync();
#pragma unroll
  for (int c = 0; c < C; ++c) {
    if (!valid[c]) continue;
#pragma unroll
    for (int jj = 0; jj < J; ++jj) {
      const int j = jj * 32 * W + 32 * w + lane;
      if (j < P) { L.beta[(size_t)lch[c] * P + j] = cur[c][jj]; L.u[(size_t)lch[c] * P + j] = ug[c * P + j]; L.z[(size_t)lch[c] * P + j] = zg[c * P + j]; }
    }
    if (w == 0 && lane == 0) { L.tau[lch[c]] = tau_c[c]; L.sig2[lch[c]] = sig2_c[c]; L.alpha[lch[c]] = al_s[g * C + c]; }
  }
  status = __reduce_or_sync(slmask, status);
  if (sl == 0 && status && cbase + dc < L.nchains) atomicOr(L.status + lch_d, status);
}

}  // namespace bb
