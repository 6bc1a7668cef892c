// bb_tri.cuh -- the triangle-mixture Gibbs sweep as one persistent kernel.
//
// One launch advances `nchains` independent chains by `niter` iterations of
//   tau -> sig2 -> omega -> u -> beta            (/root/reference/Code/C/BridgeWrapper.cpp:168-189)
// with   sample_tau_marg  BridgeRegression.cpp:453-465      sample_sig2  :436-450
//        sample_omega     :130-147     sample_u  :97-110     sample_beta  :405-433
//        rtnorm_gibbs (SVD form, the innermost loop)  :235-286.
//
// Layout of one CTA: C chains, W "compute" warps that own the coordinates j (J per lane, all C
// chains in registers) and C "draw" warps, one per chain.  Step i of the rotated sweep is
//   compute warps: bounds over j for chain c -> warp max/min -> partials in smem -> bar.arrive PART(c)
//   draw warp c  : (Philox block for trial=lane already computed) bar.sync PART(c) -> final max/min ->
//                  truncated-normal draw, 32 trials in parallel -> delta in smem -> bar.arrive DELTA(c)
//   compute warps: bar.sync DELTA(c) -> beta_j += V(j,i) delta  (and the fresh accumulation of V z)
// The C chains are staggered through the compute warps, so while draw warp c works the compute
// warps are busy with chains c+1.. ; column i of V (and of 1/|V|) is loaded once per step and
// shared by the C chains in registers; the next column is prefetched.
//
// Differences from the reference that are algebraic identities (documented in DESIGN.md):
//   * beta_j at the current z is kept incrementally instead of the O(P) dot at :258 (the author's
//     own note, Notes/bbnotes.tex:603-612); beta = V z is re-accumulated every sweep like :285;
//   * divisions by |v_ji| (:261-262) are multiplications by a precomputed 1/|v_ji|;
//   * ||y - X beta||^2 (:440-443) is r0 + sum_i d_i^2 (z_i - zc_i)^2, which never touches X.
#pragma once
#include <cfloat>
#include <math_constants.h>
#include "bb_launch.h"

namespace bb {


template <int J, int C, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) tri_sweep_kernel(const TriLaunch L) {
  extern __shared__ double smem[];
  const int P = L.m.P, ldv = L.m.ldv;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nct = blockDim.x - 32 * C;  // compute threads
  const int W = nct >> 5;
  const bool is_draw = tid >= nct;
  const int dc = is_draw ? (tid - nct) >> 5 : 0;
  const int nbar = nct + 32;

  double* z_s = smem;                 // [C][P]
  double* scr_s = z_s + C * P;        // [C][P] scratch (beta for the z initialisation)
  double* red_l = scr_s + C * P;      // [C][32]
  double* red_r = red_l + C * 32;     // [C][32]
  double* bc_s = red_r + C * 32;      // [C][4]: delta, znew | tau, sig2

  const int cbase = blockIdx.x * C;
  int lch[C]; bool valid[C];
#pragma unroll
  for (int c = 0; c < C; ++c) { valid[c] = cbase + c < L.nchains; lch[c] = valid[c] ? cbase + c : cbase; }

  const double alpha = L.h.true_alpha;  // fixed alpha (unknown alpha is handled by the host driver per iteration)
  const bool know_tau = L.h.true_tau > 0, know_sig2 = L.h.true_sig2 > 0;
  const bool do_tau = (L.stages & 1) && !know_tau, do_sig2 = (L.stages & 2) && !know_sig2;
  const bool do_omega = L.stages & 4, do_u = L.stages & 8, do_beta = L.stages & 16;

  // ---------------- load state ----------------
  double cur[C][J], uu[C][J], bb_[C][J], acc[C][J];
#pragma unroll
  for (int c = 0; c < C; ++c)
#pragma unroll
    for (int jj = 0; jj < J; ++jj) {
      const int j = tid + jj * nct;
      const bool in = !is_draw && j < P;
      cur[c][jj] = in ? L.beta[(size_t)lch[c] * P + j] : 0.0;
      uu[c][jj] = in ? L.u[(size_t)lch[c] * P + j] : 0.5;
      bb_[c][jj] = 1.0; acc[c][jj] = 0.0;
    }
  if (L.init_z) {
    for (int c = 0; c < C; ++c)
      for (int j = tid; j < P; j += blockDim.x) scr_s[c * P + j] = L.beta[(size_t)lch[c] * P + j];
    __syncthreads();
    for (int i = tid; i < P; i += blockDim.x) {
      double s[C];
#pragma unroll
      for (int c = 0; c < C; ++c) s[c] = 0.0;
      for (int j = 0; j < P; ++j) {
        const double t = __ldg(L.m.tV + i + (size_t)j * P);
#pragma unroll
        for (int c = 0; c < C; ++c) s[c] = fma(t, scr_s[c * P + j], s[c]);
      }
#pragma unroll
      for (int c = 0; c < C; ++c) z_s[c * P + i] = s[c];
    }
  } else {
    for (int c = 0; c < C; ++c)
      for (int i = tid; i < P; i += blockDim.x) z_s[c * P + i] = L.z[(size_t)lch[c] * P + i];
  }
  double tau_c[C], sig2_c[C];
#pragma unroll
  for (int c = 0; c < C; ++c) {
    tau_c[c] = know_tau ? L.h.true_tau : L.tau[lch[c]];
    sig2_c[c] = know_sig2 ? L.h.true_sig2 : L.sig2[lch[c]];
  }
  uint32_t status = 0;
  __syncthreads();

  const int nsweep = (L.h.betaburn > 0 ? L.h.betaburn : 0) + 1;

  for (int it = -L.pre_tau; it < L.niter; ++it) {
    const bool pre = it < 0;  // the tau pre-draw before burn-in
    const uint32_t t = L.iter0 + (pre ? 0u : (uint32_t)it);

    // ================= tau and sig2 (scalars per chain) =================
    if (!is_draw) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        double st = 0.0, sr = 0.0;
#pragma unroll
        for (int jj = 0; jj < J; ++jj)
          if (tid + jj * nct < P) st += exp(alpha * log(fabs(cur[c][jj])));   // BridgeRegression.cpp:459
        if (!pre && do_sig2)
          for (int i = tid; i < P; i += nct) {
            const double dz = z_s[c * P + i] - __ldg(L.m.zc + i);
            sr = fma(__ldg(L.m.d2 + i) * dz, dz, sr);
          }
        st = warp_sum(st); sr = warp_sum(sr);
        if (lane == 0) { red_l[c * 32 + warp] = st; red_r[c * 32 + warp] = sr; }
        __threadfence_block();
        bar_arrive(1 + 2 * c, nbar);
      }
#pragma unroll
      for (int c = 0; c < C; ++c) {
        bar_sync(2 + 2 * c, nbar);
        tau_c[c] = bc_s[c * 4 + 2]; sig2_c[c] = bc_s[c * 4 + 3];
      }
    } else {
      const int c = dc;
      const uint32_t gch = L.chain0 + (uint32_t)lch[c];
      bar_sync(1 + 2 * c, nbar);
      double st = lane < W ? red_l[c * 32 + lane] : 0.0, sr = lane < W ? red_r[c * 32 + lane] : 0.0;
      st = warp_sum(st); sr = warp_sum(sr);
      double tau = tau_c[c], sig2 = sig2_c[c];
      if (do_tau || (pre && !know_tau)) {  // BridgeRegression.cpp:453-465
        const double shape = L.h.nu_shape + ((double)P) / alpha;
        const double rate = L.h.nu_rate + st;
        Stream s(L.key, gch, t, pre ? ST_TAU_INIT : ST_TAU, 0);
        const double nu = gamma_rate<true>(s, shape, rate, status);
        tau = exp(-1.0 * log(nu) / alpha);
      }
      if (!pre && do_sig2) {  // BridgeRegression.cpp:436-450
        const double shape = L.h.sig2_shape + 0.5 * L.m.N;
        const double scale = L.h.sig2_scale + 0.5 * (L.m.r0z + sr);
        Stream s(L.key, gch, t, ST_SIG2, 0);
        sig2 = scale / gamma_rate<true>(s, shape, 1.0, status);
      }
      tau_c[c] = tau; sig2_c[c] = sig2;
      if (lane == 0) { bc_s[c * 4 + 2] = tau; bc_s[c * 4 + 3] = sig2; }
      __threadfence_block();
      bar_arrive(2 + 2 * c, nbar);
    }
    if (pre) continue;

    const bool store = L.sink.thin > 0 && ((it + 1) % L.sink.thin) == 0;
    const long long slot = L.sink.s0 + (it + 1) / (L.sink.thin > 0 ? L.sink.thin : 1) - 1;

    // ================= omega, u, box half-widths (elementwise) =================
    if (!is_draw) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const uint32_t gch = L.chain0 + (uint32_t)lch[c];
        const double tau = tau_c[c];
#pragma unroll
        for (int jj = 0; jj < J; ++jj) {
          const int j = tid + jj * nct;
          if (j < P) {
            const double ab = fabs(cur[c][jj]);
            // sample_omega, BridgeRegression.cpp:130-147
            double om; bool one = true;
            if (do_omega) {
              const double a_j = exp(alpha * log(ab / ((1 - uu[c][jj]) * tau)));
              const double prob = alpha / (1 + alpha * a_j);
              Stream so(L.key, gch, t, ST_OMEGA, (uint32_t)j);
              double e1, e2; so.block(1, e1, e2);
              one = so.unif(0) > prob;
              om = (one ? -log(e1) : -log(e1 * e2)) + a_j;
            } else om = L.omega[(size_t)lch[c] * P + j];
            // sample_u, BridgeRegression.cpp:97-110
            double un = uu[c][jj];
            if (do_u) {
              Stream su(L.key, gch, t, ST_U, (uint32_t)j);
              const double right = 1 - ab / tau * exp(-1.0 * log(om) / alpha);
              un = right * su.unif(0);
              uu[c][jj] = un;
            }
            // b_j, BridgeRegression.cpp:409-411
            bb_[c][jj] = (1.0 - un) * exp(log(om) / alpha) * tau;
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
    }

    // ================= beta: rotated coordinate sweeps =================
    for (int sw = 0; do_beta && sw < nsweep; ++sw) {
      if (!is_draw) {
        double v[J], w[J], vn[J], wn[J], vp[J];
#pragma unroll
        for (int jj = 0; jj < J; ++jj) {
          const bool in = tid + jj * nct < ldv;   // coordinates beyond the padded row: v = 0, 1/|v| = inf
          v[jj] = in ? __ldg(L.m.Vc + tid + jj * nct) : 0.0; w[jj] = in ? __ldg(L.m.Wc + tid + jj * nct) : CUDART_INF;
          vp[jj] = 0.0; vn[jj] = 0.0; wn[jj] = 0.0;
        }
#pragma unroll
        for (int c = 0; c < C; ++c)
#pragma unroll
          for (int jj = 0; jj < J; ++jj) acc[c][jj] = 0.0;
        for (int i = 0; i < P; ++i) {
          if (i + 1 < P) {
#pragma unroll
            for (int jj = 0; jj < J; ++jj) {
              const bool in = tid + jj * nct < ldv;
              vn[jj] = in ? __ldg(L.m.Vc + (size_t)(i + 1) * ldv + tid + jj * nct) : 0.0;
              wn[jj] = in ? __ldg(L.m.Wc + (size_t)(i + 1) * ldv + tid + jj * nct) : CUDART_INF;
            }
          }
#pragma unroll
          for (int c = 0; c < C; ++c) {
            if (i > 0) {
              bar_sync(2 + 2 * c, nbar);
              const double dl = bc_s[c * 4 + 0], zn = bc_s[c * 4 + 1];
#pragma unroll
              for (int jj = 0; jj < J; ++jj) { cur[c][jj] = fma(vp[jj], dl, cur[c][jj]); acc[c][jj] = fma(vp[jj], zn, acc[c][jj]); }
            }
            const double zi = z_s[c * P + i];
            double lm = -DBL_MAX, rm = DBL_MAX;
#pragma unroll
            for (int jj = 0; jj < J; ++jj) {   // BridgeRegression.cpp:254-265
              const double r = fma(-v[jj], zi, cur[c][jj]);
              const double dif = bb_[c][jj] - r, sum = bb_[c][jj] + r;
              const bool pos = v[jj] > 0;
              const double left = (pos ? -sum : -dif) * w[jj];
              const double right = (pos ? dif : sum) * w[jj];
              lm = fmax(lm, left); rm = fmin(rm, right);
            }
            lm = warp_max(lm); rm = warp_min(rm);
            if (lane == 0) { red_l[c * 32 + warp] = lm; red_r[c * 32 + warp] = rm; }
            __threadfence_block();
            bar_arrive(1 + 2 * c, nbar);
          }
#pragma unroll
          for (int jj = 0; jj < J; ++jj) { vp[jj] = v[jj]; v[jj] = vn[jj]; w[jj] = wn[jj]; }
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
          bar_sync(2 + 2 * c, nbar);
          const double zn = bc_s[c * 4 + 1];
#pragma unroll
          for (int jj = 0; jj < J; ++jj) cur[c][jj] = fma(vp[jj], zn, acc[c][jj]);   // beta = V z  (:285)
        }
      } else {
        const int c = dc;
        const uint32_t gch = L.chain0 + (uint32_t)lch[c];
        const double sig = sqrt(sig2_c[c]);
        for (int i = 0; i < P; ++i) {
          Stream s(L.key, gch, t, ST_BETA, (uint32_t)(sw * P + i));
          double ua, ub; s.block((uint32_t)lane, ua, ub);
          const double di = __ldg(L.m.d + i), mean = __ldg(L.m.mean + i);
          bar_sync(1 + 2 * c, nbar);
          double lm = lane < W ? red_l[c * 32 + lane] : -DBL_MAX, rm = lane < W ? red_r[c * 32 + lane] : DBL_MAX;
          lm = warp_max(lm); rm = warp_min(rm);
          const double zold = z_s[c * P + i];
          double znew;
          if (di > 1e-16) {   // BridgeRegression.cpp:268-273
            const double sd = sig / di;
            const double a = (lm - mean) / sd, b = (rm - mean) / sd;
            const TnPlan p = tn_setup(a, b);
            double x;
            if (p.mode == 5) { status |= ST_NAN; x = a + b; }
            else if (p.mode == 4) { if (a > b) status |= ST_EMPTY_INTERVAL; x = a; }
            else {
              bool acc_ = tn_eval(p, ua, ub, x);
              uint32_t mk = __ballot_sync(0xffffffffu, acc_);
              uint32_t base = 32;
              while (!mk && base < K_CAP - 32) {
                s.block(base + lane, ua, ub);
                acc_ = tn_eval(p, ua, ub, x);
                mk = __ballot_sync(0xffffffffu, acc_);
                base += 32;
              }
              if (mk) x = p.sgn * __shfl_sync(0xffffffffu, x, __ffs(mk) - 1);
              else { status |= ST_TRIAL_CAP; x = nan(""); }
            }
            znew = mean + sd * x;
            if (znew < lm) znew = lm;
            if (znew > rm) znew = rm;
          } else {            // flat(lmax, rmin), BridgeRegression.cpp:280
            const double u0 = __shfl_sync(0xffffffffu, ua, 0);
            znew = lm + (rm - lm) * u0;
          }
          if (lane == 0) { z_s[c * P + i] = znew; bc_s[c * 4 + 0] = znew - zold; bc_s[c * 4 + 1] = znew; }
          __threadfence_block();
          bar_arrive(2 + 2 * c, nbar);
        }
      }
    }

    // ================= store =================
    if (!is_draw) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        if (!valid[c]) continue;
#pragma unroll
        for (int jj = 0; jj < J; ++jj) {
          const int j = tid + jj * nct;
          if (j < P) {
            if (store) {
              const size_t o = ((size_t)lch[c] * L.sink.chain_stride + slot) * P + j;
              if (L.sink.beta) L.sink.beta[o] = cur[c][jj];
              if (L.sink.sum) { L.sink.sum[(size_t)lch[c] * P + j] += cur[c][jj]; L.sink.sumsq[(size_t)lch[c] * P + j] += cur[c][jj] * cur[c][jj]; }
            }
          }
        }
      }
    } else if (lane == 0 && valid[dc] && store) {
      const size_t o = (size_t)lch[dc] * L.sink.chain_stride + slot;
      if (L.sink.sig2) L.sink.sig2[o] = sig2_c[dc];
      if (L.sink.tau) L.sink.tau[o] = tau_c[dc];
      if (L.sink.alpha) L.sink.alpha[o] = alpha;
    }
  }

  // ---------------- write state back ----------------
  __syncthreads();
  if (!is_draw) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      if (!valid[c]) continue;
#pragma unroll
      for (int jj = 0; jj < J; ++jj) {
        const int j = tid + jj * nct;
        if (j < P) { L.beta[(size_t)lch[c] * P + j] = cur[c][jj]; L.u[(size_t)lch[c] * P + j] = uu[c][jj]; }
      }
    }
  } else if (valid[dc]) {
    if (lane == 0) {
      L.tau[lch[dc]] = tau_c[dc]; L.sig2[lch[dc]] = sig2_c[dc]; L.alpha[lch[dc]] = alpha;
      if (status) atomicOr(L.status + lch[dc], status);
    }
  }
  for (int c = 0; c < C; ++c)
    if (valid[c])
      for (int i = tid; i < P; i += blockDim.x) L.z[(size_t)lch[c] * P + i] = z_s[c * P + i];
}

}  // namespace bb
