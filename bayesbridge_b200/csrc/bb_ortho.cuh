// bb_ortho.cuh -- the `ortho=TRUE` drivers: one chain per CTA, state in shared memory.
//
//   tri : tau -> omega -> u -> sig2 -> sample_beta_ortho   (/root/reference/Code/C/BridgeWrapper.cpp:405-416;
//         sample_beta_ortho BridgeRegression.cpp:362-403: coordinate-wise truncated normals,
//         beta_j ~ TN((Xy_j - XX_sub[j] beta_-j)/XX_jj, sig2/XX_jj; [-b_j, b_j]))
//   stb : tau -> lambda -> sig2 -> sample_beta_stable_ortho (BridgeWrapper.cpp:513-522;
//         BridgeRegression.cpp:514-521: independent normals from diag(X'X))
//
// s = X'X beta is kept incrementally (s += XX[:,j] * delta after each coordinate), so that
// XX_sub[j] beta_-j = s_j - XX_jj beta_j (the reference forms it by a gemm at :382) and
// ||y - X beta||^2 = r0 - 2 d'g + d'(s - XX bc),  d = beta - bc.
#pragma once
#include "bb_launch.h"
#include "bb_retstable.cuh"
#include "bb_alpha.cuh"

namespace bb {


constexpr int ORT_T = 256;

// block-wide sum of up to 3 values; result valid in all threads
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c, double* red /*[3][8]*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
  __syncthreads();
  if (lane == 0) { red[warp] = a; red[8 + warp] = b; red[16 + warp] = c; }
  __syncthreads();
  a = b = c = 0.0;
  for (int w = 0; w < ORT_T / 32; ++w) { a += red[w]; b += red[8 + w]; c += red[16 + w]; }
}

template <bool STB>
__global__ void __launch_bounds__(ORT_T) ortho_kernel(const OrthoLaunch L) {
  extern __shared__ double smem[];
  const int P = L.m.P;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* beta_s = smem;          // [P]
  double* s_s = beta_s + P;       // [P]  X'X beta
  double* u_s = s_s + P;          // [P]
  double* b_s = u_s + P;          // [P]  box half-widths (tri) | lambda (stb)
  double* red = b_s + P;          // [24]
  double* sc = red + 24;          // tau, sig2, delta
  int* q_s = reinterpret_cast<int*>(sc + 4);   // [2][P]
  int* qn_s = q_s + 2 * P;                     // [2]
  uint32_t* kk_s = reinterpret_cast<uint32_t*>(qn_s + 2);   // [P]

  const int ch = blockIdx.x;
  if (ch >= L.nchains) return;
  const uint32_t gch = L.chain0 + (uint32_t)ch;
  const bool know_tau = L.h.true_tau > 0, know_sig2 = L.h.true_sig2 > 0, know_alpha = L.h.true_alpha > 0;
  const bool do_tau = (L.stages & 1) && !know_tau, do_sig2 = (L.stages & 2) && !know_sig2, do_alpha = (L.stages & 64) && !know_alpha;
  double alpha = know_alpha ? L.h.true_alpha : L.alpha[ch];   // unknown alpha: chain state, moved by the Metropolis step below
  const bool do_omega = L.stages & 4, do_u = L.stages & 8, do_beta = L.stages & 16, do_lambda = L.stages & 32;
  uint32_t status = 0;

  for (int j = tid; j < P; j += ORT_T) {
    beta_s[j] = L.beta[(size_t)ch * P + j];
    u_s[j] = STB ? 0.0 : L.u[(size_t)ch * P + j];
    b_s[j] = STB ? L.aux[(size_t)ch * P + j] : 0.0;
  }
  if (tid == 0) { sc[0] = know_tau ? L.h.true_tau : L.tau[ch]; sc[1] = know_sig2 ? L.h.true_sig2 : L.sig2[ch]; qn_s[0] = 0; qn_s[1] = 0; }
  __syncthreads();
  for (int k = tid; k < P; k += ORT_T) {   // s = X'X beta
    double s = 0.0;
    for (int j = 0; j < P; ++j) s = fma(__ldg(L.m.XX + k + (size_t)j * P), beta_s[j], s);
    s_s[k] = s;
  }
  __syncthreads();

  auto draw_sig2 = [&](uint32_t t) {   // BridgeRegression.cpp:436-450
    double pq = 0.0, pg = 0.0, z = 0.0;
    for (int k = tid; k < P; k += ORT_T) {
      const double dl = beta_s[k] - __ldg(L.m.bc + k);
      const double sc_k = __ldg(L.m.Xy + k) - __ldg(L.m.gc + k);   // (XX bc)_k
      pq = fma(dl, s_s[k] - sc_k, pq);
      pg = fma(dl, __ldg(L.m.gc + k), pg);
    }
    block_sum3(pq, pg, z, red);
    if (warp == 0) {
      const double rss = L.m.r0b - 2.0 * pg + pq;
      Stream s(L.key, gch, t, ST_SIG2, 0);
      const double g = gamma_rate<true>(s, L.h.sig2_shape + 0.5 * L.m.N, 1.0, status);
      if (lane == 0) sc[1] = (L.h.sig2_scale + 0.5 * rss) / g;
    }
    __syncthreads();
  };
  auto draw_tau = [&](uint32_t t, uint32_t stage) {   // BridgeRegression.cpp:453-465
    double pt = 0.0, z1 = 0.0, z2 = 0.0;
    for (int k = tid; k < P; k += ORT_T) pt += exp(alpha * log(fabs(beta_s[k])));
    block_sum3(pt, z1, z2, red);
    if (warp == 0) {
      Stream s(L.key, gch, t, stage, 0);
      const double nu = gamma_rate<true>(s, L.h.nu_shape + ((double)P) / alpha, L.h.nu_rate + pt, status);
      if (lane == 0) sc[0] = exp(-1.0 * log(nu) / alpha);
    }
    __syncthreads();
  };

  // pre-draws: tri-ortho draws sig2 then tau before burn-in (BridgeWrapper.cpp:379-380); stb-ortho none (:490)
  if (L.pre_sig2 && !know_sig2) draw_sig2(0xFFFFFFFFu);
  if (L.pre_tau && !know_tau) draw_tau(L.iter0, ST_TAU_INIT);

  for (int it = 0; it < L.niter; ++it) {
    const uint32_t t = L.iter0 + (uint32_t)it;
    const bool store = L.sink.thin > 0 && ((it + 1) % L.sink.thin) == 0;
    const long long slot = L.sink.s0 + (it + 1) / (L.sink.thin > 0 ? L.sink.thin : 1) - 1;
    if (do_tau) draw_tau(t, ST_TAU);
    const double tau = sc[0];

    if (!STB) {
      // omega, u, b  (BridgeRegression.cpp:130-147, 97-110, 390)
      for (int j = tid; j < P; j += ORT_T) {
        const double ab = fabs(beta_s[j]);
        double om; bool one = true;
        if (do_omega) {
          const double a_j = exp(alpha * log(ab / ((1 - u_s[j]) * tau)));
          const double prob = alpha / (1 + alpha * a_j);
          Stream so(L.key, gch, t, ST_OMEGA, (uint32_t)j);
          double e1, e2; so.block(1, e1, e2);
          one = so.unif(0) > prob;
          om = (one ? -log(e1) : -log(e1 * e2)) + a_j;
          if (it == L.niter - 1) { L.aux[(size_t)ch * P + j] = om; L.shape[(size_t)ch * P + j] = one ? 1.0 : 2.0; }
        } else om = L.aux[(size_t)ch * P + j];
        double un = u_s[j];
        if (do_u) {
          Stream su(L.key, gch, t, ST_U, (uint32_t)j);
          const double right = 1 - ab / tau * exp(-1.0 * log(om) / alpha);
          un = right * su.unif(0);
          u_s[j] = un;
        }
        b_s[j] = (1.0 - un) * exp(log(om) / alpha) * tau;
        if (store) {
          const size_t o = ((size_t)ch * L.sink.chain_stride + slot) * P + j;
          if (L.sink.omega) L.sink.omega[o] = om;
          if (L.sink.shape) L.sink.shape[o] = do_omega ? (one ? 1.0 : 2.0) : L.shape[(size_t)ch * P + j];
          if (L.sink.u) L.sink.u[o] = un;
        }
      }
      __syncthreads();
      if (do_sig2) draw_sig2(t);
      const double sig2 = sc[1];
      // sample_beta_ortho: sequential coordinate sweep (BridgeRegression.cpp:376-399)
      for (int j = 0; do_beta && j < P; ++j) {
        if (warp == 0) {
          const double xjj = __ldg(L.m.XX + j + (size_t)j * P);
          const double bj = beta_s[j];
          const double m = (__ldg(L.m.Xy + j) - (s_s[j] - xjj * bj)) / xjj;
          const double v = sig2 / xjj;
          Stream s(L.key, gch, t, ST_BETA, (uint32_t)j);
          const double bn = tnorm<true>(s, -1.0 * b_s[j], b_s[j], m, sqrt(v), status);
          if (lane == 0) { sc[2] = bn - bj; beta_s[j] = bn; }
        }
        __syncthreads();
        const double dl = sc[2];
        for (int k = tid; k < P; k += ORT_T) s_s[k] = fma(__ldg(L.m.XX + k + (size_t)j * P), dl, s_s[k]);
        __syncthreads();
      }
    } else {
      // lambda (BridgeRegression.cpp:506-510): retry queue + speculative look-ahead (bb_retstable.cuh)
      if (do_lambda) lambda_stage_cta<ORT_T>(tid, P, beta_s, b_s, tau, 0.5 * alpha, L.key, gch, t, nullptr, q_s, P, qn_s, kk_s, status);
      if (do_sig2) draw_sig2(t);
      const double sig2 = sc[1];
      if (do_beta) {
        // independent normals (BridgeRegression.cpp:514-521), then s = X'X beta afresh
        for (int j = tid; j < P; j += ORT_T) {
          const double uj = __ldg(L.m.XX + j + (size_t)j * P) + b_s[j] * sig2 / (tau * tau);
          const double sd = sqrt(sig2 / uj);
          const double mu = __ldg(L.m.Xy + j) / uj;
          Stream s(L.key, gch, t, ST_BETA_STB, (uint32_t)j);
          beta_s[j] = mu + sd * s.norm(0);
        }
        __syncthreads();
        for (int k = tid; k < P; k += ORT_T) {
          double s = 0.0;
          for (int j = 0; j < P; ++j) s = fma(__ldg(L.m.XX + k + (size_t)j * P), beta_s[j], s);
          s_s[k] = s;
        }
        __syncthreads();
      }
    }
    if (do_alpha) {   // BridgeRegression.cpp:477-503 (BridgeWrapper.cpp:390,412,498)
      if (warp == 0) {
        Stream sa(L.key, gch, t, ST_ALPHA, 0);
        const AlphaProposal pr = alpha_propose(sa, alpha);
        double sn = 0.0, so = 0.0;
        for (int j = lane; j < P; j += 32) alpha_terms(beta_s[j], tau, pr, sn, so);
        sn = warp_sum(sn); so = warp_sum(so);
        if (lane == 0) sc[3] = alpha_decide(sa, pr, sn, so, P, L.h.alpha_a, L.h.alpha_b);
      }
      __syncthreads();
      alpha = sc[3];
    }
    if (store) {
      for (int j = tid; j < P; j += ORT_T) {
        const size_t o = ((size_t)ch * L.sink.chain_stride + slot) * P + j;
        if (L.sink.beta) L.sink.beta[o] = beta_s[j];
        if (STB && L.sink.lambda) L.sink.lambda[o] = b_s[j];
      }
      if (tid == 0) {
        const size_t o = (size_t)ch * L.sink.chain_stride + slot;
        if (L.sink.sig2) L.sink.sig2[o] = sc[1];
        if (L.sink.tau) L.sink.tau[o] = sc[0];
        if (L.sink.alpha) L.sink.alpha[o] = alpha;
      }
    }
    __syncthreads();
  }
  for (int j = tid; j < P; j += ORT_T) {
    L.beta[(size_t)ch * P + j] = beta_s[j];
    if (STB) L.aux[(size_t)ch * P + j] = b_s[j]; else L.u[(size_t)ch * P + j] = u_s[j];
  }
  if (tid == 0) { L.tau[ch] = sc[0]; L.sig2[ch] = sc[1]; L.alpha[ch] = alpha; }
  status = __reduce_or_sync(0xffffffffu, status);
  if (lane == 0 && status) atomicOr(L.status + ch, status);
}

}  // namespace bb
