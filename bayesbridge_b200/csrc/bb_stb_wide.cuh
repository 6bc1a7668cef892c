// bb_stb_wide.cuh -- the Gaussian draw of the stable sampler for WIDE designs (P > N), opt-in.
//
// The reference draws  beta | lambda, sig2, tau ~ N(VInv^-1 X'y, sig2 VInv^-1),  VInv = X'X + D,  D = diag(lambda sig2/tau^2),
// through a P x P Cholesky (/root/reference/Code/C/BridgeRegression.cpp:552-575): P^3/3 flops per chain and iteration, which
// the authors themselves flag for P > N (Code/R/Benchmark.R:734-742).  The same law through an N x N system
// (Bhattacharya, Chakraborty & Mallick 2016, Biometrika 103(4)): with theta = beta / sig,
//     u ~ N(0, D^-1),  delta ~ N(0, I_N),  v = X u + delta,
//     M = X D^-1 X' + I_N,   w = M^-1 (y / sig - v),   theta = u + D^-1 X' w
// has theta ~ N(VInv^-1 X'y / sig, VInv^-1).  Cost N^2 P (the product, on the FP64 tensor cores) + N^3/3 instead of P^3/3:
// config c4 (N = 500, P = 2000) 5.4e8 against 2.7e9 flops, and a 2 MB instead of a 32 MB workspace per chain.
// Not the reference's arithmetic: the draws differ from sample_beta_stable's for the same normals (the LAW is the same), so
// the path is opt-in (bb_chains_option(BB_OPT_WIDE_GAUSSIAN)) and is checked distributionally and against a host
// restatement of these formulas (tests/test_gpu_wide.py), never draw for draw against the reference.
//
// Normals: item j < P of stage ST_BETA_STB is the one behind u_j, item P + n the one behind delta_n.
#pragma once
#include "bb_stb_large.cuh"

namespace bb {

constexpr int GR_T = 64;    // tile edge of the product kernel (the tiling of gram_kernel, bb_pre.cuh)
constexpr int GR_K = 32;    // k chunk
constexpr int GR_LD = 36;   // smem leading dimension (doubles)

// ---- u = D^-1/2 eps, r = y / sig - X u - delta : one CTA per chain ----
__global__ void __launch_bounds__(256) wide_draw_kernel(const StbLargeLaunch L) {
  extern __shared__ double smem[];          // [P] u
  const int P = L.m.P, N = L.m.N, tid = threadIdx.x, ch = blockIdx.x;
  if (ch >= L.nchains) return;
  const uint32_t gch = L.chain0 + (uint32_t)ch, it_ = stbl_iter(L);
  const double sig2 = L.sig2[ch], tau = L.tau[ch];
  const double f = tau * tau / sig2;        // D_j^-1 = f / lambda_j
  const double isig = 1.0 / sqrt(sig2);
  const double* lam = L.lambda + (size_t)ch * P;
  double* U = L.U + (size_t)ch * P;
  for (int j = tid; j < P; j += 256) {
    Stream s(L.key, gch, it_, ST_BETA_STB, (uint32_t)j);
    const double u = s.norm(0) * sqrt(f / lam[j]);
    smem[j] = u; U[j] = u;
  }
  __syncthreads();
  double* R = L.R + (size_t)ch * N;
  for (int n = tid; n < N; n += 256) {      // rows of X are contiguous along n: coalesced
    const double* x = L.m.X + n;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int k = 0;
    for (; k + 3 < P; k += 4) {
      s0 = fma(__ldg(x + (size_t)k * N), smem[k], s0); s1 = fma(__ldg(x + (size_t)(k + 1) * N), smem[k + 1], s1);
      s2 = fma(__ldg(x + (size_t)(k + 2) * N), smem[k + 2], s2); s3 = fma(__ldg(x + (size_t)(k + 3) * N), smem[k + 3], s3);
    }
    for (; k < P; ++k) s0 = fma(__ldg(x + (size_t)k * N), smem[k], s0);
    Stream s(L.key, gch, it_, ST_BETA_STB, (uint32_t)(P + n));
    R[n] = __ldg(L.m.y + n) * isig - ((s0 + s1) + (s2 + s3)) - s.norm(0);
  }
}

// ---- M_c = X D_c^-1 X' + I  (N x N, lower 64 x 64 tiles, diagonal tiles in full) on the FP64 tensor cores ----
// CTA = one tile (bi >= bj) of one chain, FOUR warps; warp w owns rows 16w..16w+15 of the tile (two A fragments) and all 8
// n-tiles, so that one B fragment load feeds two DMMAs: 10 shared-memory loads per 16 DMMAs instead of 9 per 8 (the 8-warp
// mapping of gram_kernel ran the shared-memory LSU at 54 % with the DMMA pipe at 50 %, profiles/r2_c4w_wide_gram.md).
// The contraction runs over the P columns of X: a chunk of 32 columns of the 64 rows of both operands is staged in shared
// memory, the left one scaled by D^-1.
constexpr int WG_T = 128;
__global__ void __launch_bounds__(WG_T) wide_gram_kernel(const StbLargeLaunch L) {
  __shared__ double As[GR_T * GR_LD];
  __shared__ double Bs[GR_T * GR_LD];
  __shared__ double dv[GR_K];
  const int P = L.m.P, N = L.m.N, ch = blockIdx.y;
  int tIdx = blockIdx.x, bi = 0;
  while ((bi + 1) * (bi + 2) / 2 <= tIdx) ++bi;
  const int bj = tIdx - bi * (bi + 1) / 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double f = L.tau[ch] * L.tau[ch] / L.sig2[ch];
  const double* lam = L.lambda + (size_t)ch * P;
  const double* X = L.m.X;
  double acc[2][8][2];
#pragma unroll
  for (int h = 0; h < 2; ++h)
#pragma unroll
    for (int t = 0; t < 8; ++t) { acc[h][t][0] = 0.0; acc[h][t][1] = 0.0; }
  for (int k0 = 0; k0 < P; k0 += GR_K) {
    if (tid < GR_K) dv[tid] = (k0 + tid < P) ? f / lam[k0 + tid] : 0.0;
    __syncthreads();
    for (int e = tid; e < GR_T * GR_K; e += WG_T) {      // lanes run along the rows: X is column-major
      const int r = e & (GR_T - 1), kk = e >> 6;
      const int k = k0 + kk, ra = bi * GR_T + r, rb = bj * GR_T + r;
      double va = 0.0, vb = 0.0;
      if (k < P) {
        if (ra < N) va = X[(size_t)k * N + ra] * dv[kk];
        if (rb < N) vb = X[(size_t)k * N + rb];
      }
      As[r * GR_LD + kk] = va;
      Bs[r * GR_LD + kk] = vb;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < GR_K; kk += 4) {
      const double a0 = As[(warp * 16 + (lane >> 2)) * GR_LD + kk + (lane & 3)];
      const double a1 = As[(warp * 16 + 8 + (lane >> 2)) * GR_LD + kk + (lane & 3)];
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const double b = Bs[(t * 8 + (lane >> 2)) * GR_LD + kk + (lane & 3)];
        dmma_m8n8k4(acc[0][t][0], acc[0][t][1], a0, b);
        dmma_m8n8k4(acc[1][t][0], acc[1][t][1], a1, b);
      }
    }
    __syncthreads();
  }
  // C fragment: row = lane/4, cols = 2*(lane%4) + {0,1};  M is stored column-major with leading dimension N
  double* M = L.A + (size_t)ch * N * N;
#pragma unroll
  for (int h = 0; h < 2; ++h)
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int r = bi * GR_T + warp * 16 + 8 * h + (lane >> 2);
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int c = bj * GR_T + t * 8 + 2 * (lane & 3) + q;
        if (r < N && c < N) M[r + (size_t)c * N] = acc[h][t][q] + (r == c ? 1.0 : 0.0);
      }
    }
}

// ---- the same product through a 3-stage cp.async ring (round 2, default): 256 threads, 16 x 32 warp tiles, the operands are
// chunks of 32 columns of X for the 64 rows of both tiles, copied k-major as they lie in memory (rows contiguous) with 16-byte
// LDGSTS and no register staging; D^-1 is applied to the right-hand fragment on its way into the DMMA instead of to the
// staged left operand.  One barrier per chunk.  (wide_gram_kernel above: LDG -> scale -> STS staging, two barriers per chunk,
// FP64 tensor pipe 50 % active.)
inline size_t wide_gram2_smem() { return (size_t)(CB_NSTAGE * CB_STAGE + CB_NSTAGE * CB_KC) * sizeof(double); }
__global__ void __launch_bounds__(256, 2) wide_gram2_kernel(const StbLargeLaunch L) {
  extern __shared__ double smem[];
  double* dvs = smem + CB_NSTAGE * CB_STAGE;           // [CB_NSTAGE][CB_KC] D^-1 of the chunk's columns
  const int P = L.m.P, N = L.m.N, ch = blockIdx.y;
  int tIdx = blockIdx.x, bi = 0;
  while ((bi + 1) * (bi + 2) / 2 <= tIdx) ++bi;
  const int bj = tIdx - bi * (bi + 1) / 2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double f = L.tau[ch] * L.tau[ch] / L.sig2[ch];
  const double* lam = L.lambda + (size_t)ch * P;
  const double* X = L.m.X;
  const int i0 = bi * SL_NB, j0 = bj * SL_NB;
  double acc[2][4][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) { acc[a][t][0] = 0.0; acc[a][t][1] = 0.0; }
  const int nchunk = (P + CB_KC - 1) / CB_KC;
  const bool v16 = (N & 1) == 0;
  auto issue = [&](int c) {
    if (c < nchunk) {
      double* st = smem + (c % CB_NSTAGE) * CB_STAGE;
      const int col = c * CB_KC, kmax = min(CB_KC, P - col);
      if (v16) { cb_copy_chunk<true>(st, X + i0 + (size_t)col * N, N, i0, tid, kmax); cb_copy_chunk<true>(st + CB_KC * SL_LD, X + j0 + (size_t)col * N, N, j0, tid, kmax); }
      else { cb_copy_chunk<false>(st, X + i0 + (size_t)col * N, N, i0, tid, kmax); cb_copy_chunk<false>(st + CB_KC * SL_LD, X + j0 + (size_t)col * N, N, j0, tid, kmax); }
      if (tid < CB_KC) dvs[(c % CB_NSTAGE) * CB_KC + tid] = tid < kmax ? f / lam[col + tid] : 0.0;
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  issue(0); issue(1);
  const int r0 = (warp & 3) * 16 + (lane >> 2), c0 = (warp >> 2) * 32 + (lane >> 2);
  for (int c = 0; c < nchunk; ++c) {
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncthreads();
    issue(c + 2);
    const double* As = smem + (c % CB_NSTAGE) * CB_STAGE;
    const double* Bs = As + CB_KC * SL_LD;
    const double* dv = dvs + (c % CB_NSTAGE) * CB_KC;
#pragma unroll 4
    for (int kk = 0; kk < CB_KC; kk += 4) {
      const double* ak = As + (kk + (lane & 3)) * SL_LD;
      const double* bk = Bs + (kk + (lane & 3)) * SL_LD;
      const double d = dv[kk + (lane & 3)];
      const double a0 = ak[r0], a1 = ak[r0 + 8];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const double b = bk[c0 + 8 * t] * d;
        dmma_m8n8k4(acc[0][t][0], acc[0][t][1], a0, b);
        dmma_m8n8k4(acc[1][t][0], acc[1][t][1], a1, b);
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  double* M = L.A + (size_t)ch * N * N;
  const int wr = (warp & 3) * 16 + (lane >> 2), wc = (warp >> 2) * 32 + 2 * (lane & 3);
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int r = i0 + wr + 8 * a;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int cc = j0 + wc + 8 * t + q;
        if (r < N && cc < N) M[r + (size_t)cc * N] = acc[a][t][q] + (r == cc ? 1.0 : 0.0);
      }
    }
}

// ---- w = M^-1 r (both substitutions with the factor of M), beta = sig (u + D^-1 X' w) : one CTA per chain ----
template <int NT> __global__ void __launch_bounds__(NT, 1024 / NT) wide_solve_kernel(const StbLargeLaunch L) {
  extern __shared__ double smem[];
  const int P = L.m.P, N = L.m.N, tid = threadIdx.x, ch = blockIdx.x;
  if (ch >= L.nchains) return;
  double* x = smem;                        // [N], then D [64][65] and ps [16][64] of the substitutions
  const double* A = L.A + (size_t)ch * N * N;
  const int nblk = (N + SL_NB - 1) / SL_NB;
  const double* Wg = L.W + (size_t)ch * nblk * SL_NB * SL_NB;
  const int lane = tid & 31, warp = tid >> 5;
  const double* R = L.R + (size_t)ch * N;
  for (int n = tid; n < N; n += NT) x[n] = R[n];
  stbl_forward<NT>(A, Wg, N, 0, N, N + SL_NB * (SL_NB + 1), tid);
  stbl_backward<NT>(A, Wg, N, 0, N, N + SL_NB * (SL_NB + 1), tid);
  __syncthreads();
  const double sig2 = L.sig2[ch], tau = L.tau[ch];
  const double f = tau * tau / sig2, sig = sqrt(sig2);
  const double* lam = L.lambda + (size_t)ch * P;
  const double* U = L.U + (size_t)ch * P;
  double* beta = L.beta + (size_t)ch * P;
  const long long slot = stbl_slot(L);
  for (int k = warp; k < P; k += NT / 32) {        // (X' w)_k: column k of X is contiguous, lanes along the rows
    const double* col = L.m.X + (size_t)k * N;
    double s = 0.0;
    for (int n = lane; n < N; n += 32) s = fma(__ldg(col + n), x[n], s);
    s = warp_sum(s);
    if (lane == 0) {
      const double b = sig * fma(f / lam[k], s, U[k]);
      beta[k] = b;
      if (slot >= 0) {
        const size_t o = ((size_t)ch * L.sink.chain_stride + slot) * P + k;
        if (L.sink.beta) L.sink.beta[o] = b;
        if (L.sink.lambda) L.sink.lambda[o] = lam[k];
      }
    }
  }
  if (tid == 0 && slot >= 0) {
    const size_t o = (size_t)ch * L.sink.chain_stride + slot;
    if (L.sink.sig2) L.sink.sig2[o] = L.sig2[ch];
    if (L.sink.tau) L.sink.tau[o] = L.tau[ch];
    if (L.sink.alpha) L.sink.alpha[o] = L.alpha[ch];
  }
}

}  // namespace bb
