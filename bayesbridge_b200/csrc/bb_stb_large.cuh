// bb_stb_large.cuh -- the stable-mixture sweep when one chain's P x P system does not fit on chip
// (P > 128): per-chain workspace in HBM, blocked right-looking Cholesky batched over chains with
// FP64 tensor-core (DMMA) trailing updates.
//
//   stbl_scalar_kernel : tau, sig2, lambda            (BridgeRegression.cpp:453-465, 436-450, 506-510)
//   stbl_build_kernel  : A_c = X'X + diag(lambda sig2/tau^2)          (:554-556)
//   stbl_update_kernel / stbl_potrf_kernel / stbl_trsm_kernel : chol(A_c), left-looking   (:559-560)
//   stbl_solve_kernel  : beta = L^-T (L^-1 Xy + sig eps)              (:563-574)
#pragma once
#include "bb_launch.h"
#include "bb_retstable.cuh"
#include "bb_alpha.cuh"

namespace bb {


constexpr int SL_NB = 64;
constexpr int SL_T = 256;
constexpr int SL_LD = 68;    // smem row stride of a staged 64 x 64 tile (k-major; 68 = 4 mod 16: the DMMA fragment loads of a half-warp hit 16 distinct 8-byte banks)

// acc(64 x 64, 8 warps x 8 rows) += A(64 x 64) B(64 x 64)' with As[k][r] = A(r,k), Bs[k][c] = B(c,k) staged in smem
__device__ __forceinline__ void tile_mma(const double* As, const double* Bs, double (&acc)[8][2], int lane, int warp) {
#pragma unroll 4
  for (int kk = 0; kk < SL_NB; kk += 4) {
    const double a = As[(kk + (lane & 3)) * SL_LD + warp * 8 + (lane >> 2)];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const double b = Bs[(kk + (lane & 3)) * SL_LD + t * 8 + (lane >> 2)];
      dmma_m8n8k4(acc[t][0], acc[t][1], a, b);
    }
  }
}

// The same product with a 16 x 32 warp tile (warp = (row group w & 3, column group w >> 2)): 2 A and 4 B fragments feed 8 DMMAs,
// 6 shared-memory loads per 8 tensor instructions instead of 9.  acc[a][t]: rows 16 (w & 3) + 8 a, columns 32 (w >> 2) + 8 t.
__device__ __forceinline__ void tile_mma_16x32(const double* As, const double* Bs, double (&acc)[2][4][2], int lane, int warp) {
  const int r0 = (warp & 3) * 16 + (lane >> 2), c0 = (warp >> 2) * 32 + (lane >> 2);
#pragma unroll 4
  for (int kk = 0; kk < SL_NB; kk += 4) {
    const double* ak = As + (kk + (lane & 3)) * SL_LD;
    const double* bk = Bs + (kk + (lane & 3)) * SL_LD;
    const double a0 = ak[r0], a1 = ak[r0 + 8];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const double b = bk[c0 + 8 * t];
      dmma_m8n8k4(acc[0][t][0], acc[0][t][1], a0, b);
      dmma_m8n8k4(acc[1][t][0], acc[1][t][1], a1, b);
    }
  }
}

// ---- d'X'Xd for the residual sum of squares (BridgeRegression.cpp:440-443 by the expansion around bc), all chains at once ----
// CTA = a strip of 32 rows of X'X x a group of 32 chains: the strip is read ONCE for the 32 chains (one CTA per chain
// walking all of X'X made 16000 dependent load+FMA steps per thread: 1.2 ms of the 1.24 ms scalar kernel at c4).
// Lane = row of the strip, warp = four chains; d is staged in chunks of 64 columns.  Q[chain][strip] = sum over the strip's
// rows of d_k (X'X d)_k: fixed-order partial sums, added up by stbl_scalar_kernel (no floating-point atomics).
constexpr int QD_R = 32, QD_J = 64;
__global__ void __launch_bounds__(256) stbl_quad_kernel(const StbLargeLaunch L, int nstrips) {
  __shared__ double ds[QD_J][33];
  const int P = L.m.P, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int strip = blockIdx.x, c0 = blockIdx.y * 32;
  const int k = strip * QD_R + lane;
  const bool kv = k < P;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  const double* xrow = L.m.XX + (kv ? k : 0);
  for (int j0 = 0; j0 < P; j0 += QD_J) {
    __syncthreads();
    for (int e = tid; e < QD_J * 32; e += 256) {
      const int jj = e & (QD_J - 1), c = e >> 6, j = j0 + jj;
      ds[jj][c] = (j < P && c0 + c < L.nchains) ? L.beta[(size_t)(c0 + c) * P + j] - __ldg(L.m.bc + j) : 0.0;
    }
    __syncthreads();
    const int nj = min(QD_J, P - j0);
#pragma unroll 8
    for (int jj = 0; jj < nj; ++jj) {
      const double x = kv ? __ldg(xrow + (size_t)(j0 + jj) * P) : 0.0;
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[q] = fma(x, ds[jj][4 * warp + q], acc[q]);
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int c = c0 + 4 * warp + q;
    double v = 0.0;
    if (kv && c < L.nchains) v = (L.beta[(size_t)c * P + k] - __ldg(L.m.bc + k)) * acc[q];
    v = warp_sum(v);
    if (lane == 0 && c < L.nchains) L.Q[(size_t)c * nstrips + strip] = v;
  }
}

// iteration index / sink slot of this launch (see StbLargeLaunch::ctl)
__device__ __forceinline__ uint32_t stbl_iter(const StbLargeLaunch& L) { return L.iter + (L.ctl ? (uint32_t)L.ctl[0] : 0u); }
__device__ __forceinline__ long long stbl_slot(const StbLargeLaunch& L) { return L.slot < 0 ? -1 : L.slot + (L.ctl ? L.ctl[1] : 0); }
__global__ void stbl_advance_kernel(long long* ctl, int stored) { ctl[0] += 1; ctl[1] += stored; }

// Phase clocks of the one-CTA-per-chain kernels (development, -DBB_STBL_PHASES): CTA 0, thread 0 stores clock64 at the marks;
// launch_stb_large_iter prints the differences of one launch.
#ifdef BB_STBL_PHASES
__device__ long long bb_stbl_ph[2][8];
#define STBL_MARK(k, i) do { __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0) bb_stbl_ph[k][i] = clock64(); } while (0)
#else
#define STBL_MARK(k, i) do { } while (0)
#endif

// ---- tau, sig2: one CTA per chain -------------------------------------------------
inline size_t stbl_scalar_smem(int P) { return ((size_t)P + 26) * sizeof(double); }
__global__ void __launch_bounds__(SL_T) stbl_scalar_kernel(const StbLargeLaunch L) {
  extern __shared__ double smem[];
  const int P = L.m.P, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* dl_s = smem;                       // [P] beta - bc
  double* red = dl_s + P;                    // [24]
  double* sc = red + 24;                     // tau, sig2
  pdl_wait();
  const int ch = blockIdx.x; if (ch >= L.nchains) return;
  STBL_MARK(0, 0);
  const uint32_t gch = L.chain0 + (uint32_t)ch, t = stbl_iter(L);
  const double alpha = L.h.true_alpha > 0 ? L.h.true_alpha : L.alpha[ch];   // unknown alpha: chain state (stbl_alpha_kernel moves it)
  const bool know_tau = L.h.true_tau > 0, know_sig2 = L.h.true_sig2 > 0;
  const bool do_tau = ((L.stages & 1) || L.pre) && !know_tau, do_sig2 = (L.stages & 2) && !know_sig2 && !L.pre;
  uint32_t status = 0;
  const double* beta = L.beta + (size_t)ch * P;
  for (int j = tid; j < P; j += SL_T) dl_s[j] = beta[j] - __ldg(L.m.bc + j);
  if (tid == 0) { sc[0] = know_tau ? L.h.true_tau : L.tau[ch]; sc[1] = know_sig2 ? L.h.true_sig2 : L.sig2[ch]; }
  __syncthreads();
  double pt = 0.0, pq = 0.0, pg = 0.0;
  for (int k = tid; k < P; k += SL_T) {
    pt += exp(alpha * log(fabs(beta[k])));
    if (do_sig2) {
      pg = fma(dl_s[k], __ldg(L.m.gc + k), pg);
    }
  }
  if (do_sig2 && L.Q == nullptr) {
    // d'X'Xd over the LOWER 32 x 32 tiles of X'X (symmetric: off-diagonal tiles count twice): 36 of 64 tiles at P = 256.  A warp
    // per tile, lane = row of the tile (coalesced), the 32 columns of the tile in flight at once.  (One thread per row walking
    // all P columns: 20 us at P = 256 as a single chain of L2 loads, 14 us with 16 loads in flight -- the CTA pulled all 512 KB
    // of X'X through its SM either way.)
    const int T = (P + 31) / 32, ntile = T * (T + 1) / 2;
    for (int t = warp; t < ntile; t += SL_T / 32) {
      int I = 0;
      while ((I + 1) * (I + 2) / 2 <= t) ++I;
      const int J = t - I * (I + 1) / 2;
      const int row = 32 * I + lane, c0 = 32 * J;
      const bool rv = row < P;
      const double* xr = L.m.XX + (rv ? row : 0) + (size_t)c0 * P;
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
      if (c0 + 32 <= P) {
#pragma unroll
        for (int c = 0; c < 32; c += 4) {
          s0 = fma(__ldg(xr + (size_t)c * P), dl_s[c0 + c], s0); s1 = fma(__ldg(xr + (size_t)(c + 1) * P), dl_s[c0 + c + 1], s1);
          s2 = fma(__ldg(xr + (size_t)(c + 2) * P), dl_s[c0 + c + 2], s2); s3 = fma(__ldg(xr + (size_t)(c + 3) * P), dl_s[c0 + c + 3], s3);
        }
      } else {
        for (int c = 0; c0 + c < P; ++c) s0 = fma(__ldg(xr + (size_t)c * P), dl_s[c0 + c], s0);
      }
      const double v = rv ? dl_s[row] * ((s0 + s1) + (s2 + s3)) : 0.0;
      pq += I == J ? v : 2.0 * v;
    }
  }
  if (do_sig2 && L.Q != nullptr) {   // d'X'Xd from stbl_quad_kernel's per-strip partial sums
    const int nstrips = (P + QD_R - 1) / QD_R;
    for (int s = tid; s < nstrips; s += SL_T) pq += L.Q[(size_t)ch * nstrips + s];
  }
  pt = warp_sum(pt); pq = warp_sum(pq); pg = warp_sum(pg);
  if (lane == 0) { red[warp] = pt; red[8 + warp] = pq; red[16 + warp] = pg; }
  __syncthreads();
  STBL_MARK(0, 1);
  if (warp == 0 && do_tau) {
    double st = lane < 8 ? red[lane] : 0.0; st = warp_sum(st);
    Stream s(L.key, gch, t, L.pre ? ST_TAU_INIT : ST_TAU, 0);
    const double nu = gamma_rate<true>(s, L.h.nu_shape + ((double)P) / alpha, L.h.nu_rate + st, status);
    if (lane == 0) sc[0] = exp(-1.0 * log(nu) / alpha);
  }
  if (warp == 1 && do_sig2) {
    double sq = lane < 8 ? red[8 + lane] : 0.0, sg = lane < 8 ? red[16 + lane] : 0.0;
    sq = warp_sum(sq); sg = warp_sum(sg);
    const double rss = L.m.r0b - 2.0 * sg + sq;
    Stream s(L.key, gch, t, ST_SIG2, 0);
    const double g = gamma_rate<true>(s, L.h.sig2_shape + 0.5 * L.m.N, 1.0, status);
    if (lane == 0) sc[1] = (L.h.sig2_scale + 0.5 * rss) / g;
  }
  __syncthreads();
  STBL_MARK(0, 2);
  pdl_trigger();
  if (tid == 0) { L.tau[ch] = sc[0]; L.sig2[ch] = sc[1]; L.alpha[ch] = alpha; }
  status = __reduce_or_sync(0xffffffffu, status);
  if (lane == 0 && status) atomicOr(L.status + ch, status);
}

// ---- lambda (BridgeRegression.cpp:506-510): nsplit CTAs per chain, each a contiguous slice of the coefficients ----
// One round of the tilted-stable sampler is ~1700 FP64 instructions per evaluation: 256 evaluations in flight keep the FP64 pipe
// of one SM busy for ~8000 cycles (-DBB_STBL_PHASES -DBB_PHASES), and with one CTA per chain the nine passes of a P = 256 chain
// took 62 of the 81 us of the scalar kernel while most SMs idled.  Split over CTAs the slices of a chain run on different SMs, and
// with fewer coefficients per CTA the speculative look-ahead (lambda_stage_cta: G offsets of a coefficient's stream evaluated at
// once) starts at the first pass.  The draws do not depend on the split (Philox item = coefficient index).
// The per-coefficient constants (RetConst) are kept in shared memory over the passes when the slice has at most SC_RC_MAXP entries.
constexpr int SC_RC_MAXP = 1024;
#ifndef BB_STBL_LAMBDA_GMAX
#define BB_STBL_LAMBDA_GMAX 32
#endif
inline size_t stbl_lambda_smem(int nmax) {
  return (nmax <= SC_RC_MAXP ? (size_t)nmax * sizeof(RetConst) : 0) + ((size_t)3 * nmax + 2) * sizeof(int);
}
__host__ __device__ inline int stbl_lambda_slice(int P, int nsplit) { return (P + nsplit - 1) / nsplit; }   // the longest slice
__global__ void __launch_bounds__(SL_T) stbl_lambda_kernel(const StbLargeLaunch L, int nsplit) {
  extern __shared__ double smem[];
  const int P = L.m.P, tid = threadIdx.x, lane = tid & 31, ch = blockIdx.y, sp = blockIdx.x;
  const int j0 = (int)((long long)P * sp / nsplit), n = (int)((long long)P * (sp + 1) / nsplit) - j0;
  const int nmax = stbl_lambda_slice(P, nsplit);
  RetConst* rc_s = nmax <= SC_RC_MAXP ? reinterpret_cast<RetConst*>(smem) : nullptr;
  int* q_s = reinterpret_cast<int*>(reinterpret_cast<char*>(smem) + (rc_s ? (size_t)nmax * sizeof(RetConst) : 0));   // [2][nmax]
  int* qn_s = q_s + 2 * nmax;
  uint32_t* kk_s = reinterpret_cast<uint32_t*>(qn_s + 2);
  if (tid == 0) { qn_s[0] = 0; qn_s[1] = 0; }
  pdl_wait();
  __syncthreads();
  STBL_MARK(0, 3);
  uint32_t status = 0;
  const double tau = L.tau[ch], alpha = L.alpha[ch];      // as left by stbl_scalar_kernel (known values included)
  lambda_stage_cta<SL_T>(tid, n, L.beta + (size_t)ch * P + j0, L.lambda + (size_t)ch * P + j0, tau, 0.5 * alpha, L.key,
                         L.chain0 + (uint32_t)ch, stbl_iter(L), rc_s, q_s, nmax, qn_s, kk_s, status, j0, BB_STBL_LAMBDA_GMAX);
  STBL_MARK(0, 4);
  pdl_trigger();
  status = __reduce_or_sync(0xffffffffu, status);
  if (lane == 0 && status) atomicOr(L.status + ch, status);
}

// ---- A_c = X'X + diag(lambda sig2 / tau^2), lower triangle (and the diagonal blocks in full) ----
__global__ void stbl_build_kernel(const StbLargeLaunch L) {
  const int P = L.m.P, ch = blockIdx.y;
  const double f = L.sig2[ch] / (L.tau[ch] * L.tau[ch]);
  double* A = L.A + (size_t)ch * P * P;
  const double* lam = L.lambda + (size_t)ch * P;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < (size_t)P * P; e += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(e % P), c = (int)(e / P);
    if (r < c && r / SL_NB != c / SL_NB) continue;
    A[e] = __ldg(L.m.XX + e) + (r == c ? lam[r] * f : 0.0);
  }
}

// ---- Cholesky of the diagonal block kb (<= 64 x 64) and the inverse of its factor, on the FP64 tensor cores ----
// History of this kernel (us per launch of 64 chains, P = 256): scaled columns, three barriers and a libm rsqrt per column 87;
// unscaled columns through shared memory 50; the block in registers (thread = row x column class), one barrier per column 52
// -- every one of its 16 conditional updates per column had become a divergent branch around a dependent LDS + DFMA, 1700
// cycles per column (scripts/mb/mb_potrf.cu); the same with unconditional loads and selects 41; this form, blocked by 8 x 8
// tiles like stb64_kernel (bb_stb64.cuh): see DESIGN.md 4.3.
constexpr int SP_LD = 68;    // row stride of the 64 x 64 arrays (4 mod 16: the DMMA fragment loads are conflict-free)
constexpr int SP_PB = 12;    // row stride of a published 8 x 8 tile / inverse diagonal block
inline size_t stbl_potrf_smem() { return (size_t)2 * SL_NB * SP_LD * sizeof(double); }

// Factor and inverse factor of one 8 x 8 diagonal tile, in registers, every lane redundantly (static indices, no shuffles on
// the dependent chain).  Columns stay UNSCALED during the elimination (multiplier = a(r,k) / a(k,k) through a reciprocal, 56
// cycles, instead of an rsqrt, 95, on the chain of the 8 pivots); the rsqrt of the pivots are taken off the chain.
// a(r,c), c <= r: on return the factor L (lower, scaled).  W = L^-1 (lower) -> Wout[8][SP_PB]: lane c < 8 computes column c
// (W(c,c) = 1/L(c,c), W(r,c) = -1/L(r,r) sum_{m=c}^{r-1} L(r,m) W(m,c); entries above the diagonal are zero), the sums of all
// rows advanced as soon as W(m,c) is known: a chain of 7 x (FMA + multiply).
__device__ __forceinline__ bool chol8_factor_inverse(double (&a)[8][8], double* __restrict__ Wout, int lane) {
  double rs[8];
  bool bad = false;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const double piv = a[k][k];
    bad = bad || !(piv > 0.0);
    const double rc = x_rcp(piv);
    rs[k] = x_rsqrt(piv);
#pragma unroll
    for (int r = k + 1; r < 8; ++r) {
      const double lip = a[r][k] * rc;
#pragma unroll
      for (int c = k + 1; c <= r; ++c) a[r][c] = fma(-lip, a[c][k], a[r][c]);
    }
  }
#pragma unroll
  for (int r = 0; r < 8; ++r)
#pragma unroll
    for (int c = 0; c <= r; ++c) a[r][c] *= rs[c];
  double w[8], sum[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) sum[r] = 0.0;
#pragma unroll
  for (int m = 0; m < 8; ++m) {
    w[m] = m == lane ? rs[m] : (m > lane ? -rs[m] * sum[m] : 0.0);
#pragma unroll
    for (int r = m + 1; r < 8; ++r) sum[r] = fma(a[r][m], w[m], sum[r]);
  }
  if (lane < 8) {
#pragma unroll
    for (int r = 0; r < 8; ++r) Wout[r * SP_PB + lane] = w[r];
  }
  return bad;
}

#ifdef BB_POTRF_PHASES
__device__ long long bb_potrf_ph[12];
#define POTRF_MARK(i) do { __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0) bb_potrf_ph[i] = clock64(); } while (0)
#else
#define POTRF_MARK(i) do { } while (0)
#endif
// The 36 lower 8 x 8 tiles of the block live in registers as DMMA accumulator fragments: tile n = i (i + 1) / 2 + j belongs to
// warp n & 7, slot n >> 3.  Block column s: publish the tiles (i,s) -> warp 0 factors the diagonal tile and inverts its factor
// -> panel L(i,s) = T(i,s) W_s' by DMMA -> trailing T(i,j) -= L(i,s) L(j,s)' by DMMA: three barriers per block column.
// Rows / columns beyond the matrix are padded with the identity.
// Where the not-yet-factored entries of the system come from.  src == nullptr: A itself (built beforehand; stbl_build_kernel,
// wide_gram_kernel, chol_lower_device).  Otherwise the entries are read from src (X'X, the same for every chain: stride 0) with
// lam[ch][r] sig2[ch] / tau[ch]^2 added on the diagonal, the first time a tile is touched -- no build pass over P x P per chain.
struct StblSrc {
  const double* src = nullptr; size_t stride = 0;
  const double *lam = nullptr, *sig2 = nullptr, *tau = nullptr;
};
// fuse_panel: the kernel also solves the FIRST tile of the panel below its block, L(kb+1,kb) = C(kb+1,kb) W_kb' (the tile every
// CTA of the next block column needs), see stbl_colblock_kernel.
#ifndef BB_POTRF_MINB
#define BB_POTRF_MINB 1
#endif
__global__ void __launch_bounds__(SL_T, BB_POTRF_MINB) stbl_potrf_kernel(double* Aall, double* Wall, int P, int kb, uint32_t* status, const StblSrc X, int fuse_panel) {
  extern __shared__ double smem[];
  double* S = smem;                        // [64][SP_LD] row-major: the factor L
  double* Wt = smem + SL_NB * SP_LD;       // [64][SP_LD] first the staged block (column-major), then W' : Wt[c][r] = W(r,c), W = L^-1
  __shared__ double pb_s[SL_NB * SP_PB];   // the published block column: rows 0..63, 8 columns
  __shared__ double wi_s[8 * 8 * SP_PB];   // the inverses of the 8 diagonal tiles of the factor
  __shared__ double tt_s[8][8 * SP_PB];    // per warp: an 8 x 8 tile, transposed (the inverse stage)
  const int ch = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* A = Aall + (size_t)ch * P * P;
  const int k0 = kb * SL_NB, nb = min(SL_NB, P - k0);
  const bool from_src = X.src != nullptr && kb == 0;  // (later diagonal blocks have been written to A by stbl_colblock_kernel)
  const double* D0 = from_src ? X.src + (size_t)ch * X.stride : A;
  const bool dadd = from_src && X.lam != nullptr;
  pdl_wait();
  // from the fourth block column on the next kernel of the chain is stbl_colblock_kernel with a main loop (>= 6 chunks) that
  // does not depend on this kernel: let it onto the SMs now (with less of a main loop its CTAs would only wait and hold
  // registers and shared memory the other sub-batches' kernels need)
  if (fuse_panel == 2 && k0 + SL_NB < P && kb >= 3) pdl_trigger();
  const double fdiag = dadd ? X.sig2[ch] / (X.tau[ch] * X.tau[ch]) : 0.0;
  POTRF_MARK(0);
  const int fr = lane >> 2, fc = 2 * (lane & 3);      // accumulator fragment: row fr, columns fc, fc + 1
  const int ar = lane >> 2, ak = lane & 3;            // A / B fragment: row ar, k = ak
  double c0[5], c1[5];
  int ti[5], tj[5];
#pragma unroll
  for (int sl = 0; sl < 5; ++sl) {                    // straight from global memory: 8 consecutive rows (64 B) per column and tile
    const int n = sl * 8 + warp;
    int i = 0;
    while ((i + 1) * (i + 2) / 2 <= n) ++i;
    ti[sl] = i; tj[sl] = n - i * (i + 1) / 2;
    if (n >= 36) { ti[sl] = 0; tj[sl] = -1; }          // no tile
    const int row = 8 * ti[sl] + fr, col = 8 * max(tj[sl], 0) + fc;
    c0[sl] = (row < nb && col < nb) ? D0[(k0 + row) + (size_t)(k0 + col) * P] : (row == col ? 1.0 : 0.0);
    c1[sl] = (row < nb && col + 1 < nb) ? D0[(k0 + row) + (size_t)(k0 + col + 1) * P] : (row == col + 1 ? 1.0 : 0.0);
    if (dadd && row < nb) {     // block column 0 straight from X'X: + lambda sig2 / tau^2 on the diagonal (BridgeRegression.cpp:554-556)
      if (row == col) c0[sl] += X.lam[(size_t)ch * P + k0 + row] * fdiag;
      if (row == col + 1) c1[sl] += X.lam[(size_t)ch * P + k0 + row] * fdiag;
    }
  }
  POTRF_MARK(1);
  bool bad = false;
  for (int s = 0; s < 8; ++s) {
    // (1) publish the tiles of block column s
#pragma unroll
    for (int sl = 0; sl < 5; ++sl)
      if (tj[sl] == s) *reinterpret_cast<double2*>(pb_s + (ti[sl] * 8 + fr) * SP_PB + fc) = make_double2(c0[sl], c1[sl]);
    __syncthreads();
    // (2) diagonal tile: factor (written to S) and inverse factor
    double a[8][8];
    if (warp == 0) {
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c <= r; ++c) a[r][c] = pb_s[(s * 8 + r) * SP_PB + c];
      bad = chol8_factor_inverse(a, wi_s + s * 8 * SP_PB, lane) || bad;
    }
    __syncthreads();
    if (warp == 0) {            // warp 0 owns no panel tile: it writes the factor of the diagonal tile meanwhile
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c <= r; ++c) S[(8 * s + r) * SP_LD + 8 * s + c] = a[r][c];
    }
    // (3) panel: L(i,s) = T(i,s) W_s'  for i > s, tile i by warp i
    if (warp > s) {
      const int i = warp;
      double d0 = 0.0, d1 = 0.0;
#pragma unroll
      for (int q = 0; q < 8; q += 4)
        dmma_m8n8k4(d0, d1, pb_s[(i * 8 + ar) * SP_PB + q + ak], wi_s[(s * 8 + ar) * SP_PB + q + ak]);
      *reinterpret_cast<double2*>(S + (8 * i + fr) * SP_LD + 8 * s + fc) = make_double2(d0, d1);
    }
    __syncthreads();
    // (4) trailing update of my tiles (i,j), i >= j > s
#pragma unroll
    for (int sl = 0; sl < 5; ++sl)
      if (tj[sl] > s) {
#pragma unroll
        for (int q = 0; q < 8; q += 4)
          dmma_m8n8k4(c0[sl], c1[sl], -S[(8 * ti[sl] + ar) * SP_LD + 8 * s + q + ak], S[(8 * tj[sl] + ar) * SP_LD + 8 * s + q + ak]);
      }
  }
  if (bad && lane == 0) atomicOr(status + ch, (uint32_t)ST_CHOL_FAIL);
  // the tile of the fused panel solve is fetched now: its L2 latency (6000 cycles when loaded where it is used) hides behind
  // the inverse stage.  Thread -> row r = tid & 63, columns (tid >> 6) + 4 q.
  const bool do_panel = fuse_panel && k0 + SL_NB < P;
  double pc[16];
  if (do_panel) {
    const int i0 = k0 + SL_NB, r = tid & (SL_NB - 1);
    const bool rv = i0 + r < P;
#pragma unroll
    for (int q = 0; q < 16; ++q) pc[q] = rv ? D0[(i0 + r) + (size_t)(k0 + (tid >> 6) + 4 * q) * P] : 0.0;
  }
  POTRF_MARK(2);
  for (int e = tid; e < SL_NB * SL_NB; e += SL_T) {
    const int r = e & (SL_NB - 1), c = e >> 6;
    if (r < nb && c <= r) A[(k0 + r) + (size_t)(k0 + c) * P] = S[r * SP_LD + c];
    Wt[c * SP_LD + r] = 0.0;
  }
  POTRF_MARK(3);
  __syncthreads();
  // W = L^-1 (lower) by 8 x 8 tiles: the diagonal tiles are the inverses kept above; block row i = 1..7, tile (i,j) by warp j < i:
  //   T(i,j) = sum_{k=j}^{i-1} L(i,k) W(k,j),   W(i,j) = -W(i,i) T(i,j)     (both on the tensor cores)
  // -- depth 7 x (at most 14 dependent DMMAs) instead of the 64 dependent rows of a plain forward substitution.
  // The panel solve below the block is a DMMA product with W' (stbl_trsm_kernel).
  {
    const int j = warp;
#pragma unroll
    for (int q = 0; q < 2; ++q) {    // diagonal tile j: Wt[8j + c][8j + r] = W_j(r,c)
      const int e = lane + 32 * q, r = e & 7, c = e >> 3;
      Wt[(8 * j + c) * SP_LD + 8 * j + r] = wi_s[(j * 8 + r) * SP_PB + c];
    }
  }
  __syncthreads();
  POTRF_MARK(4);
  for (int i = 1; i < 8; ++i) {
    const int j = warp;
    if (j < i) {
      double t0 = 0.0, t1 = 0.0;
      for (int k = j; k < i; ++k) {
#pragma unroll
        for (int q = 0; q < 8; q += 4)     // A = L(i,k)[ar][q + ak],  B(kk, n) = W(k,j)(kk, n) = Wt[8j + n][8k + kk]
          dmma_m8n8k4(t0, t1, S[(8 * i + ar) * SP_LD + 8 * k + q + ak], Wt[(8 * j + ar) * SP_LD + 8 * k + q + ak]);
      }
      double* tt = tt_s[warp];             // transposed: tt[n][kk] = T(kk, n)
      tt[fc * SP_PB + fr] = t0; tt[(fc + 1) * SP_PB + fr] = t1;
      __syncwarp();
      double w0 = 0.0, w1 = 0.0;
#pragma unroll
      for (int q = 0; q < 8; q += 4)
        dmma_m8n8k4(w0, w1, -wi_s[(i * 8 + ar) * SP_PB + q + ak], tt[ar * SP_PB + q + ak]);
      Wt[(8 * j + fc) * SP_LD + 8 * i + fr] = w0; Wt[(8 * j + fc + 1) * SP_LD + 8 * i + fr] = w1;
    }
    __syncthreads();
  }
  POTRF_MARK(5);
  pdl_trigger();
  double* W = Wall + ((size_t)ch * ((P + SL_NB - 1) / SL_NB) + kb) * SL_NB * SL_NB;   // kept per block column: the solve kernel uses them all
  for (int e = tid; e < SL_NB * SL_NB; e += SL_T) {    // W[e] = W(e % 64, e / 64); rows / columns beyond the matrix are zero
    const int r = e & (SL_NB - 1), c = e >> 6;
    W[e] = (r < nb && c < nb) ? Wt[c * SP_LD + r] : 0.0;
  }
  POTRF_MARK(6);
  if (do_panel) {
    // L(kb+1,kb) = C(kb+1,kb) W_kb'  (here nb = 64).  C staged k-major in S (the factor has been stored), warp w = rows 8w..8w+7.
    const int i0 = k0 + SL_NB, nb2 = min(SL_NB, P - i0);
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 16; ++q) S[((tid >> 6) + 4 * q) * SP_LD + (tid & (SL_NB - 1))] = pc[q];
    __syncthreads();
    POTRF_MARK(8);
    double acc[8][2];
#pragma unroll
    for (int t = 0; t < 8; ++t) { acc[t][0] = 0.0; acc[t][1] = 0.0; }
#pragma unroll 4
    for (int kk = 0; kk < SL_NB; kk += 4) {
      const double a = S[(kk + ak) * SP_LD + 8 * warp + ar];
#pragma unroll
      for (int t = 0; t < 8; ++t) dmma_m8n8k4(acc[t][0], acc[t][1], a, Wt[(kk + ak) * SP_LD + 8 * t + ar]);   // B(k,n) = W(n,k)
    }
    POTRF_MARK(9);
    if (8 * warp + fr < nb2) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        A[(i0 + 8 * warp + fr) + (size_t)(k0 + 8 * t + fc) * P] = acc[t][0];
        A[(i0 + 8 * warp + fr) + (size_t)(k0 + 8 * t + fc + 1) * P] = acc[t][1];
      }
    }
  }
  POTRF_MARK(7);
}

// ---- panel: rows below the diagonal block,  L(i,kb) = C(i,kb) W_kb'  on the FP64 tensor cores ----
// (X L_kk' = C solved as a product with the inverse factor of the 64 x 64 diagonal block: one 64^3 DMMA tile product
// per row tile instead of 64 dependent substitution steps per row.)
__global__ void __launch_bounds__(256) stbl_trsm_kernel(double* Aall, const double* Wall, int P, int kb) {
  extern __shared__ double smem[];
  double* As = smem;
  double* Bs = smem + SL_NB * SL_LD;
  const int ch = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* A = Aall + (size_t)ch * P * P;
  const double* W = Wall + ((size_t)ch * ((P + SL_NB - 1) / SL_NB) + kb) * SL_NB * SL_NB;
  const int k0 = kb * SL_NB, nb = min(SL_NB, P - k0);
  const int i0 = k0 + SL_NB + blockIdx.x * SL_NB;
  for (int e = tid; e < SL_NB * SL_NB; e += 256) {
    const int r = e % SL_NB, k = e / SL_NB;
    As[k * SL_LD + r] = (i0 + r < P && k < nb) ? A[(i0 + r) + (size_t)(k0 + k) * P] : 0.0;
    Bs[k * SL_LD + r] = W[r + k * SL_NB];          // B(k, c) = W(c, k)
  }
  __syncthreads();
  double acc[8][2];
#pragma unroll
  for (int t = 0; t < 8; ++t) { acc[t][0] = 0.0; acc[t][1] = 0.0; }
  tile_mma(As, Bs, acc, lane, warp);
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const int r = i0 + warp * 8 + (lane >> 2), c = t * 8 + 2 * (lane & 3);
    if (r < P) {
      if (c < nb) A[r + (size_t)(k0 + c) * P] = acc[t][0];
      if (c + 1 < nb) A[r + (size_t)(k0 + c + 1) * P] = acc[t][1];
    }
  }
}

// ---- block column jb, LEFT-looking, on the FP64 tensor cores:  C(i,jb) = A(i,jb) - sum_{k<jb} L(i,k) L(jb,k)'  ----
// CTA = one 64 x 64 tile (row tile i >= jb) of one chain; the sum over the jb finished block columns is
// accumulated in DMMA fragments, so every tile of the factor is read-modify-written ONCE per factorisation.
// (The right-looking form updated the whole trailing matrix after every block column: P/64 read-modify-write
// passes over a 32 MB matrix per chain, 86 GB of HBM traffic per iteration at c4 -- it ran at the speed of that
// traffic, 27 % of the DGEMM peak.  Left-looking moves 3.3x fewer bytes: each L(i,k) tile is read once per
// block column below it, the L(jb,k) tiles are shared by the CTAs of a chain through L2.)
// Panels are staged k-major in smem (ld 72: minimal 2-way pattern).
__global__ void __launch_bounds__(256, 2) stbl_update_kernel(double* Aall, int P, int jb) {
  extern __shared__ double smem[];
  double* As = smem;
  double* Bs = smem + SL_NB * SL_LD;
  const int ch = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* A = Aall + (size_t)ch * P * P;
  const int ti = jb + blockIdx.x;
  const int i0 = ti * SL_NB, j0 = jb * SL_NB;
  double acc[2][4][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) { acc[a][t][0] = 0.0; acc[a][t][1] = 0.0; }
  // thread -> 16 elements of each 64 x 64 tile: row r = tid & 63, columns k = (tid >> 6) + 4 q (coalesced along r)
  const int r = tid & (SL_NB - 1), kq = tid >> 6;
  const bool ra = i0 + r < P, rb = j0 + r < P;
  const double* pa_src = A + (i0 + r) + (size_t)kq * P;
  const double* pb_src = A + (j0 + r) + (size_t)kq * P;
  double pa[16], pb[16];
  auto fetch = [&](int kb) {   // the global loads of the NEXT step are in flight while this step's tile product runs
    const size_t off = (size_t)kb * SL_NB * P;
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      pa[q] = ra ? pa_src[off + (size_t)(4 * q) * P] : 0.0;
      pb[q] = rb ? pb_src[off + (size_t)(4 * q) * P] : 0.0;
    }
  };
  if (jb > 0) fetch(0);
  for (int kb = 0; kb < jb; ++kb) {
#pragma unroll
    for (int q = 0; q < 16; ++q) { As[(kq + 4 * q) * SL_LD + r] = pa[q]; Bs[(kq + 4 * q) * SL_LD + r] = pb[q]; }
    __syncthreads();
    if (kb + 1 < jb) fetch(kb + 1);
    tile_mma_16x32(As, Bs, acc, lane, warp);
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int rr = i0 + (warp & 3) * 16 + 8 * a + (lane >> 2), c = j0 + (warp >> 2) * 32 + t * 8 + 2 * (lane & 3);
      if (rr < P) {
        if (c < P && (ti > jb || rr >= c)) A[rr + (size_t)c * P] -= acc[a][t][0];
        if (c + 1 < P && (ti > jb || rr >= c + 1)) A[rr + (size_t)(c + 1) * P] -= acc[a][t][1];
      }
    }
}

// ---- cp.async staging of the colblock kernel: 64 rows x 32 columns of a column-major tile -> smem, k-major [32][SL_LD] ----
// (LDGSTS: no register staging and no STS instructions on the warps that run the DMMAs; rows beyond the matrix are zero-filled
//  through the src-size operand.)  V16: 16-byte copies (pairs of rows; needs an even leading dimension), else 8-byte copies.
#ifndef BB_COLBLOCK_CPASYNC
#define BB_COLBLOCK_CPASYNC 1
#endif
constexpr int CB_KC = 32;                                   // k depth of a stage
constexpr int CB_STAGE = 2 * CB_KC * SL_LD;                 // doubles per stage: A chunk, then B chunk
constexpr int CB_NSTAGE = 3;
inline size_t stbl_colblock_smem() {
  const size_t full = (size_t)2 * SL_NB * SL_LD * sizeof(double), ring = (size_t)CB_NSTAGE * CB_STAGE * sizeof(double);
  return BB_COLBLOCK_CPASYNC ? (ring > full ? ring : full) : full;
}
template <bool V16> __device__ __forceinline__ void cb_copy_chunk(double* dst, const double* src, int P, int row0, int tid, int kmax = CB_KC) {
  // src: element (row0, first column of the chunk) of the tile, leading dimension P, rows < P valid, columns < kmax valid;
  // dst: [CB_KC][SL_LD]
  if (V16) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int op = tid + 256 * q, rp = op & 31, k = op >> 5;
      const int left = k < kmax ? P - (row0 + 2 * rp) : 0;  // rows of the matrix from this pair on
      const int bytes = left >= 2 ? 16 : (left == 1 ? 8 : 0);
      const double* g = bytes ? src + 2 * rp + (size_t)k * P : src;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(smem_u32(dst + k * SL_LD + 2 * rp)), "l"(g), "r"(bytes) : "memory");
    }
  } else {
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int op = tid + 256 * q, r = op & 63, k = op >> 6;
      const int bytes = (row0 + r < P && k < kmax) ? 8 : 0;
      const double* g = bytes ? src + r + (size_t)k * P : src;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" :: "r"(smem_u32(dst + k * SL_LD + r)), "l"(g), "r"(bytes) : "memory");
    }
  }
}
// acc += A(64 x 32) B(64 x 32)' with the 16 x 32 warp tiles of tile_mma_16x32
__device__ __forceinline__ void chunk_mma_16x32(const double* As, const double* Bs, double (&acc)[2][4][2], int lane, int warp) {
  const int r0 = (warp & 3) * 16 + (lane >> 2), c0 = (warp >> 2) * 32 + (lane >> 2);
#pragma unroll 4
  for (int kk = 0; kk < CB_KC; kk += 4) {
    const double* ak = As + (kk + (lane & 3)) * SL_LD;
    const double* bk = Bs + (kk + (lane & 3)) * SL_LD;
    const double a0 = ak[r0], a1 = ak[r0 + 8];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const double b = bk[c0 + 8 * t];
      dmma_m8n8k4(acc[0][t][0], acc[0][t][1], a0, b);
      dmma_m8n8k4(acc[1][t][0], acc[1][t][1], a1, b);
    }
  }
}

// ---- block column jb >= 1 in ONE kernel (P <= 512): the panel solve of block column jb - 1 and the left-looking update ----
//   CTA = row tile ti >= jb of one chain:
//     ti > jb :  L(ti,jb-1) = C(ti,jb-1) W_{jb-1}'            (its own tile of the previous panel, written back as final)
//     C(ti,jb) = S(ti,jb) - sum_{k<jb} L(ti,k) L(jb,k)'       (the last term with the tile just solved, from shared memory)
//   L(jb,jb-1), which every CTA needs as the right-hand operand, was solved by stbl_potrf_kernel(jb-1, fuse_panel).
// Against stbl_update_kernel + stbl_trsm_kernel (+ stbl_build_kernel): the same arithmetic, one launch per block column
// instead of two, one pass less over the matrix -- at P = 256 the 64 x 64-tile kernels are bound by their latencies
// (17 us for one to three tile products of 2 us), so the SM time goes with the number of CTA launches.
__global__ void __launch_bounds__(256) stbl_colblock_v1_kernel(double* Aall, const double* Wall, int P, int jb, const StblSrc X) {
  extern __shared__ double smem[];
  double* As = smem;
  double* Bs = smem + SL_NB * SL_LD;
  const int ch = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* A = Aall + (size_t)ch * P * P;
  const double* Sm = X.src ? X.src + (size_t)ch * X.stride : A;
  const int ti = jb + blockIdx.x;
  const int i0 = ti * SL_NB, j0 = jb * SL_NB, kc = jb - 1, c0 = kc * SL_NB;
  double acc[2][4][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) { acc[a][t][0] = 0.0; acc[a][t][1] = 0.0; }
  const int r = tid & (SL_NB - 1), kq = tid >> 6;      // thread -> 16 elements of a 64 x 64 tile: row r, columns kq + 4 q
  const bool ra = i0 + r < P, rb = j0 + r < P;
  const int wr = (warp & 3) * 16 + (lane >> 2), wc = (warp >> 2) * 32 + 2 * (lane & 3);   // fragment: rows wr + 8 a, columns wc + 8 t + {0,1}
  double pa[16], pb[16];
  pdl_wait();
  // my tile of the previous panel (unsolved) and W_{jb-1}' go first: the solve heads the chain
  if (ti > jb) {
    const double* Cs = (X.src && kc == 0) ? Sm : A;
    const double* W = Wall + ((size_t)ch * ((P + SL_NB - 1) / SL_NB) + kc) * SL_NB * SL_NB;
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      pa[q] = ra ? Cs[(i0 + r) + (size_t)(c0 + kq + 4 * q) * P] : 0.0;
      pb[q] = W[r + (kq + 4 * q) * SL_NB];             // B(k, c) = W(c, k)
    }
#pragma unroll
    for (int q = 0; q < 16; ++q) { As[(kq + 4 * q) * SL_LD + r] = pa[q]; Bs[(kq + 4 * q) * SL_LD + r] = pb[q]; }
    __syncthreads();
  }
  const double* pa_src = A + (i0 + r) + (size_t)kq * P;
  const double* pb_src = A + (j0 + r) + (size_t)kq * P;
  auto fetch = [&](int kb) {
    const size_t off = (size_t)kb * SL_NB * P;
#pragma unroll
    for (int q = 0; q < 16; ++q) {
      pa[q] = ra ? pa_src[off + (size_t)(4 * q) * P] : 0.0;
      pb[q] = rb ? pb_src[off + (size_t)(4 * q) * P] : 0.0;
    }
  };
  // block column kc = jb - 1 first: right-hand tile L(jb,kc) from memory, left-hand tile = the one solved here (ti > jb) or the same tile
  fetch(kc);                                           // (for ti > jb only pb is used: in flight during the solve)
  if (ti > jb) {
    double dn[2][4][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int t = 0; t < 4; ++t) { dn[a][t][0] = 0.0; dn[a][t][1] = 0.0; }
    tile_mma_16x32(As, Bs, dn, lane, warp);            // L(ti,kc) = C(ti,kc) W'
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int rr = i0 + wr + 8 * a, c = c0 + wc + 8 * t;
        if (rr < P) { A[rr + (size_t)c * P] = dn[a][t][0]; A[rr + (size_t)(c + 1) * P] = dn[a][t][1]; }
      }
    __syncthreads();
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        As[(wc + 8 * t) * SL_LD + wr + 8 * a] = dn[a][t][0]; As[(wc + 8 * t + 1) * SL_LD + wr + 8 * a] = dn[a][t][1];
      }
  } else {
#pragma unroll
    for (int q = 0; q < 16; ++q) As[(kq + 4 * q) * SL_LD + r] = pb[q];
  }
#pragma unroll
  for (int q = 0; q < 16; ++q) Bs[(kq + 4 * q) * SL_LD + r] = pb[q];
  __syncthreads();
#if BB_COLBLOCK_CPASYNC
  tile_mma_16x32(As, Bs, acc, lane, warp);
  __syncthreads();
  // the finished block columns k < jb - 1 in chunks of 32 columns through a 3-stage cp.async ring: chunk c + 2 is in flight
  // while chunk c is multiplied, one barrier per chunk, no register staging (the LDG -> STS form ran the DMMA pipe at 71-77 %)
  {
    const int nchunk = 2 * kc;
    const bool v16 = (P & 1) == 0;
    auto issue = [&](int c) {
      if (c < nchunk) {
        double* st = smem + (c % CB_NSTAGE) * CB_STAGE;
        const size_t col = (size_t)(c >> 1) * SL_NB + (size_t)(c & 1) * CB_KC;
        if (v16) { cb_copy_chunk<true>(st, A + i0 + col * P, P, i0, tid); cb_copy_chunk<true>(st + CB_KC * SL_LD, A + j0 + col * P, P, j0, tid); }
        else { cb_copy_chunk<false>(st, A + i0 + col * P, P, i0, tid); cb_copy_chunk<false>(st + CB_KC * SL_LD, A + j0 + col * P, P, j0, tid); }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");   // (an empty group at the tail keeps the count uniform)
    };
    issue(0); issue(1);
    for (int c = 0; c < nchunk; ++c) {
      asm volatile("cp.async.wait_group 1;" ::: "memory");   // chunk c has landed (this thread's copies) ...
      __syncthreads();                                        // ... everyone's; and everyone is done with chunk c - 1
      issue(c + 2);
      const double* st = smem + (c % CB_NSTAGE) * CB_STAGE;
      chunk_mma_16x32(st, st + CB_KC * SL_LD, acc, lane, warp);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
#else
  if (kc > 0) fetch(0);
  tile_mma_16x32(As, Bs, acc, lane, warp);
  __syncthreads();
  // the finished block columns k < jb - 1
  for (int kb = 0; kb < kc; ++kb) {
#pragma unroll
    for (int q = 0; q < 16; ++q) { As[(kq + 4 * q) * SL_LD + r] = pa[q]; Bs[(kq + 4 * q) * SL_LD + r] = pb[q]; }
    __syncthreads();
    if (kb + 1 < kc) fetch(kb + 1);
    tile_mma_16x32(As, Bs, acc, lane, warp);
    __syncthreads();
  }
#endif
  pdl_trigger();
  const double fdiag = (X.lam && ti == jb) ? X.sig2[ch] / (X.tau[ch] * X.tau[ch]) : 0.0;
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int rr = i0 + wr + 8 * a, c = j0 + wc + 8 * t;
      if (rr < P) {
#pragma unroll
        for (int q = 0; q < 2; ++q)
          if (c + q < P) {
            double v = Sm[rr + (size_t)(c + q) * P];
            if (X.lam && rr == c + q) v += X.lam[(size_t)ch * P + rr] * fdiag;
            A[rr + (size_t)(c + q) * P] = v - acc[a][t][q];
          }
      }
    }
}

// ---- stbl_colblock_kernel, final form of round 2: the same work in the order  finished block columns -> panel solve -> last term.
// Everything the ring reads (block columns k <= jb - 2) was complete before stbl_potrf_kernel(jb-1) STARTED, so with
// programmatic dependent launch this kernel runs its main loop WHILE that potrf kernel (one latency-bound CTA per chain) is still
// working, and waits for it (pdl_wait) only before it touches W_{jb-1} and L(jb,jb-1).  All operands travel by cp.async (no
// register staging at all: 126 -> fewer registers live across the solve).  stbl_colblock_v1_kernel above (solve first) is kept
// for comparison (BB_STBL_COLBLOCK_V1=1).
__global__ void __launch_bounds__(256) stbl_colblock_kernel(double* Aall, const double* Wall, int P, int jb, const StblSrc X) {
  extern __shared__ double smem[];
  double* As = smem;                            // [64][SL_LD] k-major: C(ti,kc), then L(ti,kc)
  double* Bs = smem + SL_NB * SL_LD;            // [64][SL_LD]: W_kc'
  double* T3 = smem + 2 * SL_NB * SL_LD;        // [64][SL_LD]: L(jb,kc)   (the three tiles fill the ring's 3 stages exactly)
  const int ch = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* A = Aall + (size_t)ch * P * P;
  const double* Sm = X.src ? X.src + (size_t)ch * X.stride : A;
  const int ti = jb + blockIdx.x;
  const int i0 = ti * SL_NB, j0 = jb * SL_NB, kc = jb - 1, c0 = kc * SL_NB;
  const int wr = (warp & 3) * 16 + (lane >> 2), wc = (warp >> 2) * 32 + 2 * (lane & 3);   // fragment: rows wr + 8 a, columns wc + 8 t + {0,1}
  const bool v16 = (P & 1) == 0;
  double acc[2][4][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) { acc[a][t][0] = 0.0; acc[a][t][1] = 0.0; }
  // ---- the finished block columns k < jb - 1: 32-column chunks through the 3-stage cp.async ring ----
  {
    const int nchunk = 2 * kc;
    auto issue = [&](int c) {
      if (c < nchunk) {
        double* st = smem + (c % CB_NSTAGE) * CB_STAGE;
        const size_t col = (size_t)(c >> 1) * SL_NB + (size_t)(c & 1) * CB_KC;
        if (v16) { cb_copy_chunk<true>(st, A + i0 + col * P, P, i0, tid); cb_copy_chunk<true>(st + CB_KC * SL_LD, A + j0 + col * P, P, j0, tid); }
        else { cb_copy_chunk<false>(st, A + i0 + col * P, P, i0, tid); cb_copy_chunk<false>(st + CB_KC * SL_LD, A + j0 + col * P, P, j0, tid); }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");   // (an empty group at the tail keeps the count uniform)
    };
    issue(0); issue(1);
    for (int c = 0; c < nchunk; ++c) {
      asm volatile("cp.async.wait_group 1;" ::: "memory");   // chunk c has landed (this thread's copies) ...
      __syncthreads();                                        // ... everyone's; and everyone is done with chunk c - 1
      issue(c + 2);
      const double* st = smem + (c % CB_NSTAGE) * CB_STAGE;
      chunk_mma_16x32(st, st + CB_KC * SL_LD, acc, lane, warp);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                          // the ring's stages are free: they are reused as As / Bs / T3
  }
  // ---- block column kc = jb - 1: needs the potrf kernel of that block column ----
  pdl_wait();
  if (v16) { cb_copy_chunk<true>(T3, A + j0 + (size_t)c0 * P, P, j0, tid); cb_copy_chunk<true>(T3 + CB_KC * SL_LD, A + j0 + (size_t)(c0 + CB_KC) * P, P, j0, tid); }
  else { cb_copy_chunk<false>(T3, A + j0 + (size_t)c0 * P, P, j0, tid); cb_copy_chunk<false>(T3 + CB_KC * SL_LD, A + j0 + (size_t)(c0 + CB_KC) * P, P, j0, tid); }
  if (ti > jb) {
    const double* Cs = ((X.src && kc == 0) ? Sm : A) + i0 + (size_t)c0 * P;
    const double* W = Wall + ((size_t)ch * ((P + SL_NB - 1) / SL_NB) + kc) * SL_NB * SL_NB;
    if (v16) { cb_copy_chunk<true>(As, Cs, P, i0, tid); cb_copy_chunk<true>(As + CB_KC * SL_LD, Cs + (size_t)CB_KC * P, P, i0, tid); }
    else { cb_copy_chunk<false>(As, Cs, P, i0, tid); cb_copy_chunk<false>(As + CB_KC * SL_LD, Cs + (size_t)CB_KC * P, P, i0, tid); }
    // W is a full 64 x 64 block, column-major with leading dimension 64: Bs[k][c] = W(c,k) = W[c + 64 k]
    // (cb_copy_chunk's leading dimension and row limit are one parameter: a second call with the pointer advanced covers k >= 32)
    cb_copy_chunk<true>(Bs, W, SL_NB, 0, tid); cb_copy_chunk<true>(Bs + CB_KC * SL_LD, W + CB_KC * SL_NB, SL_NB, 0, tid);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (ti > jb) {
    double dn[2][4][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int t = 0; t < 4; ++t) { dn[a][t][0] = 0.0; dn[a][t][1] = 0.0; }
    tile_mma_16x32(As, Bs, dn, lane, warp);            // L(ti,kc) = C(ti,kc) W'
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int rr = i0 + wr + 8 * a, c = c0 + wc + 8 * t;
        if (rr < P) { A[rr + (size_t)c * P] = dn[a][t][0]; A[rr + (size_t)(c + 1) * P] = dn[a][t][1]; }
      }
    __syncthreads();                                   // everyone is done with C in As
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        As[(wc + 8 * t) * SL_LD + wr + 8 * a] = dn[a][t][0]; As[(wc + 8 * t + 1) * SL_LD + wr + 8 * a] = dn[a][t][1];
      }
    __syncthreads();
    tile_mma_16x32(As, T3, acc, lane, warp);           // += L(ti,kc) L(jb,kc)'
  } else {
    tile_mma_16x32(T3, T3, acc, lane, warp);
  }
  pdl_trigger();
  const double fdiag = (X.lam && ti == jb) ? X.sig2[ch] / (X.tau[ch] * X.tau[ch]) : 0.0;
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int rr = i0 + wr + 8 * a, c = j0 + wc + 8 * t;
      if (rr < P) {
#pragma unroll
        for (int q = 0; q < 2; ++q)
          if (c + q < P) {
            double v = Sm[rr + (size_t)(c + q) * P];
            if (X.lam && rr == c + q) v += X.lam[(size_t)ch * P + rr] * fdiag;
            A[rr + (size_t)(c + q) * P] = v - acc[a][t][q];
          }
      }
    }
}

// ---- the same update with a 128 x 64 CTA tile: every B fragment (the L(jb,k) panel, shared by all row tiles) feeds TWO
// DMMAs, so the shared-memory loads per tensor instruction drop from 9/8 to 10/16 -- the 64 x 64 form is bound by them.
// Warp w owns rows 16 w .. 16 w + 15 of the tile.  Launched for the row-tile PAIRS below the diagonal tile; the diagonal
// tile itself (lower triangle only) stays with the 64 x 64 kernel.
constexpr int SL_LDA = 132;  // row stride of the staged 128-row A panel (132 = 4 mod 16)
inline size_t stbl_update2_smem() { return (size_t)(SL_NB * SL_LDA + SL_NB * SL_LD) * sizeof(double); }
__global__ void __launch_bounds__(256, 2) stbl_update2_kernel(double* Aall, int P, int jb, int t0) {
  extern __shared__ double smem[];
  double* As = smem;                       // [64][SL_LDA]
  double* Bs = smem + SL_NB * SL_LDA;      // [64][SL_LD]
  const int ch = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double* A = Aall + (size_t)ch * P * P;
  const int i0 = (t0 + 2 * blockIdx.x) * SL_NB, j0 = jb * SL_NB;
  double acc[2][8][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 8; ++t) { acc[a][t][0] = 0.0; acc[a][t][1] = 0.0; }
  // thread -> A panel: row r2 = tid & 127, columns k = (tid >> 7) + 2 q (32 elements);  B panel: row r = tid & 63, k = (tid >> 6) + 4 q (16)
  const int r2 = tid & 127, kh = tid >> 7, r = tid & (SL_NB - 1), kq = tid >> 6;
  const bool ra = i0 + r2 < P, rb = j0 + r < P;
  const double* pa_src = A + (i0 + r2) + (size_t)kh * P;
  const double* pb_src = A + (j0 + r) + (size_t)kq * P;
  for (int kb = 0; kb < jb; ++kb) {
    const size_t off = (size_t)kb * SL_NB * P;
    // staged in two halves (24 loads in flight per thread): the 64 accumulator registers leave no room for all 48
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      double pa[16], pb[8];
#pragma unroll
      for (int q = 0; q < 16; ++q) pa[q] = ra ? pa_src[off + (size_t)(2 * (16 * hf + q)) * P] : 0.0;
#pragma unroll
      for (int q = 0; q < 8; ++q) pb[q] = rb ? pb_src[off + (size_t)(4 * (8 * hf + q)) * P] : 0.0;
      if (hf == 0) __syncthreads();        // the previous product is done with the panels
#pragma unroll
      for (int q = 0; q < 16; ++q) As[(kh + 2 * (16 * hf + q)) * SL_LDA + r2] = pa[q];
#pragma unroll
      for (int q = 0; q < 8; ++q) Bs[(kq + 4 * (8 * hf + q)) * SL_LD + r] = pb[q];
    }
    __syncthreads();
#pragma unroll 4
    for (int kk = 0; kk < SL_NB; kk += 4) {
      const double a0 = As[(kk + (lane & 3)) * SL_LDA + warp * 16 + (lane >> 2)];
      const double a1 = As[(kk + (lane & 3)) * SL_LDA + warp * 16 + 8 + (lane >> 2)];
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const double b = Bs[(kk + (lane & 3)) * SL_LD + t * 8 + (lane >> 2)];
        dmma_m8n8k4(acc[0][t][0], acc[0][t][1], a0, b);
        dmma_m8n8k4(acc[1][t][0], acc[1][t][1], a1, b);
      }
    }
  }
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int rr = i0 + warp * 16 + 8 * a + (lane >> 2), c = j0 + t * 8 + 2 * (lane & 3);
      if (rr < P) {
        if (c < P) A[rr + (size_t)c * P] -= acc[a][t][0];
        if (c + 1 < P) A[rr + (size_t)(c + 1) * P] -= acc[a][t][1];
      }
    }
}

// ---- the two blocked substitutions with a factor left by stbl_potrf / stbl_trsm (lower triangle of A, leading dimension P;
//      Wg = inverses of its diagonal blocks): x <- L^-1 x and x <- L^-T x, x in shared memory, one CTA of SV_T threads ----
// Threads of the solve kernels (template parameter NT): 1024 -- one CTA streams a chain's factor twice, it needs the loads in
// flight.  A 512-thread form (two CTAs per SM instead of a whole SM per chain) is kept behind BB_SV_SMALL_P (systems up to that
// size use it) but is off: measured at c5s (P = 256) 181 M draws/s against 187 M with 1024 threads.
#ifndef BB_SV_SMALL_P
#define BB_SV_SMALL_P 0
#endif
constexpr int SV_SMALL_P = BB_SV_SMALL_P;
// (x, D, ps are OFFSETS into the dynamic shared memory: handed over as generic pointers the compiler loses the address space
//  and every access becomes a generic load -- the solve kernel ran 2x slower, 51 -> 111 us at P = 256)
// W of the NEXT diagonal block travels through registers (4 values per thread): the loads are issued before the work on the
// current block and stored to shared memory after it, so that their L2 latency is off the chain of dependent blocks.
struct StblWRegs { double v[8]; };
template <int NT> __device__ __forceinline__ void stbl_fetch_w(const double* Wg, int kb, int tid, StblWRegs& w) {
  const double* W = Wg + (size_t)kb * SL_NB * SL_NB;
#pragma unroll
  for (int q = 0; q < SL_NB * SL_NB / NT; ++q) w.v[q] = W[tid + q * NT];
}
template <int NT> __device__ __forceinline__ void stbl_store_w(int D_off, int tid, const StblWRegs& w) {
  extern __shared__ double smem[];
  double* D = smem + D_off;
#pragma unroll
  for (int q = 0; q < SL_NB * SL_NB / NT; ++q) { const int e = tid + q * NT; D[(e & (SL_NB - 1)) * (SL_NB + 1) + (e >> 6)] = w.v[q]; }   // W[e] = W(e % 64, e / 64)
}
constexpr int SV_SPLIT_MAXP = 4096;   // up to here the four partial sums of the rows below a block fit in shared memory
inline size_t stbl_solve_smem(int P) { return ((size_t)P + SL_NB * (SL_NB + 1) + 16 * SL_NB + (P <= SV_SPLIT_MAXP ? 4 * (size_t)P : 0)) * sizeof(double); }
// Shared-memory layout of the solves (doubles): x [P] | D [64][65] | ps [16][64] | part [4][P].
// The diagonal blocks are applied as products with their inverses (rows / columns beyond the matrix are zero in W): a 64-term
// dot product per unknown, all unknowns of the block at once, instead of 64 dependent substitution steps with a division each.
template <int NT> __device__ __forceinline__ void stbl_forward(const double* A, const double* Wg, int P, int x_off, int D_off, int ps_off, int tid) {
  extern __shared__ double smem[];
  double* x = smem + x_off; double* D = smem + D_off; double* ps = smem + ps_off; double* part = ps + 16 * SL_NB;
  constexpr int PT = NT / SL_NB, TQ = SL_NB / PT;      // matrix-vector product: PT parts of TQ terms (16 x 4 at NT = 1024)
  constexpr int KP = NT / 256, KQ = SL_NB / KP;        // rows below the block: KP column groups of KQ columns (4 x 16 at NT = 1024)
  const int vr = tid & (SL_NB - 1), vp = tid >> 6;
  const int rs = tid & 255, kp = tid >> 8;
  StblWRegs wn;
  stbl_fetch_w<NT>(Wg, 0, tid, wn);
  // forward: L w = Xy, column-oriented
  for (int k0 = 0; k0 < P; k0 += SL_NB) {
    const int nb = min(SL_NB, P - k0);
    stbl_store_w<NT>(D_off, tid, wn);
    if (k0 + SL_NB < P) stbl_fetch_w<NT>(Wg, k0 / SL_NB + 1, tid, wn);
    __syncthreads();
    {   // w_blk = W x_blk
      double s = 0.0;
#pragma unroll
      for (int q = 0; q < TQ; ++q) { const int c = TQ * vp + q; if (c < nb) s = fma(D[vr * (SL_NB + 1) + c], x[k0 + c], s); }
      ps[vp * SL_NB + vr] = s;
    }
    __syncthreads();
    if (tid < nb) {
      double s = 0.0;
#pragma unroll
      for (int q = 0; q < PT; ++q) s += ps[q * SL_NB + tid];
      x[k0 + tid] = s;
    }
    __syncthreads();
    // rows below: x_r -= sum_k L(r,k) w_k, the 64 columns split over four thread groups (16 loads in flight per thread instead of
    // a chain of 64: with 192 rows at P = 256 only 192 threads worked and the chain of L2 latencies was 5000 cycles per block)
    if (P > SV_SPLIT_MAXP) {   // no room for the partial sums: one thread per row, the 64 columns in sequence
      for (int r = k0 + nb + tid; r < P; r += NT) {
        double s = x[r];
#pragma unroll 8
        for (int k = k0; k < k0 + nb; ++k) s = fma(-A[r + (size_t)k * P], x[k], s);
        x[r] = s;
      }
      __syncthreads();
      continue;
    }
    for (int r = k0 + nb + rs; r < P; r += 256) {
      const double* ar = A + r + (size_t)(k0 + KQ * kp) * P;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int q = 0; q < KQ; q += 2) {
        const int c = KQ * kp + q;
        const double a0 = c < nb ? ar[(size_t)q * P] : 0.0, a1 = c + 1 < nb ? ar[(size_t)(q + 1) * P] : 0.0;
        s0 = fma(a0, x[k0 + min(c, nb - 1)], s0); s1 = fma(a1, x[k0 + min(c + 1, nb - 1)], s1);
      }
      part[kp * P + r] = s0 + s1;
    }
    __syncthreads();
    for (int r = k0 + nb + tid; r < P; r += NT) {
      double sum = part[r];
#pragma unroll
      for (int g = 1; g < KP; ++g) sum += part[g * P + r];
      x[r] -= sum;
    }
    __syncthreads();
  }
}
template <int NT> __device__ __forceinline__ void stbl_backward(const double* A, const double* Wg, int P, int x_off, int D_off, int ps_off, int tid) {
  extern __shared__ double smem[];
  double* x = smem + x_off; double* D = smem + D_off; double* ps = smem + ps_off;
  const int vr = tid & (SL_NB - 1), vp = tid >> 6, lane = tid & 31, warp = tid >> 5;
  const int nblk = (P + SL_NB - 1) / SL_NB;
  // backward: L' b = x over the stored lower triangle: column k of L is contiguous, one warp per column, lanes along r
  constexpr int PT = NT / SL_NB, TQ = SL_NB / PT;
  double* pk = ps + (PT - 1) * SL_NB;     // [64] column sums (the last part of ps is free until the product below)
  StblWRegs wn;
  stbl_fetch_w<NT>(Wg, nblk - 1, tid, wn);
  for (int kbk = nblk - 1; kbk >= 0; --kbk) {
    const int k0 = kbk * SL_NB, nb = min(SL_NB, P - k0);
    stbl_store_w<NT>(D_off, tid, wn);
    if (kbk > 0) stbl_fetch_w<NT>(Wg, kbk - 1, tid, wn);
    for (int kk = warp; kk < nb; kk += NT / 32) {
      const double* col = A + (size_t)(k0 + kk) * P;
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
      int r = k0 + nb + lane;
      for (; r + 96 < P; r += 128) {
        s0 = fma(col[r], x[r], s0); s1 = fma(col[r + 32], x[r + 32], s1);
        s2 = fma(col[r + 64], x[r + 64], s2); s3 = fma(col[r + 96], x[r + 96], s3);
      }
      for (; r < P; r += 32) s0 = fma(col[r], x[r], s0);
      const double sum = warp_sum((s0 + s1) + (s2 + s3));
      if (lane == 0) pk[kk] = sum;
    }
    __syncthreads();
    if (tid < nb) x[k0 + tid] -= pk[tid];
    __syncthreads();
    {   // b_blk = W' x_blk
      double s = 0.0;
#pragma unroll
      for (int q = 0; q < TQ; ++q) { const int r = TQ * vp + q; if (r < nb) s = fma(D[r * (SL_NB + 1) + vr], x[k0 + r], s); }
      __syncthreads();               // pk (inside ps) has been read by everyone
      ps[vp * SL_NB + vr] = s;
    }
    __syncthreads();
    if (tid < nb) {
      double s = 0.0;
#pragma unroll
      for (int q = 0; q < PT; ++q) s += ps[q * SL_NB + tid];
      x[k0 + tid] = s;
    }
    __syncthreads();
  }
}

// ---- beta = L^-T (L^-1 Xy + sig eps): one CTA per chain, blocked substitution ----
template <int NT> __global__ void __launch_bounds__(NT, 1024 / NT) stbl_solve_kernel(const StbLargeLaunch L) {
  extern __shared__ double smem[];
  const int P = L.m.P, tid = threadIdx.x, ch = blockIdx.x;
  pdl_wait();
  if (ch >= L.nchains) return;
  double* x = smem;                 // [P]; then D [64][65]: inverse W = L_kk^-1 of the current diagonal block (from stbl_potrf_kernel),
                                    // then ps [16][64]: partial sums of the 64 x 64 matrix-vector products (see stbl_forward)
  const double* A = L.A + (size_t)ch * P * P;
  const int nblk = (P + SL_NB - 1) / SL_NB;
  const double* Wg = L.W + (size_t)ch * nblk * SL_NB * SL_NB;
  const uint32_t gch = L.chain0 + (uint32_t)ch, it_ = stbl_iter(L);
  const double sig = sqrt(L.sig2[ch]);
  STBL_MARK(1, 0);
  for (int j = tid; j < P; j += NT) x[j] = __ldg(L.m.Xy + j);
  const int D_off = P, ps_off = P + SL_NB * (SL_NB + 1);
  stbl_forward<NT>(A, Wg, P, 0, D_off, ps_off, tid);  // w = L^-1 Xy (the reference's first trsm, BridgeRegression.cpp:564)
  STBL_MARK(1, 1);
  for (int j = tid; j < P; j += NT) { Stream s(L.key, gch, it_, ST_BETA_STB, (uint32_t)j); x[j] = fma(sig, s.norm(0), x[j]); }
  __syncthreads();
  STBL_MARK(1, 2);
  stbl_backward<NT>(A, Wg, P, 0, D_off, ps_off, tid); // beta = L^-T (w + sig eps)  (:565, :570-574)
  STBL_MARK(1, 3);
  pdl_trigger();
  double* beta = L.beta + (size_t)ch * P;
  const long long slot = stbl_slot(L);
  for (int j = tid; j < P; j += NT) {
    beta[j] = x[j];
    if (slot >= 0) {
      const size_t o = ((size_t)ch * L.sink.chain_stride + slot) * P + j;
      if (L.sink.beta) L.sink.beta[o] = x[j];
      if (L.sink.lambda) L.sink.lambda[o] = L.lambda[(size_t)ch * P + j];
    }
  }
  if (tid == 0 && slot >= 0) {
    const size_t o = (size_t)ch * L.sink.chain_stride + slot;
    if (L.sink.sig2) L.sink.sig2[o] = L.sig2[ch];
    if (L.sink.tau) L.sink.tau[o] = L.tau[ch];
    if (L.sink.alpha) L.sink.alpha[o] = L.alpha[ch];
  }
}

// ---- alpha, only when it is unknown (BridgeRegression.cpp:477-503): one warp per chain, after the beta draw ----
__global__ void __launch_bounds__(256) stbl_alpha_kernel(const StbLargeLaunch L) {
  const int P = L.m.P, lane = threadIdx.x & 31;
  const int ch = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (ch >= L.nchains) return;
  const double* beta = L.beta + (size_t)ch * P;
  const double tau = L.h.true_tau > 0 ? L.h.true_tau : L.tau[ch];
  Stream sa(L.key, L.chain0 + (uint32_t)ch, stbl_iter(L), ST_ALPHA, 0);
  const AlphaProposal pr = alpha_propose(sa, L.alpha[ch]);
  double sn = 0.0, so = 0.0;
  for (int j = lane; j < P; j += 32) alpha_terms(beta[j], tau, pr, sn, so);
  sn = warp_sum(sn); so = warp_sum(so);
  if (lane == 0) {
    const double a = alpha_decide(sa, pr, sn, so, P, L.h.alpha_a, L.h.alpha_b);
    L.alpha[ch] = a;
    const long long slot = stbl_slot(L);
    if (slot >= 0 && L.sink.alpha) L.sink.alpha[(size_t)ch * L.sink.chain_stride + slot] = a;
  }
}

}  // namespace bb
