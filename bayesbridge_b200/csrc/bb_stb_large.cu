// bb_stb_large.cu -- host side of the large-P stable sweep (bb_stb_large.cuh): one Gibbs iteration =
// scalar stage, build, blocked left-looking Cholesky (DMMA update / potrf / trsm per block column), solve.
#include <algorithm>
#include <cstdlib>
#include "bb_stb_large.cuh"
#include "bb_stb_wide.cuh"

namespace bb {

void stb_large_prepare(int P) {
  const size_t smem_sc = stbl_scalar_smem(P);
  const size_t smem_sv = stbl_solve_smem(P);
  const size_t smem_sy = (size_t)2 * SL_NB * SL_LD * sizeof(double);
  const size_t smem_pf = stbl_potrf_smem();
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_potrf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pf));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_scalar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sc));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_lambda_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max(stbl_lambda_smem(P), stbl_lambda_smem(std::min(P, SC_RC_MAXP)))));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_solve_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sv));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_solve_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sv));

  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sy));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_trsm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sy));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_colblock_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stbl_colblock_smem()));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_colblock_v1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stbl_colblock_smem()));
  BB_CUDA_CHECK(cudaFuncSetAttribute(wide_gram2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wide_gram2_smem()));
  BB_CUDA_CHECK(cudaFuncSetAttribute(stbl_update2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stbl_update2_smem()));
  BB_CUDA_CHECK(cudaFuncSetAttribute(wide_solve_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sv));
  BB_CUDA_CHECK(cudaFuncSetAttribute(wide_solve_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sv));   // N <= P there
  if ((size_t)P * sizeof(double) > 48 * 1024) BB_CUDA_CHECK(cudaFuncSetAttribute(wide_draw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((size_t)P * sizeof(double))));
}

// Launch with programmatic stream serialisation (the kernel calls pdl_wait() before its first dependent read and pdl_trigger()
// before its last stores): the next kernel of the chain is brought onto the SMs during this one's tail instead of after it has
// drained -- the P = 256 iteration is a chain of 11 dependent kernels of 20-50 us.  BB_STBL_PDL=0 launches the ordinary way.
template <typename... KArgs, typename... Args>
static void launch_chain(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  static const bool pdl = !(std::getenv("BB_STBL_PDL") && std::getenv("BB_STBL_PDL")[0] == '0');
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cap);
  cfg.attrs = at; cfg.numAttrs = (pdl && cap == cudaStreamCaptureStatusNone) ? 1 : 0;
  BB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, KArgs(args)...));
}

// block column kb > 0 brought up to date by the 64 x 64 kernel.  BB_STBL_TILE128=1: the row tiles below the diagonal tile in
// pairs by the 128 x 64 kernel instead -- measured SLOWER at c4 (9.8 M against 10.3 M draws/s): its 64 accumulator registers
// leave no room to keep the next panels' global loads in flight during the tile product, which the 64 x 64 kernel does.
static int launch_update(double* A, int P, int kb, int nblk, int nch, cudaStream_t st) {
  static const bool tile64 = std::getenv("BB_STBL_TILE128") == nullptr;
  const size_t smem_sy = (size_t)2 * SL_NB * SL_LD * sizeof(double);
  const int below = nblk - kb - 1;
  if (tile64 || below < 2) { stbl_update_kernel<<<dim3(nblk - kb, nch), 256, smem_sy, st>>>(A, P, kb); return 1; }
  stbl_update_kernel<<<dim3(1, nch), 256, smem_sy, st>>>(A, P, kb);
  stbl_update2_kernel<<<dim3((below + 1) / 2, nch), 256, stbl_update2_smem(), st>>>(A, P, kb, kb + 1);
  return 2;
}

// Blocked Cholesky of nch P x P systems (lower triangle of A, leading dimension P; the inverses of the diagonal blocks of the
// factor go to W), left-looking: one kernel per block column (panel solve of the previous one + update, stbl_colblock_kernel)
// and one for its diagonal block.  Round 2: measured against the round-1 sequence (update / potrf / trsm kernels per block
// column after a build pass) c5s 143 -> 177 M draws/s and c4 (P = 2000, 32 block columns) 12.1 -> 15.9 M = 60 % of the DGEMM
// peak over the whole iteration.  The round-1 sequence stays reachable for comparison: BB_STBL_FUSED_MAXBLK=<n> uses it for
// systems of more than n block columns (0: always).
static const int SL_FUSED_MAXBLK = std::getenv("BB_STBL_FUSED_MAXBLK") ? std::atoi(std::getenv("BB_STBL_FUSED_MAXBLK")) : (1 << 20);
static int launch_factor(double* A, double* W, int P, int nch, uint32_t* status, const StblSrc& X, cudaStream_t st) {
  const size_t smem_sy = (size_t)2 * SL_NB * SL_LD * sizeof(double);
  const int nblk = (P + SL_NB - 1) / SL_NB;
  int n = 0;
  if (nblk <= SL_FUSED_MAXBLK) {
    for (int kb = 0; kb < nblk; ++kb) {
      if (kb > 0) {
        static const bool v1 = std::getenv("BB_STBL_COLBLOCK_V1") != nullptr;   // the solve-first form, for comparison
        if (v1) launch_chain(stbl_colblock_v1_kernel, dim3(nblk - kb, nch), dim3(256), stbl_colblock_smem(), st, A, W, P, kb, X);
        else launch_chain(stbl_colblock_kernel, dim3(nblk - kb, nch), dim3(256), stbl_colblock_smem(), st, A, W, P, kb, X);
        ++n;
      }
      static const int early = (std::getenv("BB_STBL_EARLY") && std::getenv("BB_STBL_EARLY")[0] == '1') ? 2 : 1;   // 2: potrf releases the next colblock kernel at its start
      launch_chain(stbl_potrf_kernel, dim3(nch), dim3(SL_T), stbl_potrf_smem(), st, A, W, P, kb, status, X, early); ++n;
    }
    return n;
  }
  for (int kb = 0; kb < nblk; ++kb) {   // bring block column kb up to date, factor its diagonal block, solve its panel
    if (kb > 0) n += launch_update(A, P, kb, nblk, nch, st);
    stbl_potrf_kernel<<<nch, SL_T, stbl_potrf_smem(), st>>>(A, W, P, kb, status, StblSrc{}, 0); ++n;
    const int rem = nblk - kb - 1;
    if (rem > 0) { stbl_trsm_kernel<<<dim3(rem, nch), 256, smem_sy, st>>>(A, W, P, kb); ++n; }
  }
  return n;
}

// CTAs per chain of the lambda stage.  Splitting a chain shortens its passes but evaluates more speculative rounds; measured at c5s
// (64 chains per launch, four launches in flight) 1 / 2 / 4 CTAs per chain = 157 / 154 / 148 M draws/s: the stage is bound by the
// FP64 work, so a chain is split only when the chains alone would leave most SMs idle.  Slices of at least 64 coefficients.
static int stbl_lambda_nsplit(int P, int nch) {
  static int sms = 0;
  if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
  static const int env = std::getenv("BB_STBL_LSPLIT") ? std::atoi(std::getenv("BB_STBL_LSPLIT")) : 0;
  int ns = env > 0 ? env : sms / (4 * nch);
  ns = std::min(ns, std::max(1, P / 64));
  return std::max(1, std::min(ns, 64));
}

int launch_stb_large_iter(StbLargeLaunch& L, bool beta_stage, cudaStream_t st) {
  const int P = L.m.P, nch = L.nchains;
  const size_t smem_sc = stbl_scalar_smem(P);
  const size_t smem_sv = stbl_solve_smem(P);
  const size_t smem_sy = (size_t)2 * SL_NB * SL_LD * sizeof(double);
  const int nblk = (P + SL_NB - 1) / SL_NB;
  int n = 0;
  const bool do_alpha = (L.stages & 64) && !(L.h.true_alpha > 0) && !L.pre;
  if (L.Q != nullptr && (L.stages & 2) && !(L.h.true_sig2 > 0) && !L.pre) {
    const int nstrips = (P + QD_R - 1) / QD_R;
    stbl_quad_kernel<<<dim3(nstrips, (nch + 31) / 32), 256, 0, st>>>(L, nstrips); ++n;
  }
  launch_chain(stbl_scalar_kernel, dim3(nch), dim3(SL_T), smem_sc, st, L); ++n;
  if ((L.stages & 32) && !L.pre) {
    const int nsplit = stbl_lambda_nsplit(P, nch);
    launch_chain(stbl_lambda_kernel, dim3(nsplit, nch), dim3(SL_T), stbl_lambda_smem(stbl_lambda_slice(P, nsplit)), st, L, nsplit); ++n;
  }
  if (L.pre || !beta_stage) {
    if (do_alpha) { stbl_alpha_kernel<<<(nch + 7) / 8, 256, 0, st>>>(L); ++n; }
    return n;
  }
  if (L.wide) {
    // the same draw through the N x N system (bb_stb_wide.cuh): u and the right-hand side, M = X D^-1 X' + I on the tensor
    // cores, its blocked Cholesky with the kernels below (dimension N), both substitutions and beta = sig (u + D^-1 X'w)
    const int N = L.m.N, nbn = (N + SL_NB - 1) / SL_NB, ntile = nbn * (nbn + 1) / 2;
    wide_draw_kernel<<<nch, 256, (size_t)P * sizeof(double), st>>>(L); ++n;
    static const bool gram_ldg = std::getenv("BB_WIDE_GRAM_LDG") != nullptr;   // the LDG -> STS form of round 2's first half
    if (gram_ldg) wide_gram_kernel<<<dim3(ntile, nch), WG_T, 0, st>>>(L);
    else wide_gram2_kernel<<<dim3(ntile, nch), 256, wide_gram2_smem(), st>>>(L);
    ++n;
    n += launch_factor(L.A, L.W, N, nch, L.status, StblSrc{}, st);
    if (N <= SV_SMALL_P) wide_solve_kernel<512><<<nch, 512, stbl_solve_smem(N), st>>>(L);
    else wide_solve_kernel<1024><<<nch, 1024, stbl_solve_smem(N), st>>>(L);
    ++n;
    if (do_alpha) { stbl_alpha_kernel<<<(nch + 7) / 8, 256, 0, st>>>(L); ++n; }
    return n;
  }
  if (nblk <= SL_FUSED_MAXBLK) {        // no build pass: the kernels read X'X and add the diagonal themselves
    StblSrc X; X.src = L.m.XX; X.stride = 0; X.lam = L.lambda; X.sig2 = L.sig2; X.tau = L.tau;
    n += launch_factor(L.A, L.W, P, nch, L.status, X, st);
  } else {
    stbl_build_kernel<<<dim3(64, nch), 256, 0, st>>>(L); ++n;
    n += launch_factor(L.A, L.W, P, nch, L.status, StblSrc{}, st);
  }
  if (P <= SV_SMALL_P) launch_chain(stbl_solve_kernel<512>, dim3(nch), dim3(512), smem_sv, st, L);
  else launch_chain(stbl_solve_kernel<1024>, dim3(nch), dim3(1024), smem_sv, st, L);
  ++n;
#ifdef BB_STBL_PHASES
  {
    static int calls = 0;
    if (++calls == 40) {
      cudaStreamSynchronize(st);
      long long ph[2][8]; cudaMemcpyFromSymbol(ph, bb_stbl_ph, sizeof(ph));
      fprintf(stderr, "[stbl phases] scalar: sums %lld  tau/sig2 %lld | lambda %lld | solve: forward %lld  normals %lld  backward %lld\n",
              ph[0][1] - ph[0][0], ph[0][2] - ph[0][1], ph[0][4] - ph[0][3], ph[1][1] - ph[1][0], ph[1][2] - ph[1][1], ph[1][3] - ph[1][2]);
#ifdef BB_PHASES
      unsigned long long hc[16]; cudaMemcpyFromSymbol(hc, bb_phase_cyc, sizeof(hc));
      fprintf(stderr, "[stbl lambda stage, CTA 0 thread 0, per pass] passes %llu; setup+round %.0f walk %.0f queue+barriers %.0f loop head %.0f\n", hc[12],
              (double)hc[8] / hc[12], (double)hc[9] / hc[12], (double)hc[10] / hc[12], (double)hc[11] / hc[12]);
#endif
#ifdef BB_LAMBDA_STATS
      unsigned long long q[8]; cudaMemcpyFromSymbol(q, bb_lambda_stats, sizeof(q));
      const double in = (double)q[0], out = (double)q[1], acc = (double)q[2], ev = (double)q[3], ps = (double)q[4];
      fprintf(stderr, "[stbl lambda stats] draws %.0f: sequential rounds per draw %.3f (inner %.3f, outer %.3f); evaluated per draw %.3f; passes per chain-iteration %.2f\n",
              acc, (in + out + acc) / acc, in / acc, out / acc, ev / acc, ps / (acc / P));
#endif
    }
  }
#endif
  if (do_alpha) { stbl_alpha_kernel<<<(nch + 7) / 8, 256, 0, st>>>(L); ++n; }
  return n;
}

int stb_large_advance(long long* ctl, int stored, cudaStream_t st) {
  stbl_advance_kernel<<<1, 1, 0, st>>>(ctl, stored);
  return 1;
}

// Cholesky of ONE P x P matrix with the same kernels (a batch of one chain): on return the lower triangle of A
// holds L (A = L L'); the rest of A is scratch (the callers zero it).  status: one word,
// ST_CHOL_FAIL is set when a pivot is not positive.  Used by the Cholesky-preconditioned SVD of the constructor.
// Winv: ceil(P / 64) x 64 x 64 scratch (the inverses of all diagonal blocks of the factor are kept)
int chol_lower_device(double* A, int P, double* Winv, uint32_t* status, cudaStream_t st) {
  stb_large_prepare(P);
  const int n = launch_factor(A, Winv, P, 1, status, StblSrc{}, st);
  BB_CUDA_CHECK(cudaGetLastError());
  return n;
}

}  // namespace bb
