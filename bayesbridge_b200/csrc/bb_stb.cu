// bb_stb.cu -- host launcher of the on-chip stable-mixture sweep kernel (bb_stb.cuh, P <= 128).
#include <cstdio>
#include <cstdlib>
#include "bb_stb.cuh"
#include "bb_stb64.cuh"

namespace bb {
namespace {

template <int TB>
void launch_tb(const StbLaunch& L, cudaStream_t st) {
  constexpr int PP = STB_NB * TB;
  const size_t smem = ((size_t)PP * PP + 2 * (PP + TB) + 4 * PP + 24 + 2) * sizeof(double) + (size_t)(2 * PP + 2 + PP + 2) * sizeof(int) +
                      (size_t)PP * sizeof(RetConst);
  auto kern = stb_small_kernel<TB>;
  BB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<L.nchains, STB_THREADS, smem, st>>>(L);
  check_launch();
}

template <bool TC, int NCH>
void launch64(const StbLaunch& L, cudaStream_t st) {
  const size_t smem = stb64_smem_bytes() * NCH;
  auto kern = stb64_kernel<TC, NCH>;
  BB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(L.nchains + NCH - 1) / NCH, S64_T * NCH, smem, st>>>(L);
  check_launch();
#ifdef BB_LAMBDA_STATS
  if (L.niter >= 8) {
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
    unsigned long long q[8];
    static unsigned long long qp[8] = {0};
    BB_CUDA_CHECK(cudaMemcpyFromSymbol(q, bb_lambda_stats, sizeof(q)));
    const double in = (double)(q[0] - qp[0]), out = (double)(q[1] - qp[1]), acc = (double)(q[2] - qp[2]), ev = (double)(q[3] - qp[3]), ps = (double)(q[4] - qp[4]);
    std::printf("  [lambda stats] draws %.0f: sequential rounds per draw %.3f (inner rejections %.3f, outer rejections %.3f, + the accepted one); "
                "rounds evaluated per draw (speculation included) %.3f; passes per chain-iteration %.2f\n",
                acc, (in + out + acc) / acc, in / acc, out / acc, ev / acc, ps / ((double)L.nchains / NCH * L.niter));
    for (int k = 0; k < 8; ++k) qp[k] = q[k];
  }
#endif
#ifdef BB_PHASES
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
  unsigned long long h[16];
  BB_CUDA_CHECK(cudaMemcpyFromSymbol(h, bb_phase_cyc, sizeof(h)));
  static unsigned long long prev[16] = {0};
  std::printf("  [lambda stage, CTA 0 thread 0] passes/it %.2f; cycles/it: setup+round %.0f walk %.0f queue+barriers %.0f loop head %.0f\n",
              (double)(h[12] - prev[12]) / L.niter, (double)(h[8] - prev[8]) / L.niter, (double)(h[9] - prev[9]) / L.niter,
              (double)(h[10] - prev[10]) / L.niter, (double)(h[11] - prev[11]) / L.niter);
  for (int k = 8; k < 16; ++k) prev[k] = h[k];
  const char* nm[8] = {"tau+sig2", "lambda", "tiles+normals", "factorisation", "solves", "alpha+store", "-", "loop head"};
  double tot = 0; for (int k = 0; k < 8; ++k) tot += (double)(h[k] - prev[k]);
  std::printf("  [stb64 phases, CTA 0, %d iterations]", L.niter);
  for (int k = 0; k < 8; ++k) if (h[k] != prev[k]) std::printf(" %s %.0f cyc/it (%.1f%%)", nm[k], (double)(h[k] - prev[k]) / L.niter, 100.0 * (h[k] - prev[k]) / tot);
  std::printf("\n");
  for (int k = 0; k < 8; ++k) prev[k] = h[k];
#endif
}

}  // namespace

int launch_stb_small(const StbLaunch& L, cudaStream_t st) {
  if (L.m.P <= 64) {
    static const bool scalar_path = std::getenv("BB_STB64_SCALAR") != nullptr;   // development: the pre-DMMA Cholesky
    static const int nch_env = std::getenv("BB_STB64_NCH") ? std::atoi(std::getenv("BB_STB64_NCH")) : 0;
    // two chains per CTA, two CTAs per SM: the CTAs of an SM drift apart, so one runs its latency-bound Cholesky
    // while the other fills the FP64 pipe with the lambda stage (4 chains in one CTA, in lock step: 10% slower;
    // one chain per CTA, 4 CTAs per SM: 40% slower; a forced start skew between the two CTAs changes nothing)
    const int nch = nch_env > 0 ? nch_env : (L.nchains > 2 * 148 ? 2 : 1);
    if (scalar_path) launch64<false, 4>(L, st);
    else if (nch >= 4) launch64<true, 4>(L, st);
    else if (nch >= 2) launch64<true, 2>(L, st);
    else launch64<true, 1>(L, st);
  } else launch_tb<8>(L, st);
  return 1;
}

}  // namespace bb
