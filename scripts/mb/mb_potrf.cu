// stbl_potrf_kernel alone: 64 chains x one 64 x 64 diagonal block of a P = 256 matrix, event-timed, and (built with
// -DBB_POTRF_PHASES) the clock64 marks of CTA 0: load | factor | scale | store L | W diag | W stages | store W.
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#define BB_POTRF_PHASES 1
#include "bb_stb_large.cuh"
using namespace bb;
int main() {
  const int P = 256, nch = 64;
  std::vector<double> A((size_t)nch * P * P, 0.0);
  for (int c = 0; c < nch; ++c)
    for (int i = 0; i < P; ++i)
      for (int j = 0; j < P; ++j) A[(size_t)c * P * P + i + (size_t)j * P] = (i == j ? 4.0 + 0.01 * c : 1.0 / (1.0 + abs(i - j)));
  double *dA, *dA0, *dW; uint32_t* st;
  cudaMalloc(&dA, A.size() * 8); cudaMalloc(&dA0, A.size() * 8); cudaMalloc(&dW, (size_t)nch * 4 * 64 * 64 * 8); cudaMalloc(&st, nch * 4);
  cudaMemcpy(dA0, A.data(), A.size() * 8, cudaMemcpyHostToDevice); cudaMemset(st, 0, nch * 4);
  const size_t sm = stbl_potrf_smem();
  cudaFuncSetAttribute(stbl_potrf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 5; ++rep) {
    cudaMemcpy(dA, dA0, A.size() * 8, cudaMemcpyDeviceToDevice);
    cudaEventRecord(e0);
    stbl_potrf_kernel<<<nch, SL_T, sm>>>(dA, dW, P, 0, st, StblSrc{}, 1);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long ph[12]; cudaMemcpyFromSymbol(ph, bb_potrf_ph, sizeof(ph));
    printf("rep %d: %.1f us; cycles load %lld factor %lld scale %lld storeL %lld Wdiag %lld Wstages %lld storeW %lld panel: stage %lld mma %lld store %lld  (%s)\n", rep, ms * 1e3,
           ph[1] - ph[0], ph[2] - ph[1], 0LL, ph[3] - ph[2], ph[4] - ph[3], ph[5] - ph[4], ph[6] - ph[5], ph[8] - ph[6], ph[9] - ph[8], ph[7] - ph[9], cudaGetErrorString(cudaGetLastError()));
  }
  std::vector<double> L(64 * 64); uint32_t s0; cudaMemcpy(&s0, st, 4, cudaMemcpyDeviceToHost);
  cudaMemcpy2D(L.data(), 64 * 8, dA, P * 8, 64 * 8, 64, cudaMemcpyDeviceToHost);
  double err = 0;   // L L' against the input block (chain 0)
  for (int i = 0; i < 64; ++i) for (int j = 0; j <= i; ++j) { double s = 0; for (int k = 0; k <= j; ++k) s += L[i + 64 * k] * L[j + 64 * k]; err = fmax(err, fabs(s - A[i + (size_t)j * P])); }
  std::vector<double> Wh(64 * 64); cudaMemcpy(Wh.data(), dW, 64 * 64 * 8, cudaMemcpyDeviceToHost);
  double errw = 0;   // W L against the identity
  for (int i = 0; i < 64; ++i) for (int j = 0; j < 64; ++j) { double s = 0; for (int k = 0; k < 64; ++k) s += Wh[i + 64 * k] * (k >= j ? L[k + 64 * j] : 0.0); errw = fmax(errw, fabs(s - (i == j))); }
  printf("status %u, max |LL' - A| = %.3e, max |W L - I| = %.3e\n", s0, err, errw);
  return 0;
}
