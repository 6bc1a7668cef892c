// bb_ortho.cu -- host launcher of the ortho=TRUE drivers (bb_ortho.cuh).
#include "bb_ortho.cuh"

namespace bb {

int launch_ortho(const OrthoLaunch& L, bool stable, cudaStream_t st) {
  const int P = L.m.P;
  const size_t smem = ((size_t)4 * P + 24 + 4) * sizeof(double) + ((size_t)3 * P + 2) * sizeof(int);
  if (!stable) {
    BB_CUDA_CHECK(cudaFuncSetAttribute(ortho_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ortho_kernel<false><<<L.nchains, ORT_T, smem, st>>>(L);
  } else {
    BB_CUDA_CHECK(cudaFuncSetAttribute(ortho_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ortho_kernel<true><<<L.nchains, ORT_T, smem, st>>>(L);
  }
  check_launch();
  return 1;
}

}  // namespace bb
