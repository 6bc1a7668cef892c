// bb_stb.cu -- host launcher of the on-chip stable-mixture sweep kernel (bb_stb.cuh, P <= 128).
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

}  // namespace

int launch_stb_small(const StbLaunch& L, cudaStream_t st) {
  if (L.m.P <= 64) {
    const size_t smem = stb64_smem_bytes();
    BB_CUDA_CHECK(cudaFuncSetAttribute(stb64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    stb64_kernel<<<L.nchains, S64_T, smem, st>>>(L);
    check_launch();
  } else launch_tb<8>(L, st);
  return 1;
}

}  // namespace bb
