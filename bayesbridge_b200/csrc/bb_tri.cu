// bb_tri.cu -- host launcher of the triangle-mixture sweep kernel (bb_tri.cuh).
#include <algorithm>
#include <string>
#include "bb_tri.cuh"

namespace bb {
namespace {

template <int J, int C>
void launch_tri_jc(const TriLaunch& L, int W, cudaStream_t st) {
  const int threads = 32 * (W + C);
  const size_t smem = ((size_t)2 * C * L.m.P + 2 * C * 32 + 4 * C) * sizeof(double);
  const int grid = (L.nchains + C - 1) / C;
  if (threads <= 384) {
    auto kern = tri_sweep_kernel<J, C, 384>;
    BB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, threads, smem, st>>>(L);
  } else {
    auto kern = tri_sweep_kernel<J, C, 1024>;
    BB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, threads, smem, st>>>(L);
  }
  check_launch();
}
template <int J>
void launch_tri_j(const TriLaunch& L, int W, int C, cudaStream_t st) {
  if (C >= 4) launch_tri_jc<J, 4>(L, W, st);
  else if (C >= 2) launch_tri_jc<J, 2>(L, W, st);
  else launch_tri_jc<J, 1>(L, W, st);
}

}  // namespace

int launch_tri(const TriLaunch& L, cudaStream_t st) {
  const int P = L.m.P;
  if (P > 8 * 32 * 24) throw std::string("tri: P too large for the on-chip sweep (max 6144)");
  // compute warps W and coordinates per lane J with 32*W*J >= P
  int W = std::min(8, (P + 31) / 32);
  int J = (P + 32 * W - 1) / (32 * W);
  while (J > 8) { W += 4; J = (P + 32 * W - 1) / (32 * W); }
  const int Jt = J <= 1 ? 1 : J <= 2 ? 2 : J <= 4 ? 4 : 8;
  // chains per CTA: fill the 148 SMs first, then stack chains (they share each column of V in registers)
  int C = 1;
  if (L.nchains > 148) C = 2;
  if (L.nchains > 2 * 148) C = 4;
  if (Jt == 8 && C > 2) C = 2;   // register budget
  switch (Jt) {
    case 1: launch_tri_j<1>(L, W, C, st); break;
    case 2: launch_tri_j<2>(L, W, C, st); break;
    case 4: launch_tri_j<4>(L, W, C, st); break;
    default: launch_tri_j<8>(L, W, C, st); break;
  }
  return 1;
}

}  // namespace bb
