// bb_tri.cu -- host launcher of the triangle-mixture sweep kernel (bb_tri.cuh).
#include <algorithm>
#include <string>
#include "bb_tri.cuh"

namespace bb {
namespace {

constexpr size_t SMEM_BUDGET = 220 * 1024;   // of the 227 KB a CTA may opt into on sm_100

template <int J, int W, int C, int NGMAX>
void launch_cfg(const TriLaunch& L, int NG, cudaStream_t st) {
  const int P = L.m.P, LD = L.m.ldv;
  NG = std::max(1, std::min(NG, NGMAX));
  // ring depth: everything resident when {V,1/V} fits, else as deep as the budget allows (2..8 stages)
  int S;
  while (true) {
    const size_t fixed = tri_smem_bytes(P, LD, 0, NG, W, C);
    const size_t per_stage = (size_t)LD * 16 + 16;
    if (fixed + (size_t)P * per_stage <= SMEM_BUDGET) { S = P; break; }
    S = (int)std::min<size_t>(8, fixed < SMEM_BUDGET ? (SMEM_BUDGET - fixed) / per_stage : 0);
    if (S >= 2) break;
    if (NG == 1) throw std::string("tri: P too large for the on-chip sweep");
    NG = NG / 2;
  }
  const size_t smem = tri_smem_bytes(P, LD, S, NG, W, C);
  const int CH = NG * C;
  const int grid = (L.nchains + CH - 1) / CH;
  auto kern = tri_sweep_kernel<J, W, C, NGMAX>;
  BB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<grid, 32 * NG * W, smem, st>>>(L, S);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    cudaFuncAttributes fa{};
    cudaFuncGetAttributes(&fa, kern);
    throw std::string("tri_sweep_kernel<J=") + std::to_string(J) + ",W=" + std::to_string(W) + ",C=" + std::to_string(C) + "> launch failed: " +
        cudaGetErrorString(e) + " (threads " + std::to_string(32 * NG * W) + ", dyn smem " + std::to_string(smem) + ", regs " +
        std::to_string(fa.numRegs) + ", maxThreadsPerBlock " + std::to_string(fa.maxThreadsPerBlock) + ", static smem " +
        std::to_string(fa.sharedSizeBytes) + ", local " + std::to_string(fa.localSizeBytes) + ")";
  }
}

// NGMAX (groups per CTA) bounds the thread count and with it the registers a thread may use
template <int J, int W>
void launch_jw(const TriLaunch& L, int C, int NG, cudaStream_t st) {
  constexpr int NGMAX = J >= 16 ? 1 : (W >= 4 ? 2 : 4);
  if (C >= 2 && !(J >= 16 && W >= 8)) launch_cfg<J, W, 2, NGMAX>(L, NG, st);
  else launch_cfg<J, W, 1, NGMAX>(L, NG, st);
}

}  // namespace

int launch_tri(const TriLaunch& L, cudaStream_t st) {
  if (L.m.ldv > 4096) throw std::string("tri: P too large for the on-chip sweep (max 4096)");
  // chains per SM needed to make every chain resident at once; C of them share a warp group
  const int per_sm = (L.nchains + 147) / 148;
  const int C = per_sm >= 2 ? 2 : 1;
  const int NG = (per_sm + C - 1) / C;
  const int LD = L.m.ldv;
  if (LD <= 32) launch_jw<1, 1>(L, C, NG, st);
  else if (LD <= 64) launch_jw<2, 1>(L, C, NG, st);
  else if (LD <= 128) launch_jw<4, 1>(L, C, NG, st);
  else if (LD <= 256) launch_jw<8, 1>(L, C, NG, st);
  else if (LD <= 512) launch_jw<8, 2>(L, C, NG, st);
  else if (LD <= 1024) launch_jw<8, 4>(L, C, NG, st);
  else if (LD <= 2048) launch_jw<16, 4>(L, C, NG, st);
  else launch_jw<16, 8>(L, C, NG, st);
  return 1;
}

}  // namespace bb
