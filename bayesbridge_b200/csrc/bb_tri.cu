// bb_tri.cu -- host dispatcher of the triangle-mixture sweep kernel (bb_tri.cuh): picks the number of
// warps per chain group (W), coordinates per lane (J), chains per group (C) and groups per CTA (NG).
#include <cstdlib>
#include <string>
#include "bb_launch.h"

namespace bb {

int launch_tri_w1(int J, const TriLaunch& L, int C, int NG, cudaStream_t st);
int launch_tri_w2(int J, const TriLaunch& L, int C, int NG, cudaStream_t st);
int launch_tri_w4(int J, const TriLaunch& L, int C, int NG, cudaStream_t st);
int launch_tri_w8(int J, const TriLaunch& L, int C, int NG, cudaStream_t st);

static int env_int(const char* name, int dflt) {
  const char* v = std::getenv(name);
  return (v && *v) ? std::atoi(v) : dflt;
}

int launch_tri(const TriLaunch& L, cudaStream_t st) {
  if (L.m.ldv > 4096) throw std::string("tri: P too large for the on-chip sweep (max 4096)");
  // chains per SM needed to make every chain resident at once; C of them share a warp group
  const int per_sm = (L.nchains + 147) / 148;
  // (two chains per SM run faster as two groups of one chain than as one group of two: c5 239 M -> 260 M draws/s)
  int C = per_sm >= 3 ? 2 : 1;
  // tuning overrides (development only): BB_TRI_C, BB_TRI_NG, BB_TRI_W
  C = env_int("BB_TRI_C", C);
  const int NG = env_int("BB_TRI_NG", (per_sm + C - 1) / C);
  const int Wo = env_int("BB_TRI_W", 0);
  const int LD = L.m.ldv;
  // warps per chain group: as many as keep about 8 warps (2 per SM sub-partition) on an SM, so that a
  // lightly loaded GPU (few chains) spends its idle warps on shortening each chain's step
  const int groups = (per_sm + C - 1) / C;
  int W = 1;
  while (W < 8 && groups * (W * 2) <= 8 && 32 * (W * 2) <= LD) W *= 2;
  while (32 * W * 16 < LD) W *= 2;          // at most 16 coordinates per lane
  if (Wo > 0) W = Wo;
  const int J = (LD + 32 * W - 1) / (32 * W);
  switch (W) {
    case 1: return launch_tri_w1(J, L, C, NG, st);
    case 2: return launch_tri_w2(J, L, C, NG, st);
    case 4: return launch_tri_w4(J, L, C, NG, st);
    case 8: return launch_tri_w8(J, L, C, NG, st);
    default: break;
  }
  throw std::string("tri: no kernel configuration for this P");
}

}  // namespace bb
