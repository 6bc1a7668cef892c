// Latency of warp-wide minima of doubles on B200: the transposed butterfly of bb_tri.cuh (reduce_minima<2>: 4 values,
// 7 shuffle rounds) against a two-level integer REDUX on the order-preserving image of the doubles (4 values, each
// redux.sync.min on the high word, then on the low word among the lanes that tie on the high word).
#include <cstdio>
#include <cuda_runtime.h>
#include "bb_tri.cuh"
using namespace bb;

__device__ __forceinline__ unsigned redux_min_u32(unsigned v) {
  unsigned r; asm volatile("redux.sync.min.u32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(v)); return r;
}
// order-preserving map double -> uint64 (no NaNs among the inputs)
__device__ __forceinline__ double warp_min_redux(double x) {
  long long b = __double_as_longlong(x);
  unsigned long long k = (unsigned long long)(b ^ ((b >> 63) | 0x8000000000000000LL));
  const unsigned hi = (unsigned)(k >> 32), lo = (unsigned)k;
  const unsigned mh = redux_min_u32(hi);
  const unsigned ml = redux_min_u32(hi == mh ? lo : 0xffffffffu);
  unsigned long long m = ((unsigned long long)mh << 32) | ml;
  long long r = (long long)(m ^ ((((long long)m) >> 63) ? 0x8000000000000000ULL : 0xffffffffffffffffULL));
  return __longlong_as_double(r);
}

template <int MODE> __global__ void k(double* out, long long* clk, int n) {
  const int lane = threadIdx.x & 31;
  double l[2] = {1.0 + lane * 1e-3, 2.0 - lane * 1e-3}, r[2] = {3.0 + (lane ^ 5) * 1e-3, -1.0 - (lane ^ 9) * 1e-3};
  double acc = 0.0;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < n; ++it) {
    if (MODE == 0) {
      const double keep = reduce_minima<2>(l, r, lane);
      acc += keep;
      l[0] += keep * 1e-9; l[1] -= keep * 1e-9; r[0] += keep * 1e-9; r[1] -= keep * 1e-9;   // dependent chain
    } else {
      const double a = warp_min_redux(l[0]), b = warp_min_redux(r[0]), c = warp_min_redux(l[1]), d = warp_min_redux(r[1]);
      const double keep = (lane < 8) ? a : (lane < 16 ? b : (lane < 24 ? c : d));
      acc += keep;
      l[0] += keep * 1e-9; l[1] -= keep * 1e-9; r[0] += keep * 1e-9; r[1] -= keep * 1e-9;
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
__global__ void check(double* out) {
  const int lane = threadIdx.x & 31;
  double vals[4] = {(lane * 37 % 11) - 5.0, -1e-300 * (lane + 1), (lane == 7) ? -0.0 : 0.0 + lane, 1e300 / (lane + 1)};
  for (int v = 0; v < 4; ++v) { const double a = warp_min_redux(vals[v]); const double b = warp_min(vals[v]); if (lane == 0) { out[2 * v] = a; out[2 * v + 1] = b; } }
}
int main() {
  double* o; long long* c; cudaMalloc(&o, 8 << 20); cudaMalloc(&c, 8 * 1024);
  for (int mode = 0; mode < 2; ++mode)
    for (int warps : {1, 4, 8}) {
      const int n = 20000;
      if (mode == 0) { k<0><<<1, 32 * warps>>>(o, c, n); k<0><<<1, 32 * warps>>>(o, c, n); } else { k<1><<<1, 32 * warps>>>(o, c, n); k<1><<<1, 32 * warps>>>(o, c, n); }
      long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
      printf("%-22s warps/SM %d : %6.1f cycles per 4-value warp minimum (dependent)\n", mode ? "redux (2-level u32)" : "transposed butterfly", warps, (double)h / n);
    }
  check<<<1, 32>>>(o);
  double h[8]; cudaMemcpy(h, o, sizeof(h), cudaMemcpyDeviceToHost);
  for (int v = 0; v < 4; ++v) printf("check %d: redux %.17g  shuffle %.17g %s\n", v, h[2 * v], h[2 * v + 1], h[2 * v] == h[2 * v + 1] ? "ok" : "DIFF");
  printf("status %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
