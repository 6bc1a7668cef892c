// DMMA shapes on B200: flops per clock per SM for mma.sync m8n8k4 / m16n8k4 / m16n8k8 / m16n8k16 (f64)
#include <cstdio>
#include <cuda_runtime.h>
template <int SHAPE> __global__ void k(double* out, long long* clk, int n) {
  double a[8], b[4], c[4][4];
  for (int i = 0; i < 8; ++i) a[i] = 1.0 + threadIdx.x * 1e-3 + i;
  for (int i = 0; i < 4; ++i) b[i] = 0.5 + i;
  for (int j = 0; j < 4; ++j) for (int i = 0; i < 4; ++i) c[j][i] = 0.0;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < n; ++it) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (SHAPE == 0) asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[j][0]), "+d"(c[j][1]) : "d"(a[0]), "d"(b[0]));
      if (SHAPE == 1) asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};" : "+d"(c[j][0]), "+d"(c[j][1]), "+d"(c[j][2]), "+d"(c[j][3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
      if (SHAPE == 2) asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};" : "+d"(c[j][0]), "+d"(c[j][1]), "+d"(c[j][2]), "+d"(c[j][3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
      if (SHAPE == 3) asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};" : "+d"(c[j][0]), "+d"(c[j][1]), "+d"(c[j][2]), "+d"(c[j][3]) : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
  double s = 0; for (int j = 0; j < 4; ++j) for (int i = 0; i < 4; ++i) s += c[j][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int SHAPE> void run(const char* name, double flops_per_mma) {
  double* o; long long* c; cudaMalloc(&o, 8 << 20); cudaMalloc(&c, 8 * 1024);
  for (int warps : {4, 8, 16, 32}) {
    const int n = 4000;
    k<SHAPE><<<1, 32 * warps>>>(o, c, n); k<SHAPE><<<1, 32 * warps>>>(o, c, n);
    long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("%-10s warps/SM %2d : %7.1f flop/clk/SM  (%.1f clk per mma per warp)\n", name, warps, flops_per_mma * 4.0 * n * warps / h, (double)h / (4.0 * n));
  }
}
int main() {
  run<0>("m8n8k4", 2.0 * 8 * 8 * 4); run<1>("m16n8k4", 2.0 * 16 * 8 * 4); run<2>("m16n8k8", 2.0 * 16 * 8 * 8); run<3>("m16n8k16", 2.0 * 16 * 8 * 16);
  printf("status %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
