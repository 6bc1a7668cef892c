// micro-benchmarks for the lambda stage: latency and accuracy of the FP64 primitives (libm vs bb_math.cuh) and of one
// retstable round.   build: see scripts/mb/build.sh
#include <cstdio>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#include "bb_retstable.cuh"
using namespace bb;

template <int OP> __device__ __forceinline__ double op(double x) {
  if (OP == 0) return fma(x, 1.0000001, 1e-9);
  if (OP == 1) return log(x) + 3.0;
  if (OP == 2) return exp(x * 0.3);
  if (OP == 3) return sin(x) + 1.5;
  if (OP == 4) return sqrt(x) + 1.0;
  if (OP == 5) return 1.0 / x + 0.5;
  if (OP == 6) return log1p(x) + 1.0;
  if (OP == 7) return expm1(-x) + 2.0;
  if (OP == 8) return cos(x) + 1.5;
  if (OP == 11) return x_log(x) + 3.0;
  if (OP == 12) return x_exp(x * 0.3);
  if (OP == 13) return x_sinc(x) + 1.5;
  if (OP == 14) return x_sqrt(x) + 1.0;
  if (OP == 15) return x_rcp(x) + 0.5;
  if (OP == 16) return x_log1p(x) + 1.0;
  if (OP == 17) return x_expm1(-x) + 2.0;
  if (OP == 18) return x_cos2pi(x) + 1.5;
  return x;
}
template <int OP> __global__ void lat_kernel(double x0, long long* out, double* sink, int n) {
  double x = x0 + threadIdx.x * 1e-3;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) x = op<OP>(x);
  long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = x;
}
__global__ void round_kernel(double h, double alpha, long long* out, double* sink, int n, int mode) {
  RetAlpha A; A.init(alpha);
  RetConst c; c.init(h * (1.0 + 0.01 * threadIdx.x), A);
  Stream s(RngKey{1u, 2u}, blockIdx.x, 0u, ST_LAMBDA, threadIdx.x);
  double acc = 0.0; int cnt = 0;
  long long t0 = clock64();
  for (int i = 0; i < n; ++i) {
    double o;
    cnt += mode ? retstable_round_fast(A, c, s, 3u * i, o) : retstable_round_libm(A, c, s, 3u * i, o);
    acc += o == o ? o : 0.0;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + cnt;
}
// agreement of the two rounds: codes must match, accepted values to rounding
__global__ void agree_kernel(double alpha, int n, unsigned long long* counts, double* maxrel) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  RetAlpha A; A.init(alpha);
  Stream s(RngKey{7u, 9u}, 3u, 1u, ST_LAMBDA, (uint32_t)i);
  double ua, ub; s.block(K_CAP - 1, ua, ub);
  const double h = exp(-12.0 + 24.0 * ua);            // h log-uniform in [6e-6, 1.6e5]
  RetConst c; c.init(h, A);
  double worst = 0.0; unsigned long long mism = 0, acc = 0, tot = 0, inner = 0;
  for (int r = 0; r < n; ++r) {
    double o0, o1;
    const int c0 = retstable_round_libm(A, c, s, 5u * r, o0), c1 = retstable_round_fast(A, c, s, 5u * r, o1);
    ++tot; mism += c0 != c1; inner += c0 != 0;
    if (c0 == 2 && c1 == 2) { ++acc; worst = fmax(worst, fabs(o1 - o0) / fabs(o0)); }
  }
  atomicAdd(counts + 0, tot); atomicAdd(counts + 1, mism); atomicAdd(counts + 2, acc); atomicAdd(counts + 3, inner);
  atomicMax((unsigned long long*)maxrel, (unsigned long long)__double_as_longlong(worst));
}
// accuracy of the primitives against libm
__global__ void acc_kernel(int n, double* err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  Stream s(RngKey{11u, 13u}, 0u, 0u, ST_SCALAR, (uint32_t)i);
  double e[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int r = 0; r < n; ++r) {
    double ua, ub; s.block(r, ua, ub);
    const double big = exp(-700.0 + 1400.0 * ua), ang = 4.0 * ub + 1e-9, ex = -740.0 + 1449.0 * ub, sm = exp(-40.0 + 45.0 * ua);
    auto rel = [](double a, double b) { return b == 0.0 ? fabs(a) : fabs(a - b) / fabs(b); };
    e[0] = fmax(e[0], rel(x_log(big), log(big)));
    e[0] = fmax(e[0], rel(x_log(0.5 + ub), log(0.5 + ub)));
    e[1] = fmax(e[1], rel(x_exp(ex), exp(ex)));
    e[1] = fmax(e[1], rel(x_exp(ua - 0.5), exp(ua - 0.5)));
    e[2] = fmax(e[2], rel(x_sinc(ang), sin(ang) / ang));
    e[3] = fmax(e[3], rel(x_sqrt(big), sqrt(big)));
    e[3] = fmax(e[3], rel(x_rsqrt(big), 1.0 / sqrt(big)));
    e[4] = fmax(e[4], rel(x_div(ua, big), ua / big));
    e[5] = fmax(e[5], rel(x_log1p(sm), log1p(sm)));
    e[6] = fmax(e[6], rel(x_expm1(-sm), expm1(-sm)));
    e[7] = fmax(e[7], fabs(x_cos2pi(ua) - cos(6.283185307179586 * ua)));
  }
  for (int k = 0; k < 8; ++k) atomicMax((unsigned long long*)(err + k), (unsigned long long)__double_as_longlong(e[k]));
}
template <int OP> void run_lat(const char* name, double x0) {
  long long* d; double* s; cudaMalloc(&d, 8 * 1024); cudaMalloc(&s, 8 * 1024 * 1024);
  const int n = 2000;
  for (int warps : {1, 16}) {
    lat_kernel<OP><<<1, 32 * warps>>>(x0, d, s, n);
    lat_kernel<OP><<<1, 32 * warps>>>(x0, d, s, n);
    long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%-8s warps/SM %2d : %7.1f cycles per dependent call\n", name, warps, (double)h / n);
  }
  cudaFree(d); cudaFree(s);
}
int main() {
  run_lat<0>("dfma", 1.0);
  run_lat<1>("log", 2.0); run_lat<11>("x_log", 2.0);
  run_lat<2>("exp", 1.0); run_lat<12>("x_exp", 1.0);
  run_lat<3>("sin", 1.0); run_lat<13>("x_sinc", 1.0); run_lat<8>("cos", 1.0); run_lat<18>("x_cos2pi", 0.3);
  run_lat<4>("sqrt", 2.0); run_lat<14>("x_sqrt", 2.0); run_lat<5>("div", 2.0); run_lat<15>("x_rcp", 2.0);
  run_lat<6>("log1p", 0.5); run_lat<16>("x_log1p", 0.5); run_lat<7>("expm1", 0.5); run_lat<17>("x_expm1", 0.5);
  long long* d; double* s; cudaMalloc(&d, 8 * 1024); cudaMalloc(&s, 8 * 1024 * 1024);
  for (int mode = 0; mode < 2; ++mode)
    for (double h : {0.5, 5000.0})
      for (int warps : {1, 4, 8, 16}) {
        const int n = 200;
        round_kernel<<<1, 32 * warps>>>(h, 0.25, d, s, n, mode);
        round_kernel<<<1, 32 * warps>>>(h, 0.25, d, s, n, mode);
        long long hh; cudaMemcpy(&hh, d, 8, cudaMemcpyDeviceToHost);
        printf("%s h=%g warps/SM %2d : %8.1f cycles per call\n", mode ? "round_fast" : "round_libm", h, warps, (double)hh / n);
      }
  double* err; cudaMalloc(&err, 64); cudaMemset(err, 0, 64);
  acc_kernel<<<64, 128>>>(2000, err);
  double he[8]; cudaMemcpy(he, err, 64, cudaMemcpyDeviceToHost);
  const char* nm[8] = {"x_log", "x_exp", "x_sinc", "x_sqrt/rsqrt", "x_div", "x_log1p", "x_expm1", "x_cos2pi(abs)"};
  for (int k = 0; k < 8; ++k) printf("max rel err vs libm  %-14s %.3e\n", nm[k], he[k]);
  for (double alpha : {0.0625, 0.25, 0.45}) {
    unsigned long long* cnt; cudaMalloc(&cnt, 32); cudaMemset(cnt, 0, 32);
    double* mr; cudaMalloc(&mr, 8); cudaMemset(mr, 0, 8);
    agree_kernel<<<256, 128>>>(alpha, 400, cnt, mr);
    unsigned long long hc[4]; double hm; cudaMemcpy(hc, cnt, 32, cudaMemcpyDeviceToHost); cudaMemcpy(&hm, mr, 8, cudaMemcpyDeviceToHost);
    printf("alpha=%g rounds %llu code mismatches %llu accepted %llu inner-accepted %llu max rel diff of accepted draws %.3e\n", alpha, hc[0], hc[1], hc[2], hc[3], hm);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return 0;
}
