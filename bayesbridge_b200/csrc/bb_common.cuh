// bb_common.cuh -- shared device structs and small helpers.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "bb_rng.cuh"

namespace bb {

#define BB_CUDA_CHECK(expr)                                                                    \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      std::fprintf(stderr, "bayesbridge_b200: CUDA error %s at %s:%d: %s\n", cudaGetErrorName(_e), \
                   __FILE__, __LINE__, cudaGetErrorString(_e));                               \
      throw ::bb::CudaError(_e);                                                               \
    }                                                                                          \
  } while (0)

struct CudaError { cudaError_t code; explicit CudaError(cudaError_t c) : code(c) {} };

// Chain-independent factors resident in HBM (the role of the BridgeRegression constructor,
// /root/reference/Code/C/BridgeRegression.cpp:13-69).  All column-major doubles.
struct DeviceModel {
  int N, P;
  int ldv;             // leading dimension of Vc / VQ (P rounded up to a multiple of 32)
  const double* X;     // N x P   (may be null when built from a Gram matrix only)
  const double* y;     // N
  const double* XX;    // P x P   X'X                         (:24)
  const double* Xy;    // P       X'y                         (:25)
  const double* bhat;  // P       least squares, 0 if singular (:41-42,:79-91)
  const double* tV;    // P x P   tV[i + j*P] = V(j,i)        (:47-51)  (used for z = V'beta)
  const double* Vc;    // P x ldv Vc[i*ldv + j] = V(j,i), zero padded  (sweep column i is contiguous)
  const double* VQ;    // P x (ldv+2) x 2: column i = {V(j,i), 1/V(j,i)} for j < ldv (1/V = +-inf where V is 0 or padded), then {d_i, mean_i}, {1/d_i, 0}
  const double* d;     // P       singular values (0 beyond the rank)
  const double* mean;  // P       a_i / d_i^2                  (:268)
  const double* zc;    // P       centre of the rss expansion in z-space (a_i/d_i^2 where d_i>0, else 0)
  const double* d2;    // P       d_i^2
  const double* gc;    // P       X'(y - X bc)  for the beta-space rss expansion (stable path)
  const double* bc;    // P       centre of the beta-space rss expansion (bhat, or 0 when singular)
  double r0z;          // || y - X V zc ||^2
  double r0b;          // || y - X bc ||^2
  double yy;
};

// Prior / switch block shared by all drivers (BridgeWrapper.cpp:80-109, 207-232).
struct Hyper {
  double sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b;
  double true_sig2, true_tau, true_alpha;  // > 0 means "known, keep fixed"
  int betaburn;
};

// Where one launch writes its draws.  Sample s of chain c, coordinate j goes to
// ptr[(c * chain_stride + (s0 + s)) * P + j]  (P fastest, like the .C arrays: BridgeWrapper.cpp:602-608).
struct DrawSink {
  double* beta; double* u; double* omega; double* shape; double* lambda;  // P-vectors per sample (any may be null)
  double* sig2; double* tau; double* alpha;                               // scalars per sample
  long long chain_stride;  // samples reserved per chain
  long long s0;            // first sample slot written by this launch
  int thin;                // store every thin-th iteration
  // running per-coordinate moments over the stored iterations (summary mode): [chain][P] each, may be null
  double* sum; double* sumsq;
};

// ---- mbarrier / TMA bulk-copy wrappers (PTX) ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA, 1-D); bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- named barriers (bar.sync / bar.arrive), ids 1..15; id 0 is __syncthreads ----
__device__ __forceinline__ void bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// FP64 tensor-core MMA (DMMA): D(8x8) += A(8x4, row) B(4x8, col).  Lane l holds A[l/4][l%4], B[l%4][l/4],
// C[l/4][2(l%4) + {0,1}].
__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Programmatic dependent launch (sm_90+): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may be
// scheduled while its predecessor in the stream is still running; it must call pdl_wait() before it touches anything the
// predecessor wrote (the wait returns when the predecessor grid has completed and its writes are visible).  pdl_trigger()
// tells the scheduler that the NEXT kernel of the stream may be brought onto the SMs now.  Both are no-ops in a kernel that was
// launched the ordinary way.
#ifdef BB_PDL_EARLY   // experiment: let the next kernel onto the SMs as soon as this one has started
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { }
#else
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

}  // namespace bb
