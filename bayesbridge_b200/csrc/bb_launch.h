// bb_launch.h -- launch descriptors and the host-side launchers of every kernel family.
// Each family lives in its own translation unit (bb_tri.cu, bb_stb.cu, bb_stb_large.cu,
// bb_ortho.cu, bb_pre.cu) so that nvcc can build them in parallel; bb_abi.cu holds the C ABI.
#pragma once
#include "bb_common.cuh"

namespace bb {

struct TriLaunch {
  DeviceModel m;
  Hyper h;
  RngKey key;
  int nchains;        // chains in this launch
  uint32_t chain0;    // global id of local chain 0 (Philox counter word 3)
  uint32_t iter0;     // Gibbs iteration index of the first iteration of this launch
  int niter;
  int pre_tau;        // draw tau once before the first iteration (BridgeWrapper.cpp:142)
  int init_z;         // 1: z = V'beta is formed from beta (BridgeRegression.cpp:246); 0: read from state
  int stages;         // BB_ST_* mask: which conditionals run (all = the full sweep)
  // chain state, in/out: [nchains][P] and [nchains]
  double *beta, *u, *omega, *shape, *z;
  double *sig2, *tau, *alpha;
  uint32_t* status;
  DrawSink sink;
};

struct StbLaunch {
  DeviceModel m;
  Hyper h;
  RngKey key;
  int nchains;
  uint32_t chain0;
  uint32_t iter0;
  int niter;
  int pre_tau;     // draw tau once before the first iteration (BridgeWrapper.cpp:262)
  int stages;      // BB_ST_* mask
  double *beta, *lambda;          // state [nchains][P]
  double *sig2, *tau, *alpha;     // state [nchains]
  uint32_t* status;
  DrawSink sink;
};

struct OrthoLaunch {
  DeviceModel m;
  Hyper h;
  RngKey key;
  int nchains;
  uint32_t chain0, iter0;
  int niter, pre_tau, pre_sig2, stages;
  double *beta, *u, *aux, *shape;   // aux = omega (tri) | lambda (stb)
  double *sig2, *tau, *alpha;
  uint32_t* status;
  DrawSink sink;
};

struct StbLargeLaunch {
  DeviceModel m;
  Hyper h;
  RngKey key;
  int nchains;
  uint32_t chain0, iter;   // iteration index of THIS iteration
  int pre;                 // 1: only the tau pre-draw
  int stages;
  double *beta, *lambda, *sig2, *tau, *alpha;
  uint32_t* status;
  double* A;               // [nchains][P][P] column-major workspace (lower triangle holds L)
  double* W;               // [nchains][P/64][64][64] inverses of the diagonal blocks of the factor (column-major)
  double* Q;               // [nchains][ceil(P/32)] partial sums of d'X'Xd, d = beta - bc (stbl_quad_kernel), or null
  DrawSink sink;
  long long slot;          // sink slot for this iteration, -1 = do not store
  // Replayed as a CUDA graph the launch parameters are frozen, so what changes from one iteration to the next comes from
  // device memory: ctl[0] is added to `iter`, ctl[1] to `slot` (when slot >= 0); stbl_advance_kernel moves both at the end
  // of the iteration.  null: plain launches, iter / slot as given.
  const long long* ctl;
  // wide designs, opt-in (bb_stb_wide.cuh): the Gaussian draw through the N x N system; A is then [nchains][N][N],
  // W [nchains][ceil(N/64)][64][64], U [nchains][P] (the draw u ~ N(0, D^-1)), R [nchains][N] (right-hand side)
  int wide;
  double* U;
  double* R;
};
int stb_large_advance(long long* ctl, int stored, cudaStream_t st);   // ctl[0] += 1, ctl[1] += stored

// the reference's stand-alone scalar samplers (BridgeWrapper.h:230-244), bb_scalar.cu
enum ScalarKind { SC_TEXPON_LEFT = 0, SC_TEXPON_BOTH, SC_TEXPON, SC_TNORM_LEFT, SC_TNORM_BOTH, SC_TNORM, SC_RTGAMMA, SC_RETSTABLE };
int launch_scalar(int kind, int n, const double* p0, const double* p1, const double* p2, const double* p3, double* out,
                  RngKey key, uint32_t iter, uint32_t* status, cudaStream_t st);

// ---- host launchers; each returns the number of kernel launches it issued ----
int launch_tri(const TriLaunch& L, cudaStream_t st);                       // bb_tri.cu
int launch_stb_small(const StbLaunch& L, cudaStream_t st);                 // bb_stb.cu   (P <= 128)
int launch_ortho(const OrthoLaunch& L, bool stable, cudaStream_t st);      // bb_ortho.cu
// one iteration of the large-P stable sweep (P > 128); `store_slot` < 0: no draw is stored
int launch_stb_large_iter(StbLargeLaunch& L, bool beta_stage, cudaStream_t st);   // bb_stb_large.cu
void stb_large_prepare(int P);                                             // opt-in shared memory sizes, once per P
int chol_lower_device(double* A, int P, double* Winv, uint32_t* status, cudaStream_t st);   // one matrix, lower factor in place

// ---- sum.stat on the device (bb_summary.cu) ----
int summary_device(const double* draws, int P, long long chain_stride, long long slot0, int n, int nchains, double runtime_sec,
                   double* per_chain_dev, double* pooled_dev, cudaStream_t st);

// ---- running sum.stat (bb_running.cu): chunks of stored draws folded into per-series accumulators ----
size_t running_state_doubles(int nchains, int P);
size_t running_hist_words(int P);
int running_add_device(const double* draws, int P, long long chain_stride, long long slot0, int m, int nchains, long long n_before,
                       double* state, double* range, unsigned int* hist, cudaStream_t st);
int running_end_device(const double* state, const double* range, const unsigned int* hist, int P, int nchains, long long n,
                       double runtime_sec, double* per_chain_dev, double* pooled_dev, cudaStream_t st);

// ---- one-time precompute (bb_pre.cu) ----
int gram_device(const double* X, const double* y, int N, int P, double* G, cudaStream_t st);   // G = [X y]'[X y]
int gram_device_noalloc(const double* X, const double* y, int N, int P, double* G, cudaStream_t st);   // same, partial tiles in a static __device__ array
int gram_unpack(const double* G, int P, double* XX, double* Xy, double* yy_dev, cudaStream_t st);
int jacobi_device(double* W, double* V, int M, int P, cudaStream_t st);
double resid_device(const double* X, const double* y, const double* b, int N, int P, cudaStream_t st, int* launches);
int colnorm_device(const double* W, int M, int P, double* nrm, cudaStream_t st);
int zero_upper_device(double* A, int P, cudaStream_t st);
int colscale_device(double* W, int M, int P, const double* nrm, cudaStream_t st);
int derive_device(const double* V, const int* perm, const double* dsrc, const double* Xy, int P, int ldv, double* tV, double* Vc,
                  double* VQ, double* d, double* a, double* mean, double* d2, double* zc, cudaStream_t st);
int derive_from_tv_device(const double* tV, const double* d, const double* a, int P, int ldv, double* Vc, double* VQ, double* mean,
                          double* d2, double* zc, cudaStream_t st);
int vz_device(const double* Vc, const double* z, int P, int ldv, double* out, cudaStream_t st);
int grad_device(const double* XX, const double* Xy, const double* b, int P, double* g, cudaStream_t st);

// Device memory comes from the driver's stream-ordered pool (cudaMallocAsync) with the release threshold
// lifted, so the buffers of one batched call are recycled by the next instead of paying cudaMalloc/cudaFree
// (which cost ~15 ms of a 50 ms end-to-end call).  Every free in this library happens after the work that
// used the buffer has been synchronised, so ordering frees on the per-thread stream is safe.
inline void pool_setup_once() {
  static bool done[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done[dev]) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  done[dev] = true;
}
template <typename T>
inline T* dalloc(size_t n) {
  pool_setup_once();
  T* p = nullptr;
  BB_CUDA_CHECK(cudaMallocAsync((void**)&p, (n > 0 ? n : 1) * sizeof(T), cudaStreamPerThread));
  BB_CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
  return p;
}
template <typename T>
inline void dfree(T*& p) { if (p) cudaFreeAsync((void*)p, cudaStreamPerThread); p = nullptr; }
inline void check_launch() { BB_CUDA_CHECK(cudaGetLastError()); }

}  // namespace bb
