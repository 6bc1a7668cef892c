// bb_abi.cu -- host side of libbayesbridge_b200.so: device buffers, kernel dispatch, the
// BridgeRegression-shaped sampler objects and the extern "C" boundary (include/bayesbridge_b200.h).
//
// Reference interfaces replaced (paths under /root/reference/Code/C/):
//   BridgeWrapper.h:164-226 / BridgeWrapper.cpp:572-732  -> bridge_regression, bridge_reg_stable
//   BridgeWrapper.cpp:80-537 (the four Gibbs drivers)     -> run_driver()
//   BridgeRegression.h:81-166 (class BridgeRegression)    -> struct bb_model + struct bb_chains
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/bayesbridge_b200.h"
#include "bb_launch.h"

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};
unsigned long long g_seed = 20121027ULL;
std::atomic<unsigned> g_scalar_calls{0};   // Philox `iter` word of the stand-alone scalar samplers: one per call since bb_set_seed
int g_device = 0;
int g_verbose = 0;
#ifndef BB_QUAD_MINP
#define BB_QUAD_MINP 512   // from here on d'X'Xd of the large-P stable path comes from stbl_quad_kernel (all chains of a sub-batch at once)
#endif
std::atomic<unsigned> g_scalar_status{0};  // status word (bb::Status bits) of the last stand-alone sampler call
int g_wide_default = 0;                    // bb_set_wide_gaussian: the drivers opt their chains into BB_OPT_WIDE_GAUSSIAN where it applies
int (*g_interrupt)(void) = nullptr;        // bb_set_interrupt: polled between chunks of the sampling loop on the calling thread

void set_err(const std::string& s) { g_err = s; }

using bb::dalloc;
using bb::dfree;
using bb::check_launch;

inline void launched(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// scratch device buffer that is released on every exit path (an exception between dalloc and dfree used to leak it)
template <typename T>
struct DBuf {
  T* p = nullptr;
  explicit DBuf(size_t n) : p(bb::dalloc<T>(n)) {}
  ~DBuf() { bb::dfree(p); }
  DBuf(const DBuf&) = delete;
  DBuf& operator=(const DBuf&) = delete;
  operator T*() const { return p; }
};

}  // namespace

// ===========================================================================================
// bb_model : chain-independent factors (BridgeRegression.cpp:13-69)
// ===========================================================================================
struct bb_model {
  int device = 0, N = 0, P = 0, ldv = 0;
  bool own_X = false;
  double *X = nullptr, *y = nullptr;   // may be null (Gram-only construction)
  double *XX = nullptr, *Xy = nullptr, *bhat = nullptr, *tV = nullptr, *Vc = nullptr, *VQ = nullptr;
  double *d = nullptr, *a = nullptr, *mean = nullptr, *zc = nullptr, *d2 = nullptr, *gc = nullptr, *bc = nullptr, *bz = nullptr;
  double r0z = 0, r0b = 0, yy = 0;
  bool singular = false;
  bool lazy_svd = false;   // P > N: the SVD factors (triangle sampler only) are built on first use, see ensure_svd
  std::mutex svd_mu;       // ensure_svd may be reached from several threads creating chains on one model
  int zeroed_by_gram_cut = 0;   // singular values dropped by the 1e-7 d_max cut of the Gram-eigen path although above eps d_max
  double precompute_ms = 0;
  cudaStream_t stream = nullptr;

  bb::DeviceModel view() const {
    bb::DeviceModel m;
    m.N = N; m.P = P; m.ldv = ldv; m.X = X; m.y = y; m.XX = XX; m.Xy = Xy; m.bhat = bhat; m.tV = tV; m.Vc = Vc; m.VQ = VQ;
    m.d = d; m.mean = mean; m.zc = zc; m.d2 = d2; m.gc = gc; m.bc = bc; m.r0z = r0z; m.r0b = r0b; m.yy = yy;
    return m;
  }
  ~bb_model() {
    cudaSetDevice(device);
    if (own_X) { dfree(X); dfree(y); }
    dfree(XX); dfree(Xy); dfree(bhat); dfree(tV); dfree(Vc); dfree(VQ); dfree(d); dfree(a); dfree(mean); dfree(zc); dfree(d2);
    dfree(gc); dfree(bc); dfree(bz);
    if (stream) cudaStreamDestroy(stream);
  }
};

namespace {

void model_alloc(bb_model* m) {
  const size_t P = m->P;
  m->ldv = (int)((P + 31) / 32 * 32);
  m->XX = dalloc<double>(P * P); m->Xy = dalloc<double>(P); m->bhat = dalloc<double>(P);
  m->tV = dalloc<double>(P * P); m->Vc = dalloc<double>(P * m->ldv); m->VQ = dalloc<double>(2 * P * (m->ldv + 2));
  m->d = dalloc<double>(P); m->a = dalloc<double>(P); m->mean = dalloc<double>(P); m->zc = dalloc<double>(P); m->d2 = dalloc<double>(P);
  m->gc = dalloc<double>(P); m->bc = dalloc<double>(P); m->bz = dalloc<double>(P);
}

// After tV/Vc/VQ/d/a/mean/zc/d2, XX, Xy, bhat are in place: expansion centres and residuals.
void model_finish(bb_model* m, bool have_rows, bool stable_only = false) {
  const int P = m->P; cudaStream_t st = m->stream;
  // z-space centre zc -> beta-space bz = V zc   (triangle sampler)
  if (!stable_only) launched(bb::vz_device(m->Vc, m->zc, P, m->ldv, m->bz, st));
  // beta-space centre for the stable path: bhat (0 if singular)
  BB_CUDA_CHECK(cudaMemcpyAsync(m->bc, m->bhat, P * sizeof(double), cudaMemcpyDeviceToDevice, st));
  launched(bb::grad_device(m->XX, m->Xy, m->bc, P, m->gc, st));
  if (have_rows) {
    int nl = 0;
    if (!stable_only) m->r0z = bb::resid_device(m->X, m->y, m->bz, m->N, P, st, &nl);
    m->r0b = bb::resid_device(m->X, m->y, m->bc, m->N, P, st, &nl);
    launched(nl);
  }
}

// singular values / vectors from the Jacobi result; sorts descending; fills every derived array.
void model_factor(bb_model* m, double* W, int M, bool w_is_gram) {
  const int P = m->P; cudaStream_t st = m->stream;
  DBuf<double> V((size_t)P * P);
  launched(bb::jacobi_device(W, V, M, P, st));
  DBuf<double> nrm(P);
  launched(bb::colnorm_device(W, M, P, nrm, st));
  std::vector<double> h(P);
  BB_CUDA_CHECK(cudaMemcpyAsync(h.data(), nrm, P * sizeof(double), cudaMemcpyDeviceToHost, st));
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
  // W = X V   : column norm = d_i ;   W = (X'X) V : column norm = d_i^2
  double dmax = 0;
  for (int i = 0; i < P; ++i) { if (w_is_gram) h[i] = std::sqrt(h[i]); dmax = std::max(dmax, h[i]); }
  const int rank_cap = std::min(m->N, P);
  std::vector<int> perm(P);
  for (int i = 0; i < P; ++i) perm[i] = i;
  std::stable_sort(perm.begin(), perm.end(), [&](int x, int y) { return h[x] > h[y]; });
  // numerically zero directions are exactly zero (the reference takes the flat branch when d <= 1e-16, :272-281).
  // On the Gram-eigen path singular values are only resolved to sqrt(eps) d_max: everything below 1e-7 d_max is cut.  For
  // cond(X) >= 1e7 the reference (SVD of X itself) would still draw tnorm(a_i/d_i^2, sig/d_i) in those directions; the
  // model records how many such values were cut so that the caller is told (stderr + bb_model_status) instead of
  // diverging silently -- build the model from the rows (bb_model_create, N >= P) to keep them.
  const double tol = w_is_gram ? 1e-7 * dmax : 1e-13 * dmax;
  double dmin = dmax;
  int cut = 0;
  for (int i = 0; i < P; ++i) {
    const bool beyond_rank = i >= rank_cap;
    if (beyond_rank || h[perm[i]] <= tol) {
      if (w_is_gram && !beyond_rank && h[perm[i]] > 1e-13 * dmax) ++cut;
      h[perm[i]] = 0.0;
    }
    dmin = std::min(dmin, h[perm[i]]);
  }
  m->zeroed_by_gram_cut = cut;
  if (cut > 0)
    std::fprintf(stderr, "bayesbridge_b200: warning: %d singular value(s) between 1e-13 and 1e-7 of the largest were set to zero "
                         "(factors taken from X'X, cond(X) >= 1e7); the reference's SVD of X would keep them.\n", cut);
  m->singular = !(dmin > 0.0);
  DBuf<int> d_perm(P);
  BB_CUDA_CHECK(cudaMemcpyAsync(d_perm, perm.data(), P * sizeof(int), cudaMemcpyHostToDevice, st));
  BB_CUDA_CHECK(cudaMemcpyAsync(nrm, h.data(), P * sizeof(double), cudaMemcpyHostToDevice, st));
  launched(bb::derive_device(V, d_perm, nrm, m->Xy, P, m->ldv, m->tV, m->Vc, m->VQ, m->d, m->a, m->mean, m->d2, m->zc, st));
  // least squares start: V (a / d^2), or 0 when X'X is singular (BridgeRegression.cpp:79-91)
  if (m->singular) BB_CUDA_CHECK(cudaMemsetAsync(m->bhat, 0, P * sizeof(double), st));
  else launched(bb::vz_device(m->Vc, m->zc, P, m->ldv, m->bhat, st));
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
}

// Tall designs (N >= 4 P): the factors from the Cholesky factor of the Gram matrix instead of from X itself.
// X = Q L' with L = chol(X'X), so the right singular vectors of X are the LEFT singular vectors of L and the singular
// values coincide: one-sided Jacobi on the P x P matrix L (Veselic-Hari: rotating the columns of the Cholesky factor
// converges in fewer sweeps than rotating X, and no V is accumulated -- the converged columns, normalised, ARE the
// vectors).  Each round touches 8 P^2 bytes instead of 8 (N + P) P: 11x less at c3, 390x less at c5.
// Forming X'X squares the condition number, so the result is accepted only when the factorisation succeeds and the
// singular values it delivers span less than 1e3 (relative error of d about kappa^2 eps <= 1e-10); otherwise the
// caller falls back to the rotation of X.  BB_SVD_DIRECT=1 forces the fall-back.
bool model_factor_chol(bb_model* m) {
  const int P = m->P; cudaStream_t st = m->stream;
  const char* env = std::getenv("BB_SVD_DIRECT");
  if (env && env[0] == '1') return false;
  DBuf<double> W((size_t)P * P);
  DBuf<double> Winv((size_t)((P + 63) / 64) * 64 * 64);
  DBuf<uint32_t> d_status(1);
  DBuf<double> nrm(P);
  DBuf<int> d_perm(P);
  bool ok = false;
  BB_CUDA_CHECK(cudaMemcpyAsync(W, m->XX, (size_t)P * P * sizeof(double), cudaMemcpyDeviceToDevice, st));
  BB_CUDA_CHECK(cudaMemsetAsync(d_status, 0, sizeof(uint32_t), st));
  launched(bb::chol_lower_device(W, P, Winv, d_status, st));
  uint32_t status = 0;
  BB_CUDA_CHECK(cudaMemcpyAsync(&status, d_status, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
  if (status == 0) {
    launched(bb::zero_upper_device(W, P, st));
    launched(bb::jacobi_device(W, nullptr, P, P, st));
    launched(bb::colnorm_device(W, P, P, nrm, st));
    std::vector<double> h(P);
    BB_CUDA_CHECK(cudaMemcpyAsync(h.data(), nrm, P * sizeof(double), cudaMemcpyDeviceToHost, st));
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
    double dmax = 0, dmin = 1e300;
    bool finite = true;
    for (int i = 0; i < P; ++i) { finite = finite && std::isfinite(h[i]); dmax = std::max(dmax, h[i]); dmin = std::min(dmin, h[i]); }
    if (finite && dmin > 1e-3 * dmax) {
      std::vector<int> perm(P);
      for (int i = 0; i < P; ++i) perm[i] = i;
      std::stable_sort(perm.begin(), perm.end(), [&](int x, int y) { return h[x] > h[y]; });
      launched(bb::colscale_device(W, P, P, nrm, st));
      BB_CUDA_CHECK(cudaMemcpyAsync(d_perm, perm.data(), P * sizeof(int), cudaMemcpyHostToDevice, st));
      launched(bb::derive_device(W, d_perm, nrm, m->Xy, P, m->ldv, m->tV, m->Vc, m->VQ, m->d, m->a, m->mean, m->d2, m->zc, st));
      m->singular = false;
      launched(bb::vz_device(m->Vc, m->zc, P, m->ldv, m->bhat, st));   // least squares: V (a / d^2)
      BB_CUDA_CHECK(cudaStreamSynchronize(st));
      ok = true;
    }
  }
  return ok;
}

bb_model* model_new(int N, int P, int device) {
  if (N < 1 || P < 1) { set_err("bb_model: N and P must be positive"); return nullptr; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { set_err("bb_model: no CUDA device (there is no CPU fallback)"); return nullptr; }
  if (device < 0 || device >= ndev) { set_err("bb_model: bad device index"); return nullptr; }
  BB_CUDA_CHECK(cudaSetDevice(device));
  bb_model* m = new bb_model();
  m->device = device; m->N = N; m->P = P;
  BB_CUDA_CHECK(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
  model_alloc(m);
  return m;
}

void model_take_rows(bb_model* m, const double* X, const double* y, int mem) {
  const size_t N = m->N, P = m->P;
  if (mem == BB_MEM_DEVICE) { m->X = const_cast<double*>(X); m->y = const_cast<double*>(y); m->own_X = false; return; }
  m->X = dalloc<double>(N * P); m->y = dalloc<double>(N); m->own_X = true;
  BB_CUDA_CHECK(cudaMemcpyAsync(m->X, X, N * P * sizeof(double), cudaMemcpyHostToDevice, m->stream));
  BB_CUDA_CHECK(cudaMemcpyAsync(m->y, y, N * sizeof(double), cudaMemcpyHostToDevice, m->stream));
}

void model_gram(bb_model* m) {
  const int P = m->P;
  DBuf<double> G((size_t)(P + 1) * (P + 1));
  launched(bb::gram_device(m->X, m->y, m->N, P, G, m->stream));
  DBuf<double> d_yy(1);
  launched(bb::gram_unpack(G, P, m->XX, m->Xy, d_yy, m->stream));
  BB_CUDA_CHECK(cudaMemcpyAsync(&m->yy, d_yy, sizeof(double), cudaMemcpyDeviceToHost, m->stream));
  BB_CUDA_CHECK(cudaStreamSynchronize(m->stream));
}

}  // namespace

// ===========================================================================================
// bb_chains : per-chain state in HBM + the draw buffer
// ===========================================================================================
struct bb_chains {
  const bb_model* m = nullptr;
  int device = 0;   // copy: the model may be freed before its chains
  int method = BB_TRI, ortho = 0, nchains = 0;
  unsigned long long seed = 0;
  unsigned int chain0 = 0;
  bool z_valid = false;
  double *beta = nullptr, *u = nullptr, *aux = nullptr, *shape = nullptr, *z = nullptr;
  double *sig2 = nullptr, *tau = nullptr, *alpha = nullptr;
  uint32_t* status = nullptr;
  double *A = nullptr, *W = nullptr, *Q = nullptr;   // large-P stable workspace: factors, inverse diagonal blocks, partial sums of d'X'Xd
  // running sum.stat (bb_running.cu): accumulators of one vector output, n draws per chain folded so far
  double *rs_state = nullptr, *rs_range = nullptr; unsigned int* rs_hist = nullptr; long long rs_n = 0; int rs_which = 0;
  int wide = 0;                                      // BB_OPT_WIDE_GAUSSIAN: the Gaussian draw through the N x N system (bb_stb_wide.cuh)
  double *U = nullptr, *R = nullptr;                 // its workspace: u [nchains][P], right-hand side [nchains][N]
  // draw buffer
  long long nslots = 0; int out_mask = 0;
  double *o_beta = nullptr, *o_u = nullptr, *o_omega = nullptr, *o_shape = nullptr, *o_lambda = nullptr;
  double *o_sig2 = nullptr, *o_tau = nullptr, *o_alpha = nullptr;
  // large-P stable path: one Gibbs iteration of a sub-batch (about 50 small launches at P = 256) captured once as a CUDA graph
  // and replayed; gkey = the frozen launch parameters the graph was captured with, ctl = [sub-batch][2] iteration / slot offsets
  cudaGraphExec_t gx[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  std::vector<char> gkey[8];
  int glaunches[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long* ctl = nullptr;
  cudaStream_t stream = nullptr, stream2 = nullptr, xs[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // xs: the other sub-batches of the large-P stable path
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_fork = nullptr, ev_join = nullptr, xj[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

  void free_draws() { dfree(o_beta); dfree(o_u); dfree(o_omega); dfree(o_shape); dfree(o_lambda); dfree(o_sig2); dfree(o_tau); dfree(o_alpha); nslots = 0; out_mask = 0; }
  ~bb_chains() {
    cudaSetDevice(device);
    dfree(beta); dfree(u); dfree(aux); dfree(shape); dfree(z); dfree(sig2); dfree(tau); dfree(alpha); dfree(status); dfree(A); dfree(W); dfree(Q); dfree(U); dfree(R); dfree(rs_state); dfree(rs_range); dfree(rs_hist);
    free_draws();
    for (int i = 0; i < 8; ++i) if (gx[i]) cudaGraphExecDestroy(gx[i]);
    dfree(ctl);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    if (ev_fork) cudaEventDestroy(ev_fork);
    if (ev_join) cudaEventDestroy(ev_join);
    if (stream2) cudaStreamDestroy(stream2);
    for (int i = 0; i < 7; ++i) { if (xj[i]) cudaEventDestroy(xj[i]); if (xs[i]) cudaStreamDestroy(xs[i]); }
    if (stream) cudaStreamDestroy(stream);
  }
};

namespace {

bb::Hyper to_hyper(const bb_hyper* h) {
  bb::Hyper o;
  o.sig2_shape = h->sig2_shape; o.sig2_scale = h->sig2_scale; o.nu_shape = h->nu_shape; o.nu_rate = h->nu_rate;
  o.alpha_a = h->alpha_a; o.alpha_b = h->alpha_b; o.true_sig2 = h->true_sig2; o.true_tau = h->true_tau; o.true_alpha = h->true_alpha;
  o.betaburn = h->betaburn;
  return o;
}

bb::DrawSink make_sink(const bb_chains* c, int thin, long long slot0) {
  bb::DrawSink s;
  s.beta = c->o_beta; s.u = c->o_u; s.omega = c->o_omega; s.shape = c->o_shape; s.lambda = c->o_lambda;
  s.sig2 = c->o_sig2; s.tau = c->o_tau; s.alpha = c->o_alpha;
  s.chain_stride = c->nslots; s.s0 = slot0; s.thin = (c->nslots > 0) ? thin : 0; s.sum = nullptr; s.sumsq = nullptr;
  return s;
}

__global__ void fill_kernel(double* p, size_t n, double v) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) p[e] = v;
}
__global__ void bcast_rows_kernel(double* dst, const double* src, int P, int nchains) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < (size_t)P * nchains; e += (size_t)gridDim.x * blockDim.x) dst[e] = src[e % P];
}
// copy the current state of every chain into slot `slot` of the draw buffer
__global__ void store_state_kernel(const double* src, double* dst, int width, long long stride, long long slot, int nchains) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < (size_t)width * nchains; e += (size_t)gridDim.x * blockDim.x) {
    const size_t c = e / width, j = e % width;
    dst[(c * stride + slot) * width + j] = src[e];
  }
}
// gather a slot range of the draw buffer into a dense [chain][nslots][width] array
__global__ void gather_kernel(const double* src, double* dst, int width, long long stride, long long slot0, long long ns, int nchains) {
  const size_t per = (size_t)ns * width;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < per * nchains; e += (size_t)gridDim.x * blockDim.x) {
    const size_t c = e / per, r = e % per;
    dst[e] = src[(c * stride + slot0) * width + r];
  }
}

// FP64 vector-pipe peak: 16 independent DFMA chains per thread, all SMs full (bb_measure_fp64_fma)
__global__ void __launch_bounds__(256) dfma_peak_kernel(double* out, int iters, double a, double b) {
  double x[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) x[k] = 1e-3 * (threadIdx.x + k);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 16; ++k) x[k] = fma(x[k], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 16; ++k) s += x[k];
  if (s == 123.456) out[0] = s;   // keeps the chains alive
}

void fill(double* p, size_t n, double v, cudaStream_t st) { fill_kernel<<<296, 256, 0, st>>>(p, n, v); check_launch(); launched(); }

}  // namespace

// ===========================================================================================
// run: advance all chains (the body of the four Gibbs drivers, BridgeWrapper.cpp:80-537)
// ===========================================================================================
namespace {

void chains_run(bb_chains* c, const bb_hyper* hh, unsigned iter0, int niter, int stages, int pre_tau, int pre_sig2,
                int thin, long long slot0, double* kernel_ms) {
  const bb_model* m = c->m;
  BB_CUDA_CHECK(cudaSetDevice(m->device));
  const bb::Hyper h = to_hyper(hh);
  const bb::RngKey key{(uint32_t)c->seed, (uint32_t)(c->seed >> 32)};
  cudaStream_t st = c->stream;
  if (c->nslots > 0 && thin > 0) {
    const long long last = slot0 + niter / thin - 1;
    if (slot0 < 0 || last >= c->nslots) throw std::string("bb_chains_run: draw buffer too small for slot0 + niter/thin");
  }
  BB_CUDA_CHECK(cudaEventRecord(c->ev0, st));
  const int P = m->P;
  if (c->ortho) {
    bb::OrthoLaunch L;
    L.m = m->view(); L.h = h; L.key = key; L.nchains = c->nchains; L.chain0 = c->chain0; L.iter0 = iter0; L.niter = niter;
    L.pre_tau = pre_tau; L.pre_sig2 = pre_sig2; L.stages = stages;
    L.beta = c->beta; L.u = c->u; L.aux = c->aux; L.shape = c->shape; L.sig2 = c->sig2; L.tau = c->tau; L.alpha = c->alpha; L.status = c->status;
    L.sink = make_sink(c, thin, slot0);
    launched(bb::launch_ortho(L, c->method != BB_TRI, st));
  } else if (c->method == BB_TRI) {
    bb::TriLaunch L;
    L.m = m->view(); L.h = h; L.key = key; L.nchains = c->nchains; L.chain0 = c->chain0; L.iter0 = iter0; L.niter = niter;
    L.pre_tau = pre_tau; L.init_z = c->z_valid ? 0 : 1; L.stages = stages;
    L.beta = c->beta; L.u = c->u; L.omega = c->aux; L.shape = c->shape; L.z = c->z; L.sig2 = c->sig2; L.tau = c->tau; L.alpha = c->alpha;
    L.status = c->status; L.sink = make_sink(c, thin, slot0);
    launched(bb::launch_tri(L, st));
    c->z_valid = (stages & BB_ST_BETA) != 0 || c->z_valid;
  } else if (P <= 128) {
    bb::StbLaunch L;
    L.m = m->view(); L.h = h; L.key = key; L.nchains = c->nchains; L.chain0 = c->chain0; L.iter0 = iter0; L.niter = niter;
    L.pre_tau = pre_tau; L.stages = stages;
    L.beta = c->beta; L.lambda = c->aux; L.sig2 = c->sig2; L.tau = c->tau; L.alpha = c->alpha; L.status = c->status;
    L.sink = make_sink(c, thin, slot0);
    launched(bb::launch_stb_small(L, st));
  } else {
    const int Nw = m->N;
    const size_t dimA = c->wide ? (size_t)Nw : (size_t)P;               // order of the per-chain system that is factored
    const size_t wblk = ((dimA + 63) / 64) * 64 * 64;
    if (!c->A) {
      c->A = dalloc<double>((size_t)c->nchains * dimA * dimA); c->W = dalloc<double>((size_t)c->nchains * wblk);
      if (P >= BB_QUAD_MINP && !c->Q) c->Q = dalloc<double>((size_t)c->nchains * ((P + 31) / 32));
      if (c->wide) { c->U = dalloc<double>((size_t)c->nchains * P); c->R = dalloc<double>((size_t)c->nchains * Nw); }
    }
    // (Q: the batched d'X'Xd kernel pays from P = 512 on; at P = 256 the extra launch costs more than the per-chain loop: c5s 112 -> 109 M)
    bb::StbLargeLaunch L;
    L.m = m->view(); L.h = h; L.key = key; L.nchains = c->nchains; L.chain0 = c->chain0; L.stages = stages;
    L.beta = c->beta; L.lambda = c->aux; L.sig2 = c->sig2; L.tau = c->tau; L.alpha = c->alpha; L.status = c->status;
    L.A = c->A; L.W = c->W; L.Q = c->Q; L.sink = make_sink(c, thin, slot0);
    L.wide = c->wide; L.U = c->U; L.R = c->R;
    bb::stb_large_prepare(P);
    // Sub-batches on their own streams: the low-parallelism kernels of one sub-batch (diagonal-block factorisation, panel
    // solve, substitution: one CTA per chain) run while another sub-batch's update kernel fills the SMs.  The sub-batches
    // are independent chains, so they only meet again at the end of the call.  BB_STBL_SPLIT overrides the count (1..4).
    static const int split_env = std::getenv("BB_STBL_SPLIT") ? std::atoi(std::getenv("BB_STBL_SPLIT")) : 0;
    int nsub = c->nchains >= 64 ? 4 : (c->nchains >= 16 ? 2 : 1);   // measured: c4 (128 chains) 9.2 -> 9.4 M, c5s (256 chains) 68.6 -> 71.5 M draws/s from 2 to 4
    // short systems (P <= 512: an iteration is a chain of latency-bound 64-CTA kernels): six sub-batches fill the SMs better
    // (end of round 2, c5s: 4 / 5 / 6 / 8 sub-batches = 198 / 199 / 201 / 200 M draws/s; c4 wide, N = 500: 6 and 8 lost 2 %)
    if (!c->wide && P <= 512 && c->nchains >= 96) nsub = 6;
    if (split_env >= 1 && split_env <= 8) nsub = std::min(split_env, c->nchains);
    if (nsub > 1 && !c->ev_fork) BB_CUDA_CHECK(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    for (int i = 0; i + 1 < nsub; ++i)
      if (!c->xs[i]) {
        BB_CUDA_CHECK(cudaStreamCreateWithFlags(&c->xs[i], cudaStreamNonBlocking));
        BB_CUDA_CHECK(cudaEventCreateWithFlags(&c->xj[i], cudaEventDisableTiming));
      }
    cudaStream_t sts[8] = {st, c->xs[0], c->xs[1], c->xs[2], c->xs[3], c->xs[4], c->xs[5], c->xs[6]};
    if (nsub > 1) {
      BB_CUDA_CHECK(cudaEventRecord(c->ev_fork, st));
      for (int i = 1; i < nsub; ++i) BB_CUDA_CHECK(cudaStreamWaitEvent(sts[i], c->ev_fork, 0));
    }
    auto sub = [&](int h) {
      bb::StbLargeLaunch H = L;
      const int h0 = (int)((long long)c->nchains * h / nsub), h1 = (int)((long long)c->nchains * (h + 1) / nsub);
      const size_t o1 = (size_t)h0, oP = (size_t)h0 * P;
      H.nchains = h1 - h0; H.chain0 = L.chain0 + (uint32_t)h0;
      H.beta += oP; H.lambda += oP; H.sig2 += o1; H.tau += o1; H.alpha += o1; H.status += o1;
      H.A += o1 * dimA * dimA; H.W += o1 * wblk; if (H.Q) H.Q += o1 * ((P + 31) / 32);
      if (H.wide) { H.U += oP; H.R += o1 * (size_t)Nw; }
      const size_t os = o1 * (size_t)H.sink.chain_stride;
      if (H.sink.beta) H.sink.beta += os * P;
      if (H.sink.lambda) H.sink.lambda += os * P;
      if (H.sink.sig2) H.sink.sig2 += os;
      if (H.sink.tau) H.sink.tau += os;
      if (H.sink.alpha) H.sink.alpha += os;
      if (H.sink.sum) { H.sink.sum += oP; H.sink.sumsq += oP; }
      return H;
    };
    L.ctl = nullptr;
    // Graph replay: every iteration of a run is the same sequence of launches; only the iteration index (RNG counter) and
    // the sink slot move, and those come from device memory (StbLargeLaunch::ctl).  One iteration of one sub-batch is
    // captured once -- and again whenever a frozen parameter (hyper-parameters, stages, store or not) changes -- then
    // replayed niter times: one cudaGraphLaunch per sub-batch and iteration instead of ~50 launches, which at P = 256 took
    // longer to issue from the host than to run.  thin > 1 (stores on some iterations only) and short runs launch directly.
    // MEASURED (B200, round 2): no gain -- c5s 124.6 M draws/s replayed against 125.1 M launched directly, c4 identical; at two
    // processes per box the replay was 15 % SLOWER.  The iteration is bound by the chain of ~50 dependent small kernels on the
    // GPU, not by issuing them.  The replay therefore stays opt-in (BB_STBL_GRAPH=1) for callers whose host thread is the bottleneck.
    static const bool graph_off = std::getenv("BB_STBL_GRAPH") == nullptr;
    const bool storing = c->nslots > 0 && thin == 1;
    const bool use_graph = !graph_off && niter >= 4 && (thin == 1 || thin == 0 || c->nslots == 0);
    int it_first = -pre_tau;
    if (use_graph) {
      if (!c->ctl) c->ctl = dalloc<long long>(16);
      for (int it = it_first; it < 0; ++it) {          // the tau pre-draw: plain launches
        L.pre = true; L.iter = iter0; L.slot = -1;
        for (int h = 0; h < nsub; ++h) { bb::StbLargeLaunch H = sub(h); launched(bb::launch_stb_large_iter(H, (stages & BB_ST_BETA) != 0, sts[h])); }
      }
      L.pre = false; L.iter = 0; L.slot = storing ? 0 : -1;
      const long long init[2] = {(long long)iter0, storing ? slot0 : 0};
      for (int h = 0; h < nsub; ++h) {
        bb::StbLargeLaunch H = sub(h);
        H.ctl = c->ctl + 2 * h;
        // everything a later call may change under the same chains object: priors / switches, stage mask, the draw buffer
        struct Key { bb::Hyper h; int stages, storing, nsub, pad; long long stride; double *b, *l, *s2, *ta, *al, *Q; } key;
        std::memset(&key, 0, sizeof(key));
        key.h.sig2_shape = H.h.sig2_shape; key.h.sig2_scale = H.h.sig2_scale; key.h.nu_shape = H.h.nu_shape; key.h.nu_rate = H.h.nu_rate;
        key.h.alpha_a = H.h.alpha_a; key.h.alpha_b = H.h.alpha_b; key.h.true_sig2 = H.h.true_sig2; key.h.true_tau = H.h.true_tau;
        key.h.true_alpha = H.h.true_alpha; key.h.betaburn = H.h.betaburn;
        key.stages = stages; key.storing = storing ? 1 : 0; key.nsub = nsub; key.pad = c->wide; key.stride = H.sink.chain_stride;
        key.b = H.sink.beta; key.l = H.sink.lambda; key.s2 = H.sink.sig2; key.ta = H.sink.tau; key.al = H.sink.alpha; key.Q = H.Q;
        const char* kb = reinterpret_cast<const char*>(&key);
        if (!c->gx[h] || c->gkey[h].size() != sizeof(key) || std::memcmp(c->gkey[h].data(), kb, sizeof(key)) != 0) {
          if (c->gx[h]) { cudaGraphExecDestroy(c->gx[h]); c->gx[h] = nullptr; }
          cudaGraph_t graph = nullptr;
          BB_CUDA_CHECK(cudaStreamBeginCapture(sts[h], cudaStreamCaptureModeThreadLocal));
          int nl = bb::launch_stb_large_iter(H, (stages & BB_ST_BETA) != 0, sts[h]);
          nl += bb::stb_large_advance(c->ctl + 2 * h, storing ? 1 : 0, sts[h]);
          BB_CUDA_CHECK(cudaStreamEndCapture(sts[h], &graph));
          const cudaError_t ie = cudaGraphInstantiate(&c->gx[h], graph, 0);
          cudaGraphDestroy(graph);
          BB_CUDA_CHECK(ie);
          c->gkey[h].assign(kb, kb + sizeof(key));
          c->glaunches[h] = nl;
        }
        BB_CUDA_CHECK(cudaMemcpyAsync(c->ctl + 2 * h, init, sizeof(init), cudaMemcpyHostToDevice, sts[h]));
      }
      for (int it = 0; it < niter; ++it)
        for (int h = 0; h < nsub; ++h) { BB_CUDA_CHECK(cudaGraphLaunch(c->gx[h], sts[h])); launched(c->glaunches[h]); }
      it_first = niter;
    }
    for (int it = it_first; it < niter; ++it) {
      L.pre = it < 0; L.iter = iter0 + (it < 0 ? 0 : it);
      const bool store = c->nslots > 0 && thin > 0 && it >= 0 && ((it + 1) % thin) == 0;
      L.slot = store ? slot0 + (it + 1) / thin - 1 : -1;
      for (int h = 0; h < nsub; ++h) {
        bb::StbLargeLaunch H = sub(h);
        launched(bb::launch_stb_large_iter(H, (stages & BB_ST_BETA) != 0, sts[h]));
      }
    }
    for (int i = 1; i < nsub; ++i) { BB_CUDA_CHECK(cudaEventRecord(c->xj[i - 1], sts[i])); BB_CUDA_CHECK(cudaStreamWaitEvent(st, c->xj[i - 1], 0)); }
    check_launch();
  }
  BB_CUDA_CHECK(cudaEventRecord(c->ev1, st));
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
  if (kernel_ms) { float ms = 0; BB_CUDA_CHECK(cudaEventElapsedTime(&ms, c->ev0, c->ev1)); *kernel_ms = ms; }
}

template <typename F>
int guarded(F&& f) {
  cudaGetLastError();   // drop any stale non-sticky error of an earlier call
  try { f(); return 0; }
  catch (const bb::CudaError& e) { set_err(std::string("CUDA error: ") + cudaGetErrorString(e.code)); }
  catch (const std::string& s) { set_err(s); }
  catch (const std::exception& e) { set_err(e.what()); }
  catch (...) { set_err("unknown error"); }
  return 1;
}

void ensure_svd(const bb_model* cm) {
  bb_model* m = const_cast<bb_model*>(cm);
  std::lock_guard<std::mutex> lock(m->svd_mu);   // two threads creating triangle chains on one model build the factors once
  if (!m->lazy_svd) return;
  auto t0 = std::chrono::steady_clock::now();
  const int P = m->P;
  {
    DBuf<double> W((size_t)P * P);
    BB_CUDA_CHECK(cudaMemcpyAsync(W, m->XX, (size_t)P * P * sizeof(double), cudaMemcpyDeviceToDevice, m->stream));
    model_factor(m, W, P, true);
  }
  model_finish(m, m->X != nullptr);
  BB_CUDA_CHECK(cudaStreamSynchronize(m->stream));
  m->lazy_svd = false;
  m->precompute_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

// ===========================================================================================
// extern "C" boundary
// ===========================================================================================
extern "C" {

const char* bb_last_error(void) { return g_err.c_str(); }
int bb_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }
long long bb_launch_count(void) { return g_launches.load(); }
void bb_set_seed(unsigned long long seed) { g_seed = seed; g_scalar_calls.store(0); }
void bb_set_interrupt(int (*fn)(void)) { g_interrupt = fn; }
void bb_set_wide_gaussian(int on) { g_wide_default = on ? 1 : 0; }
unsigned int bb_scalar_status(void) { return g_scalar_status.load(); }
int bb_model_status(const bb_model* m) { return m ? m->zeroed_by_gram_cut : -1; }
void bb_set_device(int device) { g_device = device; }
void bb_set_verbose(int v) { g_verbose = v; }

// Measurement aid (no counterpart in the reference): peak of the FP64 vector pipe of `device` in TFLOP/s (2 flops per
// DFMA), best of 5 launches timed with CUDA events.  bench.py uses it as the roofline denominator of the triangle sweep.
int bb_measure_fp64_fma(int device, double* tflops) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(device));
    int nsm = 0;
    BB_CUDA_CHECK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, device));
    double* d_out = dalloc<double>(1);
    cudaEvent_t e0, e1;
    BB_CUDA_CHECK(cudaEventCreate(&e0)); BB_CUDA_CHECK(cudaEventCreate(&e1));
    const int iters = 8192, grid = nsm * 8;
    double best = 1e30;
    for (int r = 0; r < 6; ++r) {
      BB_CUDA_CHECK(cudaEventRecord(e0, cudaStreamPerThread));
      dfma_peak_kernel<<<grid, 256, 0, cudaStreamPerThread>>>(d_out, iters, 0.999999, 1e-9);
      BB_CUDA_CHECK(cudaEventRecord(e1, cudaStreamPerThread));
      BB_CUDA_CHECK(cudaEventSynchronize(e1));
      check_launch(); launched();
      float ms = 0; BB_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
      if (r > 0) best = std::min(best, (double)ms);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); dfree(d_out);
    *tflops = 2.0 * 16.0 * iters * 256.0 * grid / (best * 1e-3) / 1e12;
  });
}

bb_model* bb_model_create(const double* X, const double* y, int N, int P, int device, int mem) {
  bb_model* m = nullptr;
  const int rc = guarded([&] {
    auto t0 = std::chrono::steady_clock::now();
    auto tp = t0;
    auto lap = [&](const char* what) {
      if (g_verbose < 2) return;
      cudaStreamSynchronize(m->stream);
      auto n = std::chrono::steady_clock::now();
      std::printf("  [model]  %-22s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(n - tp).count()); tp = n;
    };
    m = model_new(N, P, device);
    if (!m) throw std::string(g_err);
    lap("alloc");
    model_take_rows(m, X, y, mem);
    lap("upload X, y");
    model_gram(m);
    lap("Gram");
    // SVD factors by one-sided Jacobi: on the Cholesky factor of X'X for tall well-conditioned designs, on a copy of X
    // when N >= P otherwise (most accurate), else on X'X
    if (N >= 4 * (long long)P && model_factor_chol(m)) {
      lap("Cholesky + Jacobi");
      model_finish(m, true);
      lap("centres, residuals");
    } else if (N >= P) {
      {
        DBuf<double> W((size_t)N * P);
        BB_CUDA_CHECK(cudaMemcpyAsync(W, m->X, (size_t)N * P * sizeof(double), cudaMemcpyDeviceToDevice, m->stream));
        model_factor(m, W, N, false);
      }
      model_finish(m, true);
    } else {
      // P > N: X'X is singular by rank, so the least-squares start is 0 (BridgeRegression.cpp:79-91) and the stable
      // sampler needs nothing else.  The P x P eigen-decomposition (0.9 s at P = 2000) is only used by the rotated
      // sweep of the triangle sampler: built when such chains (or the factors) are first asked for.
      m->lazy_svd = true; m->singular = true;
      BB_CUDA_CHECK(cudaMemsetAsync(m->bhat, 0, P * sizeof(double), m->stream));
      model_finish(m, true, true);
    }
    BB_CUDA_CHECK(cudaStreamSynchronize(m->stream));
    m->precompute_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  });
  if (rc) { delete m; return nullptr; }
  return m;
}

bb_model* bb_model_create_factors(const double* X, const double* y, int N, int P, const double* tV, const double* d,
                                  const double* a, const double* bhat, int device, int mem) {
  bb_model* m = nullptr;
  const int rc = guarded([&] {
    m = model_new(N, P, device);
    if (!m) throw std::string(g_err);
    model_take_rows(m, X, y, mem);
    model_gram(m);
    cudaStream_t st = m->stream;
    BB_CUDA_CHECK(cudaMemcpyAsync(m->tV, tV, (size_t)P * P * sizeof(double), cudaMemcpyHostToDevice, st));
    BB_CUDA_CHECK(cudaMemcpyAsync(m->d, d, P * sizeof(double), cudaMemcpyHostToDevice, st));
    BB_CUDA_CHECK(cudaMemcpyAsync(m->a, a, P * sizeof(double), cudaMemcpyHostToDevice, st));
    launched(bb::derive_from_tv_device(m->tV, m->d, m->a, P, m->ldv, m->Vc, m->VQ, m->mean, m->d2, m->zc, st));
    if (bhat) BB_CUDA_CHECK(cudaMemcpyAsync(m->bhat, bhat, P * sizeof(double), cudaMemcpyHostToDevice, st));
    else launched(bb::vz_device(m->Vc, m->zc, P, m->ldv, m->bhat, st));
    model_finish(m, true);
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
  });
  if (rc) { delete m; return nullptr; }
  return m;
}

int bb_gram(const double* X_dev, const double* y_dev, int n_rows, int P, double* G_dev, int device) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(device));
    launched(bb::gram_device(X_dev, y_dev, n_rows, P, G_dev, nullptr));
  });
}

// No counterpart in the reference: bb_gram for callers that must not allocate (see include/bayesbridge_b200.h).
// Size checks come first (no CUDA call is made for a refused size); no cudaMalloc / cudaFree / pool call on this path.
int bb_gram_noalloc(const double* X_dev, const double* y_dev, int n_rows, int P, double* G_dev, int device) {
  return guarded([&] {
    if (n_rows < 1 || P < 1) throw std::string("bb_gram_noalloc: n_rows and P must be positive");
    if (P > 3071) throw std::string("bb_gram_noalloc: P = ") + std::to_string(P) + " exceeds the static scratch (P <= 3071); use bb_gram";
    BB_CUDA_CHECK(cudaSetDevice(device));
    launched(bb::gram_device_noalloc(X_dev, y_dev, n_rows, P, G_dev, nullptr));
  });
}

bb_model* bb_model_create_gram(const double* G_dev, int N, int P, int device) {
  bb_model* m = nullptr;
  const int rc = guarded([&] {
    auto t0 = std::chrono::steady_clock::now();
    m = model_new(N, P, device);
    if (!m) throw std::string(g_err);
    DBuf<double> d_yy(1);
    launched(bb::gram_unpack(G_dev, P, m->XX, m->Xy, d_yy, m->stream));
    BB_CUDA_CHECK(cudaMemcpyAsync(&m->yy, d_yy, sizeof(double), cudaMemcpyDeviceToHost, m->stream));
    if (!(N >= 4 * (long long)P && model_factor_chol(m))) {
      DBuf<double> W((size_t)P * P);
      BB_CUDA_CHECK(cudaMemcpyAsync(W, m->XX, (size_t)P * P * sizeof(double), cudaMemcpyDeviceToDevice, m->stream));
      model_factor(m, W, P, true);
    }
    model_finish(m, false);
    // without the rows: r0 from the normal equations (replaced by bb_model_set_r0 after the shard all-reduce)
    BB_CUDA_CHECK(cudaStreamSynchronize(m->stream));
    m->r0z = std::numeric_limits<double>::quiet_NaN(); m->r0b = m->r0z;
    m->precompute_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  });
  if (rc) { delete m; return nullptr; }
  return m;
}

int bb_model_resid_partial(const bb_model* m, const double* X_dev, const double* y_dev, int n_rows, double* out_dev) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(m->device));
    int nl = 0;
    const double r[2] = {bb::resid_device(X_dev, y_dev, m->bz, n_rows, m->P, m->stream, &nl), bb::resid_device(X_dev, y_dev, m->bc, n_rows, m->P, m->stream, &nl)};
    launched(nl);
    BB_CUDA_CHECK(cudaMemcpyAsync(out_dev, r, 2 * sizeof(double), cudaMemcpyHostToDevice, m->stream));
    BB_CUDA_CHECK(cudaStreamSynchronize(m->stream));   // r lives on this stack frame
  });
}
int bb_model_set_r0(bb_model* m, double r0z, double r0b) { m->r0z = r0z; m->r0b = r0b; return 0; }
void bb_model_free(bb_model* m) { delete m; }
double bb_model_precompute_ms(const bb_model* m) { return m->precompute_ms; }

int bb_model_get(const bb_model* m, double* XX, double* Xy, double* bhat, double* tV, double* d, double* a) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(m->device));
    if (tV || d || a) ensure_svd(m);
    const size_t P = m->P;
    if (XX) BB_CUDA_CHECK(cudaMemcpy(XX, m->XX, P * P * sizeof(double), cudaMemcpyDeviceToHost));
    if (Xy) BB_CUDA_CHECK(cudaMemcpy(Xy, m->Xy, P * sizeof(double), cudaMemcpyDeviceToHost));
    if (bhat) BB_CUDA_CHECK(cudaMemcpy(bhat, m->bhat, P * sizeof(double), cudaMemcpyDeviceToHost));
    if (tV) BB_CUDA_CHECK(cudaMemcpy(tV, m->tV, P * P * sizeof(double), cudaMemcpyDeviceToHost));
    if (d) BB_CUDA_CHECK(cudaMemcpy(d, m->d, P * sizeof(double), cudaMemcpyDeviceToHost));
    if (a) BB_CUDA_CHECK(cudaMemcpy(a, m->a, P * sizeof(double), cudaMemcpyDeviceToHost));
  });
}

bb_chains* bb_chains_create(const bb_model* m, int method, int ortho, int nchains, unsigned long long seed, unsigned int chain0) {
  bb_chains* c = nullptr;
  const int rc = guarded([&] {
    if (!m || nchains < 1) throw std::string("bb_chains_create: bad arguments");
    BB_CUDA_CHECK(cudaSetDevice(m->device));
    if (method == BB_TRI && !ortho) ensure_svd(m);
    c = new bb_chains();
    c->m = m; c->device = m->device; c->method = method; c->ortho = ortho; c->nchains = nchains; c->seed = seed; c->chain0 = chain0;
    BB_CUDA_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    BB_CUDA_CHECK(cudaEventCreate(&c->ev0)); BB_CUDA_CHECK(cudaEventCreate(&c->ev1));
    const size_t n = (size_t)nchains * m->P;
    c->beta = dalloc<double>(n); c->u = dalloc<double>(n); c->aux = dalloc<double>(n); c->shape = dalloc<double>(n); c->z = dalloc<double>(n);
    c->sig2 = dalloc<double>(nchains); c->tau = dalloc<double>(nchains); c->alpha = dalloc<double>(nchains);
    c->status = dalloc<uint32_t>(nchains);
    cudaStream_t st = c->stream;
    // BridgeWrapper.cpp:120-125 / 241-246 and the fills at :603-605, :686
    bcast_rows_kernel<<<296, 256, 0, st>>>(c->beta, m->bhat, m->P, nchains); check_launch(); launched();
    fill(c->u, n, 0.5, st); fill(c->aux, n, 1.0, st); fill(c->shape, n, 0.0, st); fill(c->z, n, 0.0, st);
    fill(c->sig2, nchains, 0.0, st); fill(c->tau, nchains, 0.0, st); fill(c->alpha, nchains, 0.5, st);
    BB_CUDA_CHECK(cudaMemsetAsync(c->status, 0, nchains * sizeof(uint32_t), st));
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
  });
  if (rc) { delete c; return nullptr; }
  return c;
}
void bb_chains_free(bb_chains* c) { delete c; }

int bb_chains_set(bb_chains* c, const double* beta, const double* u, const double* aux, const double* sig2, const double* tau) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    const size_t n = (size_t)c->nchains * c->m->P * sizeof(double), ns = (size_t)c->nchains * sizeof(double);
    if (beta) { BB_CUDA_CHECK(cudaMemcpy(c->beta, beta, n, cudaMemcpyHostToDevice)); c->z_valid = false; }
    if (u) BB_CUDA_CHECK(cudaMemcpy(c->u, u, n, cudaMemcpyHostToDevice));
    if (aux) BB_CUDA_CHECK(cudaMemcpy(c->aux, aux, n, cudaMemcpyHostToDevice));
    if (sig2) BB_CUDA_CHECK(cudaMemcpy(c->sig2, sig2, ns, cudaMemcpyHostToDevice));
    if (tau) BB_CUDA_CHECK(cudaMemcpy(c->tau, tau, ns, cudaMemcpyHostToDevice));
  });
}
int bb_chains_get(const bb_chains* c, double* beta, double* u, double* aux, double* shape, double* sig2, double* tau, unsigned int* status) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    const size_t n = (size_t)c->nchains * c->m->P * sizeof(double), ns = (size_t)c->nchains * sizeof(double);
    if (beta) BB_CUDA_CHECK(cudaMemcpy(beta, c->beta, n, cudaMemcpyDeviceToHost));
    if (u) BB_CUDA_CHECK(cudaMemcpy(u, c->u, n, cudaMemcpyDeviceToHost));
    if (aux) BB_CUDA_CHECK(cudaMemcpy(aux, c->aux, n, cudaMemcpyDeviceToHost));
    if (shape) BB_CUDA_CHECK(cudaMemcpy(shape, c->shape, n, cudaMemcpyDeviceToHost));
    if (sig2) BB_CUDA_CHECK(cudaMemcpy(sig2, c->sig2, ns, cudaMemcpyDeviceToHost));
    if (tau) BB_CUDA_CHECK(cudaMemcpy(tau, c->tau, ns, cudaMemcpyDeviceToHost));
    if (status) BB_CUDA_CHECK(cudaMemcpy(status, c->status, c->nchains * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  });
}

int bb_chains_alpha(bb_chains* c, const double* set, double* get) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    const size_t ns = (size_t)c->nchains * sizeof(double);
    if (set) BB_CUDA_CHECK(cudaMemcpy(c->alpha, set, ns, cudaMemcpyHostToDevice));
    if (get) BB_CUDA_CHECK(cudaMemcpy(get, c->alpha, ns, cudaMemcpyDeviceToHost));
  });
}

int bb_chains_option(bb_chains* c, int option, int value) {
  return guarded([&] {
    if (!c) throw std::string("bb_chains_option: null chains");
    if (option != BB_OPT_WIDE_GAUSSIAN) throw std::string("bb_chains_option: unknown option");
    if (value) {
      if (c->method != BB_STB || c->ortho) throw std::string("bb_chains_option: BB_OPT_WIDE_GAUSSIAN applies to the stable sampler (ortho = 0)");
      if (c->m->P <= 128) throw std::string("bb_chains_option: BB_OPT_WIDE_GAUSSIAN needs P > 128 (below that the P x P system is factored on chip)");
      if (!c->m->X) throw std::string("bb_chains_option: BB_OPT_WIDE_GAUSSIAN needs the rows of X on the device (a model built from a Gram matrix has none)");
      if (c->m->N > c->m->P) throw std::string("bb_chains_option: BB_OPT_WIDE_GAUSSIAN is for wide designs (N <= P); for N > P the P x P system is the smaller one");
    }
    if ((value != 0) != (c->wide != 0)) {     // the workspace has another shape: rebuilt on the next run
      BB_CUDA_CHECK(cudaSetDevice(c->device));
      BB_CUDA_CHECK(cudaStreamSynchronize(c->stream));
      dfree(c->A); dfree(c->W); dfree(c->U); dfree(c->R);
      c->wide = value ? 1 : 0;
    }
  });
}

int bb_chains_reserve(bb_chains* c, long long nslots, int out_mask) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    c->free_draws();
    if (nslots <= 0) return;
    const size_t nv = (size_t)c->nchains * nslots * c->m->P, ns = (size_t)c->nchains * nslots;
    if (out_mask & BB_OUT_BETA) c->o_beta = dalloc<double>(nv);
    if (out_mask & BB_OUT_U) c->o_u = dalloc<double>(nv);
    if (out_mask & BB_OUT_OMEGA) c->o_omega = dalloc<double>(nv);
    if (out_mask & BB_OUT_SHAPE) c->o_shape = dalloc<double>(nv);
    if (out_mask & BB_OUT_LAMBDA) c->o_lambda = dalloc<double>(nv);
    if (out_mask & BB_OUT_SIG2) c->o_sig2 = dalloc<double>(ns);
    if (out_mask & BB_OUT_TAU) c->o_tau = dalloc<double>(ns);
    if (out_mask & BB_OUT_ALPHA) c->o_alpha = dalloc<double>(ns);
    c->nslots = nslots; c->out_mask = out_mask;
  });
}

int bb_chains_run(bb_chains* c, const bb_hyper* h, unsigned int iter0, int niter, int stages, int pre_tau, int thin,
                  long long slot0, double* kernel_ms) {
  return guarded([&] { chains_run(c, h, iter0, niter, stages, pre_tau & 1, (pre_tau >> 1) & 1, thin, slot0, kernel_ms); });
}

int bb_chains_store_state(bb_chains* c, long long slot) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    if (slot < 0 || slot >= c->nslots) throw std::string("bb_chains_store_state: slot out of range");
    const int P = c->m->P, n = c->nchains; cudaStream_t st = c->stream;
    auto cp = [&](const double* src, double* dst, int w) {
      if (dst) { store_state_kernel<<<296, 256, 0, st>>>(src, dst, w, c->nslots, slot, n); check_launch(); launched(); }
    };
    cp(c->beta, c->o_beta, P); cp(c->u, c->o_u, P);
    cp(c->aux, c->method == BB_TRI ? c->o_omega : c->o_lambda, P); cp(c->shape, c->o_shape, P);
    cp(c->sig2, c->o_sig2, 1); cp(c->tau, c->o_tau, 1); cp(c->alpha, c->o_alpha, 1);
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
  });
}

int bb_chains_summary(const bb_chains* c, int which, long long slot0, long long nslots, double runtime_sec, double* per_chain,
                      double* pooled) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    const double* src = nullptr;
    switch (which) {
      case BB_OUT_BETA: src = c->o_beta; break;
      case BB_OUT_U: src = c->o_u; break;
      case BB_OUT_OMEGA: src = c->o_omega; break;
      case BB_OUT_SHAPE: src = c->o_shape; break;
      case BB_OUT_LAMBDA: src = c->o_lambda; break;
      default: break;
    }
    if (!src) throw std::string("bb_chains_summary: that output is not a reserved vector output");
    if (slot0 < 0 || nslots < 2 || slot0 + nslots > c->nslots) throw std::string("bb_chains_summary: slot range out of bounds");
    if (nslots > 16384) throw std::string("bb_chains_summary: at most 16384 stored draws per chain (thin the run or fetch the draws)");
    const size_t P = c->m->P, n1 = (size_t)c->nchains * 5 * P, n2 = 7 * P;
    DBuf<double> d_pc(n1);
    DBuf<double> d_po(n2);
    launched(bb::summary_device(src, (int)P, c->nslots, slot0, (int)nslots, c->nchains, runtime_sec, d_pc, d_po, c->stream));
    if (per_chain) BB_CUDA_CHECK(cudaMemcpyAsync(per_chain, d_pc, n1 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (pooled) BB_CUDA_CHECK(cudaMemcpyAsync(pooled, d_po, n2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    BB_CUDA_CHECK(cudaStreamSynchronize(c->stream));
  });
}

namespace {
const double* vector_output(const bb_chains* c, int which) {
  switch (which) {
    case BB_OUT_BETA: return c->o_beta;
    case BB_OUT_U: return c->o_u;
    case BB_OUT_OMEGA: return c->o_omega;
    case BB_OUT_SHAPE: return c->o_shape;
    case BB_OUT_LAMBDA: return c->o_lambda;
    default: return nullptr;
  }
}
}  // namespace

int bb_chains_summary_begin(bb_chains* c, int which) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    if (which != BB_OUT_BETA && which != BB_OUT_U && which != BB_OUT_OMEGA && which != BB_OUT_SHAPE && which != BB_OUT_LAMBDA)
      throw std::string("bb_chains_summary_begin: not a vector output");
    const int P = c->m->P;
    if (!c->rs_state) {
      c->rs_state = dalloc<double>(bb::running_state_doubles(c->nchains, P));
      c->rs_range = dalloc<double>((size_t)2 * P);
      c->rs_hist = dalloc<unsigned int>(bb::running_hist_words(P));
    }
    c->rs_n = 0; c->rs_which = which;
  });
}

int bb_chains_summary_add(bb_chains* c, long long slot0, long long nslots) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    if (!c->rs_state || !c->rs_which) throw std::string("bb_chains_summary_add: call bb_chains_summary_begin first");
    const double* src = vector_output(c, c->rs_which);
    if (!src) throw std::string("bb_chains_summary_add: that output is not reserved");
    if (slot0 < 0 || nslots < 1 || slot0 + nslots > c->nslots) throw std::string("bb_chains_summary_add: slot range out of bounds");
    if (c->rs_n == 0 && nslots < 2) throw std::string("bb_chains_summary_add: the first chunk needs at least 2 slots (it fixes the range of the quantile sketch)");
    long long done = 0;
    while (done < nslots) {                     // at most 4096 slots per launch (shared memory of the fold kernel)
      const int m = (int)std::min(4096LL, nslots - done);
      launched(bb::running_add_device(src, c->m->P, c->nslots, slot0 + done, m, c->nchains, c->rs_n, c->rs_state, c->rs_range, c->rs_hist, c->stream));
      c->rs_n += m; done += m;
    }
    BB_CUDA_CHECK(cudaStreamSynchronize(c->stream));
  });
}

int bb_chains_summary_end(bb_chains* c, double runtime_sec, double* per_chain, double* pooled) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    if (!c->rs_state || c->rs_n < 2) throw std::string("bb_chains_summary_end: fewer than 2 draws per chain were added");
    const size_t P = c->m->P, n1 = (size_t)c->nchains * 5 * P, n2 = 7 * P;
    DBuf<double> d_pc(n1);
    DBuf<double> d_po(n2);
    launched(bb::running_end_device(c->rs_state, c->rs_range, c->rs_hist, (int)P, c->nchains, c->rs_n, runtime_sec, d_pc, d_po, c->stream));
    if (per_chain) BB_CUDA_CHECK(cudaMemcpyAsync(per_chain, d_pc, n1 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    if (pooled) BB_CUDA_CHECK(cudaMemcpyAsync(pooled, d_po, n2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    BB_CUDA_CHECK(cudaStreamSynchronize(c->stream));
  });
}

int bb_chains_fetch(const bb_chains* c, int which, long long slot0, long long nslots, double* out) {
  return guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(c->m->device));
    const double* src = nullptr; int w = c->m->P;
    switch (which) {
      case BB_OUT_BETA: src = c->o_beta; break;
      case BB_OUT_U: src = c->o_u; break;
      case BB_OUT_OMEGA: src = c->o_omega; break;
      case BB_OUT_SHAPE: src = c->o_shape; break;
      case BB_OUT_LAMBDA: src = c->o_lambda; break;
      case BB_OUT_SIG2: src = c->o_sig2; w = 1; break;
      case BB_OUT_TAU: src = c->o_tau; w = 1; break;
      case BB_OUT_ALPHA: src = c->o_alpha; w = 1; break;
      default: break;
    }
    if (!src) throw std::string("bb_chains_fetch: that output was not reserved");
    if (slot0 < 0 || slot0 + nslots > c->nslots) throw std::string("bb_chains_fetch: slot range out of bounds");
    const size_t n = (size_t)c->nchains * nslots * w;
    if (slot0 == 0 && nslots == c->nslots) { BB_CUDA_CHECK(cudaMemcpy(out, src, n * sizeof(double), cudaMemcpyDeviceToHost)); return; }
    DBuf<double> tmp(n);
    gather_kernel<<<296, 256, 0, c->stream>>>(src, tmp, w, c->nslots, slot0, nslots, c->nchains); check_launch(); launched();
    BB_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    BB_CUDA_CHECK(cudaMemcpy(out, tmp, n * sizeof(double), cudaMemcpyDeviceToHost));
  });
}

}  // extern "C"

// ===========================================================================================
// Drivers: the schedule of BridgeWrapper.cpp:80-204 (tri), 207-313 (stb), 320-431, 434-537 (ortho)
// ===========================================================================================
namespace {

void nan_fill(double* p, size_t n) { if (p) for (size_t i = 0; i < n; ++i) p[i] = std::numeric_limits<double>::quiet_NaN(); }

// a bit-for-bit copy of a model on another device (peer copies; X, y stay behind: no kernel of the sweep reads them)
bb_model* model_clone(const bb_model* src, int device) {
  bb_model* m = model_new(src->N, src->P, device);
  if (!m) throw std::string(g_err);
  try {
    if (src->lazy_svd) throw std::string("bb_model_clone: the source model's SVD factors are not built yet");
    const size_t P = src->P, ldv = src->ldv;
    cudaStream_t st = m->stream;
    auto cp = [&](double* dst, const double* from, size_t n) {
      BB_CUDA_CHECK(cudaMemcpyPeerAsync(dst, device, from, src->device, n * sizeof(double), st));
    };
    cp(m->XX, src->XX, P * P); cp(m->Xy, src->Xy, P); cp(m->bhat, src->bhat, P); cp(m->tV, src->tV, P * P);
    cp(m->Vc, src->Vc, P * ldv); cp(m->VQ, src->VQ, 2 * P * (ldv + 2)); cp(m->d, src->d, P); cp(m->a, src->a, P);
    cp(m->mean, src->mean, P); cp(m->zc, src->zc, P); cp(m->d2, src->d2, P); cp(m->gc, src->gc, P); cp(m->bc, src->bc, P);
    cp(m->bz, src->bz, P);
    m->r0z = src->r0z; m->r0b = src->r0b; m->yy = src->yy; m->singular = src->singular; m->lazy_svd = false;
    m->zeroed_by_gram_cut = src->zeroed_by_gram_cut; m->precompute_ms = 0;
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
  } catch (...) { delete m; throw; }
  return m;
}

struct DriverOut {   // host outputs of one driver call, [nchains][nsamp][w] / [nchains][nsamp]; any may be null
  double *beta = nullptr, *aux1 = nullptr, *aux2 = nullptr, *aux3 = nullptr, *sig2 = nullptr, *tau = nullptr, *alpha = nullptr;
  unsigned int* status = nullptr;
  double *sum_pooled = nullptr, *sum_chain = nullptr;
};

// The drivers' schedule on an existing model, for chains [chain0, chain0 + nchains).  Returns the seconds of the
// post-burn-in loop INCLUDING the arrival of the last draws on the host (the reference's clock stops when its samples are
// in memory, BridgeWrapper.cpp:197-202).  poll: the calling thread polls the interrupt hook between chunks.
double run_driver_on_model(const bb_model* m, int method, int ortho, const bb_hyper* h, int nchains, unsigned long long seed,
                           unsigned int chain0, int burn, int nsamp, int thin, const DriverOut& out, bool poll) {
  if (nsamp < 1 || burn < 0 || nchains < 1) throw std::string("bridge driver: need nsamp >= 1, burn >= 0, nchains >= 1");
  if (thin < 1) thin = 1;
  const int P = m->P;
  BB_CUDA_CHECK(cudaSetDevice(m->device));
  struct Guard { bb_chains* c = nullptr; ~Guard() { delete c; } } g;
  auto tp0 = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (g_verbose >= 2) { auto n = std::chrono::steady_clock::now(); std::printf("  [driver] %-22s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(n - tp0).count()); tp0 = n; }
  };
  g.c = bb_chains_create(m, method, ortho, nchains, seed, chain0);
  if (!g.c) throw std::string(g_err);
  if (g_wide_default && method == BB_STB && !ortho && m->P > 128 && m->X && m->N <= m->P && bb_chains_option(g.c, BB_OPT_WIDE_GAUSSIAN, 1))
    throw std::string(g_err);
  // only the outputs the caller asked for are stored by the kernels (8 P bytes per chain-iteration each)
  struct Out { int which; double* host; const double* dev; int w; };
  Out outs[8] = {{BB_OUT_BETA, out.beta, nullptr, P},
                 {method == BB_TRI ? BB_OUT_U : BB_OUT_LAMBDA, out.aux1, nullptr, P},
                 {BB_OUT_OMEGA, method == BB_TRI ? out.aux2 : nullptr, nullptr, P},
                 {BB_OUT_SHAPE, method == BB_TRI ? out.aux3 : nullptr, nullptr, P},
                 {BB_OUT_SIG2, out.sig2, nullptr, 1}, {BB_OUT_TAU, out.tau, nullptr, 1}, {BB_OUT_ALPHA, out.alpha, nullptr, 1}, {0, nullptr, nullptr, 0}};
  int mask = 0;
  for (const Out& o : outs) if (o.host) mask |= o.which;
  if (mask == 0 || out.sum_pooled || out.sum_chain) mask |= BB_OUT_BETA;
  // Summary mode keeps RUNNING summaries (bb_running.cu): the draw buffer holds one chunk (at most 1024 slots) that is folded
  // into the accumulators and reused, so the run length is not bounded by device memory or by a slot limit.
  const bool running = out.sum_pooled || out.sum_chain;
  const long long cap = running ? std::min<long long>(nsamp, 1024) : nsamp;
  if (bb_chains_reserve(g.c, cap, mask)) throw std::string(g_err);
  for (Out& o : outs) {
    switch (o.host ? o.which : 0) {
      case BB_OUT_BETA: o.dev = g.c->o_beta; break;
      case BB_OUT_U: o.dev = g.c->o_u; break;
      case BB_OUT_OMEGA: o.dev = g.c->o_omega; break;
      case BB_OUT_SHAPE: o.dev = g.c->o_shape; break;
      case BB_OUT_LAMBDA: o.dev = g.c->o_lambda; break;
      case BB_OUT_SIG2: o.dev = g.c->o_sig2; break;
      case BB_OUT_TAU: o.dev = g.c->o_tau; break;
      case BB_OUT_ALPHA: o.dev = g.c->o_alpha; break;
      default: break;
    }
  }
  // The draws of a finished chunk go to the host on a copy stream while the next chunk is sampled: the host array has
  // the layout of the draw buffer, [chain][sample][w], so a slot range is one strided (2-D) copy per output.
  struct CopyStream { cudaStream_t s = nullptr; ~CopyStream() { if (s) cudaStreamDestroy(s); } } cps;
  BB_CUDA_CHECK(cudaStreamCreateWithFlags(&cps.s, cudaStreamNonBlocking));
  auto download = [&](long long s0, long long ns, long long d0 = -1) {   // host slots [s0, s0 + ns) <- device slots [d0, d0 + ns) (d0 < 0: the same)
    if (d0 < 0) d0 = s0;
    for (const Out& o : outs) {
      if (!o.dev || ns <= 0) continue;
      const size_t hpitch = (size_t)nsamp * o.w * sizeof(double), dpitch = (size_t)cap * o.w * sizeof(double);
      BB_CUDA_CHECK(cudaMemcpy2DAsync(o.host + (size_t)s0 * o.w, hpitch, o.dev + (size_t)d0 * o.w, dpitch, (size_t)ns * o.w * sizeof(double),
                                      (size_t)nchains, cudaMemcpyDeviceToHost, cps.s));
    }
  };
  auto interrupted = [&]() { return poll && g_interrupt && g_interrupt() != 0; };
  lap("chains + draw buffer");
  if (g_verbose && poll) {
    std::printf("Bridge Regression (%s):", method == BB_TRI ? "mix. of triangles" : "mix. of normals");
    if (h->true_alpha > 0) std::printf(" known alpha=%g", h->true_alpha);
    if (h->true_sig2 > 0) std::printf(", sig2=%g", h->true_sig2);
    if (h->true_tau > 0) std::printf(", tau=%g", h->true_tau);
    std::printf("\nBurn-in: %i, Num. Samples: %i, Chains: %i\n", burn, nsamp, nchains);
  }
  // burn-in, in place (tri: burn iterations, :145 ; stb: burn+1, :266), preceded by the pre-draws.  With an interrupt hook
  // the burn-in runs in slices so that the hook is polled about five times a second (the reference polls
  // R_CheckUserInterrupt every 10 iterations, BridgeWrapper.cpp:154-156,186-188).
  const int nburn = method == BB_TRI ? burn : burn + 1;
  int pre = 0;
  if (method == BB_TRI) pre = ortho ? 3 : 1;          // :142 / :379-380
  else pre = ortho ? 0 : 1;                           // :262 / none (:490)
  {
    long long done_b = 0, slice = (poll && g_interrupt) ? 16 : nburn;
    bool first = true;
    do {
      const long long n = std::min(slice, (long long)nburn - done_b);
      auto ts = std::chrono::steady_clock::now();
      chains_run(g.c, h, (unsigned)done_b, (int)n, BB_ST_ALL, first ? (pre & 1) : 0, first ? ((pre >> 1) & 1) : 0, 0, 0, nullptr);
      first = false; done_b += n;
      if (interrupted()) throw std::string("interrupted by the caller's hook (bb_set_interrupt) during burn-in");
      const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - ts).count();
      if (poll && g_interrupt && sec > 0) slice = std::max(16LL, std::min(1LL << 20, (long long)(slice * 0.2 / sec)));
    } while (done_b < nburn);
  }
  if (bb_chains_store_state(g.c, 0)) throw std::string(g_err);     // sample 0 = state after burn-in
  lap("burn-in");
  auto t0 = std::chrono::steady_clock::now();
  const long long total = (long long)(nsamp - 1) * thin;
  long long done = 0;
  // chunks: at most 8192 stored iterations between two polls, and about four chunks per call so that the download overlaps;
  // with an interrupt hook the chunk adapts to about 0.2 s of sampling
  long long chunk = std::min(8192LL, std::max(16LL, (total / thin + 3) / 4)) * thin;
  if (poll && g_interrupt) chunk = std::min(chunk, 16LL * thin);
  long long copied = 0;   // slots [0, copied) are on their way to the host
  // the sampling loop of the non-orthogonal stable driver hands (alpha_b, alpha_b) to the alpha step
  // (BridgeWrapper.cpp:294; its burn-in :272 and the other three drivers pass (alpha_a, alpha_b)): kept as written
  bb_hyper hs = *h;
  if (method != BB_TRI && !ortho) hs.alpha_a = hs.alpha_b;
  if (running && bb_chains_summary_begin(g.c, BB_OUT_BETA)) throw std::string(g_err);
  long long dev_base = 1;                         // running mode: the chunk starts behind sample 0 the first time, at slot 0 afterwards
  if (running && total == 0 && bb_chains_summary_add(g.c, 0, 1)) throw std::string(g_err);
  while (done < total) {
    long long n = std::min(chunk, total - done);
    if (running) n = std::min(n, (cap - dev_base) * thin);
    auto ts = std::chrono::steady_clock::now();
    chains_run(g.c, &hs, (unsigned)(nburn + done), (int)n, BB_ST_ALL, 0, 0, thin, running ? dev_base : 1 + done / thin, nullptr);
    done += n;
    const long long have = 1 + done / thin;
    if (running) {
      const long long cnt = have - copied;        // device slots [0, cnt) = host samples [copied, have)
      if (bb_chains_summary_add(g.c, 0, cnt)) throw std::string(g_err);
      download(copied, cnt, 0); copied = have; dev_base = 0;
      BB_CUDA_CHECK(cudaStreamSynchronize(cps.s));   // the next chunk overwrites these slots
    } else { download(copied, have - copied); copied = have; }
    if (interrupted()) {
      BB_CUDA_CHECK(cudaStreamSynchronize(cps.s));
      throw std::string("interrupted by the caller's hook (bb_set_interrupt) after ") + std::to_string(have) + " of " + std::to_string(nsamp) + " samples";
    }
    if (poll && g_interrupt) {
      const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - ts).count();
      if (sec > 0) chunk = std::max(16LL, std::min(8192LL, (long long)((chunk / thin) * 0.2 / sec))) * thin;
    }
  }
  if (!running) download(copied, nsamp - copied);
  BB_CUDA_CHECK(cudaStreamSynchronize(cps.s));    // the clock stops when the last draws are in host memory
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  lap("sampling loop + download");
  // summary mode: sum.stat of beta over the nsamp stored samples, on the device (the draws need not leave the GPU)
  if (running && bb_chains_summary_end(g.c, secs, out.sum_chain, out.sum_pooled)) throw std::string(g_err);
  if (out.status && bb_chains_get(g.c, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, out.status)) throw std::string(g_err);
  if (g_verbose && poll) std::printf("Sampling complete: %g sec. for %i iterations.\n", secs, nsamp);
  return secs;
}

void run_driver(int method, int ortho, const double* X, const double* y, int N, int P, const bb_hyper* h,
                int nchains, unsigned long long seed, unsigned int chain0, int burn, int nsamp, int thin, int device,
                double* beta, double* aux1, double* aux2, double* aux3, double* sig2, double* tau, double* alpha,
                unsigned int* status, double* runtime, double* sum_pooled = nullptr, double* sum_chain = nullptr) {
  struct Guard { bb_model* m = nullptr; ~Guard() { delete m; } } g;
  g.m = bb_model_create(X, y, N, P, device, BB_MEM_HOST);
  if (!g.m) throw std::string(g_err);
  DriverOut o;
  o.beta = beta; o.aux1 = aux1; o.aux2 = aux2; o.aux3 = aux3; o.sig2 = sig2; o.tau = tau; o.alpha = alpha; o.status = status;
  o.sum_pooled = sum_pooled; o.sum_chain = sum_chain;
  const double secs = run_driver_on_model(g.m, method, ortho, h, nchains, seed, chain0, burn, nsamp, thin, o, true);
  if (runtime) *runtime = secs;
}

// Device fan-out inside the library: the constructor runs once, on devices[0]; the other devices
// receive a bit-for-bit copy of the factors over NVLink (peer copies); device k samples the k-th contiguous block of chains
// on its own host thread.  Philox is keyed by the global chain id, so every draw equals the one-GPU run's.
void run_driver_multi(int method, int ortho, const double* X, const double* y, int N, int P, const bb_hyper* h,
                      int nchains, unsigned long long seed, unsigned int chain0, int burn, int nsamp, int thin,
                      int ndevices, const int* devices, const DriverOut& out, double* runtime) {
  int avail = 0;
  if (cudaGetDeviceCount(&avail) != cudaSuccess || avail == 0) throw std::string("bb_bridge_batched_multi: no CUDA device (there is no CPU fallback)");
  if (ndevices < 1 || ndevices > avail) throw std::string("bb_bridge_batched_multi: ndevices must be between 1 and the number of CUDA devices (") + std::to_string(avail) + ")";
  if (nchains < ndevices) ndevices = nchains;
  std::vector<int> devs(ndevices);
  for (int k = 0; k < ndevices; ++k) {
    devs[k] = devices ? devices[k] : k;
    if (devs[k] < 0 || devs[k] >= avail) throw std::string("bb_bridge_batched_multi: bad device index");
  }
  if (out.sum_pooled || out.sum_chain) throw std::string("bb_bridge_batched_multi: summary mode runs on one device (bb_bridge_batched_summary)");
  struct Models { std::vector<bb_model*> m; ~Models() { for (bb_model* x : m) delete x; } } ms;
  ms.m.assign(ndevices, nullptr);
  ms.m[0] = bb_model_create(X, y, N, P, devs[0], BB_MEM_HOST);
  if (!ms.m[0]) throw std::string(g_err);
  if (method == BB_TRI && !ortho) ensure_svd(ms.m[0]);
  else if (ms.m[0]->lazy_svd && ndevices > 1) { /* P > N, stable sampler: the clone needs no SVD factors */ }
  for (int k = 1; k < ndevices; ++k) {
    if (ms.m[0]->lazy_svd) {   // copy what exists; the stable path never touches the missing factors
      bb_model* src = ms.m[0];
      const bool lazy = src->lazy_svd; src->lazy_svd = false;
      try { ms.m[k] = model_clone(src, devs[k]); } catch (...) { src->lazy_svd = lazy; throw; }
      src->lazy_svd = lazy; ms.m[k]->lazy_svd = false;
    } else ms.m[k] = model_clone(ms.m[0], devs[k]);
  }
  std::vector<std::string> errs(ndevices);
  std::vector<double> secs(ndevices, 0.0);
  std::vector<std::thread> th;
  auto block = [&](int k, int& lo, int& n) { lo = (int)((long long)nchains * k / ndevices); n = (int)((long long)nchains * (k + 1) / ndevices) - lo; };
  auto work = [&](int k) {
    int lo, n; block(k, lo, n);
    DriverOut o = out;
    const size_t ov = (size_t)lo * nsamp * P, os = (size_t)lo * nsamp;
    if (o.beta) o.beta += ov; if (o.aux1) o.aux1 += ov; if (o.aux2) o.aux2 += ov; if (o.aux3) o.aux3 += ov;
    if (o.sig2) o.sig2 += os; if (o.tau) o.tau += os; if (o.alpha) o.alpha += os; if (o.status) o.status += lo;
    try { secs[k] = run_driver_on_model(ms.m[k], method, ortho, h, n, seed, chain0 + (unsigned)lo, burn, nsamp, thin, o, k == 0); }
    catch (const bb::CudaError& e) { errs[k] = std::string("CUDA error: ") + cudaGetErrorString(e.code); }
    catch (const std::string& e) { errs[k] = e.empty() ? std::string("error") : e; }
    catch (const std::exception& e) { errs[k] = e.what(); }
    catch (...) { errs[k] = "unknown error"; }
  };
  for (int k = 1; k < ndevices; ++k) th.emplace_back(work, k);
  work(0);                                   // device 0's block on the calling thread (it is the one that may poll the hook)
  for (std::thread& t : th) t.join();
  for (int k = 0; k < ndevices; ++k)
    if (!errs[k].empty()) throw std::string("device ") + std::to_string(devs[k]) + ": " + errs[k];
  if (runtime) *runtime = *std::max_element(secs.begin(), secs.end());
}

}  // namespace

extern "C" {

int bb_bridge_batched(int method, int ortho, const double* X, const double* y, int N, int P, const bb_hyper* h,
                      int nchains, unsigned long long seed, unsigned int chain0, int burn, int nsamp, int thin, int device,
                      double* beta, double* aux1, double* aux2, double* aux3, double* sig2, double* tau, double* alpha,
                      unsigned int* status, double* runtime) {
  return guarded([&] {
    run_driver(method, ortho, X, y, N, P, h, nchains, seed, chain0, burn, nsamp, thin, device, beta, aux1, aux2, aux3, sig2, tau,
               alpha, status, runtime);
  });
}

int bb_bridge_batched_multi(int method, int ortho, const double* X, const double* y, int N, int P, const bb_hyper* h,
                            int nchains, unsigned long long seed, unsigned int chain0, int burn, int nsamp, int thin,
                            int ndevices, const int* devices,
                            double* beta, double* aux1, double* aux2, double* aux3, double* sig2, double* tau, double* alpha,
                            unsigned int* status, double* runtime) {
  return guarded([&] {
    DriverOut o;
    o.beta = beta; o.aux1 = aux1; o.aux2 = aux2; o.aux3 = aux3; o.sig2 = sig2; o.tau = tau; o.alpha = alpha; o.status = status;
    run_driver_multi(method, ortho, X, y, N, P, h, nchains, seed, chain0, burn, nsamp, thin, ndevices, devices, o, runtime);
  });
}

bb_model* bb_model_clone(const bb_model* src, int device) {
  bb_model* m = nullptr;
  const int rc = guarded([&] {
    if (!src) throw std::string("bb_model_clone: null model");
    ensure_svd(src);
    m = model_clone(src, device);
  });
  return rc ? nullptr : m;
}

int bb_bridge_batched_summary(int method, int ortho, const double* X, const double* y, int N, int P, const bb_hyper* h,
                              int nchains, unsigned long long seed, unsigned int chain0, int burn, int nsamp, int thin, int device,
                              double* pooled, double* per_chain, double* sig2, double* tau, unsigned int* status, double* runtime) {
  return guarded([&] {
    if (!pooled && !per_chain) throw std::string("bb_bridge_batched_summary: pooled and per_chain are both NULL");
    run_driver(method, ortho, X, y, N, P, h, nchains, seed, chain0, burn, nsamp, thin, device, nullptr, nullptr, nullptr, nullptr,
               sig2, tau, nullptr, status, runtime, pooled, per_chain);
  });
}

void bridge_regression(double* betap, double* up, double* omegap, double* shapep, double* sig2p, double* taup, double* alphap,
                       const double* yp, const double* Xp, const double* sig2_shape, const double* sig2_scale,
                       const double* nu_shape, const double* nu_rate, const double* alpha_a, const double* alpha_b,
                       const double* true_sig2, const double* true_tau, const double* true_alpha, const int* P, const int* N,
                       const int* M, const int* burn, double* runtime, const int* ortho, const int* betaburn, const int* use_hmc) {
  (void)use_hmc;   // forced false in the reference (BridgeRegression.cpp:419)
  bb_hyper h{*sig2_shape, *sig2_scale, *nu_shape, *nu_rate, *alpha_a, *alpha_b, *true_sig2, *true_tau, *true_alpha, *betaburn};
  const int orth = (*reinterpret_cast<const unsigned char*>(ortho)) != 0;   // R logical: low byte, as the reference reads it
  const int rc = bb_bridge_batched(BB_TRI, orth, Xp, yp, *N, *P, &h, 1, g_seed, 0u, *burn, *M, 1, g_device, betap, up, omegap, shapep,
                                   sig2p, taup, alphap, nullptr, runtime);
  if (rc) {
    std::fprintf(stderr, "bridge_regression (bayesbridge_b200): %s\nAborting Gibbs sampler.\n", g_err.c_str());
    const size_t n = (size_t)(*P) * (*M);
    nan_fill(betap, n); nan_fill(up, n); nan_fill(omegap, n); nan_fill(shapep, n);
    nan_fill(sig2p, *M); nan_fill(taup, *M); nan_fill(alphap, *M); nan_fill(runtime, 1);
  }
}

void bridge_reg_stable(double* betap, double* lambdap, double* sig2p, double* taup, double* alphap, const double* yp,
                       const double* Xp, const double* sig2_shape, const double* sig2_scale, const double* nu_shape,
                       const double* nu_rate, const double* alpha_a, const double* alpha_b, const double* true_sig2,
                       const double* true_tau, const double* true_alpha, const int* P, const int* N, const int* M, const int* burn,
                       double* runtime, const int* ortho) {
  bb_hyper h{*sig2_shape, *sig2_scale, *nu_shape, *nu_rate, *alpha_a, *alpha_b, *true_sig2, *true_tau, *true_alpha, 0};
  const int orth = (*reinterpret_cast<const unsigned char*>(ortho)) != 0;
  const int rc = bb_bridge_batched(BB_STB, orth, Xp, yp, *N, *P, &h, 1, g_seed, 0u, *burn, *M, 1, g_device, betap, lambdap, nullptr,
                                   nullptr, sig2p, taup, alphap, nullptr, runtime);
  if (rc) {
    std::fprintf(stderr, "bridge_reg_stable (bayesbridge_b200): %s\nAborting Gibbs sampler.\n", g_err.c_str());
    const size_t n = (size_t)(*P) * (*M);
    nan_fill(betap, n); nan_fill(lambdap, n); nan_fill(sig2p, *M); nan_fill(taup, *M); nan_fill(alphap, *M); nan_fill(runtime, 1);
  }
}

// ---- the stand-alone scalar samplers (BridgeWrapper.h:230-244; bodies BridgeWrapper.cpp:762-984) ----
namespace {
void scalar_call(const char* name, int kind, int n, const double* p0, const double* p1, const double* p2, const double* p3, double* x) {
  if (n <= 0) return;
  const int rc = guarded([&] {
    BB_CUDA_CHECK(cudaSetDevice(g_device));
    cudaStream_t st = cudaStreamPerThread;
    const double* src[4] = {p0, p1, p2, p3};
    struct Bufs { double* d[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}; uint32_t* status = nullptr;
                  ~Bufs() { for (double* q : d) dfree(q); dfree(status); } } b;
    for (int k = 0; k < 4; ++k)
      if (src[k]) {
        b.d[k] = dalloc<double>((size_t)n);
        BB_CUDA_CHECK(cudaMemcpyAsync(b.d[k], src[k], (size_t)n * sizeof(double), cudaMemcpyHostToDevice, st));
      }
    b.d[4] = dalloc<double>((size_t)n);
    b.status = dalloc<uint32_t>(1);
    BB_CUDA_CHECK(cudaMemsetAsync(b.status, 0, sizeof(uint32_t), st));
    const bb::RngKey key{(uint32_t)g_seed, (uint32_t)(g_seed >> 32)};
    launched(bb::launch_scalar(kind, n, b.d[0], b.d[1], b.d[2], b.d[3], b.d[4], key, g_scalar_calls.fetch_add(1), b.status, st));
    check_launch();
    uint32_t status = 0;
    BB_CUDA_CHECK(cudaMemcpyAsync(x, b.d[4], (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
    BB_CUDA_CHECK(cudaMemcpyAsync(&status, b.status, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
    // the reference prints its own warnings (BridgeWrapper.cpp:899-934); a dropped status word would hand NaNs back silently
    if (status & bb::ST_TRIAL_CAP) std::fprintf(stderr, "%s (bayesbridge_b200): warning: a rejection sampler reached its trial cap; the affected draws are NaN.\n", name);
    if (status & bb::ST_EMPTY_INTERVAL) std::fprintf(stderr, "%s (bayesbridge_b200): warning: left > right for some items; the left bound was returned.\n", name);
    if (status & bb::ST_NAN) std::fprintf(stderr, "%s (bayesbridge_b200): warning: NaN parameters for some items.\n", name);
    g_scalar_status.store(status);
  });
  if (rc) {
    std::fprintf(stderr, "%s (bayesbridge_b200): %s\n", name, g_err.c_str());
    nan_fill(x, (size_t)n);
  }
}
}  // namespace

void rtexpon_rate_left(double* x, double* left, double* rate, int* num) {
  scalar_call("rtexpon_rate_left", bb::SC_TEXPON_LEFT, *num, left, rate, nullptr, nullptr, x);
}
void rtexpon_rate_both(double* x, double* left, double* right, double* rate, int* num) {
  scalar_call("rtexpon_rate_both", bb::SC_TEXPON_BOTH, *num, left, right, rate, nullptr, x);
}
void rtexpon_rate(double* x, double* left, double* right, double* rate, int* num) {
  scalar_call("rtexpon_rate", bb::SC_TEXPON, *num, left, right, rate, nullptr, x);
}
void rtnorm_left(double* x, double* left, double* mu, double* sig, int* num) {
  scalar_call("rtnorm_left", bb::SC_TNORM_LEFT, *num, left, mu, sig, nullptr, x);
}
void rtnorm_both(double* x, double* left, double* right, double* mu, double* sig, int* num) {
  scalar_call("rtnorm_both", bb::SC_TNORM_BOTH, *num, left, right, mu, sig, x);
}
void rtnorm(double* x, double* left, double* right, double* mu, double* sig, int* num) {
  scalar_call("rtnorm", bb::SC_TNORM, *num, left, right, mu, sig, x);
}
void rrtgamma_rate(double* x, double* scale, double* rate, double* right_t, int* num) {
  scalar_call("rrtgamma_rate", bb::SC_RTGAMMA, *num, scale, rate, right_t, nullptr, x);
}
void retstable_LD(double* x, double* alpha, double* V0, double* h, int* num) {
  scalar_call("retstable_LD", bb::SC_RETSTABLE, *num, alpha, V0, h, nullptr, x);
}

}  // extern "C"
