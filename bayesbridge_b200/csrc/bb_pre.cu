// bb_pre.cu -- host side of the one-time precompute kernels (bb_pre.cuh): Gram build, one-sided
// Jacobi SVD, derived layouts (the role of the BridgeRegression constructor, BridgeRegression.cpp:13-69).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <vector>
#include <mutex>
#include <string>
#include "bb_pre.cuh"

// Scratch of bb_gram_noalloc (the split-K partial tiles of gram_kernel): one file-scope array per GPU that loads this
// library, held until the process exits.  1184 tiles of 64 x 64 doubles = 37 MiB: enough for every P <= 3071 at any
// number of rows (slots = tiles x row splits <= 591 + tiles while the grid is being filled, = tiles beyond that).
// In HOST code this name is only a stand-in for the symbol -- the device address comes from cudaGetSymbolAddress, below.
constexpr int BB_GRAM_NOALLOC_MAX_P = 3071;
constexpr size_t BB_GRAM_NOALLOC_SLOTS = 1184;
static __device__ __align__(256) double g_gram_noalloc_scratch[BB_GRAM_NOALLOC_SLOTS * bb::GR_T * bb::GR_T];

namespace bb {

// G = [X y]'[X y] on the FP64 tensor cores.  X, y device pointers.
int gram_device(const double* X, const double* y, int N, int P, double* G, cudaStream_t st) {
  const int PA = P + 1;
  const int nt = (PA + GR_T - 1) / GR_T;
  const int ntile = nt * (nt + 1) / 2;
  // split K so that the grid covers the 148 SMs a few times over
  int nsplit = std::max(1, std::min((N + 4 * GR_K - 1) / (4 * GR_K), (148 * 4 + ntile - 1) / ntile));
  int rows = (N + nsplit - 1) / nsplit;
  rows = (rows + GR_K - 1) / GR_K * GR_K;
  nsplit = (N + rows - 1) / rows;
  if (nsplit < 1) nsplit = 1;
  double* part = dalloc<double>((size_t)nsplit * ntile * GR_T * GR_T);
  // TMA-staged operands whenever the bulk copies' alignment rules hold (N even, 16-byte aligned bases); BB_GRAM_LDG=1 forces
  // the LDG -> STS kernel (same arithmetic, same summation order: the two kernels produce identical bits)
  static const bool force_ldg = std::getenv("BB_GRAM_LDG") != nullptr;
  const bool tma_ok = !force_ldg && (N % 2 == 0) && ((uintptr_t)X % 16 == 0) && ((uintptr_t)y % 16 == 0);
  if (tma_ok) {
    static bool attr_set = false;
    if (!attr_set) { BB_CUDA_CHECK(cudaFuncSetAttribute(gram_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gram_tma_smem())); attr_set = true; }
    gram_tma_kernel<<<dim3(ntile, 1, nsplit), 288, gram_tma_smem(), st>>>(X, y, N, P, rows, part);
  } else {
    gram_kernel<<<dim3(ntile, 1, nsplit), 256, 0, st>>>(X, y, N, P, rows, part);
  }
  check_launch();
  gram_reduce_kernel<<<296, 256, 0, st>>>(part, P, nsplit, G);
  check_launch();
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
  dfree(part);
  return 2;
}

// Device address of g_gram_noalloc_scratch on the CURRENT device, looked up once per device.
static double* gram_noalloc_scratch_ptr() {
  static std::mutex mu;
  static double* ptr[64] = {};
  int dev = -1;
  BB_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) throw std::string("bb_gram_noalloc: device index out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    double* p = nullptr;
    BB_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_gram_noalloc_scratch));
    if (!p) throw std::string("bb_gram_noalloc: cudaGetSymbolAddress returned a null address");
    ptr[dev] = p;
  }
  return ptr[dev];
}

// gram_device without any allocation: the same two kernels with the same row split (so the result is bit-identical),
// the partial tiles in the static scratch.  gram_kernel writes every entry of every slot it is given before
// gram_reduce_kernel reads them, so nothing has to be cleared between calls; two calls that overlap on one GPU would race
// on the scratch.  The device must be current.
int gram_device_noalloc(const double* X, const double* y, int N, int P, double* G, cudaStream_t st) {
  if (N < 1 || P < 1) throw std::string("bb_gram_noalloc: n_rows and P must be positive");
  if (P > BB_GRAM_NOALLOC_MAX_P)
    throw std::string("bb_gram_noalloc: P = ") + std::to_string(P) + " exceeds the static scratch (P <= " +
        std::to_string(BB_GRAM_NOALLOC_MAX_P) + "); use bb_gram";
  // the split of gram_device (keep in step: tests/test_zz_no_alloc.py compares the two bit for bit)
  const int PA = P + 1;
  const int nt = (PA + GR_T - 1) / GR_T;
  const int ntile = nt * (nt + 1) / 2;
  int nsplit = std::max(1, std::min((N + 4 * GR_K - 1) / (4 * GR_K), (148 * 4 + ntile - 1) / ntile));
  int rows = (N + nsplit - 1) / nsplit;
  rows = (rows + GR_K - 1) / GR_K * GR_K;
  nsplit = (N + rows - 1) / rows;
  if (nsplit < 1) nsplit = 1;
  if ((size_t)nsplit * ntile > BB_GRAM_NOALLOC_SLOTS) throw std::string("bb_gram_noalloc: problem exceeds the static scratch; use bb_gram");
  double* part = gram_noalloc_scratch_ptr();
  gram_kernel<<<dim3(ntile, 1, nsplit), 256, 0, st>>>(X, y, N, P, rows, part);
  check_launch();
  gram_reduce_kernel<<<296, 256, 0, st>>>(part, P, nsplit, G);
  check_launch();
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
  return 2;
}

int gram_unpack(const double* G, int P, double* XX, double* Xy, double* yy_dev, cudaStream_t st) {
  gram_unpack_kernel<<<296, 256, 0, st>>>(G, P, XX, Xy, yy_dev); check_launch();
  return 1;
}

namespace {
// round-robin tournament schedule: n even, n-1 rounds of n/2 pairs
std::vector<int> tournament(int P, int& rounds, int& npairs) {
  const int n = P + (P & 1);
  rounds = n - 1; npairs = n / 2;
  std::vector<int> out((size_t)rounds * npairs * 2);
  std::vector<int> ring(n);
  for (int i = 0; i < n; ++i) ring[i] = i;
  for (int r = 0; r < rounds; ++r) {
    for (int k = 0; k < npairs; ++k) {
      int p = ring[k], q = ring[n - 1 - k];
      if (p > q) std::swap(p, q);
      if (q >= P) { p = -1; q = -1; }   // padding column
      out[((size_t)r * npairs + k) * 2] = p; out[((size_t)r * npairs + k) * 2 + 1] = q;
    }
    // rotate all but the first
    int last = ring[n - 1];
    for (int i = n - 1; i > 1; --i) ring[i] = ring[i - 1];
    ring[1] = last;
  }
  return out;
}
// the same tournament over column blocks of 8: pairs (I, J) with J = -1 for the padding block (the CTA then only
// rotates within block I)
std::vector<int> block_tournament(int P, int& rounds, int& npairs) {
  const int nb = (P + BJ_B - 1) / BJ_B;
  const int n = std::max(2, nb + (nb & 1));
  rounds = n - 1; npairs = n / 2;
  std::vector<int> out((size_t)rounds * npairs * 2);
  std::vector<int> ring(n);
  for (int i = 0; i < n; ++i) ring[i] = i;
  for (int r = 0; r < rounds; ++r) {
    for (int k = 0; k < npairs; ++k) {
      int p = ring[k], q = ring[n - 1 - k];
      if (p > q) std::swap(p, q);
      if (q >= nb) q = -1;
      if (p >= nb) p = -1;
      out[((size_t)r * npairs + k) * 2] = p; out[((size_t)r * npairs + k) * 2 + 1] = q;
    }
    int last = ring[n - 1];
    for (int i = n - 1; i > 1; --i) ring[i] = ring[i - 1];
    ring[1] = last;
  }
  return out;
}
}  // namespace

// One-sided Jacobi on W (M x P, destroyed): on return W = W0 V, V orthogonal, columns of W orthogonal.
int jacobi_device(double* W, double* V, int M, int P, cudaStream_t st) {
  int rounds, npairs, n = 0;
  // block form (one cooperative launch, 8-column blocks) unless BB_JACOBI_PAIRWISE=1 or the device cannot co-launch
  {
    const char* env = std::getenv("BB_JACOBI_PAIRWISE");
    int dev = 0, nsm = 0, per_sm = 0, coop = 0;
    BB_CUDA_CHECK(cudaGetDevice(&dev));
    BB_CUDA_CHECK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
    BB_CUDA_CHECK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    BB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jacobi_block_kernel, 256, 0));
    std::vector<int> bs = block_tournament(P, rounds, npairs);
    if (!(env && env[0] == '1') && coop && npairs <= nsm * per_sm) {
      constexpr int MAX_SWEEPS = 40;
      int* d_sched = dalloc<int>(bs.size());
      BB_CUDA_CHECK(cudaMemcpyAsync(d_sched, bs.data(), bs.size() * sizeof(int), cudaMemcpyHostToDevice, st));
      if (V) { set_identity_kernel<<<296, 256, 0, st>>>(V, P); check_launch(); ++n; }
      unsigned long long* d_convs = dalloc<unsigned long long>(MAX_SWEEPS);
      unsigned int* d_counter = dalloc<unsigned int>(1);
      const int nblocks = (P + BJ_B - 1) / BJ_B;
      unsigned int* d_ver = dalloc<unsigned int>(nblocks);
      BB_CUDA_CHECK(cudaMemsetAsync(d_convs, 0, MAX_SWEEPS * sizeof(unsigned long long), st));
      BB_CUDA_CHECK(cudaMemsetAsync(d_counter, 0, sizeof(unsigned int), st));
      BB_CUDA_CHECK(cudaMemsetAsync(d_ver, 0, nblocks * sizeof(unsigned int), st));
      // columns whose squared norm is below 1e-29 of the largest are numerically zero (the callers zero every singular
      // value below 1e-13 d_max, 1e-7 d_max on the Gram path, i.e. squared column norms below 1e-26 / 1e-28 of the largest)
      double floor2 = 0.0;
      {
        double* nrm = dalloc<double>(P);
        colnorm_kernel<<<P, 128, 0, st>>>(W, M, P, nrm); check_launch(); ++n;
        std::vector<double> h(P);
        BB_CUDA_CHECK(cudaMemcpyAsync(h.data(), nrm, P * sizeof(double), cudaMemcpyDeviceToHost, st));
        BB_CUDA_CHECK(cudaStreamSynchronize(st));
        dfree(nrm);
        double mx = 0.0; for (double v : h) mx = std::max(mx, v);
        floor2 = 1e-29 * mx * mx;
      }
      int max_sweeps = MAX_SWEEPS;
      void* args[] = {&W, &V, &M, &P, &d_sched, &rounds, &npairs, &d_convs, &d_counter, &d_ver, &floor2, &max_sweeps};
      BB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)jacobi_block_kernel, dim3(npairs), dim3(256), args, 0, st));
      ++n;
      BB_CUDA_CHECK(cudaStreamSynchronize(st));
#ifdef BB_PHASES
      {
        unsigned long long h[8];
        BB_CUDA_CHECK(cudaMemcpyFromSymbol(h, bj_phase_cyc, sizeof(h)));
        static unsigned long long prev[8] = {0};
        const char* nm[8] = {"gram", "reduce+cos", "inner sweep", "apply", "barrier+sched", "", "", ""};
        std::printf("  [jacobi phases, CTA 0, M %d P %d]", M, P);
        for (int k = 0; k < 5; ++k) std::printf(" %s %.0f kcyc", nm[k], (double)(h[k] - prev[k]) / 1e3);
        std::printf("\n");
        for (int k = 0; k < 8; ++k) prev[k] = h[k];
      }
#endif
      if (std::getenv("BB_JACOBI_VERBOSE")) {
        unsigned long long hc[MAX_SWEEPS];
        BB_CUDA_CHECK(cudaMemcpy(hc, d_convs, sizeof(hc), cudaMemcpyDeviceToHost));
        int ns = 0; while (ns < MAX_SWEEPS && hc[ns] != 0ull) ++ns;
        std::printf("  [jacobi] block form: M %d P %d, %d rounds x %d pairs, %d sweeps rotated; max |cos| per sweep:", M, P, rounds, npairs, ns);
        for (int k = 0; k < ns; ++k) { double v; std::memcpy(&v, &hc[k], sizeof(v)); std::printf(" %.1e", v); }
        std::printf("\n");
      }
      dfree(d_sched); dfree(d_convs); dfree(d_counter); dfree(d_ver);
      return n;
    }
  }
  std::vector<int> sched = tournament(P, rounds, npairs);
  int* d_sched = dalloc<int>(sched.size());
  BB_CUDA_CHECK(cudaMemcpyAsync(d_sched, sched.data(), sched.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  unsigned long long* d_conv = dalloc<unsigned long long>(1);
  if (V) { set_identity_kernel<<<296, 256, 0, st>>>(V, P); check_launch(); ++n; }
  // one cooperative launch when every pair's CTA can be resident at once (always, up to P ~ 2000)
  constexpr int MAX_SWEEPS = 40;
  int dev = 0, nsm = 0, per_sm = 0, coop = 0;
  BB_CUDA_CHECK(cudaGetDevice(&dev));
  BB_CUDA_CHECK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  BB_CUDA_CHECK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  BB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jacobi_persistent_kernel, 256, 0));
  if (coop && P > 1 && npairs <= nsm * per_sm) {
    unsigned long long* d_convs = dalloc<unsigned long long>(MAX_SWEEPS);
    unsigned int* d_counter = dalloc<unsigned int>(1);
    BB_CUDA_CHECK(cudaMemsetAsync(d_convs, 0, MAX_SWEEPS * sizeof(unsigned long long), st));
    BB_CUDA_CHECK(cudaMemsetAsync(d_counter, 0, sizeof(unsigned int), st));
    int max_sweeps = MAX_SWEEPS;
    void* args[] = {&W, &V, &M, &P, &d_sched, &rounds, &npairs, &d_convs, &d_counter, &max_sweeps};
    BB_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)jacobi_persistent_kernel, dim3(npairs), dim3(256), args, 0, st));
    ++n;
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
    dfree(d_convs); dfree(d_counter);
  } else
  for (int sweep = 0; sweep < MAX_SWEEPS && P > 1; ++sweep) {
    BB_CUDA_CHECK(cudaMemsetAsync(d_conv, 0, sizeof(unsigned long long), st));
    for (int r = 0; r < rounds; ++r) {
      jacobi_round_kernel<<<npairs, 256, 0, st>>>(W, V, M, P, d_sched + (size_t)r * npairs * 2, d_conv);
    }
    check_launch(); n += rounds;
    unsigned long long conv = 0;
    BB_CUDA_CHECK(cudaMemcpyAsync(&conv, d_conv, sizeof(conv), cudaMemcpyDeviceToHost, st));
    BB_CUDA_CHECK(cudaStreamSynchronize(st));
    if (conv == 0) break;   // no pair needed a rotation in this sweep
  }
  dfree(d_sched); dfree(d_conv);
  return n;
}

double resid_device(const double* X, const double* y, const double* b, int N, int P, cudaStream_t st, int* launches) {
  const int nb = std::max(1, std::min(592, (N + 255) / 256));
  double* part = dalloc<double>(nb);
  resid_kernel<<<nb, 256, 0, st>>>(X, y, b, N, P, part); check_launch();
  if (launches) *launches += 1;
  std::vector<double> h(nb);
  BB_CUDA_CHECK(cudaMemcpyAsync(h.data(), part, nb * sizeof(double), cudaMemcpyDeviceToHost, st));
  BB_CUDA_CHECK(cudaStreamSynchronize(st));
  dfree(part);
  double s = 0; for (double v : h) s += v;
  return s;
}

int zero_upper_device(double* A, int P, cudaStream_t st) {
  zero_upper_kernel<<<296, 256, 0, st>>>(A, P); check_launch();
  return 1;
}
int colscale_device(double* W, int M, int P, const double* nrm, cudaStream_t st) {
  colscale_kernel<<<P, 128, 0, st>>>(W, M, P, nrm); check_launch();
  return 1;
}
int colnorm_device(const double* W, int M, int P, double* nrm, cudaStream_t st) {
  colnorm_kernel<<<P, 128, 0, st>>>(W, M, P, nrm); check_launch();
  return 1;
}
int derive_device(const double* V, const int* perm, const double* dsrc, const double* Xy, int P, int ldv, double* tV, double* Vc,
                  double* VQ, double* d, double* a, double* mean, double* d2, double* zc, cudaStream_t st) {
  derive_kernel<<<P, 128, 0, st>>>(V, perm, dsrc, Xy, P, ldv, tV, Vc, VQ, d, a, mean, d2, zc); check_launch();
  return 1;
}
int derive_from_tv_device(const double* tV, const double* d, const double* a, int P, int ldv, double* Vc, double* VQ, double* mean,
                          double* d2, double* zc, cudaStream_t st) {
  derive_from_tv_kernel<<<P, 128, 0, st>>>(tV, d, a, P, ldv, Vc, VQ, mean, d2, zc); check_launch();
  return 1;
}
int vz_device(const double* Vc, const double* z, int P, int ldv, double* out, cudaStream_t st) {
  vz_kernel<<<(P + 127) / 128, 128, 0, st>>>(Vc, z, P, ldv, out); check_launch();
  return 1;
}
int grad_device(const double* XX, const double* Xy, const double* b, int P, double* g, cudaStream_t st) {
  grad_kernel<<<(P + 127) / 128, 128, 0, st>>>(XX, Xy, b, P, g); check_launch();
  return 1;
}

}  // namespace bb
