// bb_running.cu -- RUNNING sum.stat (Code/R/PublicBenchmark.R:112-134; SURVEY section 7.1 `K_out`, section 8 row f2): the
// summaries of a run of any length from a draw buffer of one chunk.  The sweep kernels store a chunk of draws, one launch
// folds the chunk into per-series accumulators, the buffer is reused.  What is kept per (chain, coefficient) series:
//
//   first      x_0, the pilot shift: everything below is accumulated on y_t = x_t - x_0 (no cancellation in the moments)
//   S1         sum_t y_t
//   C[k]       sum_{t} y_t y_{t+k}, k = 0..RS_K          (the raw lagged products of the whole series)
//   head[k]    y_k for k < RS_K, tail[k]: the last RS_K values (a ring): what the centring of C[k] needs at the two ends,
//              and what the next chunk needs to continue the lagged products across the chunk boundary
//
// from which the END step recovers exactly the quantities of the stored-draw summary (bb_summary.cu): mean, sd and the
// mean-centred autocovariances  acov_k = (1/n) sum_t (y_t - m)(y_{t+k} - m)
//                                      = (1/n) [C_k - m (2 S1 - sum of the last k - sum of the first k values) + (n - k) m^2],
// then R's ar.yw (Levinson-Durbin, order by AIC) and ESS = n var / S(0) as coda::effectiveSize defines it.  The lag window is
// floor(10 log10 n) like R's, capped at RS_K = 48 (n <= 63000 draws per chain keep R's window; longer runs use 48 lags).
//
// Quantiles (Q.1, Q.9 of sum.stat) come from a fixed-size sketch that merges over chains and GPUs: per coefficient a histogram of
// RS_NB equal bins over [m0 - 10 s0, m0 + 10 s0] (m0, s0: pooled mean and sd of the FIRST chunk) plus under / overflow counts,
// shared by all chains; the quantile is read by linear interpolation inside its bin.  Resolution 20 s0 / 2048 = 0.01 sd; the
// stored-draw summary (exact order statistics) is the check in tests/test_gpu_summary.py.
#include <cmath>
#include <string>
#include "bb_launch.h"

namespace bb {

constexpr int RS_K = 48;                       // lag capacity
constexpr int RS_NB = 2048;                    // histogram bins per coefficient
constexpr int RS_J = 4;                        // coefficients per CTA: 32-byte sectors of the draw buffer are used in full
constexpr int RS_T = 128;
constexpr int RS_STATE = 2 + (RS_K + 1) + 2 * RS_K;   // doubles per series: first, S1, C[0..K], head[K], tail[K]

size_t running_state_doubles(int nchains, int P) { return (size_t)nchains * P * RS_STATE; }
size_t running_hist_words(int P) { return (size_t)P * (RS_NB + 2); }

// ---- pooled mean / sd of the first chunk -> histogram range [lo, lo + RS_NB w) per coefficient ----
__global__ void __launch_bounds__(256) running_range_kernel(const double* __restrict__ draws, int P, long long chain_stride, long long slot0,
                                                            int m, int nchains, double* __restrict__ range /* [P][2] = lo, 1/w */) {
  const int j = blockIdx.x, tid = threadIdx.x;
  __shared__ double r1[8], r2[8];
  const long long total = (long long)nchains * m;
  const double x0 = __ldg(draws + (size_t)slot0 * P + j);
  double s = 0.0, q = 0.0;
  for (long long r = tid; r < total; r += 256) {
    const long long c = r / m, t = r - c * m;
    const double d = __ldg(draws + ((size_t)c * chain_stride + slot0 + t) * P + j) - x0;
    s += d; q = fma(d, d, q);
  }
  s = warp_sum(s); q = warp_sum(q);
  if ((tid & 31) == 0) { r1[tid >> 5] = s; r2[tid >> 5] = q; }
  __syncthreads();
  if (tid == 0) {
    s = 0.0; q = 0.0;
    for (int w = 0; w < 8; ++w) { s += r1[w]; q += r2[w]; }
    const double mean = s / total;
    double sd = total > 1 ? sqrt(fmax(q - total * mean * mean, 0.0) / (total - 1)) : 0.0;
    if (!(sd > 0.0)) sd = fmax(fabs(x0 + mean), 1.0) * 1e-12;      // a constant series: any positive width will do
    const double w = 20.0 * sd / RS_NB;
    range[2 * j] = x0 + mean - 10.0 * sd;
    range[2 * j + 1] = 1.0 / w;
  }
}

// ---- fold one chunk (slots [slot0, slot0 + m) of every chain) into the accumulators ----
// CTA = RS_J consecutive coefficients of one chain; n_before = draws already folded.
__global__ void __launch_bounds__(RS_T) running_add_kernel(const double* __restrict__ draws, int P, long long chain_stride, long long slot0,
                                                           int m, long long n_before, double* __restrict__ state,
                                                           const double* __restrict__ range, unsigned int* __restrict__ hist) {
  extern __shared__ double ys[];                 // [RS_J][RS_K + m]: the previous tail, then the chunk (shifted values)
  const int j0 = blockIdx.x * RS_J, ch = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nj = min(RS_J, P - j0);
  const int LDY = RS_K + m;
  __shared__ double first_s[RS_J];
  double* st0 = state + ((size_t)ch * P + j0) * RS_STATE;
  if (tid < nj) first_s[tid] = n_before == 0 ? __ldg(draws + ((size_t)ch * chain_stride + slot0) * P + j0 + tid) : st0[(size_t)tid * RS_STATE];
  __syncthreads();
  // previous tail (oldest first) and the chunk
  const int have = (int)(n_before < RS_K ? n_before : RS_K);     // valid entries of the tail
  for (int e = tid; e < nj * RS_K; e += RS_T) {
    const int jj = e / RS_K, k = e - jj * RS_K;                  // k-th newest is at ring position (n_before - 1 - k) mod RS_K
    double v = 0.0;
    if (k < have) v = st0[(size_t)jj * RS_STATE + 2 + (RS_K + 1) + RS_K + (int)((n_before - 1 - k) % RS_K)];
    ys[jj * LDY + RS_K - 1 - k] = v;
  }
  for (int e = tid; e < m * RS_J; e += RS_T) {
    const int t = e / RS_J, jj = e - t * RS_J;
    if (jj < nj) {
      const double x = __ldg(draws + ((size_t)ch * chain_stride + slot0 + t) * P + j0 + jj);
      ys[jj * LDY + RS_K + t] = x - first_s[jj];
      if (hist) {                                                // quantile sketch, shared by all chains
        const double u = (x - range[2 * (j0 + jj)]) * range[2 * (j0 + jj) + 1];
        const int b = u < 0.0 ? 0 : (u >= (double)RS_NB ? RS_NB + 1 : 1 + (int)u);
        atomicAdd(hist + (size_t)(j0 + jj) * (RS_NB + 2) + b, 1u);
      }
    }
  }
  __syncthreads();
  // lagged products: C[k] += sum_{t in chunk} y_{t-k} y_t  (t - k may reach into the previous tail; older than that: no term)
  for (int e = warp; e < nj * (RS_K + 1); e += RS_T / 32) {
    const int jj = e / (RS_K + 1), k = e - jj * (RS_K + 1);
    const double* y = ys + jj * LDY + RS_K;
    const int t_lo = (k > have) ? k - have : 0;                  // y_{t-k} exists for t >= k - have
    double a = 0.0;
    for (int t = t_lo + lane; t < m; t += 32) a = fma(y[t - k], y[t], a);
    a = warp_sum(a);
    if (lane == 0) st0[(size_t)jj * RS_STATE + 2 + k] += a;
  }
  // S1, first, head, tail
  for (int jj = warp; jj < nj; jj += RS_T / 32) {
    const double* y = ys + jj * LDY + RS_K;
    double s = 0.0;
    for (int t = lane; t < m; t += 32) s += y[t];
    s = warp_sum(s);
    double* st = st0 + (size_t)jj * RS_STATE;
    if (lane == 0) { if (n_before == 0) { st[0] = first_s[jj]; st[1] = s; } else st[1] += s; }
    for (int t = lane; t < m; t += 32) {
      const long long g = n_before + t;
      if (g < RS_K) st[2 + (RS_K + 1) + (int)g] = y[t];                           // head
      if (t >= m - RS_K) st[2 + (RS_K + 1) + RS_K + (int)(g % RS_K)] = y[t];       // tail ring (the last RS_K of the chunk win)
    }
  }
}

// ---- END: per-chain mean, sd, ESS from the accumulators (layout of bb_summary.cu's per_chain: [chain][5][P]) ----
__global__ void __launch_bounds__(128) running_finish_kernel(const double* __restrict__ state, int P, int nchains, long long n,
                                                            double* __restrict__ per_chain) {
  const int e = blockIdx.x * 128 + threadIdx.x;
  if (e >= nchains * P) return;
  const int ch = e / P, j = e - ch * P;
  const double* st = state + (size_t)e * RS_STATE;
  const double* C = st + 2; const double* head = C + RS_K + 1; const double* tail = head + RS_K;
  const double nn = (double)n;
  const double m = st[1] / nn;
  int K = (int)floor(10.0 * log10(nn));
  if (K > RS_K) K = RS_K;
  if (K > n - 1) K = (int)(n - 1);
  double acov[RS_K + 1];
  double sh = 0.0, stl = 0.0;                    // sums of the first k / last k values
  for (int k = 0; k <= K; ++k) {
    if (k > 0) { sh += head[k - 1]; stl += tail[(int)((n - k) % RS_K)]; }
    acov[k] = (C[k] - m * (2.0 * st[1] - stl - sh) + (nn - k) * m * m) / nn;
  }
  const double var = n > 1 ? acov[0] * nn / (nn - 1.0) : 0.0;
  double ess = 0.0;
  if (var > 0.0 && n >= 4 && acov[0] > 0.0) {    // Levinson-Durbin with AIC order selection, as summary_series_kernel
    double phi[RS_K], phin[RS_K], sp_best = 0.0;
    double v = acov[0];
    double best_aic = nn * log(v), best_v = v; int best_k = 0;
    for (int k = 1; k <= K; ++k) {
      double num = acov[k];
      for (int i = 0; i < k - 1; ++i) num -= phi[i] * acov[k - 1 - i];
      const double ref = num / v;
      for (int i = 0; i < k - 1; ++i) phin[i] = phi[i] - ref * phi[k - 2 - i];
      phin[k - 1] = ref;
      double sp = 0.0;
      for (int i = 0; i < k; ++i) { phi[i] = phin[i]; sp += phin[i]; }
      v *= (1.0 - ref * ref);
      if (!(v > 0.0)) break;
      const double aic = nn * log(v) + 2.0 * k;
      if (aic < best_aic) { best_aic = aic; best_v = v; best_k = k; sp_best = sp; }
    }
    const double vv = best_v * nn / (nn - (best_k + 1));
    const double s0 = vv / ((1.0 - sp_best) * (1.0 - sp_best));
    ess = nn * var / s0;
  }
  double* out = per_chain + (size_t)ch * 5 * P;
  const double nan_ = __longlong_as_double(0x7ff8000000000000LL);
  out[j] = st[0] + m; out[(size_t)P + j] = sqrt(var); out[2 * (size_t)P + j] = ess;
  out[3 * (size_t)P + j] = nan_; out[4 * (size_t)P + j] = nan_;      // per-chain quantiles are not kept in running mode
}

// ---- pooled Q.1 / Q.9 from the shared histogram: rank h = q (total - 1) as R's type 7, linear inside the bin ----
__global__ void __launch_bounds__(128) running_quantile_kernel(const unsigned int* __restrict__ hist, const double* __restrict__ range, int P,
                                                              double* __restrict__ pooled) {
  const int j = blockIdx.x * 128 + threadIdx.x;
  if (j >= P) return;
  const unsigned int* h = hist + (size_t)j * (RS_NB + 2);
  unsigned long long total = 0;
  for (int b = 0; b < RS_NB + 2; ++b) total += h[b];
  const double lo = range[2 * j], w = 1.0 / range[2 * j + 1];
  for (int q = 0; q < 2; ++q) {
    const double target = (double)(total - 1) * (q == 0 ? 0.1 : 0.9) + 0.5;     // position among `total` sorted values, in counts
    unsigned long long cum = 0; int b = 0;
    while (b < RS_NB + 1 && (double)(cum + h[b]) < target) { cum += h[b]; ++b; }
    double val;
    if (b == 0) val = lo;                                        // below the range: the edge (the sketch saturates)
    else if (b == RS_NB + 1) val = lo + RS_NB * w;
    else val = lo + ((b - 1) + (h[b] ? (target - (double)cum) / (double)h[b] : 0.5)) * w;
    pooled[(5 + q) * (size_t)P + j] = val;
  }
}

__global__ void running_zero_kernel(unsigned int* hist, size_t n) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) hist[e] = 0u;
}

// pooled moments / ESS over chains: summary_pool_kernel of bb_summary.cu
void summary_pool_launch(const double* per_chain, int P, int nchains, long long n, double runtime_sec, double* pooled, cudaStream_t st);

int running_add_device(const double* draws, int P, long long chain_stride, long long slot0, int m, int nchains, long long n_before,
                       double* state, double* range, unsigned int* hist, cudaStream_t st) {
  if (m < 1) return 0;
  int n = 0;
  if (n_before == 0) {
    BB_CUDA_CHECK(cudaMemsetAsync(state, 0, running_state_doubles(nchains, P) * sizeof(double), st));
    running_zero_kernel<<<296, 256, 0, st>>>(hist, running_hist_words(P)); ++n;
    running_range_kernel<<<P, 256, 0, st>>>(draws, P, chain_stride, slot0, m, nchains, range); ++n;
  }
  const size_t smem = (size_t)RS_J * (RS_K + m) * sizeof(double);
  if (smem > 200 * 1024) throw std::string("running summary: chunk too long (at most about 6000 slots per call)");
  static size_t smem_set = 0;
  if (smem > smem_set) { BB_CUDA_CHECK(cudaFuncSetAttribute(running_add_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); smem_set = smem; }
  running_add_kernel<<<dim3((P + RS_J - 1) / RS_J, nchains), RS_T, smem, st>>>(draws, P, chain_stride, slot0, m, n_before, state, range, hist); ++n;
  check_launch();
  return n;
}

int running_end_device(const double* state, const double* range, const unsigned int* hist, int P, int nchains, long long n,
                       double runtime_sec, double* per_chain_dev, double* pooled_dev, cudaStream_t st) {
  if (n < 2) throw std::string("running summary: need at least 2 draws per chain");
  running_finish_kernel<<<(nchains * P + 127) / 128, 128, 0, st>>>(state, P, nchains, n, per_chain_dev);
  summary_pool_launch(per_chain_dev, P, nchains, n, runtime_sec, pooled_dev, st);
  running_quantile_kernel<<<(P + 127) / 128, 128, 0, st>>>(hist, range, P, pooled_dev);
  check_launch();
  return 3;
}

}  // namespace bb
