// bb_summary.cu -- sum.stat on the device (Code/R/PublicBenchmark.R:112-134): mean, sd, ESS, ESS.sec, t, Q.1, Q.9 of
// the stored draws, so that a run of thousands of chains returns 7 P numbers instead of chains x samples x P draws
// (SURVEY section 8, row f2).
//
//   summary_series_kernel : one CTA per (coefficient, chain).  The series is staged in shared memory, centred, its
//                           autocovariances up to lag floor(10 log10 n) are taken, thread 0 runs Levinson-Durbin with
//                           AIC order selection (R's ar.yw, which coda::effectiveSize calls: ESS = n var / S(0),
//                           S(0) = v / (1 - sum phi)^2), then the series is sorted in place (bitonic) for the type-7
//                           quantiles R's quantile() returns.  Same definition as bayesbridge_b200/ess.py (ess_ar).
//   summary_pool_kernel   : per coefficient over chains: pooled mean / sd, ESS summed over chains.
//   summary_select_kernel : pooled Q.1 / Q.9 over ALL chains' draws of a coefficient, exact, by radix selection on the
//                           order-preserving integer image of the doubles (8 passes of 8 bits + one pass for the upper
//                           neighbour), four coefficients per CTA so that every 32-byte sector read is used in full.
#include <algorithm>
#include <cmath>
#include <string>
#include <vector>
#include "bb_launch.h"

namespace bb {

constexpr int SM_T = 128;          // threads of the series kernel
constexpr int SM_MAXLAG = 48;      // floor(10 log10 n) <= 42 for n <= 16384
constexpr int SM_MAXN = 16384;     // series length limit of the on-device summary (128 KB of shared memory)

__device__ __forceinline__ double block_sum128(double v, double* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  return red[0] + red[1] + red[2] + red[3];
}

// per_chain: [nchains][5][P] = mean, sd (n-1), ESS, Q.1, Q.9
__global__ void __launch_bounds__(SM_T) summary_series_kernel(const double* __restrict__ draws, int P, long long chain_stride,
                                                             long long slot0, int n, int npow2, double* __restrict__ per_chain) {
  extern __shared__ double xs[];          // [npow2]
  __shared__ double red[4];
  __shared__ double acov[SM_MAXLAG + 1];
  __shared__ double phi[SM_MAXLAG], phin[SM_MAXLAG], phib[SM_MAXLAG];
  __shared__ double res[2];
  const int j = blockIdx.x, ch = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double* src = draws + ((size_t)ch * chain_stride + slot0) * P + j;
  double s = 0.0;
  for (int t = tid; t < npow2; t += SM_T) {
    const double v = t < n ? __ldg(src + (size_t)t * P) : 0.0;
    xs[t] = v; s += v;
  }
  const double mean = block_sum128(s, red) / n;
  double ss = 0.0;
  for (int t = tid; t < n; t += SM_T) { const double d = xs[t] - mean; xs[t] = d; ss = fma(d, d, ss); }
  ss = block_sum128(ss, red);
  const double var = n > 1 ? ss / (n - 1) : 0.0;
  // autocovariances (divisor n, as R's acf / ar.yw)
  const int K = min(n - 1, (int)floor(10.0 * log10((double)n)));
  for (int k = warp; k <= K; k += SM_T / 32) {
    double a = 0.0;
    for (int t = lane; t + k < n; t += 32) a = fma(xs[t], xs[t + k], a);
    a = warp_sum(a);
    if (lane == 0) acov[k] = a / n;
  }
  __syncthreads();
  if (tid == 0) {
    double ess = 0.0;
    if (var > 0.0 && n >= 4 && acov[0] > 0.0) {
      // Levinson-Durbin; AIC = n log(v) + 2 order, the smallest wins (ess.py:_yule_walker)
      double v = acov[0];
      double best_aic = n * log(v), best_v = v; int best_k = 0;
      for (int k = 1; k <= K; ++k) {
        double num = acov[k];
        for (int i = 0; i < k - 1; ++i) num -= phi[i] * acov[k - 1 - i];
        const double ref = num / v;
        for (int i = 0; i < k - 1; ++i) phin[i] = phi[i] - ref * phi[k - 2 - i];
        phin[k - 1] = ref;
        for (int i = 0; i < k; ++i) phi[i] = phin[i];
        v *= (1.0 - ref * ref);
        if (!(v > 0.0)) break;
        const double aic = n * log(v) + 2.0 * k;
        if (aic < best_aic) { best_aic = aic; best_v = v; best_k = k; for (int i = 0; i < k; ++i) phib[i] = phi[i]; }
      }
      double sp = 0.0;
      for (int i = 0; i < best_k; ++i) sp += phib[i];
      const double vv = best_v * n / (n - (best_k + 1));     // ar.yw's rescaling of the innovations variance
      const double s0 = vv / ((1.0 - sp) * (1.0 - sp));
      ess = n * var / s0;
    }
    res[0] = ess;
  }
  // sort the centred series (padding = +inf) and read the type-7 quantiles
  for (int t = n + tid; t < npow2; t += SM_T) xs[t] = __longlong_as_double(0x7ff0000000000000LL);
  __syncthreads();
  for (int k = 2; k <= npow2; k <<= 1) {
    for (int jj = k >> 1; jj > 0; jj >>= 1) {
      for (int t = tid; t < npow2; t += SM_T) {
        const int p = t ^ jj;
        if (p > t) {
          const double a = xs[t], b = xs[p];
          const bool up = (t & k) == 0;
          if ((a > b) == up) { xs[t] = b; xs[p] = a; }
        }
      }
      __syncthreads();
    }
  }
  if (tid < 2) {
    const double h = (n - 1) * (tid == 0 ? 0.1 : 0.9);
    const int lo = (int)floor(h), hi = min(lo + 1, n - 1);
    const double q = xs[lo] + (h - lo) * (xs[hi] - xs[lo]);
    double* out = per_chain + (size_t)ch * 5 * P;
    out[(3 + tid) * (size_t)P + j] = mean + q;
  }
  if (tid == 0) {
    double* out = per_chain + (size_t)ch * 5 * P;
    out[j] = mean; out[(size_t)P + j] = sqrt(var); out[2 * (size_t)P + j] = res[0];
  }
}

// pooled: [7][P] = mean, sd, ESS, ESS.sec, t, Q.1, Q.9 (the last two by summary_select_kernel); one warp per coefficient
__global__ void summary_pool_kernel(const double* __restrict__ per_chain, int P, int nchains, int n, double runtime_sec,
                                    double* __restrict__ pooled) {
  const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (j >= P) return;
  double sm = 0.0, se = 0.0;
  for (int c = lane; c < nchains; c += 32) { sm += per_chain[(size_t)c * 5 * P + j]; se += per_chain[((size_t)c * 5 + 2) * P + j]; }
  sm = warp_sum(sm); se = warp_sum(se);
  const double mean = sm / nchains;
  double ss = 0.0;
  for (int c = lane; c < nchains; c += 32) {
    const double mc = per_chain[(size_t)c * 5 * P + j], sd = per_chain[((size_t)c * 5 + 1) * P + j];
    ss += (n - 1.0) * sd * sd + n * (mc - mean) * (mc - mean);
  }
  ss = warp_sum(ss);
  if (lane == 0) {
    const double tot = (double)nchains * n;
    const double sd = tot > 1 ? sqrt(ss / (tot - 1.0)) : 0.0;
    pooled[j] = mean; pooled[(size_t)P + j] = sd; pooled[2 * (size_t)P + j] = se;
    pooled[3 * (size_t)P + j] = runtime_sec > 0 ? se / runtime_sec : __longlong_as_double(0x7ff8000000000000LL);
    pooled[4 * (size_t)P + j] = sd > 0 ? mean / sd : __longlong_as_double(0x7ff8000000000000LL);
  }
}

void summary_pool_launch(const double* per_chain, int P, int nchains, long long n, double runtime_sec, double* pooled, cudaStream_t st) {
  summary_pool_kernel<<<(P + 3) / 4, 128, 0, st>>>(per_chain, P, nchains, (int)n, runtime_sec, pooled);
}

// order-preserving map double -> uint64
__device__ __forceinline__ unsigned long long dkey(double x) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double dunkey(unsigned long long k) {
  const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffULL) : ~k;
  return __longlong_as_double((long long)b);
}

constexpr int SEL_J = 4;     // coefficients per CTA
constexpr int SEL_T = 256;

// value of rank `rank` (0-based, ascending) and of rank + 1 among the nchains * n draws of each of the CTA's coefficients
__global__ void __launch_bounds__(SEL_T) summary_select_kernel(const double* __restrict__ draws, int P, long long chain_stride,
                                                             long long slot0, int n, int nchains, double* __restrict__ pooled) {
  __shared__ unsigned int hist[SEL_J][256];
  __shared__ unsigned long long prefix[SEL_J];
  __shared__ unsigned long long krem[SEL_J];
  __shared__ unsigned long long nxt[SEL_J];
  __shared__ double qv[SEL_J][2];
  const int j0 = blockIdx.x * SEL_J, tid = threadIdx.x;
  const int nj = min(SEL_J, P - j0);
  const long long total = (long long)nchains * n;
  for (int q = 0; q < 2; ++q) {
    const double h = (double)(total - 1) * (q == 0 ? 0.1 : 0.9);
    const long long rank = (long long)floor(h);
    if (tid < SEL_J) { prefix[tid] = 0ull; krem[tid] = (unsigned long long)rank; }
    for (int pass = 7; pass >= 0; --pass) {
      for (int e = tid; e < SEL_J * 256; e += SEL_T) (&hist[0][0])[e] = 0u;
      __syncthreads();
      const int shift = 8 * pass;
      for (long long r = tid; r < total; r += SEL_T) {
        const long long c = r / n, s = r - c * n;
        const double* row = draws + ((size_t)c * chain_stride + slot0 + s) * P + j0;
        for (int jj = 0; jj < nj; ++jj) {
          const unsigned long long k = dkey(__ldg(row + jj));
          const bool on = pass == 7 || (k >> (shift + 8)) == (prefix[jj] >> (shift + 8));
          // warp-aggregated: the leading bytes of a coefficient's draws are nearly all equal, one atomic per distinct bucket
          const unsigned bucket = on ? (unsigned)((k >> shift) & 255u) : 256u;
          const unsigned peers = __match_any_sync(__activemask(), bucket);
          if (on && (threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&hist[jj][bucket], (unsigned)__popc(peers));
        }
      }
      __syncthreads();
      if (tid < nj) {
        unsigned long long rem = krem[tid]; int b = 0;
        while (b < 255 && rem >= hist[tid][b]) { rem -= hist[tid][b]; ++b; }
        krem[tid] = rem; prefix[tid] |= (unsigned long long)b << shift;
      }
      __syncthreads();
    }
    // prefix = key of the rank-th value; krem = its index among equal values.  The next order statistic is the same
    // value when more equal values follow, otherwise the smallest larger key.
    if (tid < SEL_J) nxt[tid] = ~0ull;
    for (int e = tid; e < SEL_J * 256; e += SEL_T) (&hist[0][0])[e] = 0u;
    __syncthreads();
    for (long long r = tid; r < total; r += SEL_T) {
      const long long c = r / n, s = r - c * n;
      const double* row = draws + ((size_t)c * chain_stride + slot0 + s) * P + j0;
      for (int jj = 0; jj < nj; ++jj) {
        const unsigned long long k = dkey(__ldg(row + jj));
        if (k == prefix[jj]) atomicAdd(&hist[jj][0], 1u);
        else if (k > prefix[jj]) atomicMin(&nxt[jj], k);
      }
    }
    __syncthreads();
    if (tid < nj) {
      const double x0 = dunkey(prefix[tid]);
      double x1 = x0;
      if (rank + 1 < total && krem[tid] + 1 >= hist[tid][0]) x1 = dunkey(nxt[tid]);
      qv[tid][q] = x0 + (h - (double)rank) * (x1 - x0);
    }
    __syncthreads();
  }
  if (tid < nj) { pooled[5 * (size_t)P + j0 + tid] = qv[tid][0]; pooled[6 * (size_t)P + j0 + tid] = qv[tid][1]; }
}

// Host side.  draws: a vector output of the draw buffer, [nchains][chain_stride][P]; per_chain_dev [nchains][5][P] scratch
// (always needed), pooled_dev [7][P].
int summary_device(const double* draws, int P, long long chain_stride, long long slot0, int n, int nchains, double runtime_sec,
                   double* per_chain_dev, double* pooled_dev, cudaStream_t st) {
  if (n < 2) throw std::string("summary: need at least 2 stored draws per chain");
  if (n > SM_MAXN) throw std::string("summary: at most 16384 stored draws per chain (thin the run or fetch the draws)");
  int npow2 = 2; while (npow2 < n) npow2 <<= 1;
  const size_t smem = (size_t)npow2 * sizeof(double);
  BB_CUDA_CHECK(cudaFuncSetAttribute(summary_series_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  summary_series_kernel<<<dim3(P, nchains), SM_T, smem, st>>>(draws, P, chain_stride, slot0, n, npow2, per_chain_dev);
  check_launch();
  summary_pool_kernel<<<(P + 3) / 4, 128, 0, st>>>(per_chain_dev, P, nchains, n, runtime_sec, pooled_dev);
  check_launch();
  summary_select_kernel<<<(P + SEL_J - 1) / SEL_J, SEL_T, 0, st>>>(draws, P, chain_stride, slot0, n, nchains, pooled_dev);
  check_launch();
  return 3;
}

}  // namespace bb
