// bb_scalar.cu -- the stand-alone scalar samplers the reference's shared object also exports
// (/root/reference/Code/C/BridgeWrapper.h:230-244, bodies BridgeWrapper.cpp:762-984): truncated
// exponential, truncated normal, right-truncated gamma and the tilted stable law, vectorised over
// `num` parameter sets.  SURVEY.md section 8, row f3.
//
// One thread per item: item i draws from the Philox stream (key, chain 0, iter = call counter,
// stage ST_SCALAR, item i) through the same derived-draw definitions the sweep kernels use
// (bb_rng.cuh, bb_retstable.cuh), documented in DESIGN.md section 3.
#include <math_constants.h>
#include "bb_launch.h"
#include "bb_retstable.cuh"

namespace bb {

// Trial cap of the right-truncated gamma (naive rejection from the untruncated law, Notes/bbnotes.tex:792-793): when
// P(X <= right) is tiny a thread would otherwise spin through 2^24 Philox blocks (seconds).  The CPU checker carries the same cap.
constexpr uint32_t RTG_CAP = 1u << 16;

// sequential Gamma(shape, rate) starting at block k (advances k): the block order of the scalar draw spec
__device__ inline double gamma_rate_seq(const Stream& s, uint32_t& k, double shape, double rate, uint32_t& status) {
  if (shape == 1.0) return -log(s.unif(k++)) / rate;
  if (shape == 2.0) { double a, b; s.block(k++, a, b); return -log(a * b) / rate; }
  double boost = 1.0, sh = shape;
  if (shape < 1.0) { boost = pow(s.unif(k++), 1.0 / shape); sh = shape + 1.0; }
  const double d = sh - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
  while (k < K_CAP - 2) {
    const double x = s.norm(k++);
    const double u = s.unif(k++);
    double v = 1.0 + c * x;
    if (v > 0.0) {
      v = v * v * v;
      if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) return boost * d * v / rate;
    }
  }
  status |= ST_TRIAL_CAP;
  return nan("");
}

__device__ inline double tnorm_left(const Stream& s, double left, double mu, double sd, uint32_t& status) {
  const double out = mu + sd * tnorm_std_scalar(s, (left - mu) / sd, CUDART_INF, status);
  return out < left ? left : out;
}

__device__ inline double texpon_left(const Stream& s, double left, double rate) { return left + -log(s.unif(0)) / rate; }
__device__ inline double texpon_both(const Stream& s, double left, double right, double rate) {
  const double b = -expm1(rate * (left - right));   // 1 - exp(-rate (right - left))
  return left - log1p(-s.unif(0) * b) / rate;
}

__global__ void __launch_bounds__(128) scalar_kernel(const int kind, const int n, const double* __restrict__ p0,
                                                     const double* __restrict__ p1, const double* __restrict__ p2,
                                                     const double* __restrict__ p3, double* __restrict__ out,
                                                     const RngKey key, const uint32_t iter, uint32_t* status_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Stream s(key, 0u, iter, ST_SCALAR, (uint32_t)i);
  uint32_t status = 0;
  double x;
  switch (kind) {
    case SC_TEXPON_LEFT: x = texpon_left(s, p0[i], p1[i]); break;                         // BridgeWrapper.cpp:762-782
    case SC_TEXPON_BOTH: x = texpon_both(s, p0[i], p1[i], p2[i]); break;                  // :784-804
    case SC_TEXPON:                                                                       // :807-840
      x = isfinite(p1[i]) ? texpon_both(s, p0[i], p1[i], p2[i]) : texpon_left(s, p0[i], p2[i]);
      break;
    case SC_TNORM_LEFT: x = tnorm_left(s, p0[i], p1[i], p2[i], status); break;            // :846-866
    case SC_TNORM_BOTH: x = tnorm<false>(s, p0[i], p1[i], p2[i], p3[i], status); break;   // :868-888
    case SC_TNORM: {                                                                      // :890-946 (the non-R branch)
      const double l = p0[i], r = p1[i], mu = p2[i], sd = p3[i];
      if (isfinite(l) && isfinite(r)) x = tnorm<false>(s, l, r, mu, sd, status);
      else if (isfinite(l)) x = tnorm_left(s, l, mu, sd, status);
      else if (isfinite(r)) x = -1.0 * tnorm_left(s, -1.0 * r, -1.0 * mu, sd, status);
      else x = mu + sd * s.norm(0);
      break;
    }
    case SC_RTGAMMA: {                                                                    // :949-966; rejection, bbnotes.tex:792-793
      uint32_t k = 0;
      for (;;) {
        x = gamma_rate_seq(s, k, p0[i], p1[i], status);
        if (x <= p2[i] || x != x) break;
        if (k >= RTG_CAP - 64) { x = nan(""); status |= ST_TRIAL_CAP; break; }   // bounded cost: NaN + a warning from the host
      }
      break;
    }
    default: {                                                                            // SC_RETSTABLE, :968-984
      const double alpha = p0[i], V0 = p1[i], h = p2[i];
      x = retstable_scalar(s, h, alpha, status, V0);
      break;
    }
  }
  out[i] = x;
  if (status) atomicOr(status_out, status);
}

int launch_scalar(int kind, int n, const double* p0, const double* p1, const double* p2, const double* p3, double* out,
                  RngKey key, uint32_t iter, uint32_t* status, cudaStream_t st) {
  if (n <= 0) return 0;
  scalar_kernel<<<(n + 127) / 128, 128, 0, st>>>(kind, n, p0, p1, p2, p3, out, key, iter, status);
  return 1;
}

}  // namespace bb
