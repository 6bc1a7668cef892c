/*
 * bbo_port.c -- ORACLE (test infrastructure; never shipped, never on the product path).
 * Plain-C restatement of the BayesBridge Gibbs sweep; see bbo_port.h.
 * Citations are relative to /root/reference/.
 */
#include "bbo_port.h"
#include <float.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------ linear algebra */

/* upper Cholesky A = U'U in place on the upper triangle of a col-major P x P copy.
   returns 0 ok, 1 if a pivot is <= 0 or the factor is numerically rank deficient. */
static int chol_upper(double *A, int P, double sing_tol) {
  double maxdiag = 0.0, minpiv = DBL_MAX;
  for (int j = 0; j < P; ++j) if (A[j + (size_t)j * P] > maxdiag) maxdiag = A[j + (size_t)j * P];
  for (int j = 0; j < P; ++j) {
    double *cj = A + (size_t)j * P;
    for (int i = 0; i < j; ++i) { /* U(i,j) = (A(i,j) - sum_k U(k,i)U(k,j)) / U(i,i) */
      const double *ci = A + (size_t)i * P;
      double s = cj[i];
      for (int k = 0; k < i; ++k) s -= ci[k] * cj[k];
      cj[i] = s / ci[i];
    }
    double s = cj[j];
    for (int k = 0; k < j; ++k) s -= cj[k] * cj[k];
    if (!(s > 0.0)) return 1;
    if (s < minpiv) minpiv = s;
    cj[j] = sqrt(s);
  }
  if (sing_tol > 0.0 && minpiv <= sing_tol * maxdiag) return 1;
  return 0;
}
/* solve U' x = b (trans=1) or U x = b (trans=0), U upper, col-major */
static void trsv_upper(const double *U, int P, double *x, int trans) {
  if (trans) {
    for (int i = 0; i < P; ++i) {
      const double *ci = U + (size_t)i * P;
      double s = x[i];
      for (int k = 0; k < i; ++k) s -= ci[k] * x[k];
      x[i] = s / ci[i];
    }
  } else {
    for (int i = P - 1; i >= 0; --i) {
      double s = x[i];
      for (int k = i + 1; k < P; ++k) s -= U[i + (size_t)k * P] * x[k];
      x[i] = s / U[i + (size_t)i * P];
    }
  }
}

/* cyclic two-sided Jacobi on symmetric S (P x P, destroyed); eigenvectors in the columns of V */
static void sym_jacobi(double *S, double *V, int P) {
  for (int i = 0; i < P; ++i) for (int j = 0; j < P; ++j) V[i + (size_t)j * P] = (i == j);
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0, dg = 0.0;
    for (int p = 0; p < P; ++p) { dg += S[p + (size_t)p * P] * S[p + (size_t)p * P];
      for (int q = p + 1; q < P; ++q) off += S[p + (size_t)q * P] * S[p + (size_t)q * P]; }
    if (off <= 1e-30 * dg || off == 0.0) break;
    for (int p = 0; p < P - 1; ++p) for (int q = p + 1; q < P; ++q) {
      double apq = S[p + (size_t)q * P];
      if (apq == 0.0) continue;
      double app = S[p + (size_t)p * P], aqq = S[q + (size_t)q * P];
      double zeta = (aqq - app) / (2.0 * apq);
      double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
      double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
      for (int k = 0; k < P; ++k) { /* columns p,q */
        double skp = S[k + (size_t)p * P], skq = S[k + (size_t)q * P];
        S[k + (size_t)p * P] = c * skp - s * skq; S[k + (size_t)q * P] = s * skp + c * skq;
      }
      for (int k = 0; k < P; ++k) { /* rows p,q */
        double spk = S[p + (size_t)k * P], sqk = S[q + (size_t)k * P];
        S[p + (size_t)k * P] = c * spk - s * sqk; S[q + (size_t)k * P] = s * spk + c * sqk;
      }
      for (int k = 0; k < P; ++k) {
        double vkp = V[k + (size_t)p * P], vkq = V[k + (size_t)q * P];
        V[k + (size_t)p * P] = c * vkp - s * vkq; V[k + (size_t)q * P] = s * vkp + c * vkq;
      }
    }
  }
}

/* one-sided (Hestenes) Jacobi on W (N x P, N >= P): W <- X V with orthogonal columns */
static void onesided_jacobi(double *W, double *V, int N, int P) {
  for (int i = 0; i < P; ++i) for (int j = 0; j < P; ++j) V[i + (size_t)j * P] = (i == j);
  for (int sweep = 0; sweep < 60; ++sweep) {
    int rotated = 0;
    for (int p = 0; p < P - 1; ++p) for (int q = p + 1; q < P; ++q) {
      double *wp = W + (size_t)p * N, *wq = W + (size_t)q * N;
      double al = 0, be = 0, ga = 0;
      for (int k = 0; k < N; ++k) { al += wp[k] * wp[k]; be += wq[k] * wq[k]; ga += wp[k] * wq[k]; }
      if (fabs(ga) <= 1e-15 * sqrt(al * be) || ga == 0.0) continue;
      rotated = 1;
      double zeta = (be - al) / (2.0 * ga);
      double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
      double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
      for (int k = 0; k < N; ++k) { double a = wp[k], b = wq[k]; wp[k] = c * a - s * b; wq[k] = s * a + c * b; }
      double *vp = V + (size_t)p * P, *vq = V + (size_t)q * P;
      for (int k = 0; k < P; ++k) { double a = vp[k], b = vq[k]; vp[k] = c * a - s * b; vq[k] = s * a + c * b; }
    }
    if (!rotated) break;
  }
}

static int cmp_desc(const void *a, const void *b) {
  double x = ((const double *)a)[0], y = ((const double *)b)[0];
  return x < y ? 1 : (x > y ? -1 : 0);
}

static bbo_model *model_alloc(const double *X, const double *y, int N, int P) {
  bbo_model *m = (bbo_model *)calloc(1, sizeof(bbo_model));
  m->N = N; m->P = P; m->X = X; m->y = y;
  m->XX = (double *)calloc((size_t)P * P, sizeof(double));
  m->Xy = (double *)calloc(P, sizeof(double));
  m->bhat = (double *)calloc(P, sizeof(double));
  m->tV = (double *)calloc((size_t)P * P, sizeof(double));
  m->d = (double *)calloc(P, sizeof(double));
  m->a = (double *)calloc(P, sizeof(double));
  /* XX = X'X, Xy = X'y   (Code/C/BridgeRegression.cpp:24-25) */
  for (int i = 0; i < P; ++i) {
    const double *xi = X + (size_t)i * N;
    for (int j = i; j < P; ++j) {
      const double *xj = X + (size_t)j * N;
      double s = 0; for (int k = 0; k < N; ++k) s += xi[k] * xj[k];
      m->XX[i + (size_t)j * P] = s; m->XX[j + (size_t)i * P] = s;
    }
    double s = 0; for (int k = 0; k < N; ++k) s += xi[k] * y[k];
    m->Xy[i] = s;
  }
  m->yy = 0; for (int k = 0; k < N; ++k) m->yy += y[k] * y[k];
  return m;
}

/* least squares; zero vector when X'X is singular (Code/C/BridgeRegression.cpp:79-91) */
static void least_squares(bbo_model *m) {
  int P = m->P;
  double *U = (double *)malloc((size_t)P * P * sizeof(double));
  memcpy(U, m->XX, (size_t)P * P * sizeof(double));
  if (chol_upper(U, P, 1e-13)) {
    m->singular = 1; memset(m->bhat, 0, P * sizeof(double));
  } else {
    memcpy(m->bhat, m->Xy, P * sizeof(double));
    trsv_upper(U, P, m->bhat, 1); trsv_upper(U, P, m->bhat, 0);
  }
  free(U);
}

bbo_model *bbo_model_create(const double *X, const double *y, int N, int P) {
  bbo_model *m = model_alloc(X, y, N, P);
  least_squares(m);
  /* SVD X = U D V' -> tV = V', d, a = (U D)'y    (Code/C/BridgeRegression.cpp:44-54) */
  double *V = (double *)malloc((size_t)P * P * sizeof(double));
  double *key = (double *)malloc((size_t)2 * P * sizeof(double));
  double *dtmp = (double *)calloc(P, sizeof(double)), *atmp = (double *)calloc(P, sizeof(double));
  if (N >= P) {
    double *W = (double *)malloc((size_t)N * P * sizeof(double));
    memcpy(W, X, (size_t)N * P * sizeof(double));
    onesided_jacobi(W, V, N, P);
    for (int i = 0; i < P; ++i) {
      const double *w = W + (size_t)i * N; double s = 0, t = 0;
      for (int k = 0; k < N; ++k) { s += w[k] * w[k]; t += w[k] * y[k]; }
      dtmp[i] = sqrt(s); atmp[i] = t; /* a_i = d_i u_i'y = w_i'y */
    }
    free(W);
  } else {
    double *S = (double *)malloc((size_t)P * P * sizeof(double));
    memcpy(S, m->XX, (size_t)P * P * sizeof(double));
    sym_jacobi(S, V, P);
    double dmax = 0; for (int i = 0; i < P; ++i) if (S[i + (size_t)i * P] > dmax) dmax = S[i + (size_t)i * P];
    for (int i = 0; i < P; ++i) {
      double ev = S[i + (size_t)i * P];
      dtmp[i] = ev > 1e-13 * dmax ? sqrt(ev) : 0.0; /* rank min(N,P): the rest is exactly 0 */
      double t = 0; for (int k = 0; k < P; ++k) t += V[k + (size_t)i * P] * m->Xy[k];
      atmp[i] = dtmp[i] > 0 ? t : 0.0; /* a = V'X'y */
    }
    free(S);
  }
  for (int i = 0; i < P; ++i) { key[2 * i] = dtmp[i]; key[2 * i + 1] = (double)i; }
  qsort(key, P, 2 * sizeof(double), cmp_desc);
  for (int i = 0; i < P; ++i) {
    int src = (int)key[2 * i + 1];
    m->d[i] = dtmp[src]; m->a[i] = atmp[src];
    for (int j = 0; j < P; ++j) m->tV[i + (size_t)j * P] = V[j + (size_t)src * P]; /* tV(i,j) = V(j,i) */
  }
  free(V); free(key); free(dtmp); free(atmp);
  return m;
}

bbo_model *bbo_model_create_with_factors(const double *X, const double *y, int N, int P,
                                         const double *tV, const double *d, const double *a, const double *bhat) {
  bbo_model *m = model_alloc(X, y, N, P);
  if (bhat) memcpy(m->bhat, bhat, P * sizeof(double)); else least_squares(m);
  memcpy(m->tV, tV, (size_t)P * P * sizeof(double));
  memcpy(m->d, d, P * sizeof(double)); memcpy(m->a, a, P * sizeof(double));
  return m;
}

void bbo_model_free(bbo_model *m) {
  if (!m) return;
  free(m->XX); free(m->Xy); free(m->bhat); free(m->tV); free(m->d); free(m->a); free(m);
}
void bbo_model_get(const bbo_model *m, double *XX, double *Xy, double *bhat, double *tV, double *d, double *a) {
  size_t P = m->P;
  if (XX) memcpy(XX, m->XX, P * P * sizeof(double));
  if (Xy) memcpy(Xy, m->Xy, P * sizeof(double));
  if (bhat) memcpy(bhat, m->bhat, P * sizeof(double));
  if (tV) memcpy(tV, m->tV, P * P * sizeof(double));
  if (d) memcpy(d, m->d, P * sizeof(double));
  if (a) memcpy(a, m->a, P * sizeof(double));
}

/* ------------------------------------------------------------------ conditionals */

/* Code/C/BridgeRegression.cpp:453-465 */
double bbo_sample_tau_marg(const bbo_model *m, const double *beta, double alpha, double nu_shape, double nu_rate, bbo_rng *r) {
  double shape = nu_shape + ((double)m->P) / alpha;
  double rate = nu_rate;
  for (int j = 0; j < m->P; ++j) rate += exp(alpha * log(fabs(beta[j])));
  double nu = bbo_gamma_rate(r, shape, rate);
  return exp(-1.0 * log(nu) / alpha);
}

/* Code/C/BridgeRegression.cpp:436-450: residual formed directly from X, as the reference does */
double bbo_sample_sig2(const bbo_model *m, const double *beta, double sig2_shape, double sig2_scale, bbo_rng *r) {
  int N = m->N, P = m->P;
  double shape = sig2_shape + 0.5 * N;
  double *res = (double *)malloc(N * sizeof(double));
  memcpy(res, m->y, N * sizeof(double));
  for (int j = 0; j < P; ++j) {
    const double *xj = m->X + (size_t)j * N; double bj = beta[j];
    for (int k = 0; k < N; ++k) res[k] -= xj[k] * bj;
  }
  double rss = 0; for (int k = 0; k < N; ++k) rss += res[k] * res[k];
  free(res);
  double scale = sig2_scale + 0.5 * rss;
  return bbo_igamma(r, shape, scale);
}

/* Code/C/BridgeRegression.cpp:130-147; item j: unif() then gamma_rate(shape_j, 1) */
void bbo_sample_omega(const bbo_model *m, double *omega, double *shape, const double *beta, const double *u,
                      double tau, double alpha, bbo_rng *r) {
  for (int j = 0; j < m->P; ++j) {
    r->item = (uint32_t)j; r->k = 0;
    double a_j = exp(alpha * log(fabs(beta[j]) / ((1 - u[j]) * tau)));
    double prob = alpha / (1 + alpha * a_j);
    double sh, om;
    if (bbo_unif(r) > prob) { sh = 1.0; om = bbo_gamma_rate(r, 1.0, 1.0); }
    else { sh = 2.0; om = bbo_gamma_rate(r, 2.0, 1.0); }
    if (shape) shape[j] = sh;
    omega[j] = om + a_j;
  }
}

/* Code/C/BridgeRegression.cpp:97-110 */
void bbo_sample_u(const bbo_model *m, double *u, const double *beta, const double *omega, double tau, double alpha, bbo_rng *r) {
  for (int j = 0; j < m->P; ++j) {
    r->item = (uint32_t)j; r->k = 0;
    double right = 1 - fabs(beta[j]) / tau * exp(-1.0 * log(omega[j]) / alpha);
    u[j] = bbo_flat(r, 0, right);
  }
}

/* Code/C/BridgeRegression.cpp:235-286 (static rtnorm_gibbs, SVD form) */
void bbo_rtnorm_gibbs(const bbo_model *m, double *beta, const double *b, double sig2, int as_written, uint32_t item_base, bbo_rng *r) {
  int P = m->P;
  const double *tV = m->tV, *d = m->d, *a = m->a;
  double sig = sqrt(sig2);
  double *z = (double *)calloc(P, sizeof(double));
  double *cur = (double *)malloc(P * sizeof(double)); /* beta_j at the current z (incremental form) */
  for (int i = 0; i < P; ++i) { double s = 0; for (int j = 0; j < P; ++j) s += tV[i + (size_t)j * P] * beta[j]; z[i] = s; } /* :246 */
  memcpy(cur, beta, P * sizeof(double));
  for (int i = 0; i < P; ++i) {
    double lmax = -DBL_MAX, rmin = DBL_MAX;
    for (int j = 0; j < P; ++j) {
      double vji = tV[i + (size_t)j * P];
      double bj_cur;
      if (as_written) { bj_cur = 0; for (int k = 0; k < P; ++k) bj_cur += tV[k + (size_t)j * P] * z[k]; } /* :258 dot(vj,z) */
      else bj_cur = cur[j];
      double rji = bj_cur - vji * z[i];
      double dif = b[j] - rji, sum = b[j] + rji;
      double left = (vji > 0 ? -sum : -dif) / fabs(vji);
      double right = (vji > 0 ? dif : sum) / fabs(vji);
      lmax = lmax > left ? lmax : left;   /* MAX macro, BridgeRegression.h:74 */
      rmin = rmin < right ? rmin : right; /* MIN macro, :75 */
    }
    double znew;
    r->item = item_base + (uint32_t)i; r->k = 0;
    if (d[i] > 1e-16) {
      double mean = a[i] / (d[i] * d[i]);
      double sd = sig / d[i];
      znew = bbo_tnorm(r, lmax, rmin, mean, sd); /* :273 */
    } else {
      znew = bbo_flat(r, lmax, rmin); /* :280 */
    }
    if (!as_written) { double dz = znew - z[i]; for (int j = 0; j < P; ++j) cur[j] += tV[i + (size_t)j * P] * dz; }
    z[i] = znew;
  }
  for (int j = 0; j < P; ++j) { double s = 0; for (int i = 0; i < P; ++i) s += tV[i + (size_t)j * P] * z[i]; beta[j] = s; } /* :285 */
  free(z); free(cur);
}

/* Code/C/BridgeRegression.cpp:405-433 */
void bbo_sample_beta(const bbo_model *m, double *beta, const double *beta_prev, const double *u, const double *omega,
                     double sig2, double tau, double alpha, int betaburn, int as_written, uint32_t iter, bbo_rng *r) {
  int P = m->P;
  betaburn = betaburn > 0 ? betaburn : 0;
  double *b = (double *)malloc(P * sizeof(double));
  for (int j = 0; j < P; ++j) b[j] = (1.0 - u[j]) * exp(log(omega[j]) / alpha) * tau;
  if (beta != beta_prev) memcpy(beta, beta_prev, P * sizeof(double));
  bbo_rng_at(r, iter, BBO_ST_BETA, 0);
  for (int s = 0; s < betaburn + 1; ++s) /* sweep s, coordinate i -> item s*P + i */
    bbo_rtnorm_gibbs(m, beta, b, sig2, as_written, (uint32_t)s * (uint32_t)P, r);
  free(b);
}

/* Code/C/BridgeRegression.cpp:362-403 (burn=0 as every driver calls it) */
void bbo_sample_beta_ortho(const bbo_model *m, double *beta, const double *beta_prev, const double *u, const double *omega,
                           double sig2, double tau, double alpha, bbo_rng *r) {
  int P = m->P;
  if (beta != beta_prev) memcpy(beta, beta_prev, P * sizeof(double));
  for (int j = 0; j < P; ++j) {
    double xxb = 0;
    for (int k = 0; k < P; ++k) if (k != j) xxb += m->XX[j + (size_t)k * P] * beta[k]; /* XX_sub[j] * beta_{-j}, :381-382 */
    double xjj = m->XX[j + (size_t)j * P];
    double mean = (m->Xy[j] - xxb) / xjj;
    double v = sig2 / xjj;
    double b_j = (1.0 - u[j]) * exp(log(omega[j]) / alpha) * tau;
    r->item = (uint32_t)j; r->k = 0;
    beta[j] = bbo_tnorm(r, -1.0 * b_j, b_j, mean, sqrt(v));
  }
}

/* ---- Code/C/retstable.cpp:18-29, 40-47, 73-77, 94-271 ---- */
static double sinc_MM(double x) {
  double ax = fabs(x);
  if (ax < 0.006) {
    if (x == 0.) return 1;
    double x2 = x * x;
    if (ax < 2e-4) return 1. - x2 / 6.;
    else return 1. - x2 / 6. * (1 - x2 / 20.);
  }
  return sin(x) / x;
}
static double A_(double x, double alpha) {
  double Ialpha = 1. - alpha;
  return pow(Ialpha * sinc_MM(Ialpha * x), Ialpha) * pow(alpha * sinc_MM(alpha * x), alpha) / sinc_MM(x);
}
static double BdB0(double x, double alpha) {
  double Ialpha = 1. - alpha;
  double den = pow(sinc_MM(alpha * x), alpha) * pow(sinc_MM(Ialpha * x), Ialpha);
  return sinc_MM(x) / den;
}

double bbo_retstable_LD(double h, double alpha, double V0, bbo_rng *r) {
  if (alpha == 1.) return V0;
  const double c1 = sqrt(M_PI_2);
  const double c2 = 2. + c1;
  const double sqrt_pi = sqrt(M_PI);
  double b = (1. - alpha) / alpha;
  double lambda_alpha = pow(h, alpha) * V0;
  double gamma = lambda_alpha * alpha * (1. - alpha);
  double sgamma = sqrt(gamma);
  double c3 = c2 * sgamma;
  double xi = (1. + M_SQRT2 * c3) / M_PI;
  double psi = c3 * exp(-gamma * M_PI * M_PI / 8.) / sqrt_pi;
  double w1 = c1 * xi / sgamma;
  double w2 = 2. * sqrt_pi * psi;
  double w3 = xi * M_PI;
  double X, c, E;
  do {
    double U, z, Z;
    do {
      double V = bbo_unif(r);
      if (gamma >= 1) {
        if (V < w1 / (w1 + w2)) U = fabs(bbo_norm(r)) / sgamma;
        else { double W_ = bbo_unif(r); U = M_PI * (1. - W_ * W_); }
      } else {
        double W_ = bbo_unif(r);
        if (V < w3 / (w2 + w3)) U = M_PI * W_;
        else U = M_PI * (1. - W_ * W_);
      }
      double W = bbo_unif(r);
      double zeta = sqrt(BdB0(U, alpha));
      z = 1 / (1 - pow(1 + alpha * zeta / sgamma, -1 / alpha));
      double rho = M_PI * exp(-lambda_alpha * (1. - 1. / (zeta * zeta))) / ((1. + c1) * sgamma / zeta + z);
      double d = 0.;
      if (U >= 0 && gamma >= 1) d += xi * exp(-gamma * U * U / 2.);
      if (U > 0 && U < M_PI) d += psi / sqrt(M_PI - U);
      if (U >= 0 && U <= M_PI && gamma < 1) d += xi;
      rho *= d;
      Z = W * rho;
      if (r->k >= BBO_RET_CAP - 8) return NAN;
    } while (!(U < M_PI && Z <= 1.));
    double a = pow(A_(U, alpha), 1. / (1. - alpha));
    double m = pow(b / a, alpha) * lambda_alpha;
    double delta = sqrt(m * alpha / a);
    double a1 = delta * c1, a2 = delta, a3 = z / a;
    double s = a1 + a2 + a3;
    double V_ = bbo_unif(r), N_ = 0., E_ = 0.;
    if (V_ < a1 / s) { N_ = bbo_norm(r); X = m - delta * fabs(N_); }
    else if (V_ < (a1 + a2) / s) X = m + delta * bbo_unif(r);
    else { E_ = bbo_expon_rate(r, 1.0); X = m + delta + E_ * a3; }
    E = -log(Z);
    c = a * (X - m);
    c += (m != 0) ? h * (pow(X, -1. * b) - pow(m, -1. * b)) : 0.0;
    if (X < m) c -= N_ * N_ / 2.;
    else if (X > m + delta) c -= E_;
  } while (!(X >= 0 && c <= E));
  return exp(1 / alpha * log(V0) - b * log(X));
}

/* Code/C/BridgeRegression.cpp:506-510 */
void bbo_sample_lambda(const bbo_model *m, double *lambda, const double *beta, double alpha, double tau, bbo_rng *r) {
  for (int j = 0; j < m->P; ++j) {
    r->item = (uint32_t)j; r->k = 0;
    lambda[j] = 2 * bbo_retstable_LD(beta[j] * beta[j] / (tau * tau), 0.5 * alpha, 1.0, r);
  }
}

/* Code/C/BridgeRegression.cpp:552-575; returns 1 if the Cholesky fails */
int bbo_sample_beta_stable(const bbo_model *m, double *beta, const double *lambda, double sig2, double tau, bbo_rng *r) {
  int P = m->P;
  double *U = (double *)malloc((size_t)P * P * sizeof(double));
  memcpy(U, m->XX, (size_t)P * P * sizeof(double));
  for (int i = 0; i < P; ++i) U[i + (size_t)i * P] += lambda[i] * sig2 / (tau * tau);
  if (chol_upper(U, P, 0.0)) { free(U); return 1; }
  memcpy(beta, m->Xy, P * sizeof(double));
  trsv_upper(U, P, beta, 1);
  trsv_upper(U, P, beta, 0);
  double *nd = (double *)malloc(P * sizeof(double));
  for (int j = 0; j < P; ++j) { r->item = (uint32_t)j; r->k = 0; nd[j] = bbo_norm(r); }
  trsv_upper(U, P, nd, 0);
  double sig = sqrt(sig2);
  for (int i = 0; i < P; ++i) beta[i] += sig * nd[i];
  free(U); free(nd);
  return 0;
}

/* Code/C/BridgeRegression.cpp:514-521 */
void bbo_sample_beta_stable_ortho(const bbo_model *m, double *beta, const double *lambda, double sig2, double tau, bbo_rng *r) {
  int P = m->P;
  for (int i = 0; i < P; ++i) {
    double u = m->XX[i + (size_t)i * P] + lambda[i] * sig2 / (tau * tau);
    double s = sqrt(sig2 / u);
    double mean = m->Xy[i] / u;
    r->item = (uint32_t)i; r->k = 0;
    beta[i] = mean + s * bbo_norm(r);
  }
}

/* Code/C/BridgeRegression.cpp:469-503 (random-walk MH on alpha); ep = 0.1 (BridgeRegression.h:155) */
static double llh_alpha_marg(double alpha, const double *s, int P) {
  double p = (double)P;
  double llh = p * log(alpha) - p * lgamma(1.0 / alpha);
  for (int i = 0; i < P; ++i) llh -= exp(alpha * s[i]);
  return llh;
}
static double d_beta(double x, double a, double b) {
  return exp(lgamma(a + b) - lgamma(a) - lgamma(b) + (a - 1) * log(x) + (b - 1) * log1p(-x));
}
double bbo_sample_alpha_marg(const bbo_model *m, double alpha_prev, const double *beta, double tau,
                             double pr_a, double pr_b, bbo_rng *r) {
  const double ep = 0.1;
  int P = m->P;
  double *s = (double *)malloc(P * sizeof(double));
  for (int i = 0; i < P; ++i) s[i] = log(fabs(beta[i] / tau));
  double a_old = alpha_prev;
  double l_new = fmax(0.0, a_old - ep), r_new = fmin(1.0, a_old + ep);
  double d_new = r_new - l_new;
  double a_new = bbo_flat(r, l_new, r_new);
  double l_old = fmax(0.0, a_new - ep), r_old = fmin(1.0, a_new + ep);
  double d_old = r_old - l_old;
  double log_accept = llh_alpha_marg(a_new, s, P) - llh_alpha_marg(a_old, s, P)
                    + log(d_beta(a_new, pr_a, pr_b)) - log(d_beta(a_old, pr_a, pr_b))
                    + log(d_old) - log(d_new);
  free(s);
  double out = a_new;
  if (bbo_unif(r) > exp(log_accept)) out = a_old;
  return out;
}

/* ------------------------------------------------------------------ drivers */

static double now_sec(void) { return (double)clock() / CLOCKS_PER_SEC; }

/* Code/C/BridgeWrapper.cpp:80-204 (ortho=0) and :320-431 (ortho=1); buffers filled as :599-608 */
double bbo_bridge_regression(const bbo_model *m, const bbo_opts *o, int M,
                             double *beta, double *u, double *omega, double *shape,
                             double *sig2, double *tau, double *alpha) {
  int P = m->P;
  int know_sig2 = o->true_sig2 > 0, know_tau = o->true_tau > 0, know_alpha = o->true_alpha > 0;
  bbo_rng r; bbo_rng_key(&r, o->seed, o->chain);
  for (size_t i = 0; i < (size_t)P * M; ++i) { beta[i] = 0; u[i] = 0.0; omega[i] = 1.0; shape[i] = 0.0; }
  for (int i = 0; i < M; ++i) { sig2[i] = 0; tau[i] = 0; alpha[i] = 0; }
  memcpy(beta, m->bhat, P * sizeof(double));          /* :120-122 */
  for (int j = 0; j < P; ++j) u[j] = 0.5;             /* :124 */
  alpha[0] = 0.5;                                     /* :125 */
  if (know_sig2) for (int i = 0; i < M; ++i) sig2[i] = o->true_sig2;
  if (know_tau) for (int i = 0; i < M; ++i) tau[i] = o->true_tau;
  if (know_alpha) for (int i = 0; i < M; ++i) alpha[i] = o->true_alpha;
  uint32_t t = 0;
  if (o->ortho && !know_sig2) { bbo_rng_at(&r, 0xFFFFFFFFu, BBO_ST_SIG2, 0); sig2[0] = bbo_sample_sig2(m, beta, o->sig2_shape, o->sig2_scale, &r); } /* :379 */
  if (!know_tau) { bbo_rng_at(&r, 0, BBO_ST_TAU_INIT, 0); tau[0] = bbo_sample_tau_marg(m, beta, alpha[0], o->nu_shape, o->nu_rate, &r); } /* :142 */
  for (int it = 0; it < o->burn; ++it, ++t) { /* :145-157 / :383-394, in place on slice 0 */
    if (!know_tau) { bbo_rng_at(&r, t, BBO_ST_TAU, 0); tau[0] = bbo_sample_tau_marg(m, beta, alpha[0], o->nu_shape, o->nu_rate, &r); }
    if (!o->ortho && !know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[0] = bbo_sample_sig2(m, beta, o->sig2_shape, o->sig2_scale, &r); }
    bbo_rng_at(&r, t, BBO_ST_OMEGA, 0); bbo_sample_omega(m, omega, shape, beta, u, tau[0], alpha[0], &r);
    bbo_rng_at(&r, t, BBO_ST_U, 0); bbo_sample_u(m, u, beta, omega, tau[0], alpha[0], &r);
    if (o->ortho) {
      if (!know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[0] = bbo_sample_sig2(m, beta, o->sig2_shape, o->sig2_scale, &r); }
      bbo_rng_at(&r, t, BBO_ST_BETA, 0); bbo_sample_beta_ortho(m, beta, beta, u, omega, sig2[0], tau[0], alpha[0], &r);
    } else {
      bbo_sample_beta(m, beta, beta, u, omega, sig2[0], tau[0], alpha[0], o->betaburn, o->as_written, t, &r);
    }
    if (!know_alpha) { bbo_rng_at(&r, t, BBO_ST_ALPHA, 0); alpha[0] = bbo_sample_alpha_marg(m, alpha[0], beta, tau[0], o->alpha_a, o->alpha_b, &r); }
  }
  double start = now_sec();
  for (int i = 1; i < M; ++i, ++t) { /* :168-189 / :405-416 */
    double *bi = beta + (size_t)i * P, *bp = beta + (size_t)(i - 1) * P;
    double *ui = u + (size_t)i * P, *up = u + (size_t)(i - 1) * P;
    double *oi = omega + (size_t)i * P, *si = shape + (size_t)i * P;
    if (!know_tau) { bbo_rng_at(&r, t, BBO_ST_TAU, 0); tau[i] = bbo_sample_tau_marg(m, bp, alpha[i - 1], o->nu_shape, o->nu_rate, &r); }
    if (!o->ortho && !know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[i] = bbo_sample_sig2(m, bp, o->sig2_shape, o->sig2_scale, &r); }
    bbo_rng_at(&r, t, BBO_ST_OMEGA, 0); bbo_sample_omega(m, oi, si, bp, up, tau[i], alpha[i - 1], &r);
    bbo_rng_at(&r, t, BBO_ST_U, 0); bbo_sample_u(m, ui, bp, oi, tau[i], alpha[i - 1], &r);
    if (o->ortho) {
      if (!know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[i] = bbo_sample_sig2(m, bp, o->sig2_shape, o->sig2_scale, &r); }
      bbo_rng_at(&r, t, BBO_ST_BETA, 0); bbo_sample_beta_ortho(m, bi, bp, ui, oi, sig2[i], tau[i], alpha[i - 1], &r);
    } else {
      bbo_sample_beta(m, bi, bp, ui, oi, sig2[i], tau[i], alpha[i - 1], o->betaburn, o->as_written, t, &r);
    }
    if (!know_alpha) { bbo_rng_at(&r, t, BBO_ST_ALPHA, 0); alpha[i] = bbo_sample_alpha_marg(m, alpha[i - 1], bi, tau[i], o->alpha_a, o->alpha_b, &r); }
  }
  return now_sec() - start;
}

/* Code/C/BridgeWrapper.cpp:207-313 (ortho=0) and :434-537 (ortho=1); lambda filled with 1 (:686) */
double bbo_bridge_reg_stable(const bbo_model *m, const bbo_opts *o, int M,
                             double *beta, double *lambda, double *sig2, double *tau, double *alpha) {
  int P = m->P;
  int know_sig2 = o->true_sig2 > 0, know_tau = o->true_tau > 0, know_alpha = o->true_alpha > 0;
  bbo_rng r; bbo_rng_key(&r, o->seed, o->chain);
  for (size_t i = 0; i < (size_t)P * M; ++i) { beta[i] = 0; lambda[i] = 1.0; }
  for (int i = 0; i < M; ++i) { sig2[i] = 0; tau[i] = 0; alpha[i] = 0; }
  memcpy(beta, m->bhat, P * sizeof(double));
  alpha[0] = 0.5;
  if (know_sig2) for (int i = 0; i < M; ++i) sig2[i] = o->true_sig2;
  if (know_tau) for (int i = 0; i < M; ++i) tau[i] = o->true_tau;
  if (know_alpha) for (int i = 0; i < M; ++i) alpha[i] = o->true_alpha;
  uint32_t t = 0;
  if (!o->ortho && !know_tau) { bbo_rng_at(&r, 0, BBO_ST_TAU_INIT, 0); tau[0] = bbo_sample_tau_marg(m, beta, alpha[0], o->nu_shape, o->nu_rate, &r); } /* :262 */
  for (int it = 0; it < o->burn + 1; ++it, ++t) { /* :266-276 / :493-502 : burn+1 iterations */
    if (!know_tau) { bbo_rng_at(&r, t, BBO_ST_TAU, 0); tau[0] = bbo_sample_tau_marg(m, beta, alpha[0], o->nu_shape, o->nu_rate, &r); }
    if (!o->ortho && !know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[0] = bbo_sample_sig2(m, beta, o->sig2_shape, o->sig2_scale, &r); }
    bbo_rng_at(&r, t, BBO_ST_LAMBDA, 0); bbo_sample_lambda(m, lambda, beta, alpha[0], tau[0], &r);
    if (o->ortho && !know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[0] = bbo_sample_sig2(m, beta, o->sig2_shape, o->sig2_scale, &r); }
    bbo_rng_at(&r, t, BBO_ST_BETA_STB, 0);
    if (o->ortho) bbo_sample_beta_stable_ortho(m, beta, lambda, sig2[0], tau[0], &r);
    else if (bbo_sample_beta_stable(m, beta, lambda, sig2[0], tau[0], &r)) return NAN;
    if (!know_alpha) { bbo_rng_at(&r, t, BBO_ST_ALPHA, 0); alpha[0] = bbo_sample_alpha_marg(m, alpha[0], beta, tau[0], o->alpha_a, o->alpha_b, &r); }
  }
  double start = now_sec();
  for (int i = 1; i < M; ++i, ++t) { /* :287-298 / :513-522 */
    double *bi = beta + (size_t)i * P, *bp = beta + (size_t)(i - 1) * P, *li = lambda + (size_t)i * P;
    if (!know_tau) { bbo_rng_at(&r, t, BBO_ST_TAU, 0); tau[i] = bbo_sample_tau_marg(m, bp, alpha[i - 1], o->nu_shape, o->nu_rate, &r); }
    if (!o->ortho && !know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[i] = bbo_sample_sig2(m, bp, o->sig2_shape, o->sig2_scale, &r); }
    bbo_rng_at(&r, t, BBO_ST_LAMBDA, 0); bbo_sample_lambda(m, li, bp, alpha[i - 1], tau[i], &r);
    if (o->ortho && !know_sig2) { bbo_rng_at(&r, t, BBO_ST_SIG2, 0); sig2[i] = bbo_sample_sig2(m, bp, o->sig2_shape, o->sig2_scale, &r); }
    bbo_rng_at(&r, t, BBO_ST_BETA_STB, 0);
    if (o->ortho) bbo_sample_beta_stable_ortho(m, bi, li, sig2[i], tau[i], &r);
    else if (bbo_sample_beta_stable(m, bi, li, sig2[i], tau[i], &r)) return NAN;
    /* NB the reference passes (alpha_b, alpha_b) here in the non-ortho loop: BridgeWrapper.cpp:294 */
    if (!know_alpha) { bbo_rng_at(&r, t, BBO_ST_ALPHA, 0);
      alpha[i] = bbo_sample_alpha_marg(m, alpha[i - 1], bi, tau[i], o->ortho ? o->alpha_a : o->alpha_b, o->alpha_b, &r); }
  }
  return now_sec() - start;
}

/* ------------------------------------------------------------------ bulk draw probes (distribution tests) */
/* item i of (seed, chain 0, iter 0, stage SCALAR) -> out[i]; kind: 0 unif, 1 norm, 2 gamma(p0,p1),
   3 tnorm(p0,p1,p2,p3), 4 tnorm_left(p0,p1,p2), 5 texpon_left(p0,p1), 6 texpon_both(p0,p1,p2),
   7 rtgamma(p0,p1,p2), 8 retstable(h=p0, alpha=p1, V0=p2), 9 igamma(p0,p1) */
void bbo_draw_many(int kind, int n, const double *p, uint64_t seed, double *out) {
  bbo_rng r; bbo_rng_key(&r, seed, 0);
  for (int i = 0; i < n; ++i) {
    bbo_rng_at(&r, 0, BBO_ST_SCALAR, (uint32_t)i);
    switch (kind) {
      case 0: out[i] = bbo_unif(&r); break;
      case 1: out[i] = bbo_norm(&r); break;
      case 2: out[i] = bbo_gamma_rate(&r, p[0], p[1]); break;
      case 3: out[i] = bbo_tnorm(&r, p[0], p[1], p[2], p[3]); break;
      case 4: out[i] = bbo_tnorm_left(&r, p[0], p[1], p[2]); break;
      case 5: out[i] = bbo_texpon_rate_left(&r, p[0], p[1]); break;
      case 6: out[i] = bbo_texpon_rate_both(&r, p[0], p[1], p[2]); break;
      case 7: out[i] = bbo_rtgamma_rate(&r, p[0], p[1], p[2]); break;
      case 8: out[i] = bbo_retstable_LD(p[0], p[1], p[2], &r); break;
      case 9: out[i] = bbo_igamma(&r, p[0], p[1]); break;
      default: out[i] = NAN;
    }
  }
}
/* The stand-alone sampler entry points of the reference's shared object, vectorised over n parameter sets
   (Code/C/BridgeWrapper.cpp:762-984, the non-R branches for infinite bounds :830-835, :927-938).  Item i draws
   from (seed, chain 0, iter, BBO_ST_SCALAR, item i).  kind: 0 rtexpon_rate_left(left=p0, rate=p1),
   1 rtexpon_rate_both(left, right, rate), 2 rtexpon_rate(left, right, rate), 3 rtnorm_left(left, mu, sig),
   4 rtnorm_both(left, right, mu, sig), 5 rtnorm(left, right, mu, sig), 6 rrtgamma_rate(shape, rate, right),
   7 retstable_LD(alpha=p0, V0=p1, h=p2). */
void bbo_draw_vec(int kind, int n, const double *p0, const double *p1, const double *p2, const double *p3,
                  uint64_t seed, uint32_t iter, double *out) {
  bbo_rng r; bbo_rng_key(&r, seed, 0);
  for (int i = 0; i < n; ++i) {
    bbo_rng_at(&r, iter, BBO_ST_SCALAR, (uint32_t)i);
    switch (kind) {
      case 0: out[i] = bbo_texpon_rate_left(&r, p0[i], p1[i]); break;
      case 1: out[i] = bbo_texpon_rate_both(&r, p0[i], p1[i], p2[i]); break;
      case 2: out[i] = isfinite(p1[i]) ? bbo_texpon_rate_both(&r, p0[i], p1[i], p2[i]) : bbo_texpon_rate_left(&r, p0[i], p2[i]); break;
      case 3: out[i] = bbo_tnorm_left(&r, p0[i], p1[i], p2[i]); break;
      case 4: out[i] = bbo_tnorm(&r, p0[i], p1[i], p2[i], p3[i]); break;
      case 5:
        if (isfinite(p0[i]) && isfinite(p1[i])) out[i] = bbo_tnorm(&r, p0[i], p1[i], p2[i], p3[i]);
        else if (isfinite(p0[i])) out[i] = bbo_tnorm_left(&r, p0[i], p2[i], p3[i]);
        else if (isfinite(p1[i])) out[i] = -1.0 * bbo_tnorm_left(&r, -1.0 * p1[i], -1.0 * p2[i], p3[i]);
        else out[i] = p2[i] + p3[i] * bbo_norm(&r);
        break;
      case 6: out[i] = bbo_rtgamma_rate(&r, p0[i], p1[i], p2[i]); break;
      case 7: out[i] = bbo_retstable_LD(p2[i], p0[i], p1[i], &r); break;
      default: out[i] = NAN;
    }
  }
}
/* raw Philox block for known-answer tests */
void bbo_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) { bbo_philox4x32_10(ctr, key, out); }
/* single derived draws at an explicit counter position (teacher-forced per-coordinate checks) */
double bbo_tnorm_at(uint64_t seed, uint32_t chain, uint32_t iter, uint32_t stage, uint32_t item,
                    double left, double right, double mu, double sd) {
  bbo_rng r; bbo_rng_key(&r, seed, chain); bbo_rng_at(&r, iter, stage, item);
  return bbo_tnorm(&r, left, right, mu, sd);
}
double bbo_norm_at(uint64_t seed, uint32_t chain, uint32_t iter, uint32_t stage, uint32_t item) {
  bbo_rng r; bbo_rng_key(&r, seed, chain); bbo_rng_at(&r, iter, stage, item);
  return bbo_norm(&r);
}
double bbo_flat_at(uint64_t seed, uint32_t chain, uint32_t iter, uint32_t stage, uint32_t item, double a, double b) {
  bbo_rng r; bbo_rng_key(&r, seed, chain); bbo_rng_at(&r, iter, stage, item);
  return bbo_flat(&r, a, b);
}
