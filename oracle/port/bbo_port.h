/*
 * bbo_port.h -- ORACLE (test infrastructure; never shipped, never on the product path).
 *
 * Plain-C restatement of the BayesBridge Gibbs sweep.  Every function cites the reference
 * lines it follows (paths relative to /root/reference/).  Draws come from bbo_draws.h
 * (the reference's own RNG library is not in its tree: parity unpinned at the draw level).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this.
 */
#ifndef BBO_PORT_H
#define BBO_PORT_H
#include "bbo_draws.h"
#ifdef __cplusplus
extern "C" {
#endif

/* chain-independent factors: Code/C/BridgeRegression.cpp:13-69 */
typedef struct {
  int N, P;
  const double *X, *y; /* borrowed, column-major N x P */
  double *XX;          /* P x P  X'X                         (:24) */
  double *Xy;          /* P      X'y                         (:25) */
  double *bhat;        /* P      least squares, 0 if singular (:41-42, :79-91) */
  double *tV;          /* P x P  col-major, tV[i + j*P] = V(j,i)  (:47-51) */
  double *d;           /* P      singular values, 0-padded beyond min(N,P) */
  double *a;           /* P      (U diag d)' y               (:53-54) */
  double yy;
  int singular;
} bbo_model;

bbo_model *bbo_model_create(const double *X, const double *y, int N, int P);
/* same, but with externally supplied tV,d,a (e.g. the compiled reference's LAPACK SVD) */
bbo_model *bbo_model_create_with_factors(const double *X, const double *y, int N, int P,
                                         const double *tV, const double *d, const double *a,
                                         const double *bhat);
void bbo_model_free(bbo_model *m);
void bbo_model_get(const bbo_model *m, double *XX, double *Xy, double *bhat, double *tV, double *d, double *a);

/* conditionals; `r` must be positioned by the caller (bbo_rng_at) */
double bbo_sample_tau_marg(const bbo_model *m, const double *beta, double alpha, double nu_shape, double nu_rate, bbo_rng *r);
double bbo_sample_sig2(const bbo_model *m, const double *beta, double sig2_shape, double sig2_scale, bbo_rng *r);
void bbo_sample_omega(const bbo_model *m, double *omega, double *shape, const double *beta, const double *u, double tau, double alpha, bbo_rng *r);
void bbo_sample_u(const bbo_model *m, double *u, const double *beta, const double *omega, double tau, double alpha, bbo_rng *r);
/* one rotated sweep; as_written=1 recomputes beta_j by a length-P dot (O(P^3)), 0 = incremental O(P^2) */
void bbo_rtnorm_gibbs(const bbo_model *m, double *beta, const double *b, double sig2, int as_written, uint32_t item_base, bbo_rng *r);
void bbo_sample_beta(const bbo_model *m, double *beta, const double *beta_prev, const double *u, const double *omega,
                     double sig2, double tau, double alpha, int betaburn, int as_written, uint32_t iter, bbo_rng *r);
void bbo_sample_beta_ortho(const bbo_model *m, double *beta, const double *beta_prev, const double *u, const double *omega,
                           double sig2, double tau, double alpha, bbo_rng *r);
double bbo_retstable_LD(double h, double alpha, double V0, bbo_rng *r);
void bbo_sample_lambda(const bbo_model *m, double *lambda, const double *beta, double alpha, double tau, bbo_rng *r);
int bbo_sample_beta_stable(const bbo_model *m, double *beta, const double *lambda, double sig2, double tau, bbo_rng *r);
double bbo_sample_alpha_marg(const bbo_model *m, double alpha_prev, const double *beta, double tau, double pr_a, double pr_b, bbo_rng *r);
void bbo_sample_beta_stable_ortho(const bbo_model *m, double *beta, const double *lambda, double sig2, double tau, bbo_rng *r);

/* drivers (BridgeWrapper.cpp:80-204, 207-313, 320-431, 434-537); outputs P x M / M, like the .C call.
   Returns the seconds of the post-burn-in loop. */
typedef struct {
  double sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b;
  double true_sig2, true_tau, true_alpha;
  int burn, betaburn, ortho, as_written;
  uint64_t seed;
  uint32_t chain;
} bbo_opts;

double bbo_bridge_regression(const bbo_model *m, const bbo_opts *o, int M,
                             double *beta, double *u, double *omega, double *shape,
                             double *sig2, double *tau, double *alpha);
double bbo_bridge_reg_stable(const bbo_model *m, const bbo_opts *o, int M,
                             double *beta, double *lambda, double *sig2, double *tau, double *alpha);

#ifdef __cplusplus
}
#endif
#endif
