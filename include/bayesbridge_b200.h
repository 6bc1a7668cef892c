/*
 * bayesbridge_b200.h -- C ABI of libbayesbridge_b200.so
 *
 * A B200-native (sm_100a) implementation of BayesBridge's Gibbs sweep for bridge regression.
 * Plain pointers and sizes only; no C++ or torch types cross this boundary.
 * Citations are file:line under the reference tree (/root/reference/).
 *
 * Part 1 -- drop-in symbols: exactly what R's .C() binds in the reference package
 *           (Code/C/BridgeWrapper.R:170-179 and :220-228; declared extern "C" at
 *           Code/C/BridgeWrapper.h:164-226, defined at Code/C/BridgeWrapper.cpp:572-656, 659-732).
 * Part 2 -- the sampler interface of class BridgeRegression (Code/C/BridgeRegression.h:81-166)
 *           as handle-based calls: a model (chain-independent factors in HBM), a set of chains
 *           (state in HBM) and `bb_chains_run`, which advances all chains by whole iterations or
 *           by single conditionals (stage mask).
 * Part 3 -- batched extension (not in the reference): many chains, seeds, thinning.
 *
 * There is no CPU fallback: every entry point fails (error string, NaN-filled outputs) when no
 * CUDA device is usable.
 */
#ifndef BAYESBRIDGE_B200_H
#define BAYESBRIDGE_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------------------------------- */
/* Part 1: drop-in .C() entry points                                                         */
/* ---------------------------------------------------------------------------------------- */

/* Replaces bridge_regression, Code/C/BridgeWrapper.h:178-203 / BridgeWrapper.cpp:572-656.
 * Argument ORDER is ABI (R reads results positionally, BridgeWrapper.R:181-183).
 * betap,up,omegap,shapep: P*M doubles out (P fastest); sig2p,taup,alphap: M doubles out;
 * yp: N; Xp: N*P column-major; true_* > 0 means known/fixed; runtime: seconds of the
 * post-burn-in loop.  ortho/use_hmc are R logicals passed as 4-byte ints (BridgeWrapper.R:178);
 * the reference types them `const bool*` and reads the low byte -- `const int*` is binary
 * compatible and correct.  use_hmc is ignored (forced false at BridgeRegression.cpp:419). */
void bridge_regression(double *betap, double *up, double *omegap, double *shapep,
                       double *sig2p, double *taup, double *alphap,
                       const double *yp, const double *Xp,
                       const double *sig2_shape, const double *sig2_scale,
                       const double *nu_shape, const double *nu_rate,
                       const double *alpha_a, const double *alpha_b,
                       const double *true_sig2, const double *true_tau, const double *true_alpha,
                       const int *P, const int *N, const int *M, const int *burn,
                       double *runtime, const int *ortho, const int *betaburn, const int *use_hmc);

/* Replaces bridge_reg_stable, Code/C/BridgeWrapper.h:205-226 / BridgeWrapper.cpp:659-732. */
void bridge_reg_stable(double *betap, double *lambdap,
                       double *sig2p, double *taup, double *alphap,
                       const double *yp, const double *Xp,
                       const double *sig2_shape, const double *sig2_scale,
                       const double *nu_shape, const double *nu_rate,
                       const double *alpha_a, const double *alpha_b,
                       const double *true_sig2, const double *true_tau, const double *true_alpha,
                       const int *P, const int *N, const int *M, const int *burn,
                       double *runtime, const int *ortho);

/* Seed used by the two drop-in calls (the reference draws from R's global RNG between
 * GetRNGstate/PutRNGstate, BridgeWrapper.cpp:610-612,636-638; an R binding would call this with
 * one unif_rand() draw so that set.seed() governs the run).  Default 20121027. */
void bb_set_seed(unsigned long long seed);
/* Device used by the drop-in calls (default 0). */
void bb_set_device(int device);
/* Banner printing of the reference drivers (BridgeWrapper.cpp:112-117,162-163,200): 0 = quiet (default). */
void bb_set_verbose(int verbose);
/* Interrupt hook, the role of R_CheckUserInterrupt() in the reference's loops (every 10 iterations, BridgeWrapper.cpp:154-156,
 * 186-188; every 1000 rejection trials, retstable.cpp:158-166).  The sampling loop runs on the GPU in chunks; between chunks
 * (sized to about 0.2 s of work while a hook is installed) the CALLING thread calls fn(); a non-zero return aborts the run:
 * the entry point returns non-zero with bb_last_error() = "interrupted ...", the .C symbols print that and NaN-fill their
 * outputs like any other failure.  An R binding installs a function that wraps R_ToplevelExec(R_CheckUserInterrupt) (see
 * INTEGRATION.md).  NULL removes the hook. */
void bb_set_interrupt(int (*fn)(void));
/* Status word of the last stand-alone sampler call (rtnorm, rtexpon_rate, rrtgamma_rate, retstable_LD ...): bit 0 NaN
 * parameters, bit 1 empty interval (left > right), bit 2 a rejection loop reached its trial cap (those draws are NaN).
 * The same conditions are also printed to stderr, as the reference prints its own warnings (BridgeWrapper.cpp:899-934). */
unsigned int bb_scalar_status(void);

/* Stand-alone scalar samplers exported by the reference's shared object next to the two Gibbs entry points
 * (declarations BridgeWrapper.h:230-244, bodies BridgeWrapper.cpp:762-984; R callers BridgeWrapper.R:295-537).
 * Every argument is an array of length *num (R recycles before the call); x receives the draws.  Item i draws
 * from the Philox stream (seed of bb_set_seed, call counter since bb_set_seed, item i), one GPU thread per item.
 * rtexpon_rate / rtnorm take -Inf / +Inf bounds as the reference's non-R branch does (BridgeWrapper.cpp:830-835,
 * 927-938).  rrtgamma_rate: `scale` is the SHAPE of the gamma law (the reference's own argument name,
 * BridgeWrapper.cpp:949,961).  On a CUDA failure the call prints to stderr and NaN-fills x. */
void rtexpon_rate_left(double *x, double *left, double *rate, int *num);
void rtexpon_rate_both(double *x, double *left, double *right, double *rate, int *num);
void rtexpon_rate(double *x, double *left, double *right, double *rate, int *num);
void rtnorm_left(double *x, double *left, double *mu, double *sig, int *num);
void rtnorm_both(double *x, double *left, double *right, double *mu, double *sig, int *num);
void rtnorm(double *x, double *left, double *right, double *mu, double *sig, int *num);
void rrtgamma_rate(double *x, double *scale, double *rate, double *right_t, int *num);
void retstable_LD(double *x, double *alpha, double *V0, double *h, int *num);

/* ---------------------------------------------------------------------------------------- */
/* Part 2: sampler interface                                                                 */
/* ---------------------------------------------------------------------------------------- */
typedef struct bb_model bb_model;   /* role of BridgeRegression::BridgeRegression(X,y), BridgeRegression.cpp:13-69 */
typedef struct bb_chains bb_chains; /* the slices beta[i],u[i],... the drivers carry, BridgeWrapper.cpp:80-100 */

typedef struct bb_hyper {           /* BridgeWrapper.cpp:89-100 */
  double sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b;
  double true_sig2, true_tau, true_alpha;   /* > 0: known, kept fixed (BridgeWrapper.cpp:107-109) */
  int betaburn;                             /* extra rotated sweeps per iteration (BridgeRegression.cpp:421) */
} bb_hyper;

enum { BB_TRI = 0, BB_STB = 1 };
enum { BB_MEM_HOST = 0, BB_MEM_DEVICE = 1 };

/* stage mask for bb_chains_run (the conditionals of one Gibbs iteration, in driver order) */
enum {
  BB_ST_TAU = 1,     /* sample_tau_marg     BridgeRegression.cpp:453-465 */
  BB_ST_SIG2 = 2,    /* sample_sig2         :436-450 */
  BB_ST_OMEGA = 4,   /* sample_omega        :130-147   (tri) */
  BB_ST_U = 8,       /* sample_u            :97-110    (tri) */
  BB_ST_BETA = 16,   /* sample_beta :405-433 (tri) | sample_beta_stable :552-575 (stb) | the *_ortho forms */
  BB_ST_LAMBDA = 32, /* sample_lambda       :506-510   (stb) */
  BB_ST_ALPHA = 64,  /* sample_alpha_marg   :477-503   (only acts when true_alpha <= 0: alpha unknown) */
  BB_ST_ALL = 127
};

/* which per-iteration draws bb_chains_run keeps (bit mask) and bb_chains_fetch returns */
enum { BB_OUT_BETA = 1, BB_OUT_U = 2, BB_OUT_OMEGA = 4, BB_OUT_SHAPE = 8, BB_OUT_LAMBDA = 16,
       BB_OUT_SIG2 = 32, BB_OUT_TAU = 64, BB_OUT_ALPHA = 128 };

const char *bb_last_error(void);
int bb_device_count(void);

/* X: N x P column-major, y: N; `mem` says where X,y live.  Builds X'X, X'y (FP64 tensor-core Gram),
 * the SVD factors V,d,a (one-sided Jacobi) and the least-squares start on `device`. NULL on error. */
bb_model *bb_model_create(const double *X, const double *y, int N, int P, int device, int mem);
/* Same, with caller-supplied factors tV (P x P col-major, tV[i+j*P] = V(j,i)), d, a, bhat (host):
 * used to run the sampler on exactly the reference's LAPACK factors.  bhat may be NULL. */
bb_model *bb_model_create_factors(const double *X, const double *y, int N, int P,
                                  const double *tV, const double *d, const double *a, const double *bhat,
                                  int device, int mem);
/* Row-sharded construction.  bb_gram: G = [X y]'[X y] of a row shard, (P+1) x (P+1) col-major,
 * written to DEVICE memory G (to be all-reduced by the caller, e.g. NCCL).  Then
 * bb_model_create_gram builds the model from the reduced G (device memory), N = total rows. */
int bb_gram(const double *X_dev, const double *y_dev, int n_rows, int P, double *G_dev, int device);
bb_model *bb_model_create_gram(const double *G_dev, int N, int P, int device);
/* bb_gram for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * No counterpart in the reference.  Same arguments, same kernels, same
 * result bit for bit, but the split-K partial tiles live in a file-scope `static __device__` array of 37 MiB (1184
 * tiles of 64 x 64 doubles) that every GPU loading the library holds until the process exits.  LIMIT: P <= 3071 (any
 * n_rows); larger P returns 1 with bb_last_error() saying so, before any CUDA call.  The call makes no cudaMalloc /
 * cudaFree / allocating library call; inputs and the output G are the caller's device buffers.  Every call on a GPU
 * shares the one scratch array, which keeps the previous call's contents (every slot a call reads is rewritten by that
 * call first, so nothing needs clearing): calls that overlap in time on the same GPU race -- serialise them.  The first
 * call on a device also looks the array's address up (cudaGetSymbolAddress, once per device). */
int bb_gram_noalloc(const double *X_dev, const double *y_dev, int n_rows, int P, double *G_dev, int device);
/* residual sums of squares of a row shard at the two expansion centres -> out_dev[0..1] (device);
 * after the all-reduce, bb_model_set_r0 installs the totals. */
int bb_model_resid_partial(const bb_model *m, const double *X_dev, const double *y_dev, int n_rows, double *out_dev);
int bb_model_set_r0(bb_model *m, double r0z, double r0b);
void bb_model_free(bb_model *m);
/* A bit-for-bit copy of the model's factors on another device (peer copies over NVLink; the rows X, y are not copied: the
 * sweep never reads them).  Chains on the copy produce exactly the draws they would produce on the original. */
bb_model *bb_model_clone(const bb_model *src, int device);
/* 0 = clean; n > 0: n singular values between 1e-13 and 1e-7 of the largest were set to zero because the factors had to be
 * taken from X'X (bb_model_create_gram, or P > N) -- for cond(X) >= 1e7 the reference's SVD of X would keep them (also
 * printed to stderr when it happens).  Build the model from the rows (bb_model_create with N >= P) to avoid it. */
int bb_model_status(const bb_model *m);
/* copy factors to host (any pointer may be NULL) */
int bb_model_get(const bb_model *m, double *XX, double *Xy, double *bhat, double *tV, double *d, double *a);
double bb_model_precompute_ms(const bb_model *m);

/* nchains chains with the drivers' initial state: beta = least squares, u = 0.5, omega = 1,
 * lambda = 1, shape = 0 (BridgeWrapper.cpp:120-129, 241-250, 603-605, 686).  chain0 = global id of
 * the first chain: the Philox stream of a chain depends only on (seed, chain0 + local index). */
bb_chains *bb_chains_create(const bb_model *m, int method, int ortho, int nchains,
                            unsigned long long seed, unsigned int chain0);
void bb_chains_free(bb_chains *c);
/* state vectors are [nchains][P] (P fastest) / [nchains], host memory; NULL = leave untouched.
 * aux = omega (tri) or lambda (stb). */
int bb_chains_set(bb_chains *c, const double *beta, const double *u, const double *aux,
                  const double *sig2, const double *tau);
int bb_chains_get(const bb_chains *c, double *beta, double *u, double *aux, double *shape,
                  double *sig2, double *tau, unsigned int *status);
/* alpha of every chain, [nchains] host memory: `set` (may be NULL) is installed first, then the current
 * values are copied to `get` (may be NULL).  Chains start at 0.5 (BridgeWrapper.cpp:125,246,363,474); the
 * state only moves when alpha is unknown (true_alpha <= 0, BridgeWrapper.cpp:109,152,185). */
int bb_chains_alpha(bb_chains *c, const double *set, double *get);
/* Options of a set of chains.  BB_OPT_WIDE_GAUSSIAN (stable sampler, ortho = 0, P > 128, rows of X on the device): draw
 * beta | lambda, sig2, tau through the N x N system  M = X D^-1 X' + I  (Bhattacharya, Chakraborty & Mallick 2016) instead of
 * the P x P Cholesky of X'X + D (Code/C/BridgeRegression.cpp:552-575) -- the P > N remedy the reference's authors point at
 * (Code/R/Benchmark.R:734-742).  Same LAW, different arithmetic: for the same Philox stream the draws
 * differ from the reference's, so it is off by default and the per-draw parity tests do not apply to it.  N^2 P + N^3/3 flops
 * and N^2 doubles of workspace per chain instead of P^3/3 and P^2.  Returns non-zero (bb_last_error) when it does not apply. */
enum { BB_OPT_WIDE_GAUSSIAN = 1 };
int bb_chains_option(bb_chains *c, int option, int value);
/* Default for the drivers (bridge_reg_stable, bb_bridge_batched*): 1 = opt their chains into BB_OPT_WIDE_GAUSSIAN wherever it
 * applies (stable sampler, ortho = 0, P > 128, N <= P).  0 (the default) keeps the reference's P x P arithmetic. */
void bb_set_wide_gaussian(int on);
/* reserve a device buffer for `nslots` stored iterations per chain of the outputs in out_mask */
int bb_chains_reserve(bb_chains *c, long long nslots, int out_mask);
/* Advance every chain by niter iterations; iteration index of the first = iter0 (RNG counter).
 * stages: BB_ST_* mask (BB_ST_ALL = the full sweep).  pre_tau: draw tau once first
 * (BridgeWrapper.cpp:142,262).  Every thin-th iteration is stored from slot `slot0` on (thin = 0:
 * nothing stored).  kernel_ms (may be NULL) receives the device time of the launches. */
int bb_chains_run(bb_chains *c, const bb_hyper *h, unsigned int iter0, int niter, int stages,
                  int pre_tau, int thin, long long slot0, double *kernel_ms);
/* copy current state into slot `slot` of the draw buffer (sample 0 of the drivers = state after burn-in) */
int bb_chains_store_state(bb_chains *c, long long slot);
/* copy stored draws to host: out[(chain*nslots + s)*P + j] (or [chain*nslots + s] for scalars) */
int bb_chains_fetch(const bb_chains *c, int which, long long slot0, long long nslots, double *out);
/* sum.stat on the device (Code/R/PublicBenchmark.R:112-134) over slots [slot0, slot0 + nslots) of a stored VECTOR
 * output (`which` = BB_OUT_BETA, ...), 2 <= nslots <= 16384.  Host outputs, either may be NULL:
 *   per_chain[(chain*5 + k)*P + j], k = mean, sd, ESS, Q.1, Q.9 of chain `chain`'s series of coefficient j
 *   pooled[k*P + j], k = mean, sd, ESS, ESS.sec, t, Q.1, Q.9 over all chains (the columns of sum.stat): ESS is
 *   coda::effectiveSize (AR spectral estimate, order by AIC) summed over chains, ESS.sec = ESS / runtime_sec (NaN
 *   when runtime_sec <= 0), t = mean / sd, quantiles are R's type 7 over the pooled draws. */
int bb_chains_summary(const bb_chains *c, int which, long long slot0, long long nslots, double runtime_sec,
                      double *per_chain, double *pooled);
/* RUNNING sum.stat: the same table from a run of ANY length with a draw buffer of one chunk (SURVEY section 7.1 K_out).  begin
 * resets the accumulators of one vector output; add folds the stored slots [slot0, slot0 + nslots) of every chain into them, in
 * order (the caller then reuses the slots); end writes the table.  Kept per (chain, coefficient): the raw lagged products up to
 * lag 48, the first and last 48 values and the sum -- mean, sd and the autocovariances ar.yw needs are recovered exactly, so
 * ESS is coda's definition with the lag window floor(10 log10 n) capped at 48.  Quantiles come from a mergeable sketch (2048
 * bins over +-10 sd of the first chunk per coefficient, shared by all chains; resolution 0.01 sd): pooled Q.1 / Q.9 only,
 * per_chain's two quantile rows are NaN.  The first add needs >= 2 slots.  bb_bridge_batched_summary runs on this. */
int bb_chains_summary_begin(bb_chains *c, int which);
int bb_chains_summary_add(bb_chains *c, long long slot0, long long nslots);
int bb_chains_summary_end(bb_chains *c, double runtime_sec, double *per_chain, double *pooled);
/* number of kernels launched by this library in this process so far */
long long bb_launch_count(void);
/* Measurement aid, no counterpart in the reference: measured peak of the FP64 vector pipe (DFMA) of `device` in
 * TFLOP/s -- the roofline denominator of the triangle sweep, which has no contraction (bench.py). */
int bb_measure_fp64_fma(int device, double *tflops);

/* ---------------------------------------------------------------------------------------- */
/* Part 3: batched extension                                                                 */
/* ---------------------------------------------------------------------------------------- */
/* Runs nchains chains: burn iterations (stb: burn+1, BridgeWrapper.cpp:266), then nsamp stored
 * samples (sample 0 = state after burn-in, then every thin-th iteration), exactly the drivers'
 * schedule.  Host buffers: beta/aux1/aux2/aux3 are [nchains][nsamp][P], scalars [nchains][nsamp];
 * tri: aux1=u, aux2=omega, aux3=shape; stb: aux1=lambda. Any output pointer may be NULL.
 * runtime: seconds of the post-burn-in loop.  Returns 0 on success. */
int bb_bridge_batched(int method, int ortho,
                      const double *X, const double *y, int N, int P, const bb_hyper *h,
                      int nchains, unsigned long long seed, unsigned int chain0,
                      int burn, int nsamp, int thin, int device,
                      double *beta, double *aux1, double *aux2, double *aux3,
                      double *sig2, double *tau, double *alpha, unsigned int *status, double *runtime);

/* bb_bridge_batched over several GPUs of one box, inside the library: the constructor (Gram, SVD
 * factors) runs once on devices[0]; the other devices get a bit-for-bit copy of the factors by peer copies over NVLink
 * (bb_model_clone); device k samples the k-th contiguous block of chains on its own host thread and streams its draws into
 * its slice of the same host arrays.  devices: ndevices device indices, or NULL for 0 .. ndevices-1.  The Philox stream of
 * a chain depends only on (seed, chain0 + index), so the outputs equal the one-GPU call's bit for bit.  There is no
 * collective: nothing is exchanged during sampling.  runtime = the slowest device's post-burn-in loop. */
int bb_bridge_batched_multi(int method, int ortho,
                            const double *X, const double *y, int N, int P, const bb_hyper *h,
                            int nchains, unsigned long long seed, unsigned int chain0,
                            int burn, int nsamp, int thin, int ndevices, const int *devices,
                            double *beta, double *aux1, double *aux2, double *aux3,
                            double *sig2, double *tau, double *alpha, unsigned int *status, double *runtime);

/* The same run in SUMMARY mode: the draws of beta stay in HBM and sum.stat is taken on the
 * device.  pooled: [7][P] (mean, sd, ESS, ESS.sec, t, Q.1, Q.9 over all chains, ESS.sec = ESS / runtime of the
 * post-burn-in loop); per_chain: [nchains][5][P] (mean, sd, ESS, Q.1, Q.9); either may be NULL (not both).
 * sig2 / tau: [nchains][nsamp] or NULL.  nsamp >= 2, no upper limit: the summaries are running ones (bb_chains_summary_begin /
 * add / end) over a draw buffer of at most 1024 slots; per_chain's quantile rows are NaN (see there). */
int bb_bridge_batched_summary(int method, int ortho,
                              const double *X, const double *y, int N, int P, const bb_hyper *h,
                              int nchains, unsigned long long seed, unsigned int chain0,
                              int burn, int nsamp, int thin, int device,
                              double *pooled, double *per_chain, double *sig2, double *tau,
                              unsigned int *status, double *runtime);

#ifdef __cplusplus
}
#endif
#endif
