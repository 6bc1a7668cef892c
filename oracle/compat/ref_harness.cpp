// ref_harness.cpp -- ORACLE (test infrastructure; never shipped, never on the product path).
//
// C-callable probes around the UNMODIFIED reference sources
// (/root/reference/Code/C/{BridgeRegression,BridgeWrapper,retstable}.cpp) compiled against
// oracle/compat/{Matrix.h,RNG.hpp}.  Each probe positions the keyed RNG shim and calls one
// reference member function, so the CUDA sampler and the C port can be compared with the
// reference per conditional on identical Philox blocks.
#include "BridgeWrapper.h" // the reference's own header (found through -I/root/reference/Code/C)

static unsigned long long g_seed = 20121027ULL;
static unsigned int g_chain = 0;
extern "C" {
void bb_compat_set_seed(unsigned long long seed, unsigned int chain) { g_seed = seed; g_chain = chain; }
unsigned long long bb_compat_seed() { return g_seed; }
unsigned int bb_compat_chain() { return g_chain; }
}

// BridgeRegression.cpp is compiled with -Dretstable_LD=bb_compat_retstable_item so that
// sample_lambda (BridgeRegression.cpp:506-510) announces each coordinate to the keyed shim
// before entering the reference's retstable_LD (retstable.cpp:94-271).
double bb_compat_retstable_item(double h, double alpha, RNG& r, double V0) {
  r.next_item();
  return retstable_LD(h, alpha, r, V0);
}

namespace {
struct Probe : public BridgeRegression {
  Probe(const MF& X, const MF& y) : BridgeRegression(X, y) {}
  using BridgeRegression::XX; using BridgeRegression::Xy; using BridgeRegression::bhat;
  using BridgeRegression::tV; using BridgeRegression::d; using BridgeRegression::a;
  using BridgeRegression::P;
};
struct Handle { Matrix X, y; Probe* br; int N, P; };
inline MF view(double* p, int n) { return MF(p, (uint)n); }
inline MF cview(const double* p, int n) { return MF(const_cast<double*>(p), (uint)n); }
} // namespace

extern "C" {

void* ref_create(const double* X, const double* y, int N, int P) {
  Handle* h = new Handle();
  h->X = Matrix(X, (uint)N, (uint)P); h->y = Matrix(y, (uint)N, 1u); h->N = N; h->P = P;
  h->br = new Probe(h->X, h->y);
  return h;
}
void ref_free(void* hv) { Handle* h = (Handle*)hv; delete h->br; delete h; }

// d and a are zero-padded to P when P > N (the reference leaves them at length min(N,P),
// BridgeRegression.cpp:47-54, and reads past the end on the triangle path: SURVEY 3.4(v)).
void ref_get_factors(void* hv, double* XX, double* Xy, double* bhat, double* tV, double* d, double* a) {
  Handle* h = (Handle*)hv; Probe* b = h->br; size_t P = h->P;
  if (XX) std::memcpy(XX, &b->XX(0), P * P * sizeof(double));
  if (Xy) std::memcpy(Xy, &b->Xy(0), P * sizeof(double));
  if (bhat) std::memcpy(bhat, &b->bhat(0), P * sizeof(double));
  if (tV) std::memcpy(tV, &b->tV(0), P * P * sizeof(double));
  if (d) { for (size_t i = 0; i < P; ++i) d[i] = i < b->d.vol() ? b->d((uint)i) : 0.0; }
  if (a) { for (size_t i = 0; i < P; ++i) a[i] = i < b->a.vol() ? b->a((uint)i) : 0.0; }
}

double ref_sample_tau_marg(void* hv, const double* beta, double alpha, double nu_shape, double nu_rate,
                           unsigned long long seed, unsigned chain, unsigned iter, unsigned stage) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, stage, RNG::ITEM_PER_CALL);
  double tau = 0; h->br->sample_tau_marg(view(&tau, 1), cview(beta, h->P), alpha, nu_shape, nu_rate, r);
  return tau;
}
double ref_sample_sig2(void* hv, const double* beta, double sig2_shape, double sig2_scale,
                       unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_SIG2, RNG::ITEM_PER_CALL);
  double s = 0; h->br->sample_sig2(view(&s, 1), cview(beta, h->P), sig2_shape, sig2_scale, r);
  return s;
}
void ref_sample_omega(void* hv, double* omega, double* shape, const double* beta, const double* u, double tau, double alpha,
                      unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_OMEGA, RNG::ITEM_ON_UNIF);
  h->br->sample_omega(view(omega, h->P), view(shape, h->P), cview(beta, h->P), cview(u, h->P), tau, alpha, r);
}
void ref_sample_u(void* hv, double* u, const double* beta, const double* omega, double tau, double alpha,
                  unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_U, RNG::ITEM_PER_CALL);
  h->br->sample_u(view(u, h->P), cview(beta, h->P), cview(omega, h->P), tau, alpha, r);
}
void ref_sample_beta(void* hv, double* beta, const double* beta_prev, const double* u, const double* omega,
                     double sig2, double tau, double alpha, int betaburn,
                     unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_BETA, RNG::ITEM_PER_CALL);
  h->br->sample_beta(view(beta, h->P), cview(beta_prev, h->P), cview(u, h->P), cview(omega, h->P), sig2, tau, alpha, r, betaburn, false);
}
// the static SVD-form sweep with caller-supplied factors (BridgeRegression.cpp:235-286)
void ref_rtnorm_gibbs(double* beta, const double* a, const double* tV, const double* d, const double* b, double sig2, int P,
                      unsigned long long seed, unsigned chain, unsigned iter) {
  RNG r(seed, chain); r.position(iter, BBO_ST_BETA, RNG::ITEM_PER_CALL);
  BridgeRegression::rtnorm_gibbs(beta, const_cast<double*>(a), const_cast<double*>(tV), const_cast<double*>(d),
                                 const_cast<double*>(b), &sig2, &P, r);
}
void ref_sample_beta_ortho(void* hv, double* beta, const double* beta_prev, const double* u, const double* omega,
                           double sig2, double tau, double alpha,
                           unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_BETA, RNG::ITEM_PER_CALL);
  h->br->sample_beta_ortho(view(beta, h->P), cview(beta_prev, h->P), cview(u, h->P), cview(omega, h->P), sig2, tau, alpha, r, 0);
}
void ref_sample_lambda(void* hv, double* lambda, const double* beta, double alpha, double tau,
                       unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_LAMBDA, RNG::ITEM_MANUAL);
  h->br->sample_lambda(view(lambda, h->P), cview(beta, h->P), alpha, tau, r);
}
int ref_sample_beta_stable(void* hv, double* beta, const double* lambda, double alpha, double sig2, double tau,
                           unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_BETA_STB, RNG::ITEM_MANUAL);
  try { h->br->sample_beta_stable(view(beta, h->P), cview(lambda, h->P), alpha, sig2, tau, r); }
  catch (std::exception&) { return 1; }
  return 0;
}
void ref_sample_beta_stable_ortho(void* hv, double* beta, const double* lambda, double alpha, double sig2, double tau,
                                  unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_BETA_STB, RNG::ITEM_PER_CALL);
  h->br->sample_beta_stable_ortho(view(beta, h->P), cview(lambda, h->P), alpha, sig2, tau, r);
}
double ref_sample_alpha_marg(void* hv, double alpha_prev, const double* beta, double tau, double pr_a, double pr_b,
                             unsigned long long seed, unsigned chain, unsigned iter) {
  Handle* h = (Handle*)hv; RNG r(seed, chain); r.position(iter, BBO_ST_ALPHA, RNG::ITEM_MANUAL); r.next_item();
  double a_new = 0, a_old = alpha_prev;
  h->br->sample_alpha_marg(view(&a_new, 1), view(&a_old, 1), cview(beta, h->P), tau, r, pr_a, pr_b);
  return a_new;
}
// one tilted-stable draw on the blocks of (iter, LAMBDA, item)   (retstable.cpp:94-271)
double ref_retstable_LD(double hh, double alpha, double V0, unsigned long long seed, unsigned chain, unsigned iter, unsigned item) {
  RNG r(seed, chain); r.position(iter, BBO_ST_LAMBDA, RNG::ITEM_MANUAL);
  for (unsigned i = 0; i <= item; ++i) r.next_item();
  return retstable_LD(hh, alpha, r, V0);
}
double ref_sinc_MM(double x) { return sinc_MM(x); }
double ref_A_(double x, double alpha) { return A_(x, alpha); }
double ref_BdB0(double x, double alpha) { return BdB0(x, alpha); }

} // extern "C"
