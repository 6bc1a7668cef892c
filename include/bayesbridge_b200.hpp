// bayesbridge_b200.hpp -- header-only C++ face of libbayesbridge_b200.so for host C++ callers.
//
// Mirrors the two C++-level interfaces of the reference that its own test driver uses
// (/root/reference/Code/C/test_bridge.cpp:76-86):
//   * class BridgeRegression (Code/C/BridgeRegression.h:81-166): constructor = the one-time factors
//     (BridgeRegression.cpp:13-69), then one method per conditional with the reference's names
//     (sample_tau_marg :453-465, sample_sig2 :436-450, sample_omega :130-147, sample_u :97-110, sample_beta :405-433,
//     sample_lambda :506-510, sample_beta_stable :552-575, least_squares :79-91).  Here the state of ALL chains lives in
//     HBM and a method is one stage-masked launch over every chain (bb_chains_run), so the methods take no per-chain
//     arguments: read and write the state with get() / set().
//   * the four driver functions (Code/C/BridgeWrapper.h:81-157): bridge_regression, bridge_regression_stable,
//     bridge_regression_ortho, bridge_regression_stable_ortho -- same argument order, std::vector<double> in place of the
//     author's Matrix (P x 1 x M column-major == P-fastest flat arrays), return value = runtime in seconds.
// Only the C ABI (bayesbridge_b200.h) crosses the library boundary; everything here is inline.
#ifndef BAYESBRIDGE_B200_HPP
#define BAYESBRIDGE_B200_HPP

#include <cmath>
#include <stdexcept>
#include <string>
#include <vector>

#include "bayesbridge_b200.h"

namespace bayesbridge {

inline void check(int rc, const char* what) {
  if (rc) throw std::runtime_error(std::string(what) + ": " + bb_last_error());
}

inline bb_hyper make_hyper(double alpha, double sig2_shape = 0.0, double sig2_scale = 0.0, double nu_shape = 2.0, double nu_rate = 2.0,
                           double alpha_a = 1.0, double alpha_b = 1.0, double true_sig2 = 0.0, double true_tau = 0.0, int betaburn = 0) {
  bb_hyper h;
  h.sig2_shape = sig2_shape; h.sig2_scale = sig2_scale; h.nu_shape = nu_shape; h.nu_rate = nu_rate; h.alpha_a = alpha_a;
  h.alpha_b = alpha_b; h.true_sig2 = true_sig2; h.true_tau = true_tau; h.true_alpha = alpha; h.betaburn = betaburn;
  return h;
}

// state of all chains, [nchains][P] (P fastest) / [nchains]
struct State {
  std::vector<double> beta, u, aux /* omega (tri) | lambda (stb) */, shape, sig2, tau;
  std::vector<unsigned int> status;
};

class BridgeRegression {
 public:
  // BridgeRegression(const MF& X, const MF& y), BridgeRegression.cpp:13-69: X is N x P column-major
  BridgeRegression(const double* X, const double* y, int N, int P, int device = 0) : N_(N), P_(P) {
    m_ = bb_model_create(X, y, N, P, device, BB_MEM_HOST);
    if (!m_) throw std::runtime_error(std::string("BridgeRegression: ") + bb_last_error());
  }
  ~BridgeRegression() { if (c_) bb_chains_free(c_); if (m_) bb_model_free(m_); }
  BridgeRegression(const BridgeRegression&) = delete;
  BridgeRegression& operator=(const BridgeRegression&) = delete;

  int rows() const { return N_; }
  int cols() const { return P_; }

  // least_squares(Matrix& ls), BridgeRegression.cpp:79-91 (0 when X'X is singular)
  std::vector<double> least_squares() const {
    std::vector<double> b(P_);
    check(bb_model_get(m_, nullptr, nullptr, b.data(), nullptr, nullptr, nullptr), "least_squares");
    return b;
  }

  // the slices the drivers carry (BridgeWrapper.cpp:120-129, 241-250): nchains chains at the drivers' initial state
  void start_chains(int method, int nchains, unsigned long long seed, bool ortho = false, unsigned int chain0 = 0) {
    if (c_) { bb_chains_free(c_); c_ = nullptr; }
    c_ = bb_chains_create(m_, method, ortho ? 1 : 0, nchains, seed, chain0);
    if (!c_) throw std::runtime_error(std::string("start_chains: ") + bb_last_error());
    nchains_ = nchains; iter_ = 0;
  }

  // one conditional over all chains; `iter` is the Gibbs iteration index (the Philox counter word)
  void sample_tau_marg(const bb_hyper& h) { stage(h, BB_ST_TAU); }        // BridgeRegression.cpp:453-465
  void sample_sig2(const bb_hyper& h) { stage(h, BB_ST_SIG2); }          // :436-450
  void sample_omega(const bb_hyper& h) { stage(h, BB_ST_OMEGA); }        // :130-147
  void sample_u(const bb_hyper& h) { stage(h, BB_ST_U); }                // :97-110
  void sample_beta(const bb_hyper& h) { stage(h, BB_ST_BETA); }          // :405-433 / the ortho form :362-403
  void sample_lambda(const bb_hyper& h) { stage(h, BB_ST_LAMBDA); }      // :506-510
  void sample_beta_stable(const bb_hyper& h) { stage(h, BB_ST_BETA); }   // :552-575 / the ortho form :514-521
  void sample_alpha_marg(const bb_hyper& h) { stage(h, BB_ST_ALPHA); }   // :477-503 (acts when h.true_alpha <= 0)
  // whole iterations in the drivers' order; advances the iteration counter
  void sweep(const bb_hyper& h, int niter, bool pre_tau = false) {
    need_chains();
    check(bb_chains_run(c_, &h, iter_, niter, BB_ST_ALL, pre_tau ? 1 : 0, 0, 0, nullptr), "sweep");
    iter_ += (unsigned)niter;
  }
  void next_iteration() { ++iter_; }
  unsigned int iteration() const { return iter_; }

  State get() const {
    need_chains();
    State s;
    const size_t n = (size_t)nchains_ * P_;
    s.beta.resize(n); s.u.resize(n); s.aux.resize(n); s.shape.resize(n); s.sig2.resize(nchains_); s.tau.resize(nchains_);
    s.status.resize(nchains_);
    check(bb_chains_get(c_, s.beta.data(), s.u.data(), s.aux.data(), s.shape.data(), s.sig2.data(), s.tau.data(), s.status.data()), "get");
    return s;
  }
  void set(const double* beta, const double* u = nullptr, const double* aux = nullptr, const double* sig2 = nullptr,
           const double* tau = nullptr) {
    need_chains();
    check(bb_chains_set(c_, beta, u, aux, sig2, tau), "set");
  }
  bb_model* model() const { return m_; }
  bb_chains* chains() const { return c_; }

 private:
  void need_chains() const { if (!c_) throw std::runtime_error("BridgeRegression: call start_chains first"); }
  void stage(const bb_hyper& h, int mask) {
    need_chains();
    check(bb_chains_run(c_, &h, iter_, 1, mask, 0, 0, 0, nullptr), "conditional");
  }
  bb_model* m_ = nullptr;
  bb_chains* c_ = nullptr;
  int N_ = 0, P_ = 0, nchains_ = 0;
  unsigned int iter_ = 0;
};

// ---- the four drivers, Code/C/BridgeWrapper.h:81-157 (argument order kept; M = beta.size() / P samples) ----
namespace detail {
inline double tri(bool ortho, std::vector<double>& beta, std::vector<double>& u, std::vector<double>& omega, std::vector<double>& shape,
                  std::vector<double>& sig2, std::vector<double>& tau, std::vector<double>& alpha, const std::vector<double>& y,
                  const std::vector<double>& X, double sig2_shape, double sig2_scale, double nu_shape, double nu_rate, double alpha_a,
                  double alpha_b, double true_sig2, double true_tau, double true_alpha, int burn, int betaburn) {
  const int N = (int)y.size(), P = (int)(X.size() / y.size()), M = (int)(beta.size() / (size_t)P);
  const int orth = ortho ? 1 : 0, hmc = 0;
  double rt = 0.0;
  u.resize((size_t)P * M); omega.resize((size_t)P * M); shape.resize((size_t)P * M); sig2.resize(M); tau.resize(M); alpha.resize(M);
  ::bridge_regression(beta.data(), u.data(), omega.data(), shape.data(), sig2.data(), tau.data(), alpha.data(), y.data(), X.data(),
                      &sig2_shape, &sig2_scale, &nu_shape, &nu_rate, &alpha_a, &alpha_b, &true_sig2, &true_tau, &true_alpha, &P, &N, &M,
                      &burn, &rt, &orth, &betaburn, &hmc);
  return rt;
}
inline double stb(bool ortho, std::vector<double>& beta, std::vector<double>& lambda, std::vector<double>& sig2, std::vector<double>& tau,
                  std::vector<double>& alpha, const std::vector<double>& y, const std::vector<double>& X, double sig2_shape,
                  double sig2_scale, double nu_shape, double nu_rate, double alpha_a, double alpha_b, double true_sig2, double true_tau,
                  double true_alpha, int burn) {
  const int N = (int)y.size(), P = (int)(X.size() / y.size()), M = (int)(beta.size() / (size_t)P);
  const int orth = ortho ? 1 : 0;
  double rt = 0.0;
  lambda.resize((size_t)P * M); sig2.resize(M); tau.resize(M); alpha.resize(M);
  ::bridge_reg_stable(beta.data(), lambda.data(), sig2.data(), tau.data(), alpha.data(), y.data(), X.data(), &sig2_shape, &sig2_scale,
                      &nu_shape, &nu_rate, &alpha_a, &alpha_b, &true_sig2, &true_tau, &true_alpha, &P, &N, &M, &burn, &rt, &orth);
  return rt;
}
}  // namespace detail

#define BB_TRI_ARGS std::vector<double>& beta, std::vector<double>& u, std::vector<double>& omega, std::vector<double>& shape, \
  std::vector<double>& sig2, std::vector<double>& tau, std::vector<double>& alpha, const std::vector<double>& y,                 \
  const std::vector<double>& X, double sig2_shape, double sig2_scale, double nu_shape, double nu_rate, double alpha_a,          \
  double alpha_b, double true_sig2, double true_tau, double true_alpha, int burn
#define BB_STB_ARGS std::vector<double>& beta, std::vector<double>& lambda, std::vector<double>& sig2, std::vector<double>& tau, \
  std::vector<double>& alpha, const std::vector<double>& y, const std::vector<double>& X, double sig2_shape, double sig2_scale,  \
  double nu_shape, double nu_rate, double alpha_a, double alpha_b, double true_sig2, double true_tau, double true_alpha, int burn

inline double bridge_regression(BB_TRI_ARGS, int betaburn = 0) {            // BridgeWrapper.cpp:80-204
  return detail::tri(false, beta, u, omega, shape, sig2, tau, alpha, y, X, sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b,
                     true_sig2, true_tau, true_alpha, burn, betaburn);
}
inline double bridge_regression_ortho(BB_TRI_ARGS) {                        // :320-431
  return detail::tri(true, beta, u, omega, shape, sig2, tau, alpha, y, X, sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b,
                     true_sig2, true_tau, true_alpha, burn, 0);
}
inline double bridge_regression_stable(BB_STB_ARGS) {                       // :207-313
  return detail::stb(false, beta, lambda, sig2, tau, alpha, y, X, sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b, true_sig2,
                     true_tau, true_alpha, burn);
}
inline double bridge_regression_stable_ortho(BB_STB_ARGS) {                 // :434-537
  return detail::stb(true, beta, lambda, sig2, tau, alpha, y, X, sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b, true_sig2,
                     true_tau, true_alpha, burn);
}
#undef BB_TRI_ARGS
#undef BB_STB_ARGS

}  // namespace bayesbridge
#endif
