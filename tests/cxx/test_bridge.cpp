// test_bridge.cpp -- a host C++ caller of the boundary, shaped like the reference's own driver program
// (/root/reference/Code/C/test_bridge.cpp:76-93: build X, y, run one of the four drivers, print LS next to the posterior mean).
// It includes only include/bayesbridge_b200.h(pp) and links libbayesbridge_b200.so; no Python, no torch.
//
//   test_bridge [--samp M] [--burn B] [--devices G]
//
// Checks (exit code 0 only if all hold):
//   1. the four drivers (triangle / stable, ortho or not) return finite draws of the right layout;
//   2. triangle and stable posterior means agree (the reference's own acceptance test, Code/R/Trial.R:1,98-99) and both shrink
//      the least-squares fit towards zero on the null coefficients;
//   3. class BridgeRegression: the conditionals called one by one reproduce, bit for bit, the fused sweep of the same iteration;
//   4. bb_bridge_batched_multi over G devices equals bb_bridge_batched on one device bit for bit (G = 1 still goes through
//      the fan-out code);
//   5. the interrupt hook aborts a run (BridgeWrapper.cpp:154-156).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "bayesbridge_b200.hpp"

namespace bbr = bayesbridge;

static int g_fail = 0;
#define EXPECT(cond, ...) do { if (!(cond)) { ++g_fail; std::printf("FAIL %s:%d  ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); } } while (0)

static int g_polls = 0;
static int stop_after_three_polls(void) { return ++g_polls >= 3; }

static bool all_finite(const std::vector<double>& v) { for (double x : v) if (!std::isfinite(x)) return false; return true; }

static std::vector<double> col_means(const std::vector<double>& draws, int P, int M, int skip) {
  std::vector<double> m(P, 0.0);
  for (int s = skip; s < M; ++s) for (int j = 0; j < P; ++j) m[j] += draws[(size_t)s * P + j];
  for (double& x : m) x /= (M - skip);
  return m;
}

int main(int argc, char** argv) {
  int M = 3000, burn = 500, G = 1;
  for (int i = 1; i + 1 < argc; i += 2) {
    if (!std::strcmp(argv[i], "--samp")) M = std::atoi(argv[i + 1]);
    else if (!std::strcmp(argv[i], "--burn")) burn = std::atoi(argv[i + 1]);
    else if (!std::strcmp(argv[i], "--devices")) G = std::atoi(argv[i + 1]);
  }
  if (bb_device_count() < 1) { std::printf("no CUDA device: the library has no CPU fallback\n"); return 2; }
  if (G > bb_device_count()) G = bb_device_count();

  // ---- data: N x P column-major, centred unit-norm columns, 3 signals (the shape of the reference's examples) ----
  const int N = 200, P = 20;
  std::mt19937_64 gen(20121027);
  std::normal_distribution<double> nrm(0.0, 1.0);
  std::vector<double> X((size_t)N * P), y(N, 0.0), btrue(P, 0.0);
  for (int j = 0; j < P; ++j) {
    double mean = 0.0, ss = 0.0;
    for (int i = 0; i < N; ++i) { X[(size_t)j * N + i] = nrm(gen) + (j ? 0.4 * X[(size_t)(j - 1) * N + i] : 0.0); mean += X[(size_t)j * N + i]; }
    mean /= N;
    for (int i = 0; i < N; ++i) { X[(size_t)j * N + i] -= mean; ss += X[(size_t)j * N + i] * X[(size_t)j * N + i]; }
    for (int i = 0; i < N; ++i) X[(size_t)j * N + i] /= std::sqrt(ss);
  }
  btrue[1] = 6.0; btrue[7] = -4.0; btrue[13] = 3.0;
  double ymean = 0.0;
  for (int i = 0; i < N; ++i) { for (int j = 0; j < P; ++j) y[i] += X[(size_t)j * N + i] * btrue[j]; y[i] += 0.3 * nrm(gen); ymean += y[i]; }
  for (int i = 0; i < N; ++i) y[i] -= ymean / N;
  std::printf("N: %i, P: %i\n", N, P);

  const double my_alpha = 0.5;
  bb_set_seed(4242);

  // ---- 1./2. the four drivers, called as test_bridge.cpp:76-86 calls them ----
  std::vector<double> beta((size_t)P * M), u, omega, shape, sig2, tau, alpha, lambda;
  std::vector<double> mean_tri, mean_stb;
  double rt = bbr::bridge_regression(beta, u, omega, shape, sig2, tau, alpha, y, X, 0.0, 0.0, 2.0, 2.0, 1.0, 1.0, 0.0, 0.0, my_alpha, burn);
  EXPECT(rt > 0 && all_finite(beta) && all_finite(u) && all_finite(omega) && all_finite(sig2) && all_finite(tau), "bridge_regression output");
  for (size_t e = P; e < shape.size(); ++e) if (shape[e] != 1.0 && shape[e] != 2.0) { EXPECT(false, "shape not in {1,2}"); break; }
  EXPECT(alpha[M - 1] == my_alpha, "alpha is known and must be echoed");
  mean_tri = col_means(beta, P, M, 1);
  std::printf("bridge_regression              : %.3f s for %d samples\n", rt, M);

  rt = bbr::bridge_regression_stable(beta, lambda, sig2, tau, alpha, y, X, 0.0, 0.0, 2.0, 2.0, 1.0, 1.0, 0.0, 0.0, my_alpha, burn);
  EXPECT(rt > 0 && all_finite(beta) && all_finite(lambda) && all_finite(sig2) && all_finite(tau), "bridge_regression_stable output");
  for (double l : lambda) if (!(l > 0)) { EXPECT(false, "lambda must be positive"); break; }
  mean_stb = col_means(beta, P, M, 1);
  std::printf("bridge_regression_stable       : %.3f s\n", rt);

  rt = bbr::bridge_regression_ortho(beta, u, omega, shape, sig2, tau, alpha, y, X, 0.0, 0.0, 2.0, 2.0, 1.0, 1.0, 0.0, 0.0, my_alpha, burn);
  EXPECT(rt > 0 && all_finite(beta) && all_finite(sig2), "bridge_regression_ortho output");
  std::printf("bridge_regression_ortho        : %.3f s\n", rt);
  rt = bbr::bridge_regression_stable_ortho(beta, lambda, sig2, tau, alpha, y, X, 0.0, 0.0, 2.0, 2.0, 1.0, 1.0, 0.0, 0.0, my_alpha, burn);
  EXPECT(rt > 0 && all_finite(beta) && all_finite(sig2), "bridge_regression_stable_ortho output");
  std::printf("bridge_regression_stable_ortho : %.3f s\n", rt);
  std::printf("Regression done.\n");

  // ---- 3. class BridgeRegression ----
  bbr::BridgeRegression br(X.data(), y.data(), N, P, 0);
  const std::vector<double> ls = br.least_squares();
  std::printf("LS PMean(tri) PMean(stb) truth\n");
  double worst = 0.0, shrink_ls = 0.0, shrink_pm = 0.0;
  for (int j = 0; j < P; ++j) {
    std::printf("%9.4f %9.4f %9.4f %6.1f\n", ls[j], mean_tri[j], mean_stb[j], btrue[j]);
    worst = std::fmax(worst, std::fabs(mean_tri[j] - mean_stb[j]));
    if (btrue[j] == 0.0) { shrink_ls += std::fabs(ls[j]); shrink_pm += std::fabs(mean_stb[j]); }
  }
  EXPECT(worst < 0.15, "triangle and stable posterior means differ by %.3f", worst);
  EXPECT(shrink_pm < shrink_ls, "the bridge prior must shrink the null coefficients (%.3f vs LS %.3f)", shrink_pm, shrink_ls);
  {
    const bb_hyper h = bbr::make_hyper(my_alpha);
    const int nch = 3;
    // (a) the conditionals one by one: tau -> sig2 -> omega -> u -> beta, the order of BridgeWrapper.cpp:168-189
    br.start_chains(BB_TRI, nch, 99);
    br.sweep(h, 5, true);
    const bbr::State s0 = br.get();
    br.sample_tau_marg(h); br.sample_sig2(h); br.sample_omega(h); br.sample_u(h); br.sample_beta(h);
    const bbr::State a = br.get();
    // (b) the same iteration as one fused sweep from the same state
    br.start_chains(BB_TRI, nch, 99);
    br.sweep(h, 5, true);
    br.sweep(h, 1);
    const bbr::State b = br.get();
    EXPECT(a.beta == b.beta && a.u == b.u && a.aux == b.aux && a.sig2 == b.sig2 && a.tau == b.tau,
           "triangle: conditionals called one by one differ from the fused sweep");
    EXPECT(s0.beta != a.beta, "the sweep did not move beta");
    // stable: tau -> sig2 -> lambda -> beta (BridgeWrapper.cpp:287-298)
    br.start_chains(BB_STB, nch, 77);
    br.sweep(h, 4, true);
    br.sample_tau_marg(h); br.sample_sig2(h); br.sample_lambda(h); br.sample_beta_stable(h);
    const bbr::State c = br.get();
    br.start_chains(BB_STB, nch, 77);
    br.sweep(h, 4, true);
    br.sweep(h, 1);
    const bbr::State d = br.get();
    EXPECT(c.beta == d.beta && c.aux == d.aux && c.sig2 == d.sig2 && c.tau == d.tau, "stable: conditionals one by one differ from the fused sweep");
    for (unsigned st : d.status) EXPECT(st == 0, "chain status %u", st);
  }

  // ---- 4. in-library device fan-out == one device, bit for bit ----
  {
    const bb_hyper h = bbr::make_hyper(my_alpha);
    const int nch = 10, ns = 40;
    for (int method : {BB_TRI, BB_STB}) {
      std::vector<double> b1((size_t)nch * ns * P), bm(b1.size()), s1((size_t)nch * ns), sm(s1.size()), a1(b1.size()), am(b1.size());
      std::vector<unsigned> st1(nch), stm(nch);
      double r1 = 0, rm = 0;
      bbr::check(bb_bridge_batched(method, 0, X.data(), y.data(), N, P, &h, nch, 7, 100, 20, ns, 2, 0, b1.data(), a1.data(), nullptr, nullptr,
                                   s1.data(), nullptr, nullptr, st1.data(), &r1), "bb_bridge_batched");
      bbr::check(bb_bridge_batched_multi(method, 0, X.data(), y.data(), N, P, &h, nch, 7, 100, 20, ns, 2, G, nullptr, bm.data(), am.data(),
                                         nullptr, nullptr, sm.data(), nullptr, nullptr, stm.data(), &rm), "bb_bridge_batched_multi");
      EXPECT(b1 == bm && s1 == sm && a1 == am, "%s: %d-device fan-out differs from the one-device run", method == BB_TRI ? "tri" : "stb", G);
      EXPECT(rm > 0, "runtime of the fan-out run");
    }
    std::printf("bb_bridge_batched_multi over %d device(s) == bb_bridge_batched, bit for bit\n", G);
    int bad = bb_bridge_batched_multi(BB_STB, 0, X.data(), y.data(), N, P, &h, 4, 7, 0, 2, 5, 1, 999, nullptr, nullptr, nullptr, nullptr, nullptr,
                                      nullptr, nullptr, nullptr, nullptr, nullptr);
    EXPECT(bad != 0 && std::strstr(bb_last_error(), "ndevices"), "an impossible device count must be refused: %s", bb_last_error());
  }

  // ---- 5. interrupt hook ----
  {
    std::vector<double> bi((size_t)P * 200000, 0.0), s2, ta, al, la;
    bb_set_interrupt(stop_after_three_polls);
    const double r = bbr::bridge_regression_stable(bi, la, s2, ta, al, y, X, 0.0, 0.0, 2.0, 2.0, 1.0, 1.0, 0.0, 0.0, my_alpha, 100);
    bb_set_interrupt(nullptr);
    EXPECT(std::isnan(r) && std::isnan(bi[0]) && g_polls == 3, "an interrupted run must come back NaN-filled (runtime %g, polls %d)", r, g_polls);
    EXPECT(std::strstr(bb_last_error(), "interrupted") != nullptr, "error string: %s", bb_last_error());
    const double r2 = bbr::bridge_regression_stable(beta, lambda, sig2, tau, alpha, y, X, 0.0, 0.0, 2.0, 2.0, 1.0, 1.0, 0.0, 0.0, my_alpha, 10);
    EXPECT(r2 > 0 && all_finite(beta), "the library must work again after an interrupted run");
  }

  std::printf("kernel launches: %lld\n", bb_launch_count());
  std::printf(g_fail ? "test_bridge: %d check(s) FAILED\n" : "test_bridge: all checks passed\n", g_fail);
  return g_fail ? 1 : 0;
}
