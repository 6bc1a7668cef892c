// Matrix.h -- ORACLE compat shim (test infrastructure; never shipped, never on the product path).
//
// Stand-in for the reference author's un-vendored `Matrix` / `MatrixFrame` library
// (#include "Matrix.h" at /root/reference/Code/C/BridgeRegression.h:67 and BridgeWrapper.h:65;
// built from $(HOME)/Code/Matrix, /root/reference/Code/C/Makefile:4,60-64).  It provides exactly
// the surface the reference's BridgeRegression.cpp / BridgeWrapper.cpp use, so those files compile
// UNMODIFIED against it.  Semantics: column-major rows x cols x mats array of doubles; dense
// linear algebra is standard BLAS/LAPACK (scipy's bundled OpenBLAS, `scipy_`-prefixed symbols).
#ifndef BB_COMPAT_MATRIX_H
#define BB_COMPAT_MATRIX_H

#include <sys/types.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <stdexcept>
#include <algorithm>
#include <iostream>

using std::max;
using std::min;
using std::isfinite;
using std::cout;

#ifndef uint
typedef unsigned int uint;
#endif

extern "C" {
void scipy_dgemm_(const char*, const char*, const int*, const int*, const int*, const double*, const double*,
                  const int*, const double*, const int*, const double*, double*, const int*);
void scipy_dpotrf_(const char*, const int*, double*, const int*, int*);
void scipy_dtrsm_(const char*, const char*, const char*, const char*, const int*, const int*, const double*,
                  const double*, const int*, double*, const int*);
void scipy_dgesvd_(const char*, const char*, const int*, const int*, double*, const int*, double*, double*,
                   const int*, double*, const int*, double*, const int*, int*);
void scipy_dsyev_(const char*, const char*, const int*, double*, const int*, double*, double*, const int*, int*);
void scipy_dpotrs_(const char*, const int*, const int*, const double*, const int*, double*, const int*, int*);
double scipy_ddot_(const int*, const double*, const int*, const double*, const int*);
}

class Matrix;

class MatrixFrame {
 protected:
  double* p_;
  uint nr_, nc_, nm_;

 public:
  MatrixFrame() : p_(0), nr_(0), nc_(0), nm_(0) {}
  explicit MatrixFrame(double* p, uint r, uint c = 1, uint m = 1) : p_(p), nr_(r), nc_(c), nm_(m) {}
  uint rows() const { return nr_; }
  uint cols() const { return nc_; }
  uint mats() const { return nm_; }
  uint area() const { return nr_ * nc_; }
  uint vol() const { return nr_ * nc_ * nm_; }
  double& operator()(uint i) { return p_[i]; }
  const double& operator()(uint i) const { return p_[i]; }
  double& operator()(uint i, uint j) { return p_[i + (size_t)j * nr_]; }
  const double& operator()(uint i, uint j) const { return p_[i + (size_t)j * nr_]; }
  MatrixFrame operator[](uint k) const { return MatrixFrame(p_ + (size_t)k * nr_ * nc_, nr_, nc_, 1); }
  MatrixFrame col(uint j) const { return MatrixFrame(p_ + (size_t)j * nr_, nr_, 1, 1); }
  void fill(double v) { for (size_t i = 0; i < vol(); ++i) p_[i] = v; }
  // copy contents (shapes must hold the same number of elements)
  void copy(const MatrixFrame& M) {
    if (M.vol() != vol()) throw std::runtime_error("MatrixFrame::copy: size mismatch");
    std::memmove(p_, &M(0), sizeof(double) * vol());
  }
  // this(0,k) = M(row, cols(k))      (BridgeRegression.cpp:33)
  void copy(const MatrixFrame& M, uint row, const MatrixFrame& cs) {
    for (uint k = 0; k < cs.vol(); ++k) p_[k] = M(row, (uint)cs(k));
  }
  // this(k) = M(rows(k), col)        (BridgeRegression.cpp:381)
  void copy(const MatrixFrame& M, const MatrixFrame& rs, uint colj) {
    for (uint k = 0; k < rs.vol(); ++k) p_[k] = M((uint)rs(k), colj);
  }
};

typedef MatrixFrame MF;

class Matrix : public MatrixFrame {
  std::vector<double> v_;
  void bind(uint r, uint c, uint m) { nr_ = r; nc_ = c; nm_ = m; p_ = v_.empty() ? 0 : &v_[0]; }

 public:
  Matrix() { bind(0, 0, 0); }
  Matrix(uint r, uint c = 1, uint m = 1) : v_((size_t)r * c * m, 0.0) { bind(r, c, m); }
  Matrix(uint r, uint c, uint m, double f) : v_((size_t)r * c * m, f) { bind(r, c, m); }
  Matrix(int r, int c = 1, int m = 1) : v_((size_t)r * c * m, 0.0) { bind(r, c, m); }
  Matrix(int r, int c, int m, double f) : v_((size_t)r * c * m, f) { bind(r, c, m); }
  Matrix(const double* p, uint r, uint c, uint m = 1) : v_(p, p + (size_t)r * c * m) { bind(r, c, m); }
  Matrix(const MatrixFrame& M) : v_(M.vol() ? &M(0) : (const double*)0, M.vol() ? &M(0) + M.vol() : (const double*)0) {
    bind(M.rows(), M.cols(), M.mats());
  }
  Matrix(const Matrix& M) : MatrixFrame(), v_(M.v_) { bind(M.nr_, M.nc_, M.nm_); }
  // "N": natural numbers 1..n ; "W": whole numbers 0..n-1  (BridgeRegression.cpp:31-35, 605, 657)
  Matrix(const char* kind, uint n) : v_(n, 0.0) {
    bind(n, 1, 1);
    double start = (kind[0] == 'N') ? 1.0 : 0.0;
    if (kind[0] == 'I') { v_.assign((size_t)n * n, 0.0); bind(n, n, 1); for (uint i = 0; i < n; ++i) (*this)(i, i) = 1.0; return; }
    for (uint i = 0; i < n; ++i) v_[i] = start + i;
  }
  Matrix& operator=(const Matrix& M) { v_ = M.v_; bind(M.nr_, M.nc_, M.nm_); return *this; }
  void resize(uint r, uint c = 1, uint m = 1) { v_.resize((size_t)r * c * m, 0.0); bind(r, c, m); }
  void clone(const MatrixFrame& M) {
    std::vector<double> t(M.vol() ? &M(0) : (const double*)0, M.vol() ? &M(0) + M.vol() : (const double*)0);
    v_.swap(t); bind(M.rows(), M.cols(), M.mats());
  }
  void clone(const MatrixFrame& M, const MatrixFrame& rs, const MatrixFrame& cs) {
    std::vector<double> t((size_t)rs.vol() * cs.vol());
    for (uint j = 0; j < cs.vol(); ++j) for (uint i = 0; i < rs.vol(); ++i) t[i + (size_t)j * rs.vol()] = M((uint)rs(i), (uint)cs(j));
    v_.swap(t); bind(rs.vol(), cs.vol(), 1);
  }
  void clone(const MatrixFrame& M, const MatrixFrame& rs, uint colj) {
    std::vector<double> t(rs.vol());
    for (uint i = 0; i < rs.vol(); ++i) t[i] = M((uint)rs(i), colj);
    v_.swap(t); bind(rs.vol(), 1, 1);
  }
};

inline Matrix operator-(const MatrixFrame& a, const MatrixFrame& b) {
  Matrix c(a);
  for (size_t i = 0; i < c.vol(); ++i) c((uint)i) -= b((uint)i);
  return c;
}
inline std::ostream& operator<<(std::ostream& os, const MatrixFrame& M) {
  for (uint i = 0; i < M.rows(); ++i) { for (uint j = 0; j < M.cols(); ++j) os << M(i, j) << " "; os << "\n"; }
  return os;
}

// ---- BLAS / LAPACK wrappers (call sites: BridgeRegression.cpp:24-25,38-39,49-54,83,222,246,258,285,
//      382,441,443,560-570,608,631,659,669,674,698) ----

// c = alpha op(a) op(b) + beta c
inline void gemm(MatrixFrame c, const MatrixFrame& a, const MatrixFrame& b, char ta = 'N', char tb = 'N',
                 double alpha = 1.0, double beta = 0.0) {
  int m = (ta == 'N') ? a.rows() : a.cols(), k = (ta == 'N') ? a.cols() : a.rows();
  int kb = (tb == 'N') ? b.rows() : b.cols(), n = (tb == 'N') ? b.cols() : b.rows();
  if (k != kb || (int)c.rows() != m || (int)c.cols() != n) throw std::runtime_error("gemm: shapes do not conform");
  int lda = a.rows(), ldb = b.rows(), ldc = c.rows();
  scipy_dgemm_(&ta, &tb, &m, &n, &k, &alpha, &a(0), &lda, &b(0), &ldb, &beta, &c(0), &ldc);
}
inline void mult(Matrix& c, const MatrixFrame& a, const MatrixFrame& b, char ta = 'N', char tb = 'N') {
  uint m = (ta == 'N') ? a.rows() : a.cols(), n = (tb == 'N') ? b.cols() : b.rows();
  Matrix out(m, n);
  gemm(out, a, b, ta, tb);
  c = out;
}
inline double dot(const MatrixFrame& a, const MatrixFrame& b) {
  int n = a.vol(), one = 1;
  return scipy_ddot_(&n, &a(0), &one, &b(0), &one);
}
// upper/lower Cholesky; the other triangle is zeroed. Throws if not positive definite.
inline void chol(Matrix& U, const MatrixFrame& A, char uplo = 'L') {
  U.clone(A);
  int n = A.rows(), info = 0;
  scipy_dpotrf_(&uplo, &n, &U(0), &n, &info);
  if (info != 0) throw std::runtime_error("chol: matrix is not positive definite");
  for (int j = 0; j < n; ++j) for (int i = 0; i < n; ++i) if ((uplo == 'U') ? (i > j) : (i < j)) U(i, j) = 0.0;
}
inline void trsm(const MatrixFrame& A, MatrixFrame B, char uplo, char side = 'L', char ta = 'N', char diag = 'N',
                 double alpha = 1.0) {
  int m = B.rows(), n = B.cols(), lda = A.rows(), ldb = B.rows();
  scipy_dtrsm_(&side, &uplo, &ta, &diag, &m, &n, &alpha, &A(0), &lda, &B(0), &ldb);
}
// Solve A x = b for symmetric positive definite A, in place in b; A is left untouched.
// Throws when A is singular to working precision (same rule as bbo_port.c least_squares).
inline void symsolve(const MatrixFrame& A, MatrixFrame b) {
  Matrix L(A);
  int n = A.rows(), nrhs = b.cols(), info = 0; char up = 'U';
  scipy_dpotrf_(&up, &n, &L(0), &n, &info);
  if (info != 0) throw std::runtime_error("symsolve: matrix is singular");
  double maxdiag = 0, minpiv = 1e300;
  for (int i = 0; i < n; ++i) { maxdiag = std::max(maxdiag, A(i, i)); minpiv = std::min(minpiv, L(i, i) * L(i, i)); }
  if (minpiv <= 1e-13 * maxdiag) throw std::runtime_error("symsolve: matrix is singular");
  scipy_dpotrs_(&up, &n, &nrhs, &L(0), &n, &b(0), &n, &info);
}
inline void symeig_(Matrix& Q, std::vector<double>& w, const MatrixFrame& A) {
  Q.clone(A);
  int n = A.rows(), info = 0, lwork = -1; double wk; char jobz = 'V', up = 'U';
  w.assign(n, 0.0);
  scipy_dsyev_(&jobz, &up, &n, &Q(0), &n, &w[0], &wk, &lwork, &info);
  lwork = (int)wk; std::vector<double> work(lwork);
  scipy_dsyev_(&jobz, &up, &n, &Q(0), &n, &w[0], &work[0], &lwork, &info);
}
inline void symfun_(Matrix& R, const MatrixFrame& A, double power) {
  Matrix Q; std::vector<double> w; symeig_(Q, w, A);
  uint n = A.rows(); Matrix S(Q);
  for (uint j = 0; j < n; ++j) { double f = std::pow(w[j], power); for (uint i = 0; i < n; ++i) S(i, j) *= f; }
  R.resize(n, n); gemm(R, S, Q, 'N', 'T');
}
inline void symsqrt(Matrix& R, const MatrixFrame& A) { symfun_(R, A, 0.5); }
inline void syminvsqrt(Matrix& R, const MatrixFrame& A) { symfun_(R, A, -0.5); }
// X = U diag(d) tV ; jobz 'S' thin, 'A' full.
inline void svd(Matrix& U, Matrix& d, Matrix& tV, const MatrixFrame& X, char jobz = 'S') {
  Matrix Xc(X);
  int m = X.rows(), n = X.cols(), k = std::min(m, n), info = 0, lwork = -1; double wk;
  int ucols = (jobz == 'A') ? m : k, vrows = (jobz == 'A') ? n : k;
  U.resize(m, ucols); d.resize(k); tV.resize(vrows, n);
  int ldu = m, ldvt = vrows;
  scipy_dgesvd_(&jobz, &jobz, &m, &n, &Xc(0), &m, &d(0), &U(0), &ldu, &tV(0), &ldvt, &wk, &lwork, &info);
  lwork = (int)wk; std::vector<double> work(lwork);
  scipy_dgesvd_(&jobz, &jobz, &m, &n, &Xc(0), &m, &d(0), &U(0), &ldu, &tV(0), &ldvt, &work[0], &lwork, &info);
  if (info != 0) throw std::runtime_error("svd: did not converge");
}
// A <- A diag(d) over the first min(cols, len d) columns  (A = U D, Notes/bbnotes.tex:555)
inline void prodonrow(MatrixFrame A, const MatrixFrame& d) {
  uint k = std::min(A.cols(), d.vol());
  for (uint j = 0; j < A.cols(); ++j) { double f = j < k ? d(j) : 0.0; for (uint i = 0; i < A.rows(); ++i) A(i, j) *= f; }
}
inline double maxAll(const MatrixFrame& M) { double m = M(0); for (uint i = 1; i < M.vol(); ++i) m = std::max(m, M(i)); return m; }
inline double minAll(const MatrixFrame& M) { double m = M(0); for (uint i = 1; i < M.vol(); ++i) m = std::min(m, M(i)); return m; }
// conjugate gradient for SPD A (EM path only); returns iterations used
inline int cg(MatrixFrame x, const MatrixFrame& A, const MatrixFrame& b, double tol, int max_iter) {
  uint n = b.vol(); Matrix r(b), Ap(n); gemm(r, A, x, 'N', 'N', -1.0, 1.0); Matrix p(r);
  double rs = dot(r, r); int it = 0;
  while (it < max_iter && std::sqrt(rs) > tol) {
    gemm(Ap, A, p); double al = rs / dot(p, Ap);
    for (uint i = 0; i < n; ++i) { x(i) += al * p(i); r(i) -= al * Ap(i); }
    double rn = dot(r, r);
    for (uint i = 0; i < n; ++i) p(i) = r(i) + rn / rs * p(i);
    rs = rn; ++it;
  }
  return it;
}
#endif
