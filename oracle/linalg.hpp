// TEST INFRASTRUCTURE — CPU oracle for the batched DDP/FDDP path. Not part of the product.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
// build, link or call anything under oracle/.
//
// Minimal dense column-major linear algebra (stand-in for the Eigen types the reference uses:
// Eigen::VectorXd / Eigen::MatrixXd are column-major, /root/reference/include/crocoddyl/core/mathbase.hpp).
// Everything is written as plain loops; the build uses -ffp-contract=off so that no FMA
// contraction happens behind the source.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstring>
#include <vector>

namespace oracle {

typedef std::vector<double> Vec;

struct Mat {
  int r, c;
  std::vector<double> d;
  Mat() : r(0), c(0) {}
  Mat(int r_, int c_) : r(r_), c(c_), d((size_t)r_ * c_, 0.0) {}
  void resize(int r_, int c_) {
    r = r_;
    c = c_;
    d.assign((size_t)r_ * c_, 0.0);
  }
  inline double& operator()(int i, int j) { return d[(size_t)j * r + i]; }
  inline const double& operator()(int i, int j) const { return d[(size_t)j * r + i]; }
  void setZero() { std::fill(d.begin(), d.end(), 0.0); }
  void setIdentity() {
    setZero();
    for (int i = 0; i < (r < c ? r : c); ++i) (*this)(i, i) = 1.0;
  }
};

inline Mat identity(int r, int c) {
  Mat m(r, c);
  m.setIdentity();
  return m;
}

inline double dot(const Vec& a, const Vec& b) {
  double s = 0;
  for (size_t i = 0; i < a.size(); ++i) s += a[i] * b[i];
  return s;
}
inline double normInf(const Vec& a) {
  double s = 0;
  for (size_t i = 0; i < a.size(); ++i) {
    double v = std::fabs(a[i]);
    if (!(v <= s)) s = v;  // propagates NaN like Eigen's lpNorm<Infinity> (max of abs)
  }
  return s;
}
inline double normInf(const Mat& a) {
  double s = 0;
  for (size_t i = 0; i < a.d.size(); ++i) {
    double v = std::fabs(a.d[i]);
    if (!(v <= s)) s = v;
  }
  return s;
}

// y = A x
inline void gemv(const Mat& A, const Vec& x, Vec& y) {
  y.assign(A.r, 0.0);
  for (int j = 0; j < A.c; ++j)
    for (int i = 0; i < A.r; ++i) y[i] += A(i, j) * x[j];
}
// y += alpha * A x
inline void gemv_add(const Mat& A, const Vec& x, Vec& y, double alpha = 1.0) {
  for (int i = 0; i < A.r; ++i) {
    double s = 0;
    for (int j = 0; j < A.c; ++j) s += A(i, j) * x[j];
    y[i] += alpha * s;
  }
}
// y += alpha * A^T x
inline void gemtv_add(const Mat& A, const Vec& x, Vec& y, double alpha = 1.0) {
  for (int j = 0; j < A.c; ++j) {
    double s = 0;
    for (int i = 0; i < A.r; ++i) s += A(i, j) * x[i];
    y[j] += alpha * s;
  }
}
// C = A^T B
inline void gemm_tn(const Mat& A, const Mat& B, Mat& C) {
  C.resize(A.c, B.c);
  for (int j = 0; j < B.c; ++j)
    for (int i = 0; i < A.c; ++i) {
      double s = 0;
      for (int k = 0; k < A.r; ++k) s += A(k, i) * B(k, j);
      C(i, j) = s;
    }
}
// C += alpha * A B
inline void gemm_nn_add(const Mat& A, const Mat& B, Mat& C, double alpha = 1.0) {
  for (int j = 0; j < B.c; ++j)
    for (int i = 0; i < A.r; ++i) {
      double s = 0;
      for (int k = 0; k < A.c; ++k) s += A(i, k) * B(k, j);
      C(i, j) += alpha * s;
    }
}

// Lower Cholesky, unblocked, the published Eigen::LLT algorithm (Eigen/src/Cholesky/LLT.h,
// llt_inplace<Lower>::unblocked — dependency not vendored in /root/reference; call site
// src/core/solvers/ddp.cpp:298-310): pivot x = a_kk - sum l_kj^2; failure iff x <= 0 (a NaN
// pivot does NOT fail here, it propagates and is caught by raiseIfNaN in backwardPass).
struct LLT {
  Mat L;
  bool ok;
  LLT() : ok(false) {}
  bool compute(const Mat& A) {
    const int n = A.r;
    L = A;
    ok = true;
    for (int k = 0; k < n; ++k) {
      double x = L(k, k);
      for (int j = 0; j < k; ++j) x -= L(k, j) * L(k, j);
      if (x <= 0.0) {
        ok = false;
        return false;
      }
      x = std::sqrt(x);
      L(k, k) = x;
      for (int i = k + 1; i < n; ++i) {
        double s = L(i, k);
        for (int j = 0; j < k; ++j) s -= L(i, j) * L(k, j);
        L(i, k) = s / x;
      }
    }
    return true;
  }
  // in-place solve of (L L^T) X = B for the ncols columns of B (column-major, leading dim n)
  void solveInPlace(double* B, int ncols) const {
    const int n = L.r;
    for (int c = 0; c < ncols; ++c) {
      double* b = B + (size_t)c * n;
      for (int i = 0; i < n; ++i) {
        double s = b[i];
        for (int j = 0; j < i; ++j) s -= L(i, j) * b[j];
        b[i] = s / L(i, i);
      }
      for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int j = i + 1; j < n; ++j) s -= L(j, i) * b[j];
        b[i] = s / L(i, i);
      }
    }
  }
};

// Dense LU with partial pivoting (stand-in for Eigen::PartialPivLU used by
// src/core/solvers/kkt.cpp:184 `kkt_.lu().solve(-kktref_)`). Returns false if singular.
inline bool lu_solve(Mat A, Vec b, Vec& x) {
  const int n = A.r;
  std::vector<int> piv(n);
  for (int k = 0; k < n; ++k) {
    int p = k;
    double best = std::fabs(A(k, k));
    for (int i = k + 1; i < n; ++i)
      if (std::fabs(A(i, k)) > best) {
        best = std::fabs(A(i, k));
        p = i;
      }
    if (best == 0.0) return false;
    if (p != k) {
      for (int j = 0; j < n; ++j) std::swap(A(k, j), A(p, j));
      std::swap(b[k], b[p]);
    }
    for (int i = k + 1; i < n; ++i) {
      double f = A(i, k) / A(k, k);
      A(i, k) = f;
      if (f != 0.0) {
        for (int j = k + 1; j < n; ++j) A(i, j) -= f * A(k, j);
        b[i] -= f * b[k];
      }
    }
  }
  x.assign(n, 0.0);
  for (int i = n - 1; i >= 0; --i) {
    double s = b[i];
    for (int j = i + 1; j < n; ++j) s -= A(i, j) * x[j];
    x[i] = s / A(i, i);
  }
  return true;
}

}  // namespace oracle
