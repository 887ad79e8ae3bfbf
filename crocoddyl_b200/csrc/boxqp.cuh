// crocoddyl_b200 — BoxQP (projected Newton for a box-constrained QP) and the control-limited computeGains of
// SolverBoxDDP / SolverBoxFDDP, executed by ONE thread on caller-provided work arrays (registers / local memory in the
// thread-per-problem core, per-problem global scratch in the CTA-per-problem core).
//
// Reference functions restated here (paths relative to the reference tree):
//   BoxQP::solve                     src/core/solvers/box-qp.cpp:52-186
//   SolverBoxDDP::computeGains       src/core/solvers/box-ddp.cpp:46-74  (== SolverBoxFDDP::computeGains, box-fddp.cpp:46-74)
//   BoxQP(nu, 100, 0.1, 1e-5, 0.)    box-ddp.cpp:16 (maxiter, th_acceptstep, th_grad, reg of the solvers' QP)
#pragma once
#include "common.cuh"

namespace cgx {

// doubles / ints of work space for an n-variable QP
__host__ __device__ constexpr int boxqp_work_doubles(int n) { return 5 * n + 2 * n * n; }
__host__ __device__ constexpr int boxqp_work_ints(int n) { return 2 * n; }

struct BoxQPOut {
  double* x;        // [n] solution
  double* Hff_inv;  // [hdim x hdim] column-major: inverse of the free-space Hessian AS LAST COMPUTED (box-qp.cpp:104)
  int* free_idx;    // [nf]
  int* clamped_idx; // [nc]
  int nf, nc, hdim;
};

// Hff = H(free, free) (+ reg) -> unblocked lower LLT in L (ld nf), inverse in Hinv (ld nf). false: pivot <= 0.
__device__ inline bool boxqp_factor_free(int nf, const int* fidx, const double* H, int ldh, double reg, double* L,
                                         double* Hinv) {
  for (int j = 0; j < nf; ++j)
    for (int i = 0; i < nf; ++i) L[i + nf * j] = H[fidx[i] + ldh * fidx[j]] + ((i == j && reg != 0.) ? reg : 0.);
  for (int k = 0; k < nf; ++k) {
    double p = L[k + nf * k];
    for (int j = 0; j < k; ++j) p -= L[k + nf * j] * L[k + nf * j];
    if (p <= 0.) return false;
    p = sqrt(p);
    L[k + nf * k] = p;
    for (int i = k + 1; i < nf; ++i) {
      double s = L[i + nf * k];
      for (int j = 0; j < k; ++j) s -= L[i + nf * j] * L[k + nf * j];
      L[i + nf * k] = s / p;
    }
  }
  for (int c = 0; c < nf; ++c) {
    double* b = Hinv + nf * c;
    for (int i = 0; i < nf; ++i) b[i] = (i == c) ? 1. : 0.;
    for (int i = 0; i < nf; ++i) {
      double s = b[i];
      for (int j = 0; j < i; ++j) s -= L[i + nf * j] * b[j];
      b[i] = s / L[i + nf * i];
    }
    for (int i = nf - 1; i >= 0; --i) {
      double s = b[i];
      for (int j = i + 1; j < nf; ++j) s -= L[j + nf * i] * b[j];
      b[i] = s / L[i + nf * i];
    }
  }
  return true;
}

// BoxQP::solve. H: n x n column-major (leading dimension ldh), q, lb, ub, xinit: [n]. work: boxqp_work_doubles(n)
// doubles, iwork: boxqp_work_ints(n) ints. Returns 0, or 1 = "backward_error" (box-qp.cpp:112-114,152-154).
// The reference always runs `maxiter` Newton iterations unless |g|_inf <= th_grad (g includes the clamped components,
// which do not vanish at a constrained optimum). An iteration in which no step length is accepted leaves x — and so
// every later iteration — unchanged, so the loop stops there: same x, same index sets, same Hff_inv.
__device__ inline int boxqp_solve(int n, const double* H, int ldh, const double* q, const double* lb, const double* ub,
                                  const double* xinit, int maxiter, double th_acceptstep, double th_grad, double reg,
                                  int nalphas, const double* alphas, double* work, int* iwork, BoxQPOut& out) {
  double* x = work;
  double* xnew = x + n;
  double* g = xnew + n;
  double* dx = g + n;
  double* dxf = dx + n;
  double* L = dxf + n;
  double* Hinv = L + n * n;
  int* fidx = iwork;
  int* cidx = iwork + n;
  out.x = x;
  out.Hff_inv = Hinv;
  out.free_idx = fidx;
  out.clamped_idx = cidx;
  out.hdim = 0;
  for (int i = 0; i < n; ++i) x[i] = fmax(fmin(xinit[i], ub[i]), lb[i]);  // box-qp.cpp:76-78
  int nf = 0, nc = 0;
  for (int k = 0; k < maxiter; ++k) {
    nf = nc = 0;
    double gnorm = 0.;
    for (int i = 0; i < n; ++i) {  // g = q + H x   (box-qp.cpp:85-86)
      double s = 0.;
      for (int j = 0; j < n; ++j) s += H[i + ldh * j] * x[j];
      g[i] = q[i] + s;
      const double a = fabs(g[i]);
      if (!(a <= gnorm)) gnorm = a;
    }
    for (int j = 0; j < n; ++j) {
      if ((x[j] == lb[j] && g[j] > 0.) || (x[j] == ub[j] && g[j] < 0.))
        cidx[nc++] = j;
      else
        fidx[nf++] = j;
    }
    out.nf = nf;
    out.nc = nc;
    if (gnorm <= th_grad || nf == 0) {  // box-qp.cpp:103-125
      if (k == 0) {
        if (!boxqp_factor_free(nf, fidx, H, ldh, reg, L, Hinv)) return 1;
        out.hdim = nf;
      }
      return 0;
    }
    if (!boxqp_factor_free(nf, fidx, H, ldh, reg, L, Hinv)) return 1;
    out.hdim = nf;
    // Newton step in the free space (box-qp.cpp:127-170)
    for (int i = 0; i < nf; ++i) {
      double v = -q[fidx[i]];
      if (nc != 0) {
        double s = 0.;
        for (int j = 0; j < nc; ++j) s += H[fidx[i] + ldh * cidx[j]] * x[cidx[j]];
        v -= s;
      }
      dxf[i] = v;
    }
    for (int i = 0; i < nf; ++i) {
      double s = dxf[i];
      for (int j = 0; j < i; ++j) s -= L[i + nf * j] * dxf[j];
      dxf[i] = s / L[i + nf * i];
    }
    for (int i = nf - 1; i >= 0; --i) {
      double s = dxf[i];
      for (int j = i + 1; j < nf; ++j) s -= L[j + nf * i] * dxf[j];
      dxf[i] = s / L[i + nf * i];
    }
    for (int i = 0; i < n; ++i) dx[i] = 0.;
    for (int i = 0; i < nf; ++i) dx[fidx[i]] = dxf[i] - x[fidx[i]];
    // line search (box-qp.cpp:172-183)
    double xHx = 0., qx = 0.;
    for (int i = 0; i < n; ++i) {
      double s = 0.;
      for (int j = 0; j < n; ++j) s += H[i + ldh * j] * x[j];
      xHx += x[i] * s;
      qx += q[i] * x[i];
    }
    const double fold = 0.5 * xHx + qx;
    bool moved = false;
    for (int a = 0; a < nalphas; ++a) {
      const double steplength = alphas[a];
      for (int i = 0; i < n; ++i) xnew[i] = fmax(fmin(x[i] + steplength * dx[i], ub[i]), lb[i]);
      double nHn = 0., qn = 0., gd = 0.;
      for (int i = 0; i < n; ++i) {
        double s = 0.;
        for (int j = 0; j < n; ++j) s += H[i + ldh * j] * xnew[j];
        nHn += xnew[i] * s;
        qn += q[i] * xnew[i];
        gd += g[i] * (x[i] - xnew[i]);
      }
      const double fnew = 0.5 * nHn + qn;
      if (fold - fnew > th_acceptstep * gd) {
        for (int i = 0; i < n; ++i) x[i] = xnew[i];
        moved = true;
        break;
      }
    }
    if (!moved) {
      // x is a fixed point of the iteration: the remaining iterations of the reference repeat this one. What it returns
      // after maxiter iterations are the index sets / Hff_inv of an iteration identical to this one.
      return 0;
    }
  }
  // maxiter reached (box-qp.cpp:185-186): solution_.x = x_ with the index sets of the last iteration
  return 0;
}

// SolverBoxDDP::computeGains for one node by one thread (box-ddp.cpp:46-74), once Quu (with ureg), Qu, Qxu are formed
// and the caller has checked has_control_limits && is_feasible (else: the vanilla gains).
//   in : Quu (m x m, ld ldq), Qu[m] (MODIFIED: clamped entries zeroed, :70-72), Qxu (n x m, ld ldx), u[m] = us[t],
//        lb/ub[m] = u_lb/u_ub of the model, k[m] = the previous gains (warm start, :57)
//   out: K (m x n, ld ldk), k[m], Quu_inv (m x m, ld m, zero outside the free block)
// Returns 0 or 1 ("backward_error").
__device__ inline int box_compute_gains(int n, int m, const double* Quu, int ldq, double* Qu, const double* Qxu, int ldx,
                                        const double* u, const double* lb, const double* ub, double* K, int ldk,
                                        double* k, double* Quu_inv, const SolverOpts& o, double* work, int* iwork) {
  double* du_lb = work;
  double* du_ub = work + m;
  double* xinit = work + 2 * m;
  for (int i = 0; i < m; ++i) {
    du_lb[i] = lb[i] - u[i];
    du_ub[i] = ub[i] - u[i];
    xinit[i] = k[i];
  }
  BoxQPOut sol;
  // the solvers' QP: BoxQP(nu, 100, 0.1, 1e-5, 0.) with its own 10 step lengths 2^-n (box-ddp.cpp:16, box-qp.cpp:41-46)
  double qa[10];
  for (int a = 0; a < 10; ++a) qa[a] = 1. / (double)(1 << a);
  (void)o;
  if (boxqp_solve(m, Quu, ldq, Qu, du_lb, du_ub, xinit, 100, 0.1, 1e-5, 0., 10, qa, work + 3 * m, iwork, sol)) return 1;
  for (int e = 0; e < m * m; ++e) Quu_inv[e] = 0.;
  for (int i = 0; i < sol.nf; ++i)
    for (int j = 0; j < sol.nf; ++j)
      Quu_inv[sol.free_idx[i] + m * sol.free_idx[j]] = (i < sol.hdim && j < sol.hdim) ? sol.Hff_inv[i + sol.hdim * j] : 0.;
  for (int j = 0; j < n; ++j)  // K = Quu_inv Qxu^T
    for (int i = 0; i < m; ++i) {
      double s = 0.;
      for (int l = 0; l < m; ++l) s += Quu_inv[i + m * l] * Qxu[j + ldx * l];
      K[i + ldk * j] = s;
    }
  for (int i = 0; i < m; ++i) k[i] = -sol.x[i];
  for (int i = 0; i < sol.nc; ++i) Qu[sol.clamped_idx[i]] = 0.;
  return 0;
}
__host__ __device__ constexpr int box_gains_work_doubles(int m) { return 3 * m + boxqp_work_doubles(m); }
__host__ __device__ constexpr int box_gains_work_ints(int m) { return boxqp_work_ints(m); }

}  // namespace cgx
