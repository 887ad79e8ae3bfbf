// crocoddyl_b200 — thread-per-problem (TPP) DDP/FDDP core for tiny node models (nx <= ~4).
//
// One CUDA thread owns one OCP and runs the reference's whole solve() loop on the device
// (per-problem state machine, no host round trip per iteration). All per-node arrays are stored
// structure-of-arrays with the BATCH INDEX FASTEST: element i of node t of problem b of a field of
// width D lives at  ptr[(t*D + i)*Bp + b]  so that the 32 lanes of a warp (32 problems at the same
// node) touch 32 consecutive doubles = one fully used 256-byte line.
//
// Node derivatives (Fx,Fu,L**) and the gaps are generated INSIDE the backward sweep from (xs,us);
// they never touch HBM. The accepted trial becomes the candidate by flipping a per-problem buffer
// index (I_CUR) instead of copying xs_try/us_try.
//
// Reference functions restated here (paths relative to the reference tree):
//   SolverDDP::calcDiff/backwardPass/computeGains  src/core/solvers/ddp.cpp:157-253,298-310
//   SolverDDP::forwardPass/expectedImprovement/stoppingCriteria  ddp.cpp:132-155,255-296
//   SolverFDDP::forwardPass/(update)expectedImprovement  src/core/solvers/fddp.cpp:107-225
//   SolverDDP::solve / SolverFDDP::solve  ddp.cpp:41-118, fddp.cpp:19-105
#pragma once
#include "common.cuh"

namespace cgx {

struct TppCtx {
  int B, Bp, T;
  int fddp;        // 1 = SolverFDDP, 0 = SolverDDP
  int full;        // keep Qxx,Qxu,Quu,Qx,Qu per node
  double* xs[2];   // [T+1][NX][Bp] candidate / trial (selected by I_CUR)
  double* us[2];   // [T][NU][Bp]
  double* fs;      // [T+1][NX][Bp]
  double* K;       // [T][NU*NX][Bp] column-major NU x NX
  double* k;       // [T][NU][Bp]
  double* Vx;      // [T+1][NX][Bp]
  double* Vxx;     // [T+1][NX*NX][Bp]
  double* w;       // [T+1][NX][Bp]   w_t = Vxx_t fs_t  (fddp.cpp:112-118,132-144 only use Vxx through it)
  double* Qxx;     // full only
  double* Qxu;
  double* Quu;
  double* Qx;
  double* Qu;
  double* x0;      // [NX][Bp]
  double* scal;    // [S_NSLOTS][Bp]
  int* iscal;      // [I_NSLOTS][Bp]
  double* trace;   // [cap][NTRACE][Bp] or null
  int trace_cap;
  const double* mparams;  // shared model parameters
  SolverOpts o;
};

// ---------------------------------------------------------------------------------------------
// ActionModelUnicycle (reference: include/crocoddyl/core/actions/unicycle.hxx:13-73, unicycle.hpp:78-80)
struct UnicycleModel {
  static constexpr int NX = 3;
  static constexpr int NU = 2;
  double dt, w0, w1;
  __device__ void load(const double* p) {
    dt = p[0];
    w0 = p[1];
    w1 = p[2];
  }
  // calc: xnext, cost (unicycle.hxx:22-42)
  __device__ double calc(const double* x, const double* u, double* xnext) const {
    double s, c;
    sincos(x[2], &s, &c);
    xnext[0] = x[0] + c * u[0] * dt;
    xnext[1] = x[1] + s * u[0] * dt;
    xnext[2] = x[2] + u[1] * dt;
    double rr = 0.;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double r = w0 * x[i];
      rr += r * r;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const double r = w1 * u[i];
      rr += r * r;
    }
    return 0.5 * rr;
  }
  // calc + calcDiff in one go (one sincos). Matrices column-major. (unicycle.hxx:45-73)
  __device__ double calcBoth(const double* x, const double* u, double* xnext, double* Fx, double* Fu, double* Lx,
                             double* Lu, double* Lxx, double* Lxu, double* Luu) const {
    double s, c;
    sincos(x[2], &s, &c);
    xnext[0] = x[0] + c * u[0] * dt;
    xnext[1] = x[1] + s * u[0] * dt;
    xnext[2] = x[2] + u[1] * dt;
    double rr = 0.;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double r = w0 * x[i];
      rr += r * r;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const double r = w1 * u[i];
      rr += r * r;
    }
    const double wx = w0 * w0, wu = w1 * w1;
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      Fx[i] = 0.;
      Lxx[i] = 0.;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      Fu[i] = 0.;
      Lxu[i] = 0.;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) Luu[i] = 0.;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      Lx[i] = x[i] * wx;
      Lxx[i + 3 * i] = wx;
      Fx[i + 3 * i] = 1.;  // data ctor (unicycle.hpp:78-80)
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      Lu[i] = u[i] * wu;
      Luu[i + 2 * i] = wu;
    }
    Fx[0 + 3 * 2] = -s * u[0] * dt;
    Fx[1 + 3 * 2] = c * u[0] * dt;
    Fu[0 + 3 * 0] = c * dt;
    Fu[1 + 3 * 0] = s * dt;
    Fu[2 + 3 * 1] = dt;
    return 0.5 * rr;
  }
};

#define TPP_BLOCK 64
#ifdef TPP_NOINLINE
#define TPP_SWEEP_ATTR __device__ __noinline__
#else
#define TPP_SWEEP_ATTR __device__ __forceinline__
#endif
#define TPP_AT(ptr, t, D, i) (ptr)[((size_t)(t) * (D) + (i)) * (size_t)Bp + b]

// ---------------------------------------------------------------------------------------------
// problem_->calc(xs_, us_) in the reference's summation order (shooting.hxx:133-161): returns total cost.
template <class M>
TPP_SWEEP_ATTR double tpp_calc_pass(const TppCtx& c, const M& m, int b, const double* __restrict__ xs,
                                const double* __restrict__ us) {
  constexpr int NX = M::NX, NU = M::NU;
  const int Bp = c.Bp, T = c.T;
  double cost = 0.;
  double x[NX], u[NU], xn[NX];
  for (int t = 0; t < T; ++t) {
#pragma unroll
    for (int i = 0; i < NX; ++i) x[i] = TPP_AT(xs, t, NX, i);
#pragma unroll
    for (int i = 0; i < NU; ++i) u[i] = TPP_AT(us, t, NU, i);
    cost += m.calc(x, u, xn);
  }
#pragma unroll
  for (int i = 0; i < NX; ++i) x[i] = TPP_AT(xs, T, NX, i);
#pragma unroll
  for (int i = 0; i < NU; ++i) u[i] = 0.;
  cost += m.calc(x, u, xn);
  return cost;
}

// ---------------------------------------------------------------------------------------------
// SolverDDP::calcDiff (gaps) + backwardPass + computeGains, fused with the reductions of
// updateExpectedImprovement / expectedImprovement(DDP) / stoppingCriteria.
// Returns 0, or 1 = "backward_error" (LLT pivot <= 0, ddp.cpp:302-304; NaN/inf/>=1e30 in Vx/Vxx, :246-251).
// The gap term of node t+1 (Vx += Vxx fs, ddp.cpp:189-191,242-244) is applied when node t is visited,
// because fs[t+1] = xnext(xs[t],us[t]) - xs[t+1] (ddp.cpp:160-171) needs node t's calc; the operations
// and their order per node are unchanged.
template <class M>
TPP_SWEEP_ATTR int tpp_backward(const TppCtx& c, const M& m, int b, const double* __restrict__ xs,
                            const double* __restrict__ us, bool recalc, bool is_feasible, bool was_feasible,
                            double xreg, double ureg, double& dg_out, double& dq_out, double& stop_out) {
  constexpr int N = M::NX, U = M::NU;
  const int Bp = c.Bp, T = c.T;
  const bool fddp = c.fddp != 0;
  const bool full = c.full != 0;
  double dg = 0., dq = 0., stop = 0.;
  const bool has_xreg = !isnan(xreg), has_ureg = !isnan(ureg);  // ddp.cpp:185,216,226,237
  double Vxx[N * N], Vx[N], xp[N];
  double Fx[N * N], Fu[N * U], Lx[N], Lu[U], Lxx[N * N], Lxu[N * U], Luu[U * U], xn[N];
  double x[N], u[U];
  // terminal node (ddp.cpp:181-187): u = unone_ = 0 (action-base.hxx:27-37)
#pragma unroll
  for (int i = 0; i < N; ++i) xp[i] = TPP_AT(xs, T, N, i);
#pragma unroll
  for (int i = 0; i < U; ++i) u[i] = 0.;
  m.calcBoth(xp, u, xn, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu);
#pragma unroll
  for (int i = 0; i < N * N; ++i) Vxx[i] = Lxx[i];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    Vx[i] = Lx[i];
    if (has_xreg) Vxx[i + N * i] += xreg;
  }
#pragma unroll
  for (int i = 0; i < N * N; ++i) TPP_AT(c.Vxx, T, N * N, i) = Vxx[i];

  // software prefetch: node t-1's (x,u) are requested before node t's dependent Riccati chain starts
  double xq[N], uq[U];
#pragma unroll
  for (int i = 0; i < N; ++i) xq[i] = TPP_AT(xs, T - 1, N, i);
#pragma unroll
  for (int i = 0; i < U; ++i) uq[i] = TPP_AT(us, T - 1, U, i);
  for (int t = T - 1; t >= -1; --t) {
    double f[N];
    if (t >= 0) {
#pragma unroll
      for (int i = 0; i < N; ++i) x[i] = xq[i];
#pragma unroll
      for (int i = 0; i < U; ++i) u[i] = uq[i];
      if (t > 0) {
#pragma unroll
        for (int i = 0; i < N; ++i) xq[i] = TPP_AT(xs, t - 1, N, i);
#pragma unroll
        for (int i = 0; i < U; ++i) uq[i] = TPP_AT(us, t - 1, U, i);
      }
      m.calcBoth(x, u, xn, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu);
    } else {
      // gap of node 0: fs[0] = diff(xs[0], x0) = x0 - xs[0]   (ddp.cpp:162)
#pragma unroll
      for (int i = 0; i < N; ++i) xn[i] = c.x0[(size_t)i * Bp + b];
    }
    // ---- finish node t+1: gap term, Vx store ----
    if (!is_feasible) {
      if (recalc) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
          f[i] = xn[i] - xp[i];
          TPP_AT(c.fs, t + 1, N, i) = f[i];
        }
      } else {
#pragma unroll
        for (int i = 0; i < N; ++i) f[i] = TPP_AT(c.fs, t + 1, N, i);
      }
      double wv[N];
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
#pragma unroll
        for (int j = 0; j < N; ++j) s += Vxx[i + N * j] * f[j];
        wv[i] = s;
        Vx[i] += s;
        TPP_AT(c.w, t + 1, N, i) = s;
      }
      if (fddp) {  // updateExpectedImprovement gap terms (fddp.cpp:131-135,141-145)
        double a = 0., q = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) {
          a += Vx[i] * f[i];
          q += f[i] * wv[i];
        }
        dg -= a;
        dq += q;
      }
    } else if (recalc && !was_feasible) {
#pragma unroll
      for (int i = 0; i < N; ++i) TPP_AT(c.fs, t + 1, N, i) = 0.;  // ddp.cpp:172-176
    }
    {
      double nrm = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        TPP_AT(c.Vx, t + 1, N, i) = Vx[i];
        const double v = fabs(Vx[i]);
        if (!(v <= nrm)) nrm = v;
      }
      if (t + 1 < T && raise_if_nan(nrm)) return 1;  // ddp.cpp:246-248
    }
    if (t < 0) break;

    // ---- node t (ddp.cpp:202-244) ----
    double A[N * N];  // FxTVxx_p = Fx^T Vxx'
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) s += Fx[kk + N * i] * Vxx[kk + N * j];
        A[i + N * j] = s;
      }
    double Qxx[N * N], Qx[N], Qxu[N * U], Quu[U * U], Qu[U], G[U * N];  // G = FuTVxx_p = Fu^T Vxx'
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) s += A[i + N * kk] * Fx[kk + N * j];
        Qxx[i + N * j] = Lxx[i + N * j] + s;
      }
#pragma unroll
    for (int j = 0; j < N; ++j) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += Fx[i + N * j] * Vx[i];
      Qx[j] = Lx[j] + s;
    }
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < U; ++i) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) s += Fu[kk + N * i] * Vxx[kk + N * j];
        G[i + U * j] = s;
      }
#pragma unroll
    for (int j = 0; j < U; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) s += A[i + N * kk] * Fu[kk + N * j];
        Qxu[i + N * j] = Lxu[i + N * j] + s;
      }
#pragma unroll
    for (int j = 0; j < U; ++j)
#pragma unroll
      for (int i = 0; i < U; ++i) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) s += G[i + U * kk] * Fu[kk + N * j];
        Quu[i + U * j] = Luu[i + U * j] + s;
      }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += Fu[i + N * j] * Vx[i];
      Qu[j] = Lu[j] + s;
      if (has_ureg) Quu[j + U * j] += ureg;
    }
    // computeGains (ddp.cpp:298-310): unblocked lower LLT, failure iff pivot <= 0
    double L[U * U];
#pragma unroll
    for (int i = 0; i < U * U; ++i) L[i] = Quu[i];
    bool fail = false;
#pragma unroll
    for (int kk = 0; kk < U; ++kk) {
      double p = L[kk + U * kk];
#pragma unroll
      for (int j = 0; j < kk; ++j) p -= L[kk + U * j] * L[kk + U * j];
      if (p <= 0.) fail = true;
      p = sqrt(p);
      L[kk + U * kk] = p;
#pragma unroll
      for (int i = kk + 1; i < U; ++i) {
        double s = L[i + U * kk];
#pragma unroll
        for (int j = 0; j < kk; ++j) s -= L[i + U * j] * L[kk + U * j];
        L[i + U * kk] = s / p;
      }
    }
    if (fail) return 1;
    double Kg[U * N], kg[U];
    // K = Quu^-1 Qxu^T  (columns of K), k = Quu^-1 Qu
#pragma unroll
    for (int col = 0; col <= N; ++col) {
      double r[U];
#pragma unroll
      for (int i = 0; i < U; ++i) r[i] = (col < N) ? Qxu[col + N * i] : Qu[i];
#pragma unroll
      for (int i = 0; i < U; ++i) {
        double s = r[i];
#pragma unroll
        for (int j = 0; j < i; ++j) s -= L[i + U * j] * r[j];
        r[i] = s / L[i + U * i];
      }
#pragma unroll
      for (int i = U - 1; i >= 0; --i) {
        double s = r[i];
#pragma unroll
        for (int j = i + 1; j < U; ++j) s -= L[j + U * i] * r[j];
        r[i] = s / L[i + U * i];
      }
#pragma unroll
      for (int i = 0; i < U; ++i) {
        if (col < N)
          Kg[i + U * col] = r[i];
        else
          kg[i] = r[i];
      }
    }
    double Quuk[U];
#pragma unroll
    for (int i = 0; i < U; ++i) {
      double s = 0.;
#pragma unroll
      for (int j = 0; j < U; ++j) s += Quu[i + U * j] * kg[j];
      Quuk[i] = s;
    }
    // reductions: ddp.cpp:132-155 / fddp.cpp:137-140
    {
      double a = 0., q = 0., st = 0.;
#pragma unroll
      for (int i = 0; i < U; ++i) {
        a += Qu[i] * kg[i];
        q += kg[i] * Quuk[i];
        st += Qu[i] * Qu[i];
      }
      dg += a;
      dq -= q;
      stop += st;
    }
    // stores
#pragma unroll
    for (int i = 0; i < U * N; ++i) TPP_AT(c.K, t, U * N, i) = Kg[i];
#pragma unroll
    for (int i = 0; i < U; ++i) TPP_AT(c.k, t, U, i) = kg[i];
    if (full) {
#pragma unroll
      for (int i = 0; i < N * N; ++i) TPP_AT(c.Qxx, t, N * N, i) = Qxx[i];
#pragma unroll
      for (int i = 0; i < N * U; ++i) TPP_AT(c.Qxu, t, N * U, i) = Qxu[i];
#pragma unroll
      for (int i = 0; i < U * U; ++i) TPP_AT(c.Quu, t, U * U, i) = Quu[i];
#pragma unroll
      for (int i = 0; i < N; ++i) TPP_AT(c.Qx, t, N, i) = Qx[i];
#pragma unroll
      for (int i = 0; i < U; ++i) TPP_AT(c.Qu, t, U, i) = Qu[i];
    }
    // Vx = Qx + K^T Quuk - 2 K^T Qu ; Vxx = Qxx - Qxu K   (ddp.cpp:223-233)
#pragma unroll
    for (int j = 0; j < N; ++j) {
      double s1 = 0., s2 = 0.;
#pragma unroll
      for (int i = 0; i < U; ++i) {
        s1 += Kg[i + U * j] * Quuk[i];
        s2 += Kg[i + U * j] * Qu[i];
      }
      double v = Qx[j];
      if (has_ureg) {
        v += s1;
        v += -2. * s2;
      } else {
        v += -1. * s2;  // ddp.cpp:226-227
      }
      Vx[j] = v;
    }
    double Vr[N * N];
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < U; ++kk) s += Qxu[i + N * kk] * Kg[kk + U * j];
        Vr[i + N * j] = Qxx[i + N * j] + (-1.) * s;
      }
    double nrm = 0.;
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double v = 0.5 * (Vr[i + N * j] + Vr[j + N * i]);  // ddp.cpp:235
        if (i == j && has_xreg) v += xreg;                 // ddp.cpp:237-239
        Vxx[i + N * j] = v;
        TPP_AT(c.Vxx, t, N * N, i + N * j) = v;
        const double av = fabs(v);
        if (!(av <= nrm)) nrm = av;
      }
    if (raise_if_nan(nrm)) return 1;  // ddp.cpp:249-251
#pragma unroll
    for (int i = 0; i < N; ++i) xp[i] = x[i];
  }
  dg_out = dg;
  dq_out = dq;
  stop_out = stop;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// SolverFDDP::forwardPass (fddp.cpp:149-225) / SolverDDP::forwardPass (ddp.cpp:255-296, == the closed
// branch with xs_try[0] = x0) fused with SolverFDDP::expectedImprovement's gap reduction
// dv = -sum_t fs_t^T Vxx_t (xs_t - xs_try_t) = sum_t w_t^T dx_t (fddp.cpp:107-124).
// Returns 0, or 1 = "forward_error".
template <class M>
TPP_SWEEP_ATTR int tpp_forward(const TppCtx& c, const M& m, int b, const double* __restrict__ xs,
                           const double* __restrict__ us, double* __restrict__ xs_try, double* __restrict__ us_try,
                           double alpha, bool is_feasible, double& cost_try_out, double& dv_out) {
  constexpr int N = M::NX, U = M::NU;
  const int Bp = c.Bp, T = c.T;
  const bool fddp = c.fddp != 0;
  const bool closed = !fddp || is_feasible || (alpha == 1.);
  const bool use_w = fddp && !is_feasible;
  const double am1 = alpha - 1.;
  double cost_try = 0., dv = 0.;
  double xn[N], xt[N], dx[N], u[U];
#pragma unroll
  for (int i = 0; i < N; ++i) xn[i] = c.x0[(size_t)i * Bp + b];
  // software prefetch of node t+1's inputs (none of them depends on the rollout recursion)
  double qx[N], qf[N], qw[N], qu[U], qk[U], qK[U * N];
  auto fetch = [&](int t) {
#pragma unroll
    for (int i = 0; i < N; ++i) qx[i] = TPP_AT(xs, t, N, i);
    if (!closed) {
#pragma unroll
      for (int i = 0; i < N; ++i) qf[i] = TPP_AT(c.fs, t, N, i);
    }
    if (use_w) {
#pragma unroll
      for (int i = 0; i < N; ++i) qw[i] = TPP_AT(c.w, t, N, i);
    }
    if (t < T) {
#pragma unroll
      for (int i = 0; i < U; ++i) {
        qu[i] = TPP_AT(us, t, U, i);
        qk[i] = TPP_AT(c.k, t, U, i);
      }
#pragma unroll
      for (int i = 0; i < U * N; ++i) qK[i] = TPP_AT(c.K, t, U * N, i);
    }
  };
  fetch(0);
  for (int t = 0; t <= T; ++t) {
    double cx[N], cw[N], cu[U], ck[U], cK[U * N];
    if (closed) {
#pragma unroll
      for (int i = 0; i < N; ++i) xt[i] = xn[i];
    } else {
#pragma unroll
      for (int i = 0; i < N; ++i) xt[i] = xn[i] + qf[i] * am1;  // integrate(xnext, fs*(a-1))
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
      cx[i] = qx[i];
      cw[i] = qw[i];
    }
#pragma unroll
    for (int i = 0; i < U; ++i) {
      cu[i] = qu[i];
      ck[i] = qk[i];
    }
#pragma unroll
    for (int i = 0; i < U * N; ++i) cK[i] = qK[i];
    if (t < T) fetch(t + 1);
#pragma unroll
    for (int i = 0; i < N; ++i) {
      dx[i] = xt[i] - cx[i];
      TPP_AT(xs_try, t, N, i) = xt[i];
    }
    if (use_w) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) s += cw[i] * dx[i];
      dv += s;
    }
    if (t == T) break;
#pragma unroll
    for (int i = 0; i < U; ++i) {
      double s = 0.;
#pragma unroll
      for (int j = 0; j < N; ++j) s += cK[i + U * j] * dx[j];
      u[i] = cu[i] - ck[i] * alpha - s;
      TPP_AT(us_try, t, U, i) = u[i];
    }
    cost_try += m.calc(xt, u, xn);
    double nrm = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const double v = fabs(xn[i]);
      if (!(v <= nrm)) nrm = v;
    }
    if (raise_if_nan(cost_try) || raise_if_nan(nrm)) return 1;
  }
#pragma unroll
  for (int i = 0; i < U; ++i) u[i] = 0.;
  cost_try += m.calc(xt, u, xn);
  if (raise_if_nan(cost_try)) return 1;
  cost_try_out = cost_try;
  dv_out = dv;
  return 0;
}

// acceptance tests of the line search (ddp.cpp:84-93 / fddp.cpp:64-80)
__device__ inline bool accept_step(const SolverOpts& o, bool fddp, bool is_feasible, double d0, double dV,
                                   double dVexp) {
  if (fddp) {
    if (dVexp >= 0) return d0 < o.th_grad || dV > o.th_acceptstep * dVexp;
    return dV > o.th_acceptnegstep * dVexp;
  }
  if (dVexp >= 0) return d0 < o.th_grad || !is_feasible || dV > o.th_acceptstep * dVexp;
  return false;
}

#define TPP_S(slot) c.scal[(size_t)(slot) * Bp + b]
#define TPP_I(slot) c.iscal[(size_t)(slot) * Bp + b]

// ---------------------------------------------------------------------------------------------
// Whole SolverDDP::solve / SolverFDDP::solve loop for one problem per thread (ddp.cpp:41-118, fddp.cpp:19-105).
template <class M>
__global__ void tpp_solve_kernel(const __grid_constant__ TppCtx c, int maxiter, int is_feasible0, double reginit) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const int Bp = c.Bp;
  const bool fddp = c.fddp != 0;
  M m;
  m.load(c.mparams);
  int cur = TPP_I(I_CUR);
  bool feas = is_feasible0 != 0, wasfeas = false;
  double xreg = isnan(reginit) ? c.o.regmin : reginit, ureg = xreg;
  double cost = TPP_S(S_COST), stop = TPP_S(S_STOP), d0 = TPP_S(S_D0), d1 = TPP_S(S_D1);
  double steplength = TPP_S(S_STEP), dV = TPP_S(S_DV), dVexp = TPP_S(S_DVEXP), cost_try = TPP_S(S_COST_TRY);
  double dg = 0., dq = 0.;
  int nbwd = 0, nfwd = 0, nrec = 0, converged = 0, iter = 0;
  bool recalc = true, need_gaps = true, done = false;
  for (iter = 0; iter < maxiter && !done; ++iter) {
    const double* xs = c.xs[cur];
    const double* us = c.us[cur];
    int retries = 0;
    if (recalc && iter == 0) {  // ddp.cpp:158
      cost = tpp_calc_pass(c, m, b, xs, us);
      ++nfwd;
    }
    if (recalc) need_gaps = true;
    while (true) {  // computeDirection with regularisation retry (ddp.cpp:55-68)
      const int st = tpp_backward(c, m, b, xs, us, need_gaps, feas, wasfeas, xreg, ureg, dg, dq, stop);
      ++nbwd;
      if (st == 0) {
        need_gaps = false;
        break;
      }
      recalc = false;
      reg_increase(c.o, xreg, ureg);
      ++retries;
      if (xreg == c.o.regmax) {
        done = true;
        break;
      }
    }
    if (done) break;  // return false (iter_ keeps its value)
    if (!fddp) {      // SolverDDP::expectedImprovement before the line search (ddp.cpp:71)
      d0 = dg;
      d1 = dq;
    }
    recalc = false;
    int accepted = -1;
    for (int a = 0; a < c.o.nalphas; ++a) {
      steplength = c.o.alphas[a];
      double dv = 0.;
      const int st = tpp_forward(c, m, b, xs, us, c.xs[cur ^ 1], c.us[cur ^ 1], steplength, feas, cost_try, dv);
      ++nfwd;
      if (st) continue;
      dV = cost - cost_try;
      if (fddp) {
        d0 = dg + dv;
        d1 = dq - 2 * dv;
      }
      dVexp = steplength * (d0 + 0.5 * steplength * d1);
      if (accept_step(c.o, fddp, feas, d0, dV, dVexp)) {
        wasfeas = feas;
        feas = fddp ? (wasfeas || (steplength == 1.)) : true;
        cost = cost_try;
        cur ^= 1;
        recalc = true;
        accepted = a;
        break;
      }
    }
    if (steplength > c.o.th_stepdec) reg_decrease(c.o, xreg, ureg);
    if (steplength <= c.o.th_stepinc) {
      reg_increase(c.o, xreg, ureg);
      if (xreg == c.o.regmax) break;  // return false
    }
    if (c.trace && nrec < c.trace_cap) {
      double* tr = c.trace + (size_t)nrec * NTRACE * Bp + b;
      tr[0 * (size_t)Bp] = iter;
      tr[1 * (size_t)Bp] = cost;
      tr[2 * (size_t)Bp] = stop;
      tr[3 * (size_t)Bp] = d0;
      tr[4 * (size_t)Bp] = d1;
      tr[5 * (size_t)Bp] = xreg;
      tr[6 * (size_t)Bp] = ureg;
      tr[7 * (size_t)Bp] = steplength;
      tr[8 * (size_t)Bp] = feas ? 1. : 0.;
      tr[9 * (size_t)Bp] = dV;
      tr[10 * (size_t)Bp] = dVexp;
      tr[11 * (size_t)Bp] = accepted;
      tr[12 * (size_t)Bp] = retries;
    }
    ++nrec;
    if (wasfeas && stop < c.o.th_stop) {
      converged = 1;
      break;  // return true with iter_ = iter
    }
  }
  TPP_S(S_COST) = cost;
  TPP_S(S_STOP) = stop;
  TPP_S(S_D0) = d0;
  TPP_S(S_D1) = d1;
  TPP_S(S_XREG) = xreg;
  TPP_S(S_UREG) = ureg;
  TPP_S(S_STEP) = steplength;
  TPP_S(S_DV) = dV;
  TPP_S(S_DVEXP) = dVexp;
  TPP_S(S_COST_TRY) = cost_try;
  TPP_S(S_DG) = dg;
  TPP_S(S_DQ) = dq;
  TPP_I(I_ITER) = iter;
  TPP_I(I_FEAS) = feas;
  TPP_I(I_WASFEAS) = wasfeas;
  TPP_I(I_CONVERGED) = converged;
  TPP_I(I_CUR) = cur;
  TPP_I(I_NREC) = nrec;
  TPP_I(I_NBWD) = nbwd;
  TPP_I(I_NFWD) = nfwd;
}

// ---------------------------------------------------------------------------------------------
// Step-wise kernels (same device functions, state kept in scal/iscal between calls).
// computeDirection(recalc) (ddp.cpp:120-125): calcDiff (calc + gaps, ddp.cpp:157-178; the step-wise API
// always runs with iter_ == 0, so problem_->calc is evaluated) + backwardPass.
template <class M>
__global__ void __launch_bounds__(64) tpp_direction_kernel(const __grid_constant__ TppCtx c, int recalc) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const int Bp = c.Bp;
  M m;
  m.load(c.mparams);
  const int cur = TPP_I(I_CUR);
  const bool feas = TPP_I(I_FEAS) != 0, wasfeas = TPP_I(I_WASFEAS) != 0;
  if (recalc) TPP_S(S_COST) = tpp_calc_pass(c, m, b, c.xs[cur], c.us[cur]);
  double dg = 0., dq = 0., stop = 0.;
  const int st = tpp_backward(c, m, b, c.xs[cur], c.us[cur], recalc != 0, feas, wasfeas, TPP_S(S_XREG),
                              TPP_S(S_UREG), dg, dq, stop);
  TPP_I(I_BSTATUS) = st;
  if (st == 0) {
    TPP_S(S_DG) = dg;
    TPP_S(S_DQ) = dq;
    TPP_S(S_STOP) = stop;
    if (!c.fddp) {
      TPP_S(S_D0) = dg;
      TPP_S(S_D1) = dq;
    }
  }
}

// tryStep(alpha) (ddp.cpp:127-130)
template <class M>
__global__ void __launch_bounds__(64) tpp_trystep_kernel(const __grid_constant__ TppCtx c, double alpha) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const int Bp = c.Bp;
  M m;
  m.load(c.mparams);
  const int cur = TPP_I(I_CUR);
  const bool feas = TPP_I(I_FEAS) != 0;
  double cost_try = 0., dv = 0.;
  const int st = tpp_forward(c, m, b, c.xs[cur], c.us[cur], c.xs[cur ^ 1], c.us[cur ^ 1], alpha, feas, cost_try, dv);
  TPP_I(I_FSTATUS) = st;
  if (st == 0) {
    TPP_S(S_COST_TRY) = cost_try;
    TPP_S(S_DV) = TPP_S(S_COST) - cost_try;
    TPP_S(S_DVGAP) = dv;
    TPP_S(S_STEP) = alpha;
    if (c.fddp) {
      TPP_S(S_D0) = TPP_S(S_DG) + dv;
      TPP_S(S_D1) = TPP_S(S_DQ) - 2 * dv;
    }
  }
}

// ShootingProblem::calc / calcDiff for one (problem, node) per thread, results in the host ("user") layout
// [B][T+1][size] (shooting.hxx:133-195). xs/us in SoA layout.
template <class M>
__global__ void tpp_node_data_kernel(TppCtx c, const double* __restrict__ xs, const double* __restrict__ us,
                                     int diff, double* cost, double* xnext, double* Fx, double* Fu, double* Lx,
                                     double* Lu, double* Lxx, double* Lxu, double* Luu) {
  constexpr int N = M::NX, U = M::NU;
  const int Bp = c.Bp, T = c.T;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int t = blockIdx.y;
  if (b >= c.B) return;
  M m;
  m.load(c.mparams);
  double x[N], u[U], xn[N], fx[N * N], fu[N * U], lx[N], lu[U], lxx[N * N], lxu[N * U], luu[U * U];
#pragma unroll
  for (int i = 0; i < N; ++i) x[i] = TPP_AT(xs, t, N, i);
#pragma unroll
  for (int i = 0; i < U; ++i) u[i] = (t < T) ? TPP_AT(us, t < T ? t : 0, U, i) : 0.;
  const size_t node = (size_t)b * (T + 1) + t;
  if (!diff) {
    cost[node] = m.calc(x, u, xn);
    for (int i = 0; i < N; ++i) xnext[node * N + i] = xn[i];
  } else {
    m.calcBoth(x, u, xn, fx, fu, lx, lu, lxx, lxu, luu);
    for (int i = 0; i < N * N; ++i) Fx[node * N * N + i] = fx[i];
    for (int i = 0; i < N * U; ++i) Fu[node * N * U + i] = fu[i];
    for (int i = 0; i < N; ++i) Lx[node * N + i] = lx[i];
    for (int i = 0; i < U; ++i) Lu[node * U + i] = lu[i];
    for (int i = 0; i < N * N; ++i) Lxx[node * N * N + i] = lxx[i];
    for (int i = 0; i < N * U; ++i) Lxu[node * N * U + i] = lxu[i];
    for (int i = 0; i < U * U; ++i) Luu[node * U * U + i] = luu[i];
  }
}

// ShootingProblem::rollout (shooting.hxx:198-223); xs/us SoA
template <class M>
__global__ void tpp_rollout_kernel(TppCtx c, const double* __restrict__ us, double* __restrict__ xs) {
  constexpr int N = M::NX, U = M::NU;
  const int Bp = c.Bp, T = c.T;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  M m;
  m.load(c.mparams);
  double x[N], u[U], xn[N];
#pragma unroll
  for (int i = 0; i < N; ++i) x[i] = c.x0[(size_t)i * Bp + b];
  for (int t = 0; t < T; ++t) {
#pragma unroll
    for (int i = 0; i < N; ++i) TPP_AT(xs, t, N, i) = x[i];
#pragma unroll
    for (int i = 0; i < U; ++i) u[i] = TPP_AT(us, t, U, i);
    m.calc(x, u, xn);
#pragma unroll
    for (int i = 0; i < N; ++i) x[i] = xn[i];
  }
#pragma unroll
  for (int i = 0; i < N; ++i) TPP_AT(xs, T, N, i) = x[i];
}

// ---------------------------------------------------------------------------------------------
// Layout conversion between the host layout [B][W] (W = nt*D contiguous doubles per problem) and the
// batch-fastest SoA [W][Bp], through a 32x33 shared-memory tile so both sides are coalesced.
// to_soa != 0: user -> soa ; else soa -> user. Problems b0..b0+nb-1 (user buffer holds nb problems).
// If cur != null, soa0/soa1 is picked per problem by cur[b] ^ flip (double-buffered candidate/trial).
__global__ void tpp_transpose_kernel(double* user, double* soa0, double* soa1, const int* cur, int flip, int W, int Bp,
                                     int b0, int nb, int to_soa) {
  __shared__ double tile[32][33];
  const int bx = blockIdx.x * 32;  // problem offset within the range
  const int wy = blockIdx.y * 32;  // element offset
  const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
  if (to_soa) {
    for (int r = ty; r < 32; r += 8) {
      const int bl = bx + r, wi = wy + tx;
      if (bl < nb && wi < W) tile[r][tx] = user[(size_t)bl * W + wi];
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
      const int wi = wy + r, bl = bx + tx;
      if (bl < nb && wi < W) {
        const int b = b0 + bl;
        double* dst = (cur && ((cur[b] ^ flip) & 1)) ? soa1 : soa0;
        dst[(size_t)wi * Bp + b] = tile[tx][r];
      }
    }
  } else {
    for (int r = ty; r < 32; r += 8) {
      const int wi = wy + r, bl = bx + tx;
      if (bl < nb && wi < W) {
        const int b = b0 + bl;
        const double* src = (cur && ((cur[b] ^ flip) & 1)) ? soa1 : soa0;
        tile[tx][r] = src[(size_t)wi * Bp + b];
      }
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
      const int bl = bx + r, wi = wy + tx;
      if (bl < nb && wi < W) user[(size_t)bl * W + wi] = tile[r][tx];
    }
  }
}

// setCandidate bookkeeping (solver-base.cpp:42-67) + reset of the step-wise state
__global__ void tpp_reset_kernel(TppCtx c, int is_feasible) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.Bp) return;
  const int Bp = c.Bp;
  TPP_I(I_ITER) = 0;
  TPP_I(I_FEAS) = is_feasible;
  TPP_I(I_WASFEAS) = 0;
  TPP_I(I_CONVERGED) = 0;
  TPP_I(I_CUR) = 0;
  TPP_I(I_NREC) = 0;
  TPP_I(I_BSTATUS) = 0;
  TPP_I(I_FSTATUS) = 0;
  TPP_I(I_NBWD) = 0;
  TPP_I(I_NFWD) = 0;
  TPP_S(S_DVGAP) = 0.;
}

}  // namespace cgx
