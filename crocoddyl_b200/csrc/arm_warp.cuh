// crocoddyl_b200 — warp-cooperative node calc for the arm rollout (SolverFDDP::forwardPass, fddp.cpp:168,200):
// xnext, cost = IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics).calc(x, u) by ONE WARP.
// The rollout is serial in t, so the latency of one ABA evaluation is the critical path of the whole forward pass
// (measured on B200: 17.6 us per node with ABA on one lane). Here
//   lanes 0..NV-1 : joint placements, body inertias, bias terms of their own joint (parallel over joints)
//   lanes 0..5    : the articulated-inertia sweep, lane r holding ROW r of the 6x6 inertia. Rows transform like
//                   forces (row . X^-1 = actF(X,row)^T), the result of X* Ia X^-1 is symmetric, so one 6x6 transpose
//                   through shared memory per joint turns "columns of X* M" back into the rows of the parent inertia
//   lane NV       : frame placement, log6 and the cost terms, concurrently with the sweep
//   all lanes     : the velocity / acceleration recursions redundantly (no communication)
// Same algorithm and conventions as rb::aba (rigid_body.cuh); summation order inside the 6x6 products differs from
// the single-lane version at rounding level only.
#pragma once
#include "rigid_body.cuh"

namespace cgx {

template <int NV>
struct ArmWarpSmem {
  double X[NV][12];    // liMi: R (row-major 9), p (3)
  double c[NV][6];     // bias acceleration v_i x S_i qd_i
  double pa0[NV][6];   // v_i x* (I_i v_i)
  double Ib[NV][36];   // body inertia as a 6x6 (row-major, [lin; ang])
  double U[NV][6];
  double Du[NV][2];    // Dinv, u
  double Tm[6][7];     // transpose buffer (padded)
  double pv[6];
  double cost;
};

__device__ __forceinline__ double sel6(const rb::Sp& s, int r) {
  return r < 3 ? (r == 0 ? s.lin.x : (r == 1 ? s.lin.y : s.lin.z)) : (r == 3 ? s.ang.x : (r == 4 ? s.ang.y : s.ang.z));
}
__device__ __forceinline__ rb::SE3 ld_se3(const double* p) {
  rb::SE3 M;
#pragma unroll
  for (int i = 0; i < 9; ++i) M.R.m[i] = p[i];
  M.p = rb::V3{p[9], p[10], p[11]};
  return M;
}
__device__ __forceinline__ rb::Sp ld_sp(const double* p) {
  return rb::Sp{rb::V3{p[0], p[1], p[2]}, rb::V3{p[3], p[4], p[5]}};
}

// called by the 32 lanes of one warp; x (2 NV), u (NV or nullptr = zero control of the terminal node) and xnext in
// shared memory; *cost_out written by one lane. W: the warp's scratch.
template <int NV>
__device__ void arm_calc_warp(const double* params, bool terminal, const double* x, const double* u, double* xnext,
                              double* cost_out, ArmWarpSmem<NV>& W) {
  using namespace rb;
  static_assert(NV + 1 <= 32 && NV >= 1, "one lane per joint plus one");
  const int lane = threadIdx.x & 31;
  const ArmParams P{params};
  // S0: per joint (lane i): placement, body inertia
  if (lane < NV) {
    const SE3 X = P.joint_transform(lane, x[lane]);
#pragma unroll
    for (int e = 0; e < 9; ++e) W.X[lane][e] = X.R.m[e];
    W.X[lane][9] = X.p.x;
    W.X[lane][10] = X.p.y;
    W.X[lane][11] = X.p.z;
    const I6 Y = i6_from(P.inertia(lane));
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int cc = 0; cc < 3; ++cc) {
        W.Ib[lane][6 * r + cc] = Y.A(r, cc);
        W.Ib[lane][6 * r + 3 + cc] = Y.B(r, cc);
        W.Ib[lane][6 * (3 + r) + cc] = Y.B(cc, r);
        W.Ib[lane][6 * (3 + r) + 3 + cc] = Y.C(r, cc);
      }
  }
  __syncwarp();
  // S1: velocity recursion, redundantly on every lane; lane i keeps v_i
  Sp vmine = sp_zero();
  {
    Sp v = sp_zero();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const SE3 X = ld_se3(W.X[i]);
      const Sp vj{V3{0., 0., 0.}, x[NV + i] * P.axis(i)};
      v = (i == 0) ? vj : actInvM(X, v) + vj;
      if (lane == i) vmine = v;
    }
  }
  // S2: lane i: bias terms of joint i; lane NV: frame placement, residual and the differential cost
  if (lane < NV) {
    const V3 ax = P.axis(lane);
    const Sp ci = crossM_S(vmine, ax, x[NV + lane]);
    const Sp p0 = crossF(vmine, mulI(P.inertia(lane), vmine));
    W.c[lane][0] = ci.lin.x; W.c[lane][1] = ci.lin.y; W.c[lane][2] = ci.lin.z;
    W.c[lane][3] = ci.ang.x; W.c[lane][4] = ci.ang.y; W.c[lane][5] = ci.ang.z;
    W.pa0[lane][0] = p0.lin.x; W.pa0[lane][1] = p0.lin.y; W.pa0[lane][2] = p0.lin.z;
    W.pa0[lane][3] = p0.ang.x; W.pa0[lane][4] = p0.ang.y; W.pa0[lane][5] = p0.ang.z;
  } else if (lane == NV) {
    const int fp = P.frame_parent();
    SE3 oM = ld_se3(W.X[0]);
#pragma unroll
    for (int i = 1; i < NV; ++i)
      if (i <= fp) oM = compose(oM, ld_se3(W.X[i]));
    const FrameRes fr = frame_residual_at(P, compose(oM, P.frame_placement()));
    W.cost = diff_cost<NV>(P, fr, x, u);
  }
  __syncwarp();
  // S3: articulated-inertia sweep on lanes 0..5 (lane r = row r); the other lanes skip it
  constexpr unsigned M6 = 0x3fu;
  if (lane < 6) {
    const int r = lane;
    double Ir[6] = {0., 0., 0., 0., 0., 0.};  // row r of the inertia propagated from the child
    double par = 0.;                          // component r of the propagated bias force
#pragma unroll
    for (int i = NV - 1; i >= 0; --i) {
      const V3 ax = P.axis(i);
      double Ia[6];
#pragma unroll
      for (int cc = 0; cc < 6; ++cc) Ia[cc] = W.Ib[i][6 * r + cc] + Ir[cc];
      const double pa = W.pa0[i][r] + par;
      const double Ur = Ia[3] * ax.x + Ia[4] * ax.y + Ia[5] * ax.z;
      double Uall[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) Uall[k] = __shfl_sync(M6, Ur, k);
      const double pa3 = __shfl_sync(M6, pa, 3), pa4 = __shfl_sync(M6, pa, 4), pa5 = __shfl_sync(M6, pa, 5);
      const double D = ax.x * Uall[3] + ax.y * Uall[4] + ax.z * Uall[5] + P.armature(i);  // rb::aba
      const double Dinv = 1. / D;
      const double ui = (u ? u[i] : 0.) - (ax.x * pa3 + ax.y * pa4 + ax.z * pa5);
      W.U[i][r] = Ur;
      if (r == 0) {
        W.Du[i][0] = Dinv;
        W.Du[i][1] = ui;
      }
      if (i > 0) {
        double cv[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) cv[k] = W.c[i][k];
        double s = 0.;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          Ia[k] -= Ur * Dinv * Uall[k];
          s += Ia[k] * cv[k];
        }
        const double pr = pa + s + (Dinv * ui) * Ur;
        const SE3 X = ld_se3(W.X[i]);
        const Sp Mrow = actF(X, Sp{V3{Ia[0], Ia[1], Ia[2]}, V3{Ia[3], Ia[4], Ia[5]}});  // row . X^-1
        W.Tm[r][0] = Mrow.lin.x; W.Tm[r][1] = Mrow.lin.y; W.Tm[r][2] = Mrow.lin.z;
        W.Tm[r][3] = Mrow.ang.x; W.Tm[r][4] = Mrow.ang.y; W.Tm[r][5] = Mrow.ang.z;
        W.pv[r] = pr;
        __syncwarp(M6);
        const Sp col{V3{W.Tm[0][r], W.Tm[1][r], W.Tm[2][r]}, V3{W.Tm[3][r], W.Tm[4][r], W.Tm[5][r]}};
        const Sp Ycol = actF(X, col);  // column r of X* M = row r of the (symmetric) parent contribution
        Ir[0] = Ycol.lin.x; Ir[1] = Ycol.lin.y; Ir[2] = Ycol.lin.z;
        Ir[3] = Ycol.ang.x; Ir[4] = Ycol.ang.y; Ir[5] = Ycol.ang.z;
        par = sel6(actF(X, ld_sp(W.pv)), r);
        __syncwarp(M6);  // Tm / pv are rewritten by the next joint
      }
    }
  }
  __syncwarp();
  // S4: acceleration recursion, redundantly on every lane; lane i writes joint i of xnext (euler.hxx:65-74)
  double dt = P.dt();  // the terminal node has its own block (dt = 0 by default, benchmark/factory/arm.hpp:89-90)
  if (dt < 0.) dt = 1e-3;
  {
    const V3 g = P.gravity();
    Sp a{V3{-g.x, -g.y, -g.z}, V3{0., 0., 0.}};
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const SE3 X = ld_se3(W.X[i]);
      const Sp ai = actInvM(X, a) + ld_sp(W.c[i]);
      const double qdd = W.Du[i][0] * (W.Du[i][1] - dotS(ld_sp(W.U[i]), ai));
      a = ai + Sp{V3{0., 0., 0.}, qdd * P.axis(i)};
      if (lane == i) {
        if (dt != 0.) {
          const double dq = x[NV + i] * dt + qdd * (dt * dt);
          const double dv = qdd * dt;
          xnext[i] = x[i] + dq;
          xnext[NV + i] = x[NV + i] + dv;
        } else {
          xnext[i] = x[i];
          xnext[NV + i] = x[NV + i];
        }
      }
    }
  }
  if (lane == NV) *cost_out = (dt != 0.) ? dt * W.cost : W.cost;
}

}  // namespace cgx
