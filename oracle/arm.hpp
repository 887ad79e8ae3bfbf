// TEST INFRASTRUCTURE — CPU oracle (see linalg.hpp header). Not part of the product.
//
// IntegratedActionModelEuler( DifferentialActionModelFreeFwdDynamics ) on a fixed-base revolute chain, with
// ActuationModelFull and CostModelSum{gripperPose: FramePlacement, uReg: Control, xReg: State}, restated from
//   include/crocoddyl/core/integrator/euler.hxx:17-131
//   include/crocoddyl/multibody/actions/free-fwddyn.hxx:44-169 (ABA branch, with_armature_ == true)
//   include/crocoddyl/multibody/actuations/full.hpp:38-63 (tau = u, dtau_du = I, dtau_dx = 0)
//   include/crocoddyl/multibody/costs/cost-sum.hxx:89-160 (std::map => name order gripperPose < uReg < xReg)
//   include/crocoddyl/multibody/costs/{state.hxx:130-169, control.hxx:56-87, frame-placement.hxx:52-86}
//   include/crocoddyl/core/activations/quadratic.hpp:34-60
//   benchmark/factory/arm.hpp:60-90 (running model dt, terminal model = the SAME differential model with dt = 0)
// The rigid-body arithmetic (Pinocchio in the reference) is in rbd.hpp — parity unpinned, see its header.
//
// Parameter block (doubles), ARM_HEADER + ARM_PER_JOINT * nv:
//   [0] dt  [1..3] gravity  [4] w_gripperPose  [5] w_xReg  [6] w_uReg  [7] frame parent joint (0-based)
//   [8..19] frame placement in its parent joint: R (row-major 9), p (3)     [20..31] Mref: R (9), p (3)
//   [32] goal cost: 0 CostModelFramePlacement, 1 CostModelFrameTranslation, 2 CostModelFrameRotation  [33..35] reserved
//   per joint: axis (3), placement in parent R (9) p (3), mass, COM lever (3), inertia about COM Ixx Ixy Ixz Iyy Iyz Izz,
//              armature (free-fwddyn.hxx:66-74; any non-zero entry selects the Minv branch), weights of the state
//              cost for q_i and v_i (ActivationModelWeightedQuad, weighted-quadratic.hpp:39-49; 1 = ActivationModelQuad)
#pragma once
#include "models.hpp"
#include "rbd.hpp"

namespace oracle {

constexpr int ARM_HEADER = 36, ARM_PER_JOINT = 28;
inline long arm_param_count(int nv) { return ARM_HEADER + (long)ARM_PER_JOINT * nv; }

template <typename S>
inline rbd::SE3<S> read_se3(const double* p) {
  rbd::SE3<S> M;
  for (int i = 0; i < 9; ++i) M.R.m[i] = S(p[i]);
  M.p = rbd::V3<S>{S(p[9]), S(p[10]), S(p[11])};
  return M;
}

template <typename S>
struct ArmParams {
  rbd::Chain<S> chain;
  S dt, w_gp, w_x, w_u;
  rbd::SE3<S> Mref;
  int goal = 0;
  bool has_armature = false;
  S armature[rbd::MAXNV], wstate[2 * rbd::MAXNV];
  void parse(const double* p, long nparams, int nv) {
    if (nv < 1 || nv > rbd::MAXNV) throw std::invalid_argument("arm: unsupported nv");
    if (nparams != arm_param_count(nv)) throw std::invalid_argument("arm params: wrong count");
    dt = S(p[0]);
    chain.nv = nv;
    chain.gravity = rbd::V3<S>{S(p[1]), S(p[2]), S(p[3])};
    w_gp = S(p[4]);
    w_x = S(p[5]);
    w_u = S(p[6]);
    chain.frame_parent = (int)p[7];
    if (chain.frame_parent < 0 || chain.frame_parent >= nv) throw std::invalid_argument("arm: frame parent out of range");
    chain.frame_placement = read_se3<S>(p + 8);
    Mref = read_se3<S>(p + 20);
    goal = (int)p[32];
    if (goal < 0 || goal > 2) throw std::invalid_argument("arm: unknown goal cost");
    has_armature = false;
    for (int i = 0; i < nv; ++i) {
      const double* j = p + ARM_HEADER + ARM_PER_JOINT * i;
      chain.axis[i] = rbd::V3<S>{S(j[0]), S(j[1]), S(j[2])};
      chain.placement[i] = read_se3<S>(j + 3);
      chain.inertia[i].mass = S(j[15]);
      chain.inertia[i].c = rbd::V3<S>{S(j[16]), S(j[17]), S(j[18])};
      rbd::M3<S>& I = chain.inertia[i].I;
      I(0, 0) = S(j[19]);
      I(0, 1) = I(1, 0) = S(j[20]);
      I(0, 2) = I(2, 0) = S(j[21]);
      I(1, 1) = S(j[22]);
      I(1, 2) = I(2, 1) = S(j[23]);
      I(2, 2) = S(j[24]);
      armature[i] = S(j[25]);
      if (j[25] != 0.) has_armature = true;
      wstate[i] = S(j[26]);
      wstate[nv + i] = S(j[27]);
    }
  }
};

// differential model pieces shared by calc / calcDiff
template <typename S>
struct ArmEval {
  // calc: acceleration and differential cost (free-fwddyn.hxx:44-80)
  static S calc(const ArmParams<S>& P, rbd::ChainData<S>& d, const S* x, const S* u, S* a, rbd::SE3<S>& rMf, S* r6) {
    const int nv = P.chain.nv;
    if (!P.has_armature) {
      rbd::aba(P.chain, d, x, x + nv, u, a);  // tau = u (full.hpp:44) ; fills oMi (updateGlobalPlacements)
    } else {
      // free-fwddyn.hxx:66-74: M.diagonal() += armature; Minv; xout = Minv (tau - nle)
      std::vector<S> M(nv * nv), Minv(nv * nv), nle(nv), zero(nv, S(0));
      rbd::crba(P.chain, d, x, M.data());
      for (int i = 0; i < nv; ++i) M[i + nv * i] += P.armature[i];
      if (!rbd::spdInverse(M.data(), nv, Minv.data())) throw std::runtime_error("M + armature is not positive definite");
      rbd::rnea(P.chain, d, x, x + nv, zero.data(), nle.data());
      for (int i = 0; i < nv; ++i) {
        S s = S(0);
        for (int k = 0; k < nv; ++k) s += Minv[i + nv * k] * (u[k] - nle[k]);
        a[i] = s;
      }
      rbd::aba(P.chain, d, x, x + nv, u, zero.data());  // only for the placements oMi it leaves in d
    }
    const rbd::SE3<S> oMf = rbd::framePlacement(P.chain, d);
    for (int i = 0; i < 6; ++i) r6[i] = S(0);
    if (P.goal == 1) {  // frame-translation.hxx:56: r = oMf.translation - xref
      rMf = oMf;
      r6[0] = oMf.p.x - P.Mref.p.x;
      r6[1] = oMf.p.y - P.Mref.p.y;
      r6[2] = oMf.p.z - P.Mref.p.z;
    } else if (P.goal == 2) {  // frame-rotation.hxx:57-58: r = log3(oRf_inv_ * oRf)
      rMf = rbd::actInv(P.Mref, oMf);
      S theta;
      const rbd::V3<S> w = rbd::log3(rMf.R, theta);
      r6[3] = w.x;
      r6[4] = w.y;
      r6[5] = w.z;
    } else {
      rMf = rbd::actInv(P.Mref, oMf);  // oMf_inv_ * oMf   (frame-placement.hxx:58)
      rbd::toArr(rbd::log6(rMf), r6);
    }
    S c_gp = S(0), c_u = S(0), c_x = S(0);
    for (int i = 0; i < 6; ++i) c_gp += r6[i] * r6[i];
    for (int i = 0; i < nv; ++i) c_u += u[i] * u[i];
    for (int i = 0; i < 2 * nv; ++i) c_x += x[i] * (P.wstate[i] * x[i]);  // weighted-quadratic.hpp:47-48
    S cost = S(0);  // cost-sum.hxx:103-118, name order
    cost += P.w_gp * (S(0.5) * c_gp);
    cost += P.w_u * (S(0.5) * c_u);
    cost += P.w_x * (S(0.5) * c_x);
    return cost;
  }
};

struct ActionModelArm : ActionModel {
  ArmParams<double> P;
  int nv;
  bool terminal;  // dt = 0  => enable_integration_ = false (euler.hxx:32-34)
  ActionModelArm(const double* params, long nparams, int nv_, bool terminal_)
      : ActionModel(2 * nv_, nv_, 2 * nv_ + nv_ + 6), nv(nv_), terminal(terminal_) {
    P.parse(params, nparams, nv_);
    if (terminal) P.dt = 0.;
    if (P.dt < 0.) P.dt = 1e-3;  // euler.hxx:27-31
  }
  void calc(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    rbd::ChainData<double> cd;
    double a[rbd::MAXNV], r6[6];
    rbd::SE3<double> rMf;
    const double cost = ArmEval<double>::calc(P, cd, x.data(), u.data(), a, rMf, r6);
    const double dt = P.dt;
    if (dt != 0.) {  // euler.hxx:65-69
      for (int i = 0; i < nv; ++i) {
        const double dq = x[nv + i] * dt + a[i] * (dt * dt);
        const double dv = a[i] * dt;
        d.xnext[i] = x[i] + dq;
        d.xnext[nv + i] = x[nv + i] + dv;
      }
      d.cost = dt * cost;
    } else {  // euler.hxx:70-74
      d.xnext = x;
      d.cost = cost;
    }
  }
  void calcDiff(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    const int n = 2 * nv;
    rbd::ChainData<double> cd;
    double a[rbd::MAXNV], r6[6];
    rbd::SE3<double> rMf;
    ArmEval<double>::calc(P, cd, x.data(), u.data(), a, rMf, r6);  // the data calc left behind (same x, u)
    // dynamics derivatives (free-fwddyn.hxx:103-107): da_dx = [ddq_dq ddq_dv], da_du = Minv
    std::vector<double> dq(nv * nv), dv(nv * nv), Minv(nv * nv);
    if (!P.has_armature) {
      rbd::abaDerivatives(P.chain, cd, x.data(), x.data() + nv, a, dq.data(), dv.data(), Minv.data());
    } else {  // free-fwddyn.hxx:108-114: Fx = Minv (-dtau_dq | -dtau_dv) at xout, Fu = Minv, Minv = (M + armature)^-1
      std::vector<double> M(nv * nv), tq(nv * nv), tv(nv * nv);
      rbd::crba(P.chain, cd, x.data(), M.data());
      for (int i = 0; i < nv; ++i) M[i + nv * i] += P.armature[i];
      if (!rbd::spdInverse(M.data(), nv, Minv.data())) throw std::runtime_error("M + armature is not positive definite");
      rbd::rneaDerivatives(P.chain, cd, x.data(), x.data() + nv, a, tq.data(), tv.data());
      for (int j = 0; j < nv; ++j)
        for (int i = 0; i < nv; ++i) {
          double s1 = 0, s2 = 0;
          for (int k = 0; k < nv; ++k) {
            s1 += Minv[i + nv * k] * tq[k + nv * j];
            s2 += Minv[i + nv * k] * tv[k + nv * j];
          }
          dq[i + nv * j] = -s1;
          dv[i + nv * j] = -s2;
        }
    }
    // cost derivatives (frame-placement.hxx:68-86): J = Jlog6(rMf) * fJf
    double rJf[36];
    std::vector<double> fJf(6 * nv), J(6 * nv);
    {
      std::vector<double> scratch(nv);
      rbd::aba(P.chain, cd, x.data(), x.data() + nv, u.data(), scratch.data());  // restore liMi/oMi of the forward dynamics
    }
    rbd::frameJacobianLocal(P.chain, cd, fJf.data());
    for (int e = 0; e < 36; ++e) rJf[e] = 0.;
    if (P.goal == 1) {  // frame-translation.hxx:71-72: J = oRf * fJf.topRows(3)   (rMf holds oMf here)
      for (int i = 0; i < 3; ++i)
        for (int k = 0; k < 3; ++k) rJf[i + 6 * k] = rMf.R(i, k);
    } else if (P.goal == 2) {  // frame-rotation.hxx:76-78: J = Jlog3(rRf) * fJf.bottomRows(3)
      double theta;
      const rbd::V3<double> w = rbd::log3(rMf.R, theta);
      const rbd::M3<double> J3 = rbd::Jlog3(theta, w);
      for (int i = 0; i < 3; ++i)
        for (int k = 0; k < 3; ++k) rJf[(3 + i) + 6 * (3 + k)] = J3(i, k);
    } else {
      rbd::Jlog6(rMf, rJf);
    }
    for (int j = 0; j < nv; ++j)
      for (int i = 0; i < 6; ++i) {
        double s = 0;
        for (int k = 0; k < 6; ++k) s += rJf[i + 6 * k] * fJf[k + 6 * j];
        J[i + 6 * j] = s;
      }
    Vec Lx(n, 0.), Lu(nv, 0.);
    Mat Lxx(n, n), Luu(nv, nv);
    // cost-sum.hxx:140-158 in name order: gripperPose, uReg, xReg
    for (int j = 0; j < nv; ++j) {
      double s = 0;
      for (int k = 0; k < 6; ++k) s += J[k + 6 * j] * r6[k];
      Lx[j] += P.w_gp * s;
    }
    for (int j = 0; j < nv; ++j)
      for (int i = 0; i < nv; ++i) {
        double s = 0;
        for (int k = 0; k < 6; ++k) s += J[k + 6 * i] * J[k + 6 * j];
        Lxx(i, j) += P.w_gp * s;
      }
    for (int i = 0; i < nv; ++i) {
      Lu[i] += P.w_u * u[i];
      Luu(i, i) += P.w_u * 1.;
    }
    for (int i = 0; i < n; ++i) {  // state.hxx:150-160 with Ar = w o r, Arr = diag(w) (weighted-quadratic.hpp:59-63)
      Lx[i] += P.w_x * (P.wstate[i] * x[i]);
      Lxx(i, i) += P.w_x * P.wstate[i];
    }
    const double dt = P.dt;
    d.Fx.setZero();
    d.Fu.setZero();
    d.Lxu.setZero();
    if (dt != 0.) {  // euler.hxx:100-121
      const double dt2 = dt * dt;
      for (int j = 0; j < nv; ++j)
        for (int i = 0; i < nv; ++i) {
          d.Fx(i, j) = dq[i + nv * j] * dt2;
          d.Fx(i, nv + j) = dv[i + nv * j] * dt2;
          d.Fx(nv + i, j) = dq[i + nv * j] * dt;
          d.Fx(nv + i, nv + j) = dv[i + nv * j] * dt;
          d.Fu(i, j) = Minv[i + nv * j] * dt2;
          d.Fu(nv + i, j) = Minv[i + nv * j] * dt;
        }
      for (int i = 0; i < nv; ++i) d.Fx(i, nv + i) += dt;
      for (int i = 0; i < n; ++i) d.Fx(i, i) += 1.;  // Jintegrate(first, addto) on a vector space
      for (int i = 0; i < n; ++i) d.Lx[i] = dt * Lx[i];
      for (int i = 0; i < nv; ++i) d.Lu[i] = dt * Lu[i];
      for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i) d.Lxx(i, j) = dt * Lxx(i, j);
      for (int j = 0; j < nv; ++j)
        for (int i = 0; i < nv; ++i) d.Luu(i, j) = dt * Luu(i, j);
    } else {  // euler.hxx:122-130
      for (int i = 0; i < n; ++i) d.Fx(i, i) = 1.;
      d.Lx = Lx;
      d.Lu = Lu;
      d.Lxx = Lxx;
      d.Luu = Luu;
    }
  }
  // quasiStatic (free-fwddyn.hxx:146-169): u = pinv(dtau_du = I) * rnea(q, 0, 0)
  void quasiStatic(const Vec& x, Vec& u) const {
    rbd::ChainData<double> cd;
    std::vector<double> z(nv, 0.);
    u.assign(nv, 0.);
    rbd::rnea(P.chain, cd, x.data(), z.data(), z.data(), u.data());
  }
};

inline void make_arm_models(const double* params, long nparams, int nx, int nu, std::shared_ptr<ActionModel>& run,
                            std::shared_ptr<ActionModel>& term) {
  if (nx != 2 * nu) throw std::invalid_argument("arm: nx must be 2 nv");
  run = std::make_shared<ActionModelArm>(params, nparams, nu, false);
  term = std::make_shared<ActionModelArm>(params, nparams, nu, true);
}

// Rigid-body primitives by name, for the self-consistency tests. in/out layouts (nv from the params):
//   "aba"   in q,v,tau -> out qdd(nv)                "rnea" in q,v,a -> out tau(nv)        "crba" in q -> M(nv*nv)
//   "aba_derivatives" in q,v,tau -> ddq_dq, ddq_dv, Minv (3 nv*nv)      "rnea_derivatives" in q,v,a -> dtau_dq, dtau_dv
//   "frame" in q -> oMf R(9 row-major), p(3), fJf (6*nv col-major)       "log6" in R(9),p(3) -> 6
//   "jlog6" in R(9),p(3) -> 36 col-major                                  "quasistatic" in q -> u(nv)
inline int arm_op(const double* params, long nparams, const char* op_, const double* in, double* out) {
  const int nv = (int)((nparams - ARM_HEADER) / ARM_PER_JOINT);
  ArmParams<double> P;
  P.parse(params, nparams, nv);
  rbd::ChainData<double> d;
  const std::string op(op_);
  if (op == "aba") {
    rbd::aba(P.chain, d, in, in + nv, in + 2 * nv, out);
    return nv;
  }
  if (op == "rnea" || op == "quasistatic") {
    std::vector<double> z(nv, 0.);
    if (op == "rnea")
      rbd::rnea(P.chain, d, in, in + nv, in + 2 * nv, out);
    else
      rbd::rnea(P.chain, d, in, z.data(), z.data(), out);
    return nv;
  }
  if (op == "crba") {
    rbd::crba(P.chain, d, in, out);
    return nv * nv;
  }
  if (op == "aba_derivatives") {
    std::vector<double> a(nv);
    rbd::aba(P.chain, d, in, in + nv, in + 2 * nv, a.data());
    rbd::abaDerivatives(P.chain, d, in, in + nv, a.data(), out, out + nv * nv, out + 2 * nv * nv);
    return 3 * nv * nv;
  }
  if (op == "rnea_derivatives") {
    rbd::rneaDerivatives(P.chain, d, in, in + nv, in + 2 * nv, out, out + nv * nv);
    return 2 * nv * nv;
  }
  if (op == "frame") {
    std::vector<double> z(nv, 0.), a(nv);
    rbd::aba(P.chain, d, in, z.data(), z.data(), a.data());
    const rbd::SE3<double> M = rbd::framePlacement(P.chain, d);
    for (int i = 0; i < 9; ++i) out[i] = M.R.m[i];
    out[9] = M.p.x;
    out[10] = M.p.y;
    out[11] = M.p.z;
    rbd::frameJacobianLocal(P.chain, d, out + 12);
    return 12 + 6 * nv;
  }
  if (op == "log6") {
    rbd::toArr(rbd::log6(read_se3<double>(in)), out);
    return 6;
  }
  if (op == "jlog6") {
    rbd::Jlog6(read_se3<double>(in), out);
    return 36;
  }
  return -1;
}

}  // namespace oracle
