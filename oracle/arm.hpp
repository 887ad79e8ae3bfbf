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
//   per joint: axis (3), placement in parent R (9) p (3), mass, COM lever (3), inertia about COM Ixx Ixy Ixz Iyy Iyz Izz
#pragma once
#include "models.hpp"
#include "rbd.hpp"

namespace oracle {

constexpr int ARM_HEADER = 32, ARM_PER_JOINT = 25;
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
    }
  }
};

// differential model pieces shared by calc / calcDiff
template <typename S>
struct ArmEval {
  // calc: acceleration and differential cost (free-fwddyn.hxx:44-80)
  static S calc(const ArmParams<S>& P, rbd::ChainData<S>& d, const S* x, const S* u, S* a, rbd::SE3<S>& rMf, S* r6) {
    const int nv = P.chain.nv;
    rbd::aba(P.chain, d, x, x + nv, u, a);  // tau = u (full.hpp:44) ; fills oMi (updateGlobalPlacements)
    const rbd::SE3<S> oMf = rbd::framePlacement(P.chain, d);
    rMf = rbd::actInv(P.Mref, oMf);  // oMf_inv_ * oMf   (frame-placement.hxx:58)
    rbd::toArr(rbd::log6(rMf), r6);
    S c_gp = S(0), c_u = S(0), c_x = S(0);
    for (int i = 0; i < 6; ++i) c_gp += r6[i] * r6[i];
    for (int i = 0; i < nv; ++i) c_u += u[i] * u[i];
    for (int i = 0; i < 2 * nv; ++i) c_x += x[i] * x[i];
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
    rbd::abaDerivatives(P.chain, cd, x.data(), x.data() + nv, a, dq.data(), dv.data(), Minv.data());
    // cost derivatives (frame-placement.hxx:68-86): J = Jlog6(rMf) * fJf
    double rJf[36];
    std::vector<double> fJf(6 * nv), J(6 * nv);
    rbd::Jlog6(rMf, rJf);
    rbd::aba(P.chain, cd, x.data(), x.data() + nv, u.data(), a);  // restore liMi/oMi of the forward dynamics
    rbd::frameJacobianLocal(P.chain, cd, fJf.data());
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
    for (int i = 0; i < n; ++i) {
      Lx[i] += P.w_x * x[i];
      Lxx(i, i) += P.w_x * 1.;
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
