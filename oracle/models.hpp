// TEST INFRASTRUCTURE — CPU oracle (see linalg.hpp header). Not part of the product.
//
// Node plugin API + the vector-space action models, restated from the reference:
//   ActionModelAbstract / ActionDataAbstract  include/crocoddyl/core/action-base.hpp:34-142
//   StateVector (diff = x1-x0, integrate = x+dx) include/crocoddyl/core/states/euclidean.hxx:28-61
//   ActionModelLQR      include/crocoddyl/core/actions/lqr.hxx:14-70, lqr.hpp:88-95
//   ActionModelUnicycle include/crocoddyl/core/actions/unicycle.hxx:13-73, unicycle.hpp:78-80
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include "linalg.hpp"

namespace oracle {

struct ActionModel;

// action-base.hpp:101-142
struct ActionData {
  double cost;
  Vec xnext, r, Lx, Lu;
  Mat Fx, Fu, Lxx, Lxu, Luu;
  ActionData(int nx, int ndx, int nu, int nr)
      : cost(0.), xnext(nx, 0.), r(nr, 0.), Lx(ndx, 0.), Lu(nu, 0.), Fx(ndx, ndx), Fu(ndx, nu), Lxx(ndx, ndx),
        Lxu(ndx, nu), Luu(nu, nu) {}
  virtual ~ActionData() {}
};

// action-base.hpp:34-98. All states on this path are vector spaces (StateVector, or StateMultibody
// restricted to fixed-base revolute chains, multibody/states/multibody.hxx:69-70,89-90), so the
// manifold operations are restated once here.
struct ActionModel {
  int nx, ndx, nu, nr;
  Vec unone;
  // control limits, action-base.hxx:19-22,117-140: +-inf by default; has_control_limits = some finite entry in BOTH
  Vec u_lb, u_ub;
  bool has_control_limits;
  ActionModel(int nx_, int nu_, int nr_)
      : nx(nx_), ndx(nx_), nu(nu_), nr(nr_), unone(nu_, 0.), u_lb(nu_, -INFINITY), u_ub(nu_, INFINITY),
        has_control_limits(false) {}
  virtual ~ActionModel() {}
  void set_u_lb(const Vec& v) {
    if ((int)v.size() != nu) throw std::invalid_argument("Invalid argument: lower bound has wrong dimension");
    u_lb = v;
    update_has_control_limits();
  }
  void set_u_ub(const Vec& v) {
    if ((int)v.size() != nu) throw std::invalid_argument("Invalid argument: upper bound has wrong dimension");
    u_ub = v;
    update_has_control_limits();
  }
  void update_has_control_limits() {  // action-base.hxx:138-140
    bool a = false, b = false;
    for (int i = 0; i < nu; ++i) {
      a = a || std::isfinite(u_lb[i]);
      b = b || std::isfinite(u_ub[i]);
    }
    has_control_limits = a && b;
  }
  virtual void calc(ActionData& d, const Vec& x, const Vec& u) = 0;
  virtual void calcDiff(ActionData& d, const Vec& x, const Vec& u) = 0;
  virtual std::shared_ptr<ActionData> createData() { return std::make_shared<ActionData>(nx, ndx, nu, nr); }
  // action-base.hxx:27-37: x-only overloads forward unone_ (zeros)
  void calc(ActionData& d, const Vec& x) { calc(d, x, unone); }
  void calcDiff(ActionData& d, const Vec& x) { calcDiff(d, x, unone); }
  // euclidean.hxx:28-35 zero(); :44-52 diff; :55-61 integrate
  Vec zero() const { return Vec(nx, 0.); }
  void diff(const Vec& x0, const Vec& x1, Vec& dx) const {
    dx.resize(ndx);
    for (int i = 0; i < ndx; ++i) dx[i] = x1[i] - x0[i];
  }
  void integrate(const Vec& x, const Vec& dx, Vec& xout) const {
    xout.resize(nx);
    for (int i = 0; i < nx; ++i) xout[i] = x[i] + dx[i];
  }
  void check(const Vec& x, const Vec& u) const {
    if ((int)x.size() != nx)
      throw std::invalid_argument("Invalid argument: x has wrong dimension (it should be " + std::to_string(nx) + ")");
    if ((int)u.size() != nu)
      throw std::invalid_argument("Invalid argument: u has wrong dimension (it should be " + std::to_string(nu) + ")");
  }
};

// Sensitivity experiment only (liboracle_perturb.so, -DORACLE_PERTURB_SINCOS): move the result of sin / cos by one
// ulp for about a quarter of the arguments (chosen by a hash of the argument's bits). Models two correct libm's that
// round differently (glibc on the host vs CUDA's sin/cos on the device, both < 1-2 ulp) and measures how many solves
// change their iteration count because of that alone (tools/sincos_sensitivity.py, profiles/r02_sincos_sensitivity.json).
inline double libm_result(double v, double arg) {
#ifdef ORACLE_PERTURB_SINCOS
  uint64_t h;
  std::memcpy(&h, &arg, 8);
  h ^= h >> 33;
  h *= 0xff51afd7ed558ccdULL;
  h ^= h >> 33;
  if ((h & 7) == 0) return std::nextafter(v, INFINITY);
  if ((h & 7) == 1) return std::nextafter(v, -INFINITY);
#else
  (void)arg;
#endif
  return v;
}

// ---------------------------------------------------------------------------------------------
// ActionModelLQR — lqr.hxx:14-70
struct ActionModelLQR : ActionModel {
  bool drift_free;
  Mat Fx, Fu, Lxx, Lxu, Luu;
  Vec f0, lx, lu;
  ActionModelLQR(int nx_, int nu_, bool drift_free_ = true) : ActionModel(nx_, nu_, 0), drift_free(drift_free_) {
    Fx = identity(nx_, nx_);   // lqr.hxx:17
    Fu = identity(nx_, nu_);   // :18
    f0.assign(nx_, 1.);        // :19
    Lxx = identity(nx_, nx_);  // :20
    Lxu = identity(nx_, nu_);  // :21
    Luu = identity(nu_, nu_);  // :22
    lx.assign(nx_, 1.);        // :23
    lu.assign(nu_, 1.);        // :24
  }
  void calc(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    // xnext = Fx x + Fu u (+ f0)   lqr.hxx:42-46
    Vec a, b;
    gemv(Fx, x, a);
    gemv(Fu, u, b);
    for (int i = 0; i < nx; ++i) d.xnext[i] = drift_free ? (a[i] + b[i]) : (a[i] + b[i] + f0[i]);
    // cost = .5 x'Lxx x + .5 u'Luu u + x'Lxu u + lx'x + lu'u   lqr.hxx:47-48
    Vec Lxx_x, Luu_u, Lxu_u;
    gemv(Lxx, x, Lxx_x);
    gemv(Luu, u, Luu_u);
    gemv(Lxu, u, Lxu_u);
    d.cost = 0.5 * dot(x, Lxx_x) + 0.5 * dot(u, Luu_u) + dot(x, Lxu_u) + dot(lx, x) + dot(lu, u);
  }
  void calcDiff(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    // Lx = lx + Lxx x + Lxu u ; Lu = lu + Lxu' x + Luu u   lqr.hxx:63-64
    Vec a, b;
    gemv(Lxx, x, a);
    gemv(Lxu, u, b);
    for (int i = 0; i < nx; ++i) d.Lx[i] = lx[i] + a[i] + b[i];
    Vec c(nu, 0.), e;
    gemtv_add(Lxu, x, c);
    gemv(Luu, u, e);
    for (int i = 0; i < nu; ++i) d.Lu[i] = lu[i] + c[i] + e[i];
    d.Fx = Fx;  // lqr.hxx:65-69 (copies)
    d.Fu = Fu;
    d.Lxx = Lxx;
    d.Lxu = Lxu;
    d.Luu = Luu;
  }
};

// ---------------------------------------------------------------------------------------------
// IntegratedActionModelEuler(DifferentialActionModelLQR(nq, nu, drift_free), dt) — diff-lqr.hxx:14-88 evaluated first,
// then the semi-implicit Euler step of euler.hxx:40-131 on a vector-space state (Jintegrate = identity). dt = 0 is the
// reference's enable_integration_ = false branch (euler.hxx:32-34,70-74,122-130).
struct ActionModelEulerDiffLQR : ActionModel {
  int nq;
  bool drift_free;
  double dt;
  Mat Fq, Fv, Fu, Lxx, Lxu, Luu;
  Vec f0, lx, lu;
  ActionModelEulerDiffLQR(int nq_, int nu_, bool drift_free_, double dt_)
      : ActionModel(2 * nq_, nu_, 0), nq(nq_), drift_free(drift_free_), dt(dt_) {
    Fq = identity(nq, nq);     // diff-lqr.hxx:19
    Fv = identity(nq, nq);     // :20
    Fu = identity(nq, nu_);    // :21
    f0.assign(nq, 1.);         // :22
    Lxx = identity(nx, nx);    // :23
    Lxu = identity(nx, nu_);   // :24
    Luu = identity(nu_, nu_);  // :25
    lx.assign(nx, 1.);         // :26
    lu.assign(nu_, 1.);        // :27
  }
  // differential model: xout and cost (diff-lqr.hxx:48-57)
  double diff_calc(const Vec& x, const Vec& u, Vec& xout) const {
    Vec q(x.begin(), x.begin() + nq), v(x.begin() + nq, x.end());
    Vec a, b, c;
    gemv(Fq, q, a);
    gemv(Fv, v, b);
    gemv(Fu, u, c);
    xout.assign(nq, 0.);
    for (int i = 0; i < nq; ++i) xout[i] = drift_free ? (a[i] + b[i] + c[i]) : (a[i] + b[i] + c[i] + f0[i]);
    Vec Lxx_x, Luu_u, Lxu_u;
    gemv(Lxx, x, Lxx_x);
    gemv(Luu, u, Luu_u);
    gemv(Lxu, u, Lxu_u);
    return 0.5 * dot(x, Lxx_x) + 0.5 * dot(u, Luu_u) + dot(x, Lxu_u) + dot(lx, x) + dot(lu, u);
  }
  void calc(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    Vec acc;
    const double cost = diff_calc(x, u, acc);
    if (dt != 0.) {  // euler.hxx:65-69
      const double dt2 = dt * dt;
      for (int i = 0; i < nq; ++i) {
        const double dq = x[nq + i] * dt + acc[i] * dt2;
        const double dv = acc[i] * dt;
        d.xnext[i] = x[i] + dq;
        d.xnext[nq + i] = x[nq + i] + dv;
      }
      d.cost = dt * cost;
    } else {  // euler.hxx:70-74
      d.xnext = x;
      d.cost = cost;
    }
  }
  void calcDiff(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    // differential derivatives (diff-lqr.hxx:76-84)
    Vec a, b, Lx(nx), Lu(nu);
    gemv(Lxx, x, a);
    gemv(Lxu, u, b);
    for (int i = 0; i < nx; ++i) Lx[i] = lx[i] + a[i] + b[i];
    Vec c(nu, 0.), e;
    gemtv_add(Lxu, x, c);
    gemv(Luu, u, e);
    for (int i = 0; i < nu; ++i) Lu[i] = lu[i] + c[i] + e[i];
    d.Fx.setZero();
    d.Fu.setZero();
    if (dt != 0.) {  // euler.hxx:100-121
      const double dt2 = dt * dt;
      for (int j = 0; j < nq; ++j)
        for (int i = 0; i < nq; ++i) {
          d.Fx(i, j) = Fq(i, j) * dt2;
          d.Fx(i, nq + j) = Fv(i, j) * dt2;
          d.Fx(nq + i, j) = Fq(i, j) * dt;
          d.Fx(nq + i, nq + j) = Fv(i, j) * dt;
        }
      for (int j = 0; j < nu; ++j)
        for (int i = 0; i < nq; ++i) {
          d.Fu(i, j) = Fu(i, j) * dt2;
          d.Fu(nq + i, j) = Fu(i, j) * dt;
        }
      for (int i = 0; i < nq; ++i) d.Fx(i, nq + i) += dt;
      for (int i = 0; i < nx; ++i) d.Fx(i, i) += 1.;  // Jintegrate(first, addto) on a vector space
      for (int i = 0; i < nx; ++i) d.Lx[i] = dt * Lx[i];
      for (int i = 0; i < nu; ++i) d.Lu[i] = dt * Lu[i];
      for (int j = 0; j < nx; ++j)
        for (int i = 0; i < nx; ++i) d.Lxx(i, j) = dt * Lxx(i, j);
      for (int j = 0; j < nu; ++j)
        for (int i = 0; i < nx; ++i) d.Lxu(i, j) = dt * Lxu(i, j);
      for (int j = 0; j < nu; ++j)
        for (int i = 0; i < nu; ++i) d.Luu(i, j) = dt * Luu(i, j);
    } else {  // euler.hxx:122-130
      for (int i = 0; i < nx; ++i) d.Fx(i, i) = 1.;
      d.Lx = Lx;
      d.Lu = Lu;
      d.Lxx = Lxx;
      d.Lxu = Lxu;
      d.Luu = Luu;
    }
  }
};

// ---------------------------------------------------------------------------------------------
// ActionModelUnicycle — unicycle.hxx:13-73; data ctor sets Fx = I once (unicycle.hpp:78-80)
struct ActionModelUnicycle : ActionModel {
  double dt;
  double cost_weights[2];
  ActionModelUnicycle() : ActionModel(3, 2, 5), dt(0.1) {
    cost_weights[0] = 10.;
    cost_weights[1] = 1.;
  }
  std::shared_ptr<ActionData> createData() override {
    std::shared_ptr<ActionData> d = std::make_shared<ActionData>(nx, ndx, nu, nr);
    d->Fx.setIdentity();
    return d;
  }
  void calc(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    const double c = libm_result(std::cos(x[2]), x[2]);
    const double s = libm_result(std::sin(x[2]), x[2]);
    d.xnext[0] = x[0] + c * u[0] * dt;  // unicycle.hxx:38
    d.xnext[1] = x[1] + s * u[0] * dt;
    d.xnext[2] = x[2] + u[1] * dt;
    for (int i = 0; i < 3; ++i) d.r[i] = cost_weights[0] * x[i];      // :39
    for (int i = 0; i < 2; ++i) d.r[3 + i] = cost_weights[1] * u[i];  // :40
    double rr = 0;
    for (int i = 0; i < 5; ++i) rr += d.r[i] * d.r[i];
    d.cost = 0.5 * rr;  // :41
  }
  void calcDiff(ActionData& d, const Vec& x, const Vec& u) override {
    check(x, u);
    const double w_x = cost_weights[0] * cost_weights[0];
    const double w_u = cost_weights[1] * cost_weights[1];
    for (int i = 0; i < 3; ++i) d.Lx[i] = x[i] * w_x;  // :61
    for (int i = 0; i < 2; ++i) d.Lu[i] = u[i] * w_u;  // :62
    for (int i = 0; i < 3; ++i) d.Lxx(i, i) = w_x;     // :63
    for (int i = 0; i < 2; ++i) d.Luu(i, i) = w_u;     // :64
    const double c = libm_result(std::cos(x[2]), x[2]);
    const double s = libm_result(std::sin(x[2]), x[2]);
    d.Fx(0, 2) = -s * u[0] * dt;  // :69-73
    d.Fx(1, 2) = c * u[0] * dt;
    d.Fu(0, 0) = c * dt;
    d.Fu(1, 0) = s * dt;
    d.Fu(2, 1) = dt;
  }
};

}  // namespace oracle
