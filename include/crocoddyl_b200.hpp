// crocoddyl_b200 — C++ facade over the C ABI (crocoddyl_b200.h) with the reference's class and method names:
//   the node plugin interface  ActionModelAbstract / ActionDataAbstract / StateVector
//                              (include/crocoddyl/core/action-base.hpp:34-142, core/states/euclidean.hxx:28-61),
//   ShootingProblem(x0, running_models, terminal_model)  (include/crocoddyl/core/optctrl/shooting.hpp:48-207),
//   SolverDDP / SolverFDDP / SolverBoxDDP / SolverBoxFDDP  (core/solvers/ddp.hpp:48-302, fddp.hpp:50-100, box-ddp.hpp,
//                              box-fddp.hpp), crocoddyl::Exception (core/utils/exception.hpp).
// Header-only; link with -lcrocoddyl_b200. Eigen and Boost are not required: vectors are std::vector<double> (matrices
// column-major, as Eigen::MatrixXd), shared ownership is std::shared_ptr. A problem is a batch of B OCPs (x0 holds B
// initial states); with B = 1 the calls read like the reference's (benchmark/unicycle-optctrl.cpp:25-47).
//
// What a node model is here: the reference lets a user derive ActionModelAbstract and write calc/calcDiff in C++. On
// this path calc/calcDiff run inside CUDA kernels, so the models that can be put into a ShootingProblem are the
// families the device implements (ActionModelUnicycle, ActionModelLQR, ActionModelArm = IntegratedActionModelEuler o
// DifferentialActionModelFreeFwdDynamics on a revolute chain); each exposes the reference's virtual interface
// (calc(data, x, u), calcDiff, createData, quasiStatic, get_nu, get_state, control limits) and evaluates it on the GPU.
#ifndef CROCODDYL_B200_HPP_
#define CROCODDYL_B200_HPP_

#include <cmath>
#include <cstddef>
#include <exception>
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "crocoddyl_b200.h"

namespace crocoddyl_b200 {

typedef std::vector<double> VectorXd;  // one vector / one column-major matrix
typedef std::vector<double> MatrixXd;

class Exception : public std::exception {
 public:
  Exception(int code, const std::string& msg) : code_(code), msg_(msg) {}
  const char* what() const noexcept override { return msg_.c_str(); }
  int code() const { return code_; }

 private:
  int code_;
  std::string msg_;
};
inline void check(int status) {
  if (status < 0) throw Exception(status, cgx_last_error());
}

// StateVector (core/states/euclidean.hxx:28-61): the state of every model on this path
class StateVector {
 public:
  explicit StateVector(std::size_t nx, std::size_t nq = 0) : nx_(nx), nq_(nq ? nq : nx / 2) {}
  std::size_t get_nx() const { return nx_; }
  std::size_t get_ndx() const { return nx_; }
  std::size_t get_nq() const { return nq_; }
  std::size_t get_nv() const { return nx_ - nq_; }
  VectorXd zero() const { return VectorXd(nx_, 0.); }
  VectorXd diff(const VectorXd& x0, const VectorXd& x1) const {
    VectorXd d(nx_);
    for (std::size_t i = 0; i < nx_; ++i) d[i] = x1[i] - x0[i];
    return d;
  }
  VectorXd integrate(const VectorXd& x, const VectorXd& dx) const {
    VectorXd o(nx_);
    for (std::size_t i = 0; i < nx_; ++i) o[i] = x[i] + dx[i];
    return o;
  }

 private:
  std::size_t nx_, nq_;
};
typedef StateVector StateAbstract;

class ActionModelAbstract;

// ActionDataAbstract (core/action-base.hpp:101-142)
struct ActionDataAbstract {
  explicit ActionDataAbstract(const ActionModelAbstract* model);
  double cost;
  VectorXd xnext, r;
  MatrixXd Fx, Fu;
  VectorXd Lx, Lu;
  MatrixXd Lxx, Lxu, Luu;
};

// ActionModelAbstract (core/action-base.hpp:34-98, action-base.hxx)
class ActionModelAbstract {
 public:
  ActionModelAbstract(std::shared_ptr<StateAbstract> state, std::size_t nu, std::size_t nr = 0)
      : nu_(nu), nr_(nr), state_(state), unone_(nu, 0.), u_lb_(nu, -std::numeric_limits<double>::infinity()),
        u_ub_(nu, std::numeric_limits<double>::infinity()), has_control_limits_(false), node_(nullptr) {}
  virtual ~ActionModelAbstract() {
    if (node_) cgx_problem_destroy(node_);
  }
  ActionModelAbstract(const ActionModelAbstract&) = delete;
  ActionModelAbstract& operator=(const ActionModelAbstract&) = delete;

  // device mapping of the model family (crocoddyl_b200.h: cgx_model_kind and the parameter-block layouts)
  virtual int kind() const = 0;
  virtual std::vector<double> params() const = 0;

  // calc / calcDiff of ONE node, evaluated on the GPU through a one-node problem (action-base.hpp:50-73)
  virtual void calc(const std::shared_ptr<ActionDataAbstract>& data, const VectorXd& x, const VectorXd& u) {
    eval(data, x, &u, false);
  }
  virtual void calcDiff(const std::shared_ptr<ActionDataAbstract>& data, const VectorXd& x, const VectorXd& u) {
    eval(data, x, &u, true);
  }
  // x-only overloads forward unone_ (action-base.hxx:27-37) — how a terminal node is evaluated
  void calc(const std::shared_ptr<ActionDataAbstract>& data, const VectorXd& x) { eval(data, x, nullptr, false); }
  void calcDiff(const std::shared_ptr<ActionDataAbstract>& data, const VectorXd& x) { eval(data, x, nullptr, true); }
  virtual std::shared_ptr<ActionDataAbstract> createData() { return std::make_shared<ActionDataAbstract>(this); }
  virtual bool checkData(const std::shared_ptr<ActionDataAbstract>& data) { return (bool)data; }
  // quasiStatic (action-base.hxx:39-83; closed form for the arm, free-fwddyn.hxx:146-169)
  virtual void quasiStatic(const std::shared_ptr<ActionDataAbstract>&, VectorXd& u, const VectorXd& x,
                           std::size_t /*maxiter*/ = 100, double /*tol*/ = 1e-9) {
    if (u.size() != nu_) throw Exception(CGX_ERR_INVALID, "Invalid argument: u has wrong dimension (it should be " + std::to_string(nu_) + ")");
    node_problem(x);
    check(cgx_problem_quasi_static(node_, x.data(), u.data()));
  }
  VectorXd quasiStatic_x(const std::shared_ptr<ActionDataAbstract>& data, const VectorXd& x, std::size_t maxiter = 100,
                         double tol = 1e-9) {
    VectorXd u(nu_, 0.);
    quasiStatic(data, u, x, maxiter, tol);
    return u;
  }
  const std::size_t& get_nu() const { return nu_; }
  const std::size_t& get_nr() const { return nr_; }
  const std::shared_ptr<StateAbstract>& get_state() const { return state_; }
  const VectorXd& get_u_lb() const { return u_lb_; }
  const VectorXd& get_u_ub() const { return u_ub_; }
  bool get_has_control_limits() const { return has_control_limits_; }
  void set_u_lb(const VectorXd& u_lb) {
    if (u_lb.size() != nu_) throw Exception(CGX_ERR_INVALID, "Invalid argument: lower bound has wrong dimension (it should be " + std::to_string(nu_) + ")");
    u_lb_ = u_lb;
    update_has_control_limits();
  }
  void set_u_ub(const VectorXd& u_ub) {
    if (u_ub.size() != nu_) throw Exception(CGX_ERR_INVALID, "Invalid argument: upper bound has wrong dimension (it should be " + std::to_string(nu_) + ")");
    u_ub_ = u_ub;
    update_has_control_limits();
  }

 protected:
  void update_has_control_limits() {  // action-base.hxx:142-144
    bool a = false, b = false;
    for (std::size_t i = 0; i < nu_; ++i) {
      a = a || std::isfinite(u_lb_[i]);
      b = b || std::isfinite(u_ub_[i]);
    }
    has_control_limits_ = a && b;
  }
  void invalidate() {  // parameters changed: the cached one-node problem is stale
    if (node_) cgx_problem_destroy(node_);
    node_ = nullptr;
  }
  void node_problem(const VectorXd& x) {
    const std::size_t nx = state_->get_nx();
    if (x.size() != nx) throw Exception(CGX_ERR_INVALID, "Invalid argument: x has wrong dimension (it should be " + std::to_string(nx) + ")");
    if (!node_) {
      const std::vector<double> p = params();
      check(cgx_problem_create(&node_, kind(), 1, 1, (int)nx, (int)nu_, p.data(), (long)p.size(), 0, x.data()));
      check(cgx_problem_set_terminal_params(node_, p.data(), (long)p.size()));  // running = terminal = this model
    }
  }
  void eval(const std::shared_ptr<ActionDataAbstract>& d, const VectorXd& x, const VectorXd* u, bool diff) {
    if (!d) throw Exception(CGX_ERR_INVALID, "Invalid argument: data is null");
    if (u && u->size() != nu_) throw Exception(CGX_ERR_INVALID, "Invalid argument: u has wrong dimension (it should be " + std::to_string(nu_) + ")");
    node_problem(x);
    const std::size_t nx = state_->get_nx();
    VectorXd xs(2 * nx), us(nu_ ? nu_ : 1, 0.);
    for (std::size_t i = 0; i < nx; ++i) xs[i] = xs[nx + i] = x[i];
    if (u) us = *u;
    const int t = u ? 0 : 1;  // the x-only overload is the terminal node of the one-node problem
    double c = 0.;
    check(cgx_problem_calc(node_, xs.data(), us.data(), &c));
    check(cgx_problem_get_data(node_, "cost", 0, t, &d->cost));
    check(cgx_problem_get_data(node_, "xnext", 0, t, d->xnext.data()));
    // the cost residual, where the reference's model writes it (ActionModelUnicycle::calc, unicycle.hxx:39-40)
    if (kind() == CGX_MODEL_UNICYCLE && d->r.size() == 5) check(cgx_problem_get_data(node_, "r", 0, t, d->r.data()));
    if (diff) {
      check(cgx_problem_calc_diff(node_, xs.data(), us.data(), &c));
      check(cgx_problem_get_data(node_, "Fx", 0, t, d->Fx.data()));
      check(cgx_problem_get_data(node_, "Fu", 0, t, d->Fu.data()));
      check(cgx_problem_get_data(node_, "Lx", 0, t, d->Lx.data()));
      check(cgx_problem_get_data(node_, "Lu", 0, t, d->Lu.data()));
      check(cgx_problem_get_data(node_, "Lxx", 0, t, d->Lxx.data()));
      check(cgx_problem_get_data(node_, "Lxu", 0, t, d->Lxu.data()));
      check(cgx_problem_get_data(node_, "Luu", 0, t, d->Luu.data()));
    }
  }
  std::size_t nu_, nr_;
  std::shared_ptr<StateAbstract> state_;
  VectorXd unone_, u_lb_, u_ub_;
  bool has_control_limits_;
  cgx_problem_t node_;
};

inline ActionDataAbstract::ActionDataAbstract(const ActionModelAbstract* m)
    : cost(0.), xnext(m->get_state()->get_nx(), 0.), r(m->get_nr(), 0.),
      Fx(m->get_state()->get_ndx() * m->get_state()->get_ndx(), 0.), Fu(m->get_state()->get_ndx() * m->get_nu(), 0.),
      Lx(m->get_state()->get_ndx(), 0.), Lu(m->get_nu(), 0.), Lxx(m->get_state()->get_ndx() * m->get_state()->get_ndx(), 0.),
      Lxu(m->get_state()->get_ndx() * m->get_nu(), 0.), Luu(m->get_nu() * m->get_nu(), 0.) {}

// ActionModelUnicycle — include/crocoddyl/core/actions/unicycle.hxx:13-73 (dt = 0.1, cost weights (10, 1))
class ActionModelUnicycle : public ActionModelAbstract {
 public:
  ActionModelUnicycle() : ActionModelAbstract(std::make_shared<StateVector>(3), 2, 5), dt_(0.1), w_{10., 1.} {}
  int kind() const override { return CGX_MODEL_UNICYCLE; }
  std::vector<double> params() const override { return {dt_, w_[0], w_[1]}; }
  const double* get_cost_weights() const { return w_; }
  void set_cost_weights(double wx, double wu) {
    w_[0] = wx;
    w_[1] = wu;
    invalidate();
  }
  double get_dt() const { return dt_; }
  void set_dt(double dt) {
    dt_ = dt;
    invalidate();
  }

 private:
  double dt_, w_[2];
};

// ActionModelLQR(nx, nu, drift_free) — include/crocoddyl/core/actions/lqr.hxx:14-25 (identity / ones defaults) with the
// setters of lqr.hpp:51-75 (matrices column-major)
class ActionModelLQR : public ActionModelAbstract {
 public:
  ActionModelLQR(std::size_t nx, std::size_t nu, bool drift_free = true)
      : ActionModelAbstract(std::make_shared<StateVector>(nx), nu, 0), drift_free_(drift_free), Fx_(eye(nx, nx)), Fu_(eye(nx, nu)),
        f0_(nx, 1.), Lxx_(eye(nx, nx)), Lxu_(eye(nx, nu)), Luu_(eye(nu, nu)), lx_(nx, 1.), lu_(nu, 1.) {}
  int kind() const override { return CGX_MODEL_LQR; }
  std::vector<double> params() const override {
    std::vector<double> p(1, drift_free_ ? 1. : 0.);
    for (const VectorXd* v : {&Fx_, &Fu_, &f0_, &Lxx_, &Lxu_, &Luu_, &lx_, &lu_}) p.insert(p.end(), v->begin(), v->end());
    return p;
  }
  const MatrixXd& get_Fx() const { return Fx_; }
  const MatrixXd& get_Fu() const { return Fu_; }
  const VectorXd& get_f0() const { return f0_; }
  const VectorXd& get_lx() const { return lx_; }
  const VectorXd& get_lu() const { return lu_; }
  const MatrixXd& get_Lxx() const { return Lxx_; }
  const MatrixXd& get_Lxu() const { return Lxu_; }
  const MatrixXd& get_Luu() const { return Luu_; }
  void set_Fx(const MatrixXd& v) { set(Fx_, v, "Fx"); }
  void set_Fu(const MatrixXd& v) { set(Fu_, v, "Fu"); }
  void set_f0(const VectorXd& v) { set(f0_, v, "f0"); }
  void set_lx(const VectorXd& v) { set(lx_, v, "lx"); }
  void set_lu(const VectorXd& v) { set(lu_, v, "lu"); }
  void set_Lxx(const MatrixXd& v) { set(Lxx_, v, "Lxx"); }
  void set_Lxu(const MatrixXd& v) { set(Lxu_, v, "Lxu"); }
  void set_Luu(const MatrixXd& v) { set(Luu_, v, "Luu"); }

 private:
  static MatrixXd eye(std::size_t r, std::size_t c) {
    MatrixXd m(r * c, 0.);
    for (std::size_t i = 0; i < (r < c ? r : c); ++i) m[i + r * i] = 1.;
    return m;
  }
  void set(VectorXd& dst, const VectorXd& v, const char* name) {  // lqr.hxx:128-203
    if (v.size() != dst.size()) throw Exception(CGX_ERR_INVALID, std::string("Invalid argument: ") + name + " has wrong dimension");
    dst = v;
    invalidate();
  }
  bool drift_free_;
  MatrixXd Fx_, Fu_;
  VectorXd f0_;
  MatrixXd Lxx_, Lxu_, Luu_;
  VectorXd lx_, lu_;
};

// DifferentialActionModelLQR(nq, nu, drift_free) — include/crocoddyl/core/actions/diff-lqr.hxx:14-28 (identity / ones
// defaults): a = Fq q + Fv v + Fu u (+ f0), cost 1/2 [x,u]^T [Lxx Lxu; Lxu^T Luu] [x,u] + lx^T x + lu^T u. Matrices
// column-major. Put into a problem through IntegratedActionModelEuler below.
class DifferentialActionModelLQR {
 public:
  DifferentialActionModelLQR(std::size_t nq, std::size_t nu, bool drift_free = true)
      : nq_(nq), nu_(nu), drift_free_(drift_free), Fq(eye(nq, nq)), Fv(eye(nq, nq)), Fu(eye(nq, nu)), f0(nq, 1.),
        Lxx(eye(2 * nq, 2 * nq)), Lxu(eye(2 * nq, nu)), Luu(eye(nu, nu)), lx(2 * nq, 1.), lu(nu, 1.) {}
  std::size_t get_nq() const { return nq_; }
  std::size_t get_nu() const { return nu_; }
  bool get_drift_free() const { return drift_free_; }
  MatrixXd Fq, Fv, Fu;
  VectorXd f0;
  MatrixXd Lxx, Lxu, Luu;
  VectorXd lx, lu;

 private:
  static MatrixXd eye(std::size_t r, std::size_t c) {
    MatrixXd m(r * c, 0.);
    for (std::size_t i = 0; i < (r < c ? r : c); ++i) m[i + r * i] = 1.;
    return m;
  }
  std::size_t nq_, nu_;
  bool drift_free_;
};

// IntegratedActionModelEuler(DifferentialActionModelLQR, dt) (core/integrator/euler.hxx:41-131 over diff-lqr.hxx:48-84):
// the semi-implicit Euler step of a linear differential model is linear in (x, u), so the node runs on the LQR kernels
// with the discrete matrices below, built in the reference's operation order (da * dt^2, da * dt, + dt on the top-right
// diagonal, + identity; costs scaled by dt; dt = 0: identity step, unscaled cost). A negative dt falls back to 1e-3
// (euler.hxx:27-31).
inline std::shared_ptr<ActionModelLQR> IntegratedActionModelEuler(const DifferentialActionModelLQR& d, double dt = 1e-3) {
  if (dt < 0.) dt = 1e-3;
  const std::size_t nq = d.get_nq(), nu = d.get_nu(), nx = 2 * nq;
  auto m = std::make_shared<ActionModelLQR>(nx, nu, d.get_drift_free() || dt == 0.);
  MatrixXd Fx(nx * nx, 0.), Fu(nx * nu, 0.);
  VectorXd f0(nx, 0.);
  const double dt2 = dt * dt, sc = dt != 0. ? dt : 1.;
  if (dt != 0.) {
    for (std::size_t j = 0; j < nq; ++j)
      for (std::size_t i = 0; i < nq; ++i) {
        Fx[i + nx * j] = d.Fq[i + nq * j] * dt2;
        Fx[i + nx * (nq + j)] = d.Fv[i + nq * j] * dt2;
        Fx[(nq + i) + nx * j] = d.Fq[i + nq * j] * dt;
        Fx[(nq + i) + nx * (nq + j)] = d.Fv[i + nq * j] * dt;
      }
    for (std::size_t i = 0; i < nq; ++i) Fx[i + nx * (nq + i)] += dt;
    for (std::size_t i = 0; i < nx; ++i) Fx[i + nx * i] += 1.;
    for (std::size_t j = 0; j < nu; ++j)
      for (std::size_t i = 0; i < nq; ++i) {
        Fu[i + nx * j] = d.Fu[i + nq * j] * dt2;
        Fu[(nq + i) + nx * j] = d.Fu[i + nq * j] * dt;
      }
    for (std::size_t i = 0; i < nq; ++i) {
      f0[i] = d.f0[i] * dt2;
      f0[nq + i] = d.f0[i] * dt;
    }
  } else {
    for (std::size_t i = 0; i < nx; ++i) Fx[i + nx * i] = 1.;
  }
  auto scaled = [sc](VectorXd v) {
    for (double& e : v) e *= sc;
    return v;
  };
  m->set_Fx(Fx);
  m->set_Fu(Fu);
  m->set_f0(f0);
  m->set_Lxx(scaled(d.Lxx));
  m->set_Lxu(scaled(d.Lxu));
  m->set_Luu(scaled(d.Luu));
  m->set_lx(scaled(d.lx));
  m->set_lu(scaled(d.lu));
  return m;
}

// IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics(StateMultibody, ActuationModelFull, CostModelSum{
// gripperPose, xReg, uReg}), dt) on a fixed-base revolute chain (core/integrator/euler.hxx, multibody/actions/
// free-fwddyn.hxx, benchmark/factory/arm.hpp:60-90): the parameter block of cgx_arm_param_count(nv) doubles
class ActionModelArm : public ActionModelAbstract {
 public:
  ActionModelArm(std::size_t nv, const std::vector<double>& params)
      : ActionModelAbstract(std::make_shared<StateVector>(2 * nv, nv), nv, 6 + 3 * nv), p_(params) {
    if ((long)params.size() != cgx_arm_param_count((int)nv)) throw Exception(CGX_ERR_INVALID, "Invalid argument: arm parameter block has wrong size");
  }
  int kind() const override { return CGX_MODEL_ARM; }
  std::vector<double> params() const override { return p_; }
  double get_dt() const { return p_[0]; }
  // the same differential model integrated with another dt (dt = 0: a terminal knot, benchmark/factory/arm.hpp:89-90),
  // optionally with other cost weights (examples/arm_manipulation.py:41-57: terminal cost = gripperPose only)
  std::shared_ptr<ActionModelArm> with(double dt, double w_gripperPose, double w_xReg, double w_uReg) const {
    std::vector<double> p = p_;
    p[0] = dt;
    p[4] = w_gripperPose;
    p[5] = w_xReg;
    p[6] = w_uReg;
    return std::make_shared<ActionModelArm>(get_nu(), p);
  }
  std::shared_ptr<ActionModelArm> terminal() const { return with(0., p_[4], p_[5], p_[6]); }

 private:
  std::vector<double> p_;
};

class ShootingProblem {
 public:
  // ShootingProblem(x0, running_models, terminal_model) (shooting.hpp:48-63). x0 holds B initial states ([B][nx]); ONE
  // running model object is aliased into all T nodes (std::vector<...>(N, model), benchmark/unicycle-optctrl.cpp:28);
  // the terminal model may be another object of the same family and dimensions.
  ShootingProblem(const VectorXd& x0, const std::vector<std::shared_ptr<ActionModelAbstract> >& running_models,
                  std::shared_ptr<ActionModelAbstract> terminal_model)
      : running_models_(running_models), terminal_model_(terminal_model), h_(nullptr) {
    if (running_models.empty() || !terminal_model) throw Exception(CGX_ERR_INVALID, "Invalid argument: at least one running model and a terminal model are needed");
    const std::shared_ptr<ActionModelAbstract>& m = running_models[0];
    for (const auto& r : running_models)
      if (r != m) throw Exception(CGX_ERR_UNSUPPORTED, "only problems that alias one running model into every node are supported");
    if (terminal_model->kind() != m->kind() || terminal_model->get_state()->get_nx() != m->get_state()->get_nx() ||
        terminal_model->get_nu() != m->get_nu())
      throw Exception(CGX_ERR_UNSUPPORTED, "the terminal model must be of the running model's family and dimensions");
    nx_ = (int)m->get_state()->get_nx();
    nu_ = (int)m->get_nu();
    if (nx_ <= 0 || x0.size() % nx_ != 0 || x0.empty())
      throw Exception(CGX_ERR_INVALID, "Invalid argument: x0 has wrong dimension (it should be " + std::to_string(nx_) + ")");
    B_ = (int)(x0.size() / nx_);
    T_ = (int)running_models.size();
    const std::vector<double> p = m->params(), pt = terminal_model->params();
    check(cgx_problem_create(&h_, m->kind(), B_, T_, nx_, nu_, p.data(), (long)p.size(), 0, x0.data()));
    check(cgx_problem_set_terminal_params(h_, pt.data(), (long)pt.size()));
    if (m->get_has_control_limits() || any_finite(m->get_u_lb()) || any_finite(m->get_u_ub()))
      check(cgx_problem_set_u_bounds(h_, m->get_u_lb().data(), m->get_u_ub().data(), 0));
  }
  // T copies of one model, the model also at the terminal node (benchmark/unicycle-optctrl.cpp:28-32)
  ShootingProblem(const VectorXd& x0, std::size_t T, std::shared_ptr<ActionModelAbstract> model)
      : ShootingProblem(x0, std::vector<std::shared_ptr<ActionModelAbstract> >(T, model), model) {}
  ~ShootingProblem() {
    if (h_) cgx_problem_destroy(h_);
  }
  ShootingProblem(const ShootingProblem&) = delete;
  ShootingProblem& operator=(const ShootingProblem&) = delete;

  std::size_t get_T() const { return (std::size_t)T_; }
  std::size_t get_nx() const { return (std::size_t)nx_; }
  std::size_t get_ndx() const { return (std::size_t)nx_; }
  std::size_t get_nu_max() const { return (std::size_t)nu_; }
  std::size_t get_batch() const { return (std::size_t)B_; }
  const std::vector<std::shared_ptr<ActionModelAbstract> >& get_runningModels() const { return running_models_; }
  const std::shared_ptr<ActionModelAbstract>& get_terminalModel() const { return terminal_model_; }
  VectorXd get_x0() const {
    VectorXd x((std::size_t)B_ * nx_);
    check(cgx_problem_get_x0(h_, x.data()));
    return x;
  }
  void set_x0(const VectorXd& x0s) {
    if (x0s.size() != (std::size_t)B_ * nx_) throw Exception(CGX_ERR_INVALID, "Invalid argument: x0 has wrong dimension");
    check(cgx_problem_set_x0(h_, x0s.data()));
  }
  // calc / calcDiff return the total cost of every problem; xs = [B][T+1][nx], us = [B][T][nu] (shooting.hxx:133-195)
  std::vector<double> calc(const VectorXd& xs, const VectorXd& us) {
    check_traj(xs, us);
    std::vector<double> c(B_);
    check(cgx_problem_calc(h_, xs.data(), us.data(), c.data()));
    return c;
  }
  std::vector<double> calcDiff(const VectorXd& xs, const VectorXd& us) {
    check_traj(xs, us);
    std::vector<double> c(B_);
    check(cgx_problem_calc_diff(h_, xs.data(), us.data(), c.data()));
    return c;
  }
  VectorXd rollout(const VectorXd& us) {
    if (us.size() != (std::size_t)B_ * T_ * nu_) throw Exception(CGX_ERR_INVALID, "Invalid argument: us has wrong dimension");
    VectorXd xs((std::size_t)B_ * (T_ + 1) * nx_);
    check(cgx_problem_rollout(h_, us.data(), xs.data()));
    return xs;
  }
  VectorXd quasiStatic(const VectorXd& xs_point) {
    VectorXd us((std::size_t)B_ * nu_);
    check(cgx_problem_quasi_static(h_, xs_point.data(), us.data()));
    return us;
  }
  // runningDatas[t] / terminalData (t == T) of problem b after calc / calcDiff (shooting.hpp:170-181)
  std::shared_ptr<ActionDataAbstract> get_data(int b, int t) {
    std::shared_ptr<ActionDataAbstract> d = (t == T_ ? terminal_model_ : running_models_[0])->createData();
    check(cgx_problem_get_data(h_, "cost", b, t, &d->cost));
    check(cgx_problem_get_data(h_, "xnext", b, t, d->xnext.data()));
    check(cgx_problem_get_data(h_, "Fx", b, t, d->Fx.data()));
    check(cgx_problem_get_data(h_, "Fu", b, t, d->Fu.data()));
    check(cgx_problem_get_data(h_, "Lx", b, t, d->Lx.data()));
    check(cgx_problem_get_data(h_, "Lu", b, t, d->Lu.data()));
    check(cgx_problem_get_data(h_, "Lxx", b, t, d->Lxx.data()));
    check(cgx_problem_get_data(h_, "Lxu", b, t, d->Lxu.data()));
    check(cgx_problem_get_data(h_, "Luu", b, t, d->Luu.data()));
    return d;
  }
  cgx_problem_t handle() const { return h_; }

 private:
  static bool any_finite(const VectorXd& v) {
    for (double e : v)
      if (std::isfinite(e)) return true;
    return false;
  }
  void check_traj(const VectorXd& xs, const VectorXd& us) const {
    if (xs.size() != (std::size_t)B_ * (T_ + 1) * nx_)
      throw Exception(CGX_ERR_INVALID, "Invalid argument: xs has wrong dimension (it should be " + std::to_string(T_ + 1) + ")");
    if (us.size() != (std::size_t)B_ * T_ * nu_)
      throw Exception(CGX_ERR_INVALID, "Invalid argument: us has wrong dimension (it should be " + std::to_string(T_) + ")");
  }
  std::vector<std::shared_ptr<ActionModelAbstract> > running_models_;
  std::shared_ptr<ActionModelAbstract> terminal_model_;
  int B_, T_, nx_, nu_;
  cgx_problem_t h_;
};

class SolverDDP {
 public:
  explicit SolverDDP(std::shared_ptr<ShootingProblem> problem) : SolverDDP(problem, CGX_SOLVER_DDP) {}
  virtual ~SolverDDP() {
    if (h_) cgx_solver_destroy(h_);
  }
  SolverDDP(const SolverDDP&) = delete;
  SolverDDP& operator=(const SolverDDP&) = delete;

  // bool solve(init_xs, init_us, maxiter = 100, is_feasible = false, regInit = 1e-9) — ddp.cpp:41-118 / fddp.cpp:19-105.
  // Empty init_xs / init_us mean the reference's defaults (zero states / zero controls, solver-base.cpp:46-60).
  // Returns the bool of problem 0; converged() holds all B.
  bool solve(const VectorXd& init_xs = VectorXd(), const VectorXd& init_us = VectorXd(), std::size_t maxiter = 100,
             bool is_feasible = false, double regInit = 1e-9) {
    const std::size_t B = problem_->get_batch(), T = problem_->get_T(), nx = problem_->get_nx(), nu = problem_->get_nu_max();
    if (!init_xs.empty() && init_xs.size() != B * (T + 1) * nx)
      throw Exception(CGX_ERR_INVALID, "Warm start state has wrong dimension, got " + std::to_string(init_xs.size() / (B * nx)) +
                                           " expecting " + std::to_string(T + 1));
    if (!init_us.empty() && init_us.size() != B * T * nu)
      throw Exception(CGX_ERR_INVALID, "Warm start control has wrong dimension, got " + std::to_string(init_us.size() / (B * nu)) +
                                           " expecting " + std::to_string(T));
    converged_.assign(B, 0);
    check(cgx_solver_solve(h_, init_xs.empty() ? nullptr : init_xs.data(), init_us.empty() ? nullptr : init_us.data(), (int)maxiter,
                           is_feasible ? 1 : 0, regInit, converged_.data()));
    return converged_[0] != 0;
  }
  // solve(...) followed by get_xs() / get_us() as one call (cgx_solver_solve_fetch): xs_out = [B][T+1][nx],
  // us_out = [B][T][nu] are resized and filled; on the unicycle backend the copies overlap the solve.
  bool solve(const VectorXd& init_xs, const VectorXd& init_us, VectorXd& xs_out, VectorXd& us_out, std::size_t maxiter = 100,
             bool is_feasible = false, double regInit = 1e-9) {
    const std::size_t B = problem_->get_batch(), T = problem_->get_T(), nx = problem_->get_nx(), nu = problem_->get_nu_max();
    if (init_xs.size() != B * (T + 1) * nx || init_us.size() != B * T * nu)
      throw Exception(CGX_ERR_INVALID, "Warm start has wrong dimension");
    xs_out.resize(B * (T + 1) * nx);
    us_out.resize(B * T * nu);
    converged_.assign(B, 0);
    check(cgx_solver_solve_fetch(h_, init_xs.data(), init_us.data(), (int)maxiter, is_feasible ? 1 : 0, regInit,
                                 converged_.data(), xs_out.data(), us_out.data()));
    return converged_[0] != 0;
  }
  const std::vector<unsigned char>& converged() const { return converged_; }
  void setCandidate(const VectorXd& xs, const VectorXd& us, bool is_feasible = false) {
    check(cgx_solver_set_candidate(h_, xs.empty() ? nullptr : xs.data(), us.empty() ? nullptr : us.data(), is_feasible ? 1 : 0));
  }
  void computeDirection(bool recalc = true) { check(cgx_solver_compute_direction(h_, recalc ? 1 : 0, nullptr)); }
  std::vector<double> tryStep(double steplength = 1.) {
    std::vector<double> dV(problem_->get_batch());
    check(cgx_solver_try_step(h_, steplength, dV.data(), nullptr));
    return dV;
  }
  std::vector<double> expectedImprovement() {
    std::vector<double> d(2 * problem_->get_batch());
    check(cgx_solver_expected_improvement(h_, d.data()));
    return d;
  }
  std::vector<double> stoppingCriteria() {
    std::vector<double> s(problem_->get_batch());
    check(cgx_solver_stopping_criteria(h_, s.data()));
    return s;
  }
  // accessors: whole batch, [B][nt][size] (solver-base.hpp:162-167, ddp.hpp:191-236)
  std::vector<double> get_xs() const { return field("xs"); }
  std::vector<double> get_us() const { return field("us"); }
  std::vector<double> get_fs() const { return field("fs"); }
  std::vector<double> get_K() const { return field("K"); }
  std::vector<double> get_k() const { return field("k"); }
  std::vector<double> get_Vx() const { return field("Vx"); }
  std::vector<double> get_Vxx() const { return field("Vxx"); }
  std::vector<double> get_Qxx() const { return field("Qxx"); }
  std::vector<double> get_Qxu() const { return field("Qxu"); }
  std::vector<double> get_Quu() const { return field("Quu"); }
  std::vector<double> get_Qx() const { return field("Qx"); }
  std::vector<double> get_Qu() const { return field("Qu"); }
  // one node of one problem, e.g. get("K", b, t): nu x ndx column-major
  std::vector<double> get(const std::string& name, int b, int t) const {
    std::vector<double> out(problem_->get_nx() * problem_->get_nx() + 1);
    const int n = cgx_solver_get(h_, name.c_str(), b, t, out.data());
    check(n);
    out.resize((std::size_t)n);
    return out;
  }
  // scalars of problem b (solver-base.hpp:172-227): one row of CGX_NSCALARS doubles per call
  double get_cost(int b = 0) const { return scalar(b, 0); }
  double get_stop(int b = 0) const { return scalar(b, 1); }
  std::vector<double> get_d(int b = 0) const { return {scalar(b, 2), scalar(b, 3)}; }
  double get_xreg(int b = 0) const { return scalar(b, 4); }
  double get_ureg(int b = 0) const { return scalar(b, 5); }
  double get_steplength(int b = 0) const { return scalar(b, 6); }
  double get_dV(int b = 0) const { return scalar(b, 7); }
  double get_dVexp(int b = 0) const { return scalar(b, 8); }
  std::size_t get_iter(int b = 0) const { return (std::size_t)scalar(b, 9); }
  bool get_is_feasible(int b = 0) const { return scalar(b, 10) != 0.; }
  // thresholds and regularisation (ddp.hpp:156-271, solver-base.hpp:217-257, fddp.hpp:87-92)
  double get_param(const std::string& name) const {
    double v = 0.;
    check(cgx_solver_get_param(h_, name.c_str(), &v));
    return v;
  }
  void set_param(const std::string& name, double v) { check(cgx_solver_set_param(h_, name.c_str(), v)); }
  double get_th_stop() const { return get_param("th_stop"); }
  void set_th_stop(double v) { set_param("th_stop", v); }
  double get_th_acceptstep() const { return get_param("th_acceptstep"); }
  void set_th_acceptstep(double v) { set_param("th_acceptstep", v); }
  double get_th_grad() const { return get_param("th_grad"); }
  void set_th_grad(double v) { set_param("th_grad", v); }
  double get_th_stepdec() const { return get_param("th_stepdec"); }
  void set_th_stepdec(double v) { set_param("th_stepdec", v); }
  double get_th_stepinc() const { return get_param("th_stepinc"); }
  void set_th_stepinc(double v) { set_param("th_stepinc", v); }
  double get_regfactor() const { return get_param("regfactor"); }
  void set_regfactor(double v) { set_param("regfactor", v); }
  double get_regmin() const { return get_param("regmin"); }
  void set_regmin(double v) { set_param("regmin", v); }
  double get_regmax() const { return get_param("regmax"); }
  void set_regmax(double v) { set_param("regmax", v); }
  void set_xreg(double v) { set_param("xreg", v); }
  void set_ureg(double v) { set_param("ureg", v); }
  std::vector<double> get_alphas() const {
    std::vector<double> a(16);
    int n = 0;
    check(cgx_solver_get_alphas(h_, a.data(), &n));
    a.resize((std::size_t)n);
    return a;
  }
  void set_alphas(const std::vector<double>& a) { check(cgx_solver_set_alphas(h_, a.data(), (int)a.size())); }
  // setCallbacks: fn(user, record) once per recorded iteration of every problem, after the device loop finished
  void setCallbacks(cgx_callback_fn fn, void* user, int max_records = 100) {
    check(cgx_solver_enable_trace(h_, fn ? 1 : 0, max_records));
    check(cgx_solver_set_callback(h_, fn, user));
  }
  const std::shared_ptr<ShootingProblem>& get_problem() const { return problem_; }
  cgx_solver_t handle() const { return h_; }

 protected:
  SolverDDP(std::shared_ptr<ShootingProblem> problem, int kind) : problem_(problem), h_(nullptr) {
    if (!problem_) throw Exception(CGX_ERR_INVALID, "Invalid argument: null problem");
    check(cgx_solver_create(&h_, problem_->handle(), kind));
  }
  std::vector<double> field(const char* name) const {
    const std::size_t B = problem_->get_batch(), T = problem_->get_T(), n = problem_->get_nx(), m = problem_->get_nu_max();
    const std::string s(name);
    std::size_t per = 0;
    if (s == "xs" || s == "fs" || s == "Vx") per = (T + 1) * n;
    else if (s == "Vxx") per = (T + 1) * n * n;
    else if (s == "us" || s == "k" || s == "Qu") per = T * m;
    else if (s == "K" || s == "Qxu") per = T * m * n;
    else if (s == "Qxx") per = T * n * n;
    else if (s == "Quu" || s == "Quu_inv") per = T * m * m;
    else if (s == "Qx") per = T * n;
    std::vector<double> out(B * per);
    check(cgx_solver_get_all(h_, name, out.data()));
    return out;
  }
  double scalar(int b, int col) const {
    double row[CGX_NSCALARS];
    check(cgx_solver_get_scalars_row(h_, b, row));
    return row[col];
  }
  std::shared_ptr<ShootingProblem> problem_;
  cgx_solver_t h_;
  std::vector<unsigned char> converged_;
};

class SolverFDDP : public SolverDDP {
 public:
  explicit SolverFDDP(std::shared_ptr<ShootingProblem> problem) : SolverDDP(problem, CGX_SOLVER_FDDP) {}
  double get_th_acceptnegstep() const { return get_param("th_acceptnegstep"); }
  void set_th_acceptnegstep(double v) { set_param("th_acceptnegstep", v); }

 protected:
  SolverFDDP(std::shared_ptr<ShootingProblem> problem, int kind) : SolverDDP(problem, kind) {}
};

// SolverBoxDDP / SolverBoxFDDP (box-ddp.hpp, box-fddp.hpp): control-limited; read u_lb / u_ub of the running model
class SolverBoxDDP : public SolverDDP {
 public:
  explicit SolverBoxDDP(std::shared_ptr<ShootingProblem> problem) : SolverDDP(problem, CGX_SOLVER_BOXDDP) {}
  std::vector<double> get_Quu_inv() const { return field("Quu_inv"); }
};
class SolverBoxFDDP : public SolverFDDP {
 public:
  explicit SolverBoxFDDP(std::shared_ptr<ShootingProblem> problem) : SolverFDDP(problem, CGX_SOLVER_BOXFDDP) {}
  std::vector<double> get_Quu_inv() const { return field("Quu_inv"); }
};

}  // namespace crocoddyl_b200
#endif  // CROCODDYL_B200_HPP_
