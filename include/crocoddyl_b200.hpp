// crocoddyl_b200 — C++ facade over the C ABI (crocoddyl_b200.h) with the reference's class and method names for the
// solve path: ShootingProblem (include/crocoddyl/core/optctrl/shooting.hpp:48-207), SolverDDP / SolverFDDP
// (include/crocoddyl/core/solvers/ddp.hpp:48-302, fddp.hpp:50-100), crocoddyl::Exception (core/utils/exception.hpp).
// Header-only; link with -lcrocoddyl_b200. Eigen and Boost are not required: vectors are std::vector<double>
// (matrices column-major, as Eigen::MatrixXd), shared ownership is std::shared_ptr. A problem is a batch of B OCPs;
// with B = 1 the calls read like the reference's (benchmark/unicycle-optctrl.cpp:25-47).
#ifndef CROCODDYL_B200_HPP_
#define CROCODDYL_B200_HPP_

#include <cmath>
#include <cstddef>
#include <exception>
#include <memory>
#include <string>
#include <vector>

#include "crocoddyl_b200.h"

namespace crocoddyl_b200 {

typedef std::vector<double> VectorXd;                 // one vector / one column-major matrix
typedef std::vector<std::vector<double> > Trajectory;  // std::vector<Eigen::VectorXd> of the reference

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

// Node models of the path (parameter blocks documented in crocoddyl_b200.h)
struct ActionModel {
  int kind, nx, nu;
  std::vector<double> params;
};
// ActionModelUnicycle() — include/crocoddyl/core/actions/unicycle.hxx:13-16 (dt = 0.1, cost weights (10, 1))
inline ActionModel ActionModelUnicycle(double dt = 0.1, double wx = 10., double wu = 1.) {
  return ActionModel{CGX_MODEL_UNICYCLE, 3, 2, {dt, wx, wu}};
}
// ActionModelLQR(nx, nu, drift_free) — include/crocoddyl/core/actions/lqr.hxx:14-25 (identity / ones defaults)
inline ActionModel ActionModelLQR(int nx, int nu, bool drift_free = true) {
  ActionModel m{CGX_MODEL_LQR, nx, nu, {}};
  std::vector<double>& p = m.params;
  p.push_back(drift_free ? 1. : 0.);
  auto eye = [&p](int r, int c) {
    for (int j = 0; j < c; ++j)
      for (int i = 0; i < r; ++i) p.push_back(i == j ? 1. : 0.);
  };
  eye(nx, nx);                      // Fx
  eye(nx, nu);                      // Fu
  p.insert(p.end(), nx, 1.);        // f0
  eye(nx, nx);                      // Lxx
  eye(nx, nu);                      // Lxu
  eye(nu, nu);                      // Luu
  p.insert(p.end(), nx, 1.);        // lx
  p.insert(p.end(), nu, 1.);        // lu
  return m;
}
// Euler o FreeFwdDynamics on a revolute chain: pass the parameter block of cgx_arm_param_count(nv) doubles
inline ActionModel ActionModelArm(int nv, const std::vector<double>& params) {
  if ((long)params.size() != cgx_arm_param_count(nv)) throw Exception(CGX_ERR_INVALID, "Invalid argument: arm parameter block has wrong size");
  return ActionModel{CGX_MODEL_ARM, 2 * nv, nv, params};
}

class ShootingProblem {
 public:
  // ShootingProblem(x0, running_models, terminal_model): the path aliases ONE model into every node, so the
  // running models are given by the horizon length T. x0s holds B initial states ([B][nx]).
  ShootingProblem(const std::vector<double>& x0s, std::size_t T, const ActionModel& model) : model_(model), h_(nullptr) {
    if (model.nx <= 0 || x0s.size() % model.nx != 0 || x0s.empty())
      throw Exception(CGX_ERR_INVALID, "Invalid argument: x0 has wrong dimension (it should be " + std::to_string(model.nx) + ")");
    B_ = (int)(x0s.size() / model.nx);
    T_ = (int)T;
    check(cgx_problem_create(&h_, model.kind, B_, T_, model.nx, model.nu, model.params.data(), (long)model.params.size(), 0,
                             x0s.data()));
  }
  ~ShootingProblem() {
    if (h_) cgx_problem_destroy(h_);
  }
  ShootingProblem(const ShootingProblem&) = delete;
  ShootingProblem& operator=(const ShootingProblem&) = delete;

  std::size_t get_T() const { return (std::size_t)T_; }
  std::size_t get_nx() const { return (std::size_t)model_.nx; }
  std::size_t get_ndx() const { return (std::size_t)model_.nx; }
  std::size_t get_nu_max() const { return (std::size_t)model_.nu; }
  std::size_t get_batch() const { return (std::size_t)B_; }
  std::vector<double> get_x0() const {
    std::vector<double> x((std::size_t)B_ * model_.nx);
    check(cgx_problem_get_x0(h_, x.data()));
    return x;
  }
  void set_x0(const std::vector<double>& x0s) {
    if (x0s.size() != (std::size_t)B_ * model_.nx) throw Exception(CGX_ERR_INVALID, "Invalid argument: x0 has wrong dimension");
    check(cgx_problem_set_x0(h_, x0s.data()));
  }
  // calc / calcDiff return the total cost of every problem; xs = [B][T+1][nx], us = [B][T][nu]
  std::vector<double> calc(const std::vector<double>& xs, const std::vector<double>& us) {
    check_traj(xs, us);
    std::vector<double> c(B_);
    check(cgx_problem_calc(h_, xs.data(), us.data(), c.data()));
    return c;
  }
  std::vector<double> calcDiff(const std::vector<double>& xs, const std::vector<double>& us) {
    check_traj(xs, us);
    std::vector<double> c(B_);
    check(cgx_problem_calc_diff(h_, xs.data(), us.data(), c.data()));
    return c;
  }
  std::vector<double> rollout(const std::vector<double>& us) {
    if (us.size() != (std::size_t)B_ * T_ * model_.nu) throw Exception(CGX_ERR_INVALID, "Invalid argument: us has wrong dimension");
    std::vector<double> xs((std::size_t)B_ * (T_ + 1) * model_.nx);
    check(cgx_problem_rollout(h_, us.data(), xs.data()));
    return xs;
  }
  std::vector<double> quasiStatic(const std::vector<double>& xs_point) {
    std::vector<double> us((std::size_t)B_ * model_.nu);
    check(cgx_problem_quasi_static(h_, xs_point.data(), us.data()));
    return us;
  }
  cgx_problem_t handle() const { return h_; }

 private:
  void check_traj(const std::vector<double>& xs, const std::vector<double>& us) const {
    if (xs.size() != (std::size_t)B_ * (T_ + 1) * model_.nx)
      throw Exception(CGX_ERR_INVALID, "Invalid argument: xs has wrong dimension (it should be " + std::to_string(T_ + 1) + ")");
    if (us.size() != (std::size_t)B_ * T_ * model_.nu)
      throw Exception(CGX_ERR_INVALID, "Invalid argument: us has wrong dimension (it should be " + std::to_string(T_) + ")");
  }
  ActionModel model_;
  int B_, T_;
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
  bool solve(const std::vector<double>& init_xs = std::vector<double>(), const std::vector<double>& init_us = std::vector<double>(),
             std::size_t maxiter = 100, bool is_feasible = false, double regInit = 1e-9) {
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
  bool solve(const std::vector<double>& init_xs, const std::vector<double>& init_us, std::vector<double>& xs_out,
             std::vector<double>& us_out, std::size_t maxiter = 100, bool is_feasible = false, double regInit = 1e-9) {
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
  void setCandidate(const std::vector<double>& xs, const std::vector<double>& us, bool is_feasible = false) {
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
  // scalars of problem b (solver-base.hpp:172-227)
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
    else if (s == "Quu") per = T * m * m;
    else if (s == "Qx") per = T * n;
    std::vector<double> out(B * per);
    check(cgx_solver_get_all(h_, name, out.data()));
    return out;
  }
  double scalar(int b, int col) const {
    std::vector<double> sc((std::size_t)CGX_NSCALARS * problem_->get_batch());
    check(cgx_solver_get_scalars(h_, sc.data()));
    return sc[(std::size_t)b * CGX_NSCALARS + col];
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
};

}  // namespace crocoddyl_b200
#endif  // CROCODDYL_B200_HPP_
