// TEST INFRASTRUCTURE — CPU oracle (see linalg.hpp header). Not part of the product.
//
// Control-limited solvers, restated from the reference:
//   BoxQP (projected Newton, Bertsekas 82)   src/core/solvers/box-qp.cpp:15-186, include/crocoddyl/core/solvers/box-qp.hpp
//   SolverBoxDDP                             src/core/solvers/box-ddp.cpp:15-138
//   SolverBoxFDDP                            src/core/solvers/box-fddp.cpp:15-166
// Pinned by the structures of unittest/test_boxqp.cpp (unbounded problem = Newton step, negative-definite Hessian
// throws, proximity to scipy's bounded least squares) in tests/test_box_oracle_cpu.py.
#pragma once
#include "solvers.hpp"

namespace oracle {

// box-qp.hpp:30-52
struct BoxQPSolution {
  Mat Hff_inv;
  Vec x;
  std::vector<int> free_idx, clamped_idx;
};

struct BoxQP {
  int nx, maxiter;
  double th_acceptstep, th_grad, reg;
  std::vector<double> alphas;
  BoxQPSolution solution;
  int iters;  // bookkeeping: Newton iterations of the last solve

  // box-qp.cpp:15-48
  BoxQP(int nx_, int maxiter_ = 100, double th_acceptstep_ = 0.1, double th_grad_ = 1e-9, double reg_ = 1e-9)
      : nx(nx_), maxiter(maxiter_), th_acceptstep(th_acceptstep_), th_grad(th_grad_), reg(reg_), iters(0) {
    alphas.resize(10);
    for (int n = 0; n < 10; ++n) alphas[n] = 1. / std::pow(2., (double)n);
  }

  // Hff = H(free, free) (+ reg), its Cholesky factor and inverse; box-qp.cpp:106-121 / 148-160
  void factor_free(const Mat& H, LLT& llt) {
    const int nf = (int)solution.free_idx.size();
    Mat Hff(nf, nf);
    for (int i = 0; i < nf; ++i)
      for (int j = 0; j < nf; ++j) Hff(i, j) = H(solution.free_idx[i], solution.free_idx[j]);
    if (reg != 0.)
      for (int i = 0; i < nf; ++i) Hff(i, i) += reg;
    if (nf > 0 && !llt.compute(Hff)) throw SolveError("backward_error");
    if (nf == 0) llt.L = Mat(0, 0);
    solution.Hff_inv = identity(nf, nf);
    if (nf > 0) llt.solveInPlace(solution.Hff_inv.d.data(), nf);
  }

  // box-qp.cpp:52-186
  const BoxQPSolution& solve(const Mat& H, const Vec& q, const Vec& lb, const Vec& ub, const Vec& xinit) {
    if (H.r != nx || H.c != nx) throw std::invalid_argument("Invalid argument: H has wrong dimension");
    if ((int)q.size() != nx || (int)lb.size() != nx || (int)ub.size() != nx || (int)xinit.size() != nx)
      throw std::invalid_argument("Invalid argument: q / lb / ub / xinit has wrong dimension");
    Vec x(nx), xnew(nx), g(nx), dx(nx), Hx;
    for (int i = 0; i < nx; ++i) x[i] = std::max(std::min(xinit[i], ub[i]), lb[i]);  // :76-78
    LLT llt;
    iters = 0;
    for (int k = 0; k < maxiter; ++k) {
      solution.clamped_idx.clear();
      solution.free_idx.clear();
      // g = q + H x   :85-86
      g = q;
      gemv_add(H, x, g);
      for (int j = 0; j < nx; ++j) {
        if ((x[j] == lb[j] && g[j] > 0.) || (x[j] == ub[j] && g[j] < 0.))
          solution.clamped_idx.push_back(j);
        else
          solution.free_idx.push_back(j);
      }
      const int nf = (int)solution.free_idx.size(), nc = (int)solution.clamped_idx.size();
      if (normInf(g) <= th_grad || nf == 0) {  // :103
        if (k == 0) factor_free(H, llt);       // otherwise Hff_inv is the one of the previous iteration (:104-122)
        solution.x = x;
        return solution;
      }
      iters = k + 1;
      // Newton step in the free space  :127-170
      factor_free(H, llt);
      Vec dxf(nf);
      for (int i = 0; i < nf; ++i) dxf[i] = -q[solution.free_idx[i]];
      if (nc != 0)
        for (int i = 0; i < nf; ++i) {
          double s = 0;
          for (int j = 0; j < nc; ++j) s += H(solution.free_idx[i], solution.clamped_idx[j]) * x[solution.clamped_idx[j]];
          dxf[i] -= s;
        }
      llt.solveInPlace(dxf.data(), 1);
      for (int i = 0; i < nf; ++i) dxf[i] -= x[solution.free_idx[i]];
      std::fill(dx.begin(), dx.end(), 0.);
      for (int i = 0; i < nf; ++i) dx[solution.free_idx[i]] = dxf[i];
      // line search  :172-183
      gemv(H, x, Hx);
      const double fold = 0.5 * dot(x, Hx) + dot(q, x);
      for (size_t a = 0; a < alphas.size(); ++a) {
        const double steplength = alphas[a];
        for (int i = 0; i < nx; ++i) xnew[i] = std::max(std::min(x[i] + steplength * dx[i], ub[i]), lb[i]);
        gemv(H, xnew, Hx);
        const double fnew = 0.5 * dot(xnew, Hx) + dot(q, xnew);
        double gd = 0;
        for (int i = 0; i < nx; ++i) gd += g[i] * (x[i] - xnew[i]);
        if (fold - fnew > th_acceptstep * gd) {
          x = xnew;
          break;
        }
      }
    }
    solution.x = x;
    return solution;
  }
};

// box-ddp.cpp:46-74 == box-fddp.cpp:46-74: computeGains through the box QP on the control step.
// Hff_inv(i, j) is read with the dimensions it was last computed with (box-qp.cpp:104: after the first Newton iteration
// it belongs to the previous iteration's free set); entries outside that matrix do not exist in the reference
// (out-of-range Eigen access) and are taken as zero here.
template <class Base>
struct BoxGains : Base {
  BoxQP qp;
  std::vector<Mat> Quu_inv;
  Vec du_lb, du_ub;
  std::vector<int> qp_iters;  // bookkeeping
  explicit BoxGains(const std::shared_ptr<ShootingProblem>& p)
      : Base(p), qp(p->running_models[0]->nu, 100, 0.1, 1e-5, 0.) {
    const int T = p->T, nu = p->nu_max();
    Quu_inv.assign(T, Mat(nu, nu));
    du_lb.assign(nu, 0.);
    du_ub.assign(nu, 0.);
    qp_iters.assign(T, 0);
    this->th_stop = 5e-5;  // box-ddp.cpp:24-29
  }
  void computeGains(int t) override {
    ActionModel& m = *this->problem->running_models[t];
    const int nu = m.nu;
    if (nu > 0) {
      if (!m.has_control_limits || !this->is_feasible) {
        Base::computeGains(t);  // vanilla DDP
        return;
      }
      for (int i = 0; i < nu; ++i) {
        du_lb[i] = m.u_lb[i] - this->us[t][i];
        du_ub[i] = m.u_ub[i] - this->us[t][i];
      }
      const BoxQPSolution& sol = qp.solve(this->Quu[t], this->Qu[t], du_lb, du_ub, this->k[t]);
      qp_iters[t] = qp.iters;
      Quu_inv[t].setZero();
      const int nf = (int)sol.free_idx.size();
      for (int i = 0; i < nf; ++i)
        for (int j = 0; j < nf; ++j)
          Quu_inv[t](sol.free_idx[i], sol.free_idx[j]) = (i < sol.Hff_inv.r && j < sol.Hff_inv.c) ? sol.Hff_inv(i, j) : 0.;
      // K = Quu_inv Qxu^T ; k = -x
      const int ndx = this->problem->ndx();
      for (int j = 0; j < ndx; ++j)
        for (int i = 0; i < nu; ++i) {
          double s = 0;
          for (int l = 0; l < nu; ++l) s += Quu_inv[t](i, l) * this->Qxu[t](j, l);
          this->K[t](i, j) = s;
        }
      for (int i = 0; i < nu; ++i) this->k[t][i] = -sol.x[i];
      for (size_t i = 0; i < sol.clamped_idx.size(); ++i) this->Qu[t][sol.clamped_idx[i]] = 0.;
    }
  }
  void clamp(const ActionModel& m, Vec& u) const {
    if (m.has_control_limits)
      for (int i = 0; i < m.nu; ++i) u[i] = std::min(std::max(u[i], m.u_lb[i]), m.u_ub[i]);
  }
};

// box-ddp.cpp:76-134
struct SolverBoxDDP : BoxGains<SolverDDP> {
  explicit SolverBoxDDP(const std::shared_ptr<ShootingProblem>& p) : BoxGains<SolverDDP>(p) {}
  void forwardPass(double steplength_) override {
    if (steplength_ > 1. || steplength_ < 0.)
      throw std::invalid_argument("Invalid argument: invalid step length, value is between 0. to 1.");
    cost_try = 0.;
    xnext = problem->x0;
    const int T = problem->T;
    for (int t = 0; t < T; ++t) {
      ActionModel& m = *problem->running_models[t];
      ActionData& dt = *problem->running_datas[t];
      xs_try[t] = xnext;
      m.diff(xs[t], xs_try[t], dx[t]);
      if (m.nu != 0) {
        Vec Kdx(m.nu, 0.);
        gemv_add(K[t], dx[t], Kdx);
        for (int i = 0; i < m.nu; ++i) us_try[t][i] = us[t][i] - k[t][i] * steplength_ - Kdx[i];
        clamp(m, us_try[t]);
        m.calc(dt, xs_try[t], us_try[t]);
      } else {
        m.calc(dt, xs_try[t]);
      }
      xnext = dt.xnext;
      cost_try += dt.cost;
      if (raiseIfNaN(cost_try)) throw SolveError("forward_error");
      if (raiseIfNaN(normInf(xnext))) throw SolveError("forward_error");
    }
    ActionModel& m = *problem->terminal_model;
    if (is_feasible || steplength_ == 1) {
      xs_try.back() = xnext;
    } else {  // only the terminal node keeps its gap in this solver (box-ddp.cpp:121-125)
      Vec scaled(fs.back().size());
      for (size_t i = 0; i < scaled.size(); ++i) scaled[i] = fs.back()[i] * (steplength_ - 1);
      m.integrate(xnext, scaled, xs_try.back());
    }
    m.calc(*problem->terminal_data, xs_try.back());
    cost_try += problem->terminal_data->cost;
    if (raiseIfNaN(cost_try)) throw SolveError("forward_error");
  }
};

// box-fddp.cpp:76-162: SolverFDDP::forwardPass with the control clamp
struct SolverBoxFDDP : BoxGains<SolverFDDP> {
  explicit SolverBoxFDDP(const std::shared_ptr<ShootingProblem>& p) : BoxGains<SolverFDDP>(p) {}
  void forwardPass(double steplength_) override {
    if (steplength_ > 1. || steplength_ < 0.)
      throw std::invalid_argument("Invalid argument: invalid step length, value is between 0. to 1.");
    cost_try = 0.;
    xnext = problem->x0;
    const int T = problem->T;
    const bool closed = is_feasible || (steplength_ == 1);
    Vec scaled;
    for (int t = 0; t < T; ++t) {
      ActionModel& m = *problem->running_models[t];
      ActionData& dt = *problem->running_datas[t];
      if (closed) {
        xs_try[t] = xnext;
      } else {
        scaled.resize(fs[t].size());
        for (size_t i = 0; i < fs[t].size(); ++i) scaled[i] = fs[t][i] * (steplength_ - 1);
        m.integrate(xnext, scaled, xs_try[t]);
      }
      m.diff(xs[t], xs_try[t], dx[t]);
      if (m.nu != 0) {
        Vec Kdx(m.nu, 0.);
        gemv_add(K[t], dx[t], Kdx);
        for (int i = 0; i < m.nu; ++i) us_try[t][i] = us[t][i] - k[t][i] * steplength_ - Kdx[i];
        clamp(m, us_try[t]);
        m.calc(dt, xs_try[t], us_try[t]);
      } else {
        m.calc(dt, xs_try[t]);
      }
      xnext = dt.xnext;
      cost_try += dt.cost;
      if (raiseIfNaN(cost_try)) throw SolveError("forward_error");
      if (raiseIfNaN(normInf(xnext))) throw SolveError("forward_error");
    }
    ActionModel& m = *problem->terminal_model;
    if (closed) {
      xs_try.back() = xnext;
    } else {
      scaled.resize(fs.back().size());
      for (size_t i = 0; i < fs.back().size(); ++i) scaled[i] = fs.back()[i] * (steplength_ - 1);
      m.integrate(xnext, scaled, xs_try.back());
    }
    m.calc(*problem->terminal_data, xs_try.back());
    cost_try += problem->terminal_data->cost;
    if (raiseIfNaN(cost_try)) throw SolveError("forward_error");
  }
};

}  // namespace oracle
