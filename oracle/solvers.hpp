// TEST INFRASTRUCTURE — CPU oracle (see linalg.hpp header). Not part of the product.
//
// Restatement of the reference solve path, function by function:
//   ShootingProblem::calc/calcDiff/rollout   include/crocoddyl/core/optctrl/shooting.hxx:133-223
//   SolverAbstract ctor / setCandidate       src/core/solver-base.cpp:14-67 ; raiseIfNaN :175-181
//   SolverDDP                                src/core/solvers/ddp.cpp (all)
//   SolverFDDP                               src/core/solvers/fddp.cpp (all)
//   SolverKKT (independent cross-check)      src/core/solvers/kkt.cpp:33-227
//   per-iteration record = CallbackVerbose   src/core/utils/callbacks.cpp:37-46
#pragma once
#include <functional>
#include <limits>
#include "models.hpp"

namespace oracle {

struct SolveError : std::runtime_error {
  explicit SolveError(const char* w) : std::runtime_error(w) {}
};

// solver-base.cpp:175-181
inline bool raiseIfNaN(double value) { return std::isnan(value) || std::isinf(value) || value >= 1e30; }

// shooting.hpp:48-63, shooting.hxx:133-223
struct ShootingProblem {
  Vec x0;
  std::vector<std::shared_ptr<ActionModel> > running_models;
  std::shared_ptr<ActionModel> terminal_model;
  std::vector<std::shared_ptr<ActionData> > running_datas;
  std::shared_ptr<ActionData> terminal_data;
  int T;
  double cost;
  ShootingProblem(const Vec& x0_, const std::vector<std::shared_ptr<ActionModel> >& running,
                  const std::shared_ptr<ActionModel>& terminal)
      : x0(x0_), running_models(running), terminal_model(terminal), T((int)running.size()), cost(0.) {
    if ((int)x0.size() != terminal->nx) throw std::invalid_argument("Invalid argument: x0 has wrong dimension");
    for (int i = 0; i < T; ++i) running_datas.push_back(running_models[i]->createData());  // shooting.hxx:358-364
    terminal_data = terminal_model->createData();
  }
  int nx() const { return terminal_model->nx; }
  int ndx() const { return terminal_model->ndx; }
  int nu_max() const {
    int m = 0;
    for (int i = 0; i < T; ++i) m = running_models[i]->nu > m ? running_models[i]->nu : m;
    return m;
  }
  double calc(const std::vector<Vec>& xs, const std::vector<Vec>& us) {
    if ((int)xs.size() != T + 1) throw std::invalid_argument("Invalid argument: xs has wrong dimension");
    if ((int)us.size() != T) throw std::invalid_argument("Invalid argument: us has wrong dimension");
    for (int i = 0; i < T; ++i) {
      if (running_models[i]->nu != 0)
        running_models[i]->calc(*running_datas[i], xs[i], us[i]);
      else
        running_models[i]->calc(*running_datas[i], xs[i]);
    }
    terminal_model->calc(*terminal_data, xs.back());
    cost = 0.;
    for (int i = 0; i < T; ++i) cost += running_datas[i]->cost;  // shooting.hxx:155-159
    cost += terminal_data->cost;
    return cost;
  }
  double calcDiff(const std::vector<Vec>& xs, const std::vector<Vec>& us) {
    if ((int)xs.size() != T + 1) throw std::invalid_argument("Invalid argument: xs has wrong dimension");
    if ((int)us.size() != T) throw std::invalid_argument("Invalid argument: us has wrong dimension");
    for (int i = 0; i < T; ++i) {
      if (running_models[i]->nu != 0)
        running_models[i]->calcDiff(*running_datas[i], xs[i], us[i]);
      else
        running_models[i]->calcDiff(*running_datas[i], xs[i]);
    }
    terminal_model->calcDiff(*terminal_data, xs.back());
    cost = 0.;
    for (int i = 0; i < T; ++i) cost += running_datas[i]->cost;  // shooting.hxx:188-192
    cost += terminal_data->cost;
    return cost;
  }
  void rollout(const std::vector<Vec>& us, std::vector<Vec>& xs) {
    if ((int)xs.size() != T + 1) throw std::invalid_argument("Invalid argument: xs has wrong dimension");
    if ((int)us.size() != T) throw std::invalid_argument("Invalid argument: us has wrong dimension");
    xs[0] = x0;
    for (int i = 0; i < T; ++i) {
      if (running_models[i]->nu != 0)
        running_models[i]->calc(*running_datas[i], xs[i], us[i]);
      else
        running_models[i]->calc(*running_datas[i], xs[i]);
      xs[i + 1] = running_datas[i]->xnext;
    }
    terminal_model->calc(*terminal_data, xs.back());
  }
};

// One row of CallbackVerbose (callbacks.cpp:37-46): iter cost stop grad(-d[1]) xreg ureg step feas (+ dVexp dV)
struct IterRecord {
  double iter, cost, stop, d0, d1, xreg, ureg, steplength, is_feasible, dV, dVexp;
};

struct SolverAbstract {
  std::shared_ptr<ShootingProblem> problem;
  std::vector<Vec> xs, us;
  bool is_feasible;
  double cost, stop;
  double d[2];
  double xreg, ureg, steplength, dV, dVexp, th_acceptstep, th_stop;
  int iter;
  std::function<void(SolverAbstract&)> callback;
  std::vector<IterRecord> trace;  // filled every iteration (the reference's callback hook point)

  explicit SolverAbstract(const std::shared_ptr<ShootingProblem>& p)
      : problem(p), is_feasible(false), cost(0.), stop(0.), xreg(NAN), ureg(NAN), steplength(1.), dV(0.), dVexp(0.),
        th_acceptstep(0.1), th_stop(1e-9), iter(0) {
    d[0] = d[1] = 0.;
    const int T = p->T;
    xs.resize(T + 1);
    us.resize(T);
    for (int t = 0; t < T; ++t) {
      xs[t] = p->running_models[t]->zero();
      us[t] = Vec(p->nu_max(), 0.);
    }
    xs.back() = p->terminal_model->zero();
  }
  virtual ~SolverAbstract() {}
  // solver-base.cpp:42-67
  void setCandidate(const std::vector<Vec>& xs_warm, const std::vector<Vec>& us_warm, bool feasible) {
    const int T = problem->T;
    if (xs_warm.size() == 0) {
      for (int t = 0; t < T; ++t) xs[t] = problem->running_models[t]->zero();
      xs.back() = problem->terminal_model->zero();
    } else {
      if ((int)xs_warm.size() != T + 1) throw std::invalid_argument("Warm start state has wrong dimension");
      if (&xs_warm != &xs) xs = xs_warm;
    }
    if (us_warm.size() == 0) {
      for (int t = 0; t < T; ++t) us[t] = Vec(problem->nu_max(), 0.);
    } else {
      if ((int)us_warm.size() != T) throw std::invalid_argument("Warm start control has wrong dimension");
      if (&us_warm != &us) us = us_warm;
    }
    is_feasible = feasible;
  }
  void record() {
    IterRecord r = {(double)iter, cost, stop, d[0], d[1], xreg, ureg, steplength, is_feasible ? 1. : 0., dV, dVexp};
    trace.push_back(r);
    if (callback) callback(*this);
  }
};

// ---------------------------------------------------------------------------------------------
struct SolverDDP : SolverAbstract {
  double regfactor, regmin, regmax, cost_try, th_grad, th_stepdec, th_stepinc;
  bool was_feasible;
  std::vector<Vec> xs_try, us_try, dx;
  std::vector<Mat> Vxx, Qxx, Qxu, Quu, K, FuTVxx_p;
  std::vector<Vec> Vx, Qx, Qu, k, fs, Quuk;
  std::vector<LLT> Quu_llt;
  Mat FxTVxx_p;
  Vec fTVxx_p, xnext;
  std::vector<double> alphas;
  // bookkeeping for the accept/reject parity artefact: number of backward-pass retries and the
  // index of the accepted alpha (-1 = none) for each iteration
  std::vector<int> trace_alpha_idx, trace_retries;

  explicit SolverDDP(const std::shared_ptr<ShootingProblem>& p)
      : SolverAbstract(p), regfactor(10.), regmin(1e-9), regmax(1e9), cost_try(0.), th_grad(1e-12), th_stepdec(0.5),
        th_stepinc(0.01), was_feasible(false) {
    allocateData();
    alphas.resize(10);
    for (int n = 0; n < 10; ++n) alphas[n] = 1. / std::pow(2., (double)n);  // ddp.cpp:27-31
    if (th_stepinc < alphas[9]) th_stepinc = alphas[9];                    // ddp.cpp:32-36
  }
  // ddp.cpp:328-384
  void allocateData() {
    const int T = problem->T, ndx = problem->ndx(), nu = problem->nu_max();
    Vxx.assign(T + 1, Mat(ndx, ndx));
    Vx.assign(T + 1, Vec(ndx, 0.));
    Qxx.assign(T, Mat(ndx, ndx));
    Qxu.assign(T, Mat(ndx, nu));
    Quu.assign(T, Mat(nu, nu));
    Qx.assign(T, Vec(ndx, 0.));
    Qu.assign(T, Vec(nu, 0.));
    K.assign(T, Mat(nu, ndx));
    k.assign(T, Vec(nu, 0.));
    fs.assign(T + 1, Vec(ndx, 0.));
    xs_try.assign(T + 1, Vec(problem->nx(), NAN));
    xs_try[0] = problem->x0;
    us_try.assign(T, Vec(nu, NAN));
    dx.assign(T + 1, Vec(ndx, 0.));
    FuTVxx_p.assign(T, Mat(nu, ndx));
    Quu_llt.assign(T, LLT());
    Quuk.assign(T, Vec(nu, 0.));
    FxTVxx_p.resize(ndx, ndx);
    fTVxx_p.assign(ndx, 0.);
    xnext.assign(problem->nx(), 0.);
  }

  // ddp.cpp:41-118
  virtual bool solve(const std::vector<Vec>& init_xs, const std::vector<Vec>& init_us, int maxiter = 100,
                     bool is_feasible_ = false, double reginit = 1e-9) {
    trace.clear();
    trace_alpha_idx.clear();
    trace_retries.clear();
    xs_try[0] = problem->x0;
    setCandidate(init_xs, init_us, is_feasible_);
    if (std::isnan(reginit)) {
      xreg = regmin;
      ureg = regmin;
    } else {
      xreg = reginit;
      ureg = reginit;
    }
    was_feasible = false;
    bool recalcDiff = true;
    for (iter = 0; iter < maxiter; ++iter) {
      int retries = 0;
      while (true) {
        try {
          computeDirection(recalcDiff);
        } catch (SolveError&) {
          recalcDiff = false;
          increaseRegularization();
          ++retries;
          if (xreg == regmax) return false;
          continue;
        }
        break;
      }
      expectedImprovement();
      recalcDiff = false;
      int accepted = -1;
      for (size_t a = 0; a < alphas.size(); ++a) {
        steplength = alphas[a];
        try {
          dV = tryStep(steplength);
        } catch (SolveError&) {
          continue;
        }
        dVexp = steplength * (d[0] + 0.5 * steplength * d[1]);
        if (dVexp >= 0) {
          if (d[0] < th_grad || !is_feasible || dV > th_acceptstep * dVexp) {
            was_feasible = is_feasible;
            setCandidate(xs_try, us_try, true);
            cost = cost_try;
            recalcDiff = true;
            accepted = (int)a;
            break;
          }
        }
      }
      trace_alpha_idx.push_back(accepted);
      trace_retries.push_back(retries);
      if (steplength > th_stepdec) decreaseRegularization();
      if (steplength <= th_stepinc) {
        increaseRegularization();
        if (xreg == regmax) return false;
      }
      stoppingCriteria();
      record();
      if (was_feasible && stop < th_stop) return true;
    }
    return false;
  }

  // ddp.cpp:120-125
  virtual void computeDirection(bool recalcDiff) {
    if (recalcDiff) calcDiff();
    backwardPass();
  }
  // ddp.cpp:127-130
  virtual double tryStep(double steplength_) {
    forwardPass(steplength_);
    return cost - cost_try;
  }
  // ddp.cpp:132-142
  double stoppingCriteria() {
    stop = 0.;
    const int T = problem->T;
    for (int t = 0; t < T; ++t)
      if (problem->running_models[t]->nu != 0) stop += dot(Qu[t], Qu[t]);
    return stop;
  }
  // ddp.cpp:144-155
  virtual const double* expectedImprovement() {
    d[0] = d[1] = 0.;
    const int T = problem->T;
    for (int t = 0; t < T; ++t)
      if (problem->running_models[t]->nu != 0) {
        d[0] += dot(Qu[t], k[t]);
        d[1] -= dot(k[t], Quuk[t]);
      }
    return d;
  }
  // ddp.cpp:157-178
  double calcDiff() {
    if (iter == 0) problem->calc(xs, us);
    cost = problem->calcDiff(xs, us);
    const int T = problem->T;
    if (!is_feasible) {
      problem->running_models[0]->diff(xs[0], problem->x0, fs[0]);
      for (int t = 0; t < T; ++t) problem->running_models[t]->diff(xs[t + 1], problem->running_datas[t]->xnext, fs[t + 1]);
    } else if (!was_feasible) {
      for (size_t t = 0; t < fs.size(); ++t) std::fill(fs[t].begin(), fs[t].end(), 0.);
    }
    return cost;
  }
  // ddp.cpp:180-253
  void backwardPass() {
    const ActionData& d_T = *problem->terminal_data;
    const int T = problem->T;
    Vxx[T] = d_T.Lxx;
    Vx[T] = d_T.Lx;
    const int ndx = problem->ndx();
    if (!std::isnan(xreg))
      for (int i = 0; i < ndx; ++i) Vxx[T](i, i) += xreg;
    if (!is_feasible) gemv_add(Vxx[T], fs[T], Vx[T]);

    for (int t = T - 1; t >= 0; --t) {
      const ActionModel& m = *problem->running_models[t];
      const ActionData& dt = *problem->running_datas[t];
      const Mat& Vxx_p = Vxx[t + 1];
      const Vec& Vx_p = Vx[t + 1];
      const int nu = m.nu;

      Qxx[t] = dt.Lxx;
      Qx[t] = dt.Lx;
      gemm_tn(dt.Fx, Vxx_p, FxTVxx_p);       // :204
      gemm_nn_add(FxTVxx_p, dt.Fx, Qxx[t]);  // :205
      gemtv_add(dt.Fx, Vx_p, Qx[t]);         // :206
      if (nu != 0) {
        Qxu[t] = dt.Lxu;
        Quu[t] = dt.Luu;
        Qu[t] = dt.Lu;
        gemm_tn(dt.Fu, Vxx_p, FuTVxx_p[t]);       // :211
        gemm_nn_add(FxTVxx_p, dt.Fu, Qxu[t]);     // :212
        gemm_nn_add(FuTVxx_p[t], dt.Fu, Quu[t]);  // :213
        gemtv_add(dt.Fu, Vx_p, Qu[t]);            // :214
        if (!std::isnan(ureg))
          for (int i = 0; i < nu; ++i) Quu[t](i, i) += ureg;
      }

      computeGains(t);

      Vx[t] = Qx[t];
      Vxx[t] = Qxx[t];
      if (nu != 0) {
        if (std::isnan(ureg)) {
          gemtv_add(K[t], Qu[t], Vx[t], -1.);
        } else {
          gemv(Quu[t], k[t], Quuk[t]);         // :229
          gemtv_add(K[t], Quuk[t], Vx[t]);     // :230
          gemtv_add(K[t], Qu[t], Vx[t], -2.);  // :231
        }
        gemm_nn_add(Qxu[t], K[t], Vxx[t], -1.);  // :233
      }
      // :235  Vxx = 0.5 (Vxx + Vxx^T)
      for (int j = 0; j < ndx; ++j)
        for (int i = j + 1; i < ndx; ++i) {
          const double s = 0.5 * (Vxx[t](i, j) + Vxx[t](j, i));
          Vxx[t](i, j) = s;
          Vxx[t](j, i) = s;
        }
      for (int i = 0; i < ndx; ++i) Vxx[t](i, i) = 0.5 * (Vxx[t](i, i) + Vxx[t](i, i));
      if (!std::isnan(xreg))
        for (int i = 0; i < ndx; ++i) Vxx[t](i, i) += xreg;
      if (!is_feasible) gemv_add(Vxx[t], fs[t], Vx[t]);  // :242-244
      if (raiseIfNaN(normInf(Vx[t]))) throw SolveError("backward_error");
      if (raiseIfNaN(normInf(Vxx[t]))) throw SolveError("backward_error");
    }
  }
  // ddp.cpp:255-296
  virtual void forwardPass(double steplength_) {
    if (steplength_ > 1. || steplength_ < 0.)
      throw std::invalid_argument("Invalid argument: invalid step length, value is between 0. to 1.");
    cost_try = 0.;
    const int T = problem->T;
    for (int t = 0; t < T; ++t) {
      ActionModel& m = *problem->running_models[t];
      ActionData& dt = *problem->running_datas[t];
      m.diff(xs[t], xs_try[t], dx[t]);
      if (m.nu != 0) {
        for (int i = 0; i < m.nu; ++i) us_try[t][i] = us[t][i];
        for (int i = 0; i < m.nu; ++i) us_try[t][i] -= k[t][i] * steplength_;
        gemv_add(K[t], dx[t], us_try[t], -1.);
        m.calc(dt, xs_try[t], us_try[t]);
      } else {
        m.calc(dt, xs_try[t]);
      }
      xs_try[t + 1] = dt.xnext;
      cost_try += dt.cost;
      if (raiseIfNaN(cost_try)) throw SolveError("forward_error");
      if (raiseIfNaN(normInf(xs_try[t + 1]))) throw SolveError("forward_error");
    }
    problem->terminal_model->calc(*problem->terminal_data, xs_try.back());
    cost_try += problem->terminal_data->cost;
    if (raiseIfNaN(cost_try)) throw SolveError("forward_error");
  }
  // ddp.cpp:298-310 (virtual: SolverBoxDDP / SolverBoxFDDP replace it, box-ddp.cpp:46-74)
  virtual void computeGains(int t) {
    const int nu = problem->running_models[t]->nu;
    if (nu > 0) {
      if (!Quu_llt[t].compute(Quu[t])) throw SolveError("backward_error");
      const int ndx = problem->ndx();
      for (int j = 0; j < ndx; ++j)
        for (int i = 0; i < nu; ++i) K[t](i, j) = Qxu[t](j, i);
      Quu_llt[t].solveInPlace(K[t].d.data(), ndx);
      k[t] = Qu[t];
      Quu_llt[t].solveInPlace(k[t].data(), 1);
    }
  }
  // ddp.cpp:312-326
  void increaseRegularization() {
    xreg *= regfactor;
    if (xreg > regmax) xreg = regmax;
    ureg = xreg;
  }
  void decreaseRegularization() {
    xreg /= regfactor;
    if (xreg < regmin) xreg = regmin;
    ureg = xreg;
  }
};

// ---------------------------------------------------------------------------------------------
struct SolverFDDP : SolverDDP {
  double dg, dq, dv, th_acceptnegstep;
  explicit SolverFDDP(const std::shared_ptr<ShootingProblem>& p)
      : SolverDDP(p), dg(0), dq(0), dv(0), th_acceptnegstep(2) {}

  // fddp.cpp:19-105
  bool solve(const std::vector<Vec>& init_xs, const std::vector<Vec>& init_us, int maxiter = 100,
             bool is_feasible_ = false, double reginit = 1e-9) override {
    trace.clear();
    trace_alpha_idx.clear();
    trace_retries.clear();
    xs_try[0] = problem->x0;
    setCandidate(init_xs, init_us, is_feasible_);
    if (std::isnan(reginit)) {
      xreg = regmin;
      ureg = regmin;
    } else {
      xreg = reginit;
      ureg = reginit;
    }
    was_feasible = false;
    bool recalcDiff = true;
    for (iter = 0; iter < maxiter; ++iter) {
      int retries = 0;
      while (true) {
        try {
          computeDirection(recalcDiff);
        } catch (SolveError&) {
          recalcDiff = false;
          increaseRegularization();
          ++retries;
          if (xreg == regmax) return false;
          continue;
        }
        break;
      }
      updateExpectedImprovement();
      recalcDiff = false;
      int accepted = -1;
      for (size_t a = 0; a < alphas.size(); ++a) {
        steplength = alphas[a];
        try {
          dV = tryStep(steplength);
        } catch (SolveError&) {
          continue;
        }
        expectedImprovement();
        dVexp = steplength * (d[0] + 0.5 * steplength * d[1]);
        if (dVexp >= 0) {  // descend direction
          if (d[0] < th_grad || dV > th_acceptstep * dVexp) {
            was_feasible = is_feasible;
            setCandidate(xs_try, us_try, was_feasible || (steplength == 1));
            cost = cost_try;
            recalcDiff = true;
            accepted = (int)a;
            break;
          }
        } else {  // reducing the gaps by allowing a small increment in the cost value
          if (dV > th_acceptnegstep * dVexp) {
            was_feasible = is_feasible;
            setCandidate(xs_try, us_try, was_feasible || (steplength == 1));
            cost = cost_try;
            recalcDiff = true;
            accepted = (int)a;
            break;
          }
        }
      }
      trace_alpha_idx.push_back(accepted);
      trace_retries.push_back(retries);
      if (steplength > th_stepdec) decreaseRegularization();
      if (steplength <= th_stepinc) {
        increaseRegularization();
        if (xreg == regmax) return false;
      }
      stoppingCriteria();
      record();
      if (was_feasible && stop < th_stop) return true;
    }
    return false;
  }
  // fddp.cpp:107-124
  const double* expectedImprovement() override {
    dv = 0;
    const int T = problem->T;
    if (!is_feasible) {
      problem->terminal_model->diff(xs_try.back(), xs.back(), dx.back());
      gemv(Vxx.back(), dx.back(), fTVxx_p);
      dv -= dot(fs.back(), fTVxx_p);
      for (int t = 0; t < T; ++t) {
        problem->running_models[t]->diff(xs_try[t], xs[t], dx[t]);
        gemv(Vxx[t], dx[t], fTVxx_p);
        dv -= dot(fs[t], fTVxx_p);
      }
    }
    d[0] = dg + dv;
    d[1] = dq - 2 * dv;
    return d;
  }
  // fddp.cpp:126-147
  void updateExpectedImprovement() {
    dg = 0;
    dq = 0;
    const int T = problem->T;
    if (!is_feasible) {
      dg -= dot(Vx.back(), fs.back());
      gemv(Vxx.back(), fs.back(), fTVxx_p);
      dq += dot(fs.back(), fTVxx_p);
    }
    for (int t = 0; t < T; ++t) {
      if (problem->running_models[t]->nu != 0) {
        dg += dot(Qu[t], k[t]);
        dq -= dot(k[t], Quuk[t]);
      }
      if (!is_feasible) {
        dg -= dot(Vx[t], fs[t]);
        gemv(Vxx[t], fs[t], fTVxx_p);
        dq += dot(fs[t], fTVxx_p);
      }
    }
  }
  // fddp.cpp:149-225
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
        // us_try = us - k*alpha - K dx   (one Eigen expression, fddp.cpp:167/199)
        Vec Kdx(m.nu, 0.);
        gemv_add(K[t], dx[t], Kdx);
        for (int i = 0; i < m.nu; ++i) us_try[t][i] = us[t][i] - k[t][i] * steplength_ - Kdx[i];
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

// ---------------------------------------------------------------------------------------------
// SolverKKT — kkt.cpp:33-227. Dense KKT system solved by LU; used only as the independent
// cross-check the reference's own test uses (unittest/test_solvers.cpp:65-110).
struct SolverKKT : SolverAbstract {
  double th_grad, cost_try;
  bool was_feasible;
  int ndx_tot, nu_tot;
  Mat kkt;
  Vec kktref, primaldual, primal, dual;
  std::vector<Vec> dxs, dus, lambdas, xs_try, us_try;
  std::vector<double> alphas;
  explicit SolverKKT(const std::shared_ptr<ShootingProblem>& p)
      : SolverAbstract(p), th_grad(1e-12), cost_try(0.), was_feasible(false) {
    xreg = 0.;
    ureg = 0.;
    alphas.resize(10);
    for (int n = 0; n < 10; ++n) alphas[n] = 1. / std::pow(2., (double)n);
    const int T = p->T, ndx = p->ndx(), nu = p->nu_max();
    ndx_tot = (T + 1) * ndx;
    nu_tot = 0;
    for (int t = 0; t < T; ++t) nu_tot += p->running_models[t]->nu;
    kkt.resize(2 * ndx_tot + nu_tot, 2 * ndx_tot + nu_tot);
    kktref.assign(2 * ndx_tot + nu_tot, 0.);
    dxs.assign(T + 1, Vec(ndx, 0.));
    dus.assign(T, Vec(nu, 0.));
    lambdas.assign(T + 1, Vec(ndx, 0.));
    xs_try.assign(T + 1, Vec(p->nx(), 0.));
    us_try.assign(T, Vec(nu, 0.));
  }
  // kkt.cpp:177-221
  double calc() {
    cost = problem->calc(xs, us);
    cost = problem->calcDiff(xs, us);
    const int T = problem->T;
    const int cx0 = problem->running_models[0]->ndx;
    const int off = ndx_tot + nu_tot;
    kkt.setZero();
    for (int i = 0; i < ndx_tot; ++i) kkt(off + i, i) = 1.;
    int ix = 0, iu = 0;
    for (int t = 0; t < T; ++t) {
      const ActionModel& m = *problem->running_models[t];
      const ActionData& dt = *problem->running_datas[t];
      const int ndxi = m.ndx, nui = m.nu;
      if (t == 0)
        for (int i = 0; i < ndxi; ++i) kktref[off + i] = xs[0][i] - problem->x0[i];  // diff(x0, xs[0])
      for (int j = 0; j < ndxi; ++j)
        for (int i = 0; i < ndxi; ++i) {
          kkt(ix + i, ix + j) = dt.Lxx(i, j);
          kkt(off + cx0 + ix + i, ix + j) = -dt.Fx(i, j);
        }
      for (int j = 0; j < nui; ++j)
        for (int i = 0; i < ndxi; ++i) {
          kkt(ix + i, ndx_tot + iu + j) = dt.Lxu(i, j);
          kkt(ndx_tot + iu + j, ix + i) = dt.Lxu(i, j);
          kkt(off + cx0 + ix + i, ndx_tot + iu + j) = -dt.Fu(i, j);
        }
      for (int j = 0; j < nui; ++j)
        for (int i = 0; i < nui; ++i) kkt(ndx_tot + iu + i, ndx_tot + iu + j) = dt.Luu(i, j);
      for (int i = 0; i < ndxi; ++i) kktref[ix + i] = dt.Lx[i];
      for (int i = 0; i < nui; ++i) kktref[ndx_tot + iu + i] = dt.Lu[i];
      for (int i = 0; i < ndxi; ++i) kktref[off + cx0 + ix + i] = xs[t + 1][i] - dt.xnext[i];  // diff(xnext, xs[t+1])
      ix += ndxi;
      iu += nui;
    }
    const ActionData& df = *problem->terminal_data;
    const int ndxf = problem->terminal_model->ndx;
    for (int j = 0; j < ndxf; ++j)
      for (int i = 0; i < ndxf; ++i) kkt(ix + i, ix + j) = df.Lxx(i, j);
    for (int i = 0; i < ndxf; ++i) kktref[ix + i] = df.Lx[i];
    for (int i = 0; i < off; ++i)
      for (int j = 0; j < ndx_tot; ++j) kkt(i, off + j) = kkt(off + j, i);
    return cost;
  }
  // kkt.cpp:223-227 + :87-112
  void computeDirection(bool recalc) {
    if (recalc) calc();
    Vec rhs(kktref.size());
    for (size_t i = 0; i < rhs.size(); ++i) rhs[i] = -kktref[i];
    if (!lu_solve(kkt, rhs, primaldual)) throw SolveError("kkt singular");
    const int T = problem->T, ndx = problem->ndx();
    int ix = 0, iu = 0;
    for (int t = 0; t < T; ++t) {
      const int nui = problem->running_models[t]->nu;
      for (int i = 0; i < ndx; ++i) dxs[t][i] = primaldual[ix + i];
      for (int i = 0; i < nui; ++i) dus[t][i] = primaldual[ndx_tot + iu + i];
      for (int i = 0; i < ndx; ++i) lambdas[t][i] = primaldual[ndx_tot + nu_tot + ix + i];
      ix += ndx;
      iu += nui;
    }
    for (int i = 0; i < ndx; ++i) dxs[T][i] = primaldual[ix + i];
    for (int i = 0; i < ndx; ++i) lambdas[T][i] = primaldual[ndx_tot + nu_tot + ix + i];
  }
  // kkt.cpp:114-131
  double tryStep(double a) {
    const int T = problem->T;
    for (int t = 0; t < T; ++t) {
      for (size_t i = 0; i < xs[t].size(); ++i) xs_try[t][i] = xs[t][i] + a * dxs[t][i];
      for (size_t i = 0; i < us[t].size(); ++i) us_try[t][i] = us[t][i] + a * dus[t][i];
    }
    for (size_t i = 0; i < xs[T].size(); ++i) xs_try[T][i] = xs[T][i] + a * dxs[T][i];
    cost_try = problem->calc(xs_try, us_try);
    return cost - cost_try;
  }
  // kkt.cpp:33-85 (stopping criterion simplified to the primal/dual residual norm; the cross-check only
  // needs xs/us of the converged solution)
  bool solve(const std::vector<Vec>& init_xs, const std::vector<Vec>& init_us, int maxiter = 100,
             bool is_feasible_ = false) {
    setCandidate(init_xs, init_us, is_feasible_);
    for (iter = 0; iter < maxiter; ++iter) {
      computeDirection(true);
      const int np = ndx_tot + nu_tot;
      d[0] = 0;
      for (int i = 0; i < np; ++i) d[0] -= kktref[i] * primaldual[i];
      d[1] = 0;
      for (int i = 0; i < np; ++i) {
        double s = 0;
        for (int j = 0; j < np; ++j) s += kkt(i, j) * primaldual[j];
        d[1] -= s * primaldual[i];
      }
      for (size_t a = 0; a < alphas.size(); ++a) {
        steplength = alphas[a];
        dV = tryStep(steplength);
        dVexp = steplength * d[0] + 0.5 * steplength * steplength * d[1];
        if (d[0] < th_grad || !is_feasible || dV > th_acceptstep * dVexp) {
          was_feasible = is_feasible;
          setCandidate(xs_try, us_try, true);
          cost = cost_try;
          break;
        }
      }
      // stop = |primal step|^2 (converged when the Newton step vanishes)
      stop = 0;
      for (int i = 0; i < np; ++i) stop += primaldual[i] * primaldual[i];
      if (was_feasible && stop < th_stop) return true;
    }
    return false;
  }
};

}  // namespace oracle
