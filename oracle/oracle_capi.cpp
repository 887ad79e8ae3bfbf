// TEST INFRASTRUCTURE — CPU oracle (see linalg.hpp header). Not part of the product.
//
// extern "C" surface over the oracle so that tests/ (ctypes) and bench.py's cpu_baseline /
// --impl reference leg can drive a batch of independent OCPs. The batch loop is an OpenMP
// `parallel for schedule(dynamic)` over problems (one ShootingProblem + solver + datas per
// problem, models read-only), which is the CPU-baseline plan of BASELINE.md §2.
#include <omp.h>
#include <time.h>
#include <cstdio>
#include <cstring>
#include <string>
#include "arm.hpp"
#include "box.hpp"
#include "solvers.hpp"

using namespace oracle;

namespace {

enum { KIND_LQR = 0, KIND_UNICYCLE = 1, KIND_ARM = 2, KIND_EULER_DIFFLQR = 3 };
enum { SOLVER_DDP = 0, SOLVER_FDDP = 1, SOLVER_BOXDDP = 2, SOLVER_BOXFDDP = 3 };

struct Batch {
  int kind, solver_kind, B, T, nx, nu;
  std::vector<std::shared_ptr<ShootingProblem> > problems;
  std::vector<std::shared_ptr<SolverDDP> > solvers;
  std::vector<int> converged;
  std::string error;
};

size_t lqr_param_count(int n, int m) { return 1 + 2 * (size_t)n * n + 2 * (size_t)n * m + (size_t)m * m + 2 * n + m; }

std::shared_ptr<ActionModel> make_lqr(int n, int m, const double* p) {
  std::shared_ptr<ActionModelLQR> M = std::make_shared<ActionModelLQR>(n, m, p[0] != 0.0);
  const double* q = p + 1;
  auto take = [&](double* dst, size_t cnt) {
    std::memcpy(dst, q, cnt * sizeof(double));
    q += cnt;
  };
  take(M->Fx.d.data(), (size_t)n * n);
  take(M->Fu.d.data(), (size_t)n * m);
  take(M->f0.data(), n);
  take(M->Lxx.d.data(), (size_t)n * n);
  take(M->Lxu.d.data(), (size_t)n * m);
  take(M->Luu.d.data(), (size_t)m * m);
  take(M->lx.data(), n);
  take(M->lu.data(), m);
  return M;
}

// [dt, drift_free, Fq(nq*nq), Fv(nq*nq), Fu(nq*m), f0(nq), Lxx(n*n), Lxu(n*m), Luu(m*m), lx(n), lu(m)], column-major
size_t difflqr_param_count(int n, int m) {
  const size_t nq = n / 2;
  return 2 + 2 * nq * nq + nq * m + nq + (size_t)n * n + (size_t)n * m + (size_t)m * m + n + m;
}
std::shared_ptr<ActionModel> make_euler_difflqr(int n, int m, const double* p) {
  const int nq = n / 2;
  std::shared_ptr<ActionModelEulerDiffLQR> M = std::make_shared<ActionModelEulerDiffLQR>(nq, m, p[1] != 0.0, p[0]);
  const double* q = p + 2;
  auto take = [&](double* dst, size_t cnt) {
    std::memcpy(dst, q, cnt * sizeof(double));
    q += cnt;
  };
  take(M->Fq.d.data(), (size_t)nq * nq);
  take(M->Fv.d.data(), (size_t)nq * nq);
  take(M->Fu.d.data(), (size_t)nq * m);
  take(M->f0.data(), nq);
  take(M->Lxx.d.data(), (size_t)n * n);
  take(M->Lxu.d.data(), (size_t)n * m);
  take(M->Luu.d.data(), (size_t)m * m);
  take(M->lx.data(), n);
  take(M->lu.data(), m);
  return M;
}

std::vector<Vec> unpack(const double* p, int count, int len) {
  std::vector<Vec> out;
  if (!p) return out;
  out.resize(count);
  for (int i = 0; i < count; ++i) out[i].assign(p + (size_t)i * len, p + (size_t)(i + 1) * len);
  return out;
}

int copy_mat(const Mat& m, double* out) {
  std::memcpy(out, m.d.data(), m.d.size() * sizeof(double));
  return (int)m.d.size();
}
int copy_vec(const Vec& v, double* out) {
  std::memcpy(out, v.data(), v.size() * sizeof(double));
  return (int)v.size();
}

}  // namespace

extern "C" {

// params layout:
//   LQR      : [drift_free, Fx(n*n), Fu(n*m), f0(n), Lxx(n*n), Lxu(n*m), Luu(m*m), lx(n), lu(m)]  column-major
//   UNICYCLE : [dt, w_x, w_u]
//   EULER_DIFFLQR : [dt, drift_free, Fq, Fv, Fu, f0, Lxx, Lxu, Luu, lx, lu]  (difflqr_param_count)
//   ARM      : see arm.hpp (arm_param_count)
// per_problem_params != 0: params holds B consecutive blocks; else one shared block.
// tparams (may be NULL): parameter block(s) of the TERMINAL model, same layout / count / per-problem-ness as params
// (ShootingProblem(x0, running_models, terminal_model) with a terminal model of its own, shooting.hpp:48-63). NULL: the
// running model is aliased into the terminal node (arm: the running differential model with dt = 0, arm.hpp:89-90).
void* orc_batch_create2(int kind, int solver_kind, int B, int T, int nx, int nu, const double* params, long nparams,
                        int per_problem_params, const double* x0s, const double* tparams);
void* orc_batch_create(int kind, int solver_kind, int B, int T, int nx, int nu, const double* params, long nparams,
                       int per_problem_params, const double* x0s) {
  return orc_batch_create2(kind, solver_kind, B, T, nx, nu, params, nparams, per_problem_params, x0s, nullptr);
}
void* orc_batch_create2(int kind, int solver_kind, int B, int T, int nx, int nu, const double* params, long nparams,
                        int per_problem_params, const double* x0s, const double* tparams) {
  Batch* bt = new Batch();
  bt->kind = kind;
  bt->solver_kind = solver_kind;
  bt->B = B;
  bt->T = T;
  bt->nx = nx;
  bt->nu = nu;
  bt->converged.assign(B, 0);
  try {
    std::shared_ptr<ActionModel> shared_run, shared_term;
    for (int b = 0; b < B; ++b) {
      const double* p = params + (per_problem_params ? (size_t)b * nparams : 0);
      std::shared_ptr<ActionModel> run, term;
      if (b == 0 || per_problem_params) {
        if (kind == KIND_LQR) {
          if ((size_t)nparams != lqr_param_count(nx, nu)) throw std::invalid_argument("LQR params: wrong count");
          run = term = make_lqr(nx, nu, p);
          if (tparams) term = make_lqr(nx, nu, tparams + (per_problem_params ? (size_t)b * nparams : 0));
        } else if (kind == KIND_UNICYCLE) {
          if (nparams != 3) throw std::invalid_argument("unicycle params: wrong count");
          std::shared_ptr<ActionModelUnicycle> u = std::make_shared<ActionModelUnicycle>();
          u->dt = p[0];
          u->cost_weights[0] = p[1];
          u->cost_weights[1] = p[2];
          run = term = u;
          if (tparams) {
            const double* q = tparams + (per_problem_params ? (size_t)b * nparams : 0);
            std::shared_ptr<ActionModelUnicycle> ut = std::make_shared<ActionModelUnicycle>();
            ut->dt = q[0];
            ut->cost_weights[0] = q[1];
            ut->cost_weights[1] = q[2];
            term = ut;
          }
        } else if (kind == KIND_EULER_DIFFLQR) {
          if (nx % 2 || (size_t)nparams != difflqr_param_count(nx, nu)) throw std::invalid_argument("DiffLQR params: wrong count");
          run = term = make_euler_difflqr(nx, nu, p);
          if (tparams) term = make_euler_difflqr(nx, nu, tparams + (per_problem_params ? (size_t)b * nparams : 0));
        } else if (kind == KIND_ARM) {
          make_arm_models(p, nparams, nx, nu, run, term);
          if (tparams) {  // the terminal block is used as given (its own dt, weights, frame reference)
            std::shared_ptr<ActionModel> unused;
            make_arm_models(tparams + (per_problem_params ? (size_t)b * nparams : 0), nparams, nx, nu, term, unused);
          }
        } else {
          throw std::invalid_argument("unknown model kind");
        }
        shared_run = run;
        shared_term = term;
      } else {
        run = shared_run;
        term = shared_term;
      }
      Vec x0(x0s + (size_t)b * nx, x0s + (size_t)(b + 1) * nx);
      std::vector<std::shared_ptr<ActionModel> > running(T, run);
      std::shared_ptr<ShootingProblem> prob = std::make_shared<ShootingProblem>(x0, running, term);
      bt->problems.push_back(prob);
      if (solver_kind == SOLVER_FDDP)
        bt->solvers.push_back(std::make_shared<SolverFDDP>(prob));
      else if (solver_kind == SOLVER_BOXDDP)
        bt->solvers.push_back(std::make_shared<SolverBoxDDP>(prob));
      else if (solver_kind == SOLVER_BOXFDDP)
        bt->solvers.push_back(std::make_shared<SolverBoxFDDP>(prob));
      else
        bt->solvers.push_back(std::make_shared<SolverDDP>(prob));
    }
  } catch (std::exception& e) {
    std::fprintf(stderr, "orc_batch_create: %s\n", e.what());
    delete bt;
    return nullptr;
  }
  return bt;
}

void orc_batch_destroy(void* h) { delete (Batch*)h; }

// ActionModelAbstract::set_u_lb / set_u_ub (action-base.hxx:117-140) on the running model(s) of every problem.
// lb/ub: nu doubles shared by all problems (per_problem == 0) or B blocks. +-inf = unbounded.
int orc_batch_set_u_bounds(void* h, const double* lb, const double* ub, int per_problem) {
  Batch* bt = (Batch*)h;
  const int nu = bt->nu;
  for (int b = 0; b < bt->B; ++b) {
    const double* l = lb + (per_problem ? (size_t)b * nu : 0);
    const double* u = ub + (per_problem ? (size_t)b * nu : 0);
    ShootingProblem& p = *bt->problems[b];
    if (per_problem && b > 0 && p.running_models[0] == bt->problems[0]->running_models[0]) return -1;  // shared model
    ActionModel* last = nullptr;
    for (auto& m : p.running_models) {
      if (m.get() == last) continue;
      last = m.get();
      m->set_u_lb(Vec(l, l + nu));
      m->set_u_ub(Vec(u, u + nu));
    }
  }
  return 0;
}

// BoxQP::solve (box-qp.cpp:52-186) stand-alone: H (n x n column-major), q, lb, ub, xinit -> x, Hff_inv (nf x nf
// column-major, nf returned through *nf_out), free/clamped flags (clamped[i] = 1). Returns the number of Newton
// iterations, or -1 on "backward_error".
int orc_boxqp_solve(int n, const double* H, const double* q, const double* lb, const double* ub, const double* xinit,
                    int maxiter, double th_acceptstep, double th_grad, double reg, double* x_out, double* Hff_inv_out,
                    int* nf_out, int* clamped_out) {
  BoxQP qp(n, maxiter, th_acceptstep, th_grad, reg);
  Mat Hm(n, n);
  std::memcpy(Hm.d.data(), H, sizeof(double) * n * n);
  try {
    const BoxQPSolution& sol = qp.solve(Hm, Vec(q, q + n), Vec(lb, lb + n), Vec(ub, ub + n), Vec(xinit, xinit + n));
    std::memcpy(x_out, sol.x.data(), sizeof(double) * n);
    if (Hff_inv_out) std::memcpy(Hff_inv_out, sol.Hff_inv.d.data(), sizeof(double) * sol.Hff_inv.d.size());
    if (nf_out) *nf_out = sol.Hff_inv.r;
    if (clamped_out) {
      for (int i = 0; i < n; ++i) clamped_out[i] = 0;
      for (int i : sol.clamped_idx) clamped_out[i] = 1;
    }
  } catch (SolveError&) {
    return -1;
  }
  return qp.iters;
}

// thresholds: name in {th_acceptstep, th_stop, th_grad, th_stepdec, th_stepinc, th_acceptnegstep, regfactor, regmin, regmax}
int orc_batch_set(void* h, const char* name, double v) {
  Batch* bt = (Batch*)h;
  std::string n(name);
  for (auto& s : bt->solvers) {
    if (n == "th_acceptstep") s->th_acceptstep = v;
    else if (n == "th_stop") s->th_stop = v;
    else if (n == "th_grad") s->th_grad = v;
    else if (n == "th_stepdec") s->th_stepdec = v;
    else if (n == "th_stepinc") s->th_stepinc = v;
    else if (n == "regfactor") s->regfactor = v;
    else if (n == "regmin") s->regmin = v;
    else if (n == "regmax") s->regmax = v;
    else if (n == "xreg") { s->xreg = v; }
    else if (n == "ureg") { s->ureg = v; }
    else if (n == "th_acceptnegstep") {
      SolverFDDP* f = dynamic_cast<SolverFDDP*>(s.get());
      if (!f) return -1;
      f->th_acceptnegstep = v;
    } else return -1;
  }
  return 0;
}

// SolverDDP::set_alphas (ddp.cpp:444-462): the list replaces alphas_; thresholds are left alone
int orc_batch_set_alphas(void* h, const double* alphas, int n) {
  Batch* bt = (Batch*)h;
  if (n <= 0) return -1;
  for (auto& s : bt->solvers) s->alphas.assign(alphas, alphas + n);
  return 0;
}

// init_xs: B*(T+1)*nx or NULL (=> state zero, solver-base.cpp:46-50); init_us: B*T*nu or NULL (zeros).
// Returns number of converged problems; converged[b] readable via orc_batch_scalars. seconds_out (optional)
// receives the CLOCK_MONOTONIC wall time of the parallel region (reference clock: core/utils/timer.hpp:17-43).
int orc_batch_solve(void* h, const double* init_xs, const double* init_us, int maxiter, int is_feasible,
                    double reginit, int nthreads, double* seconds_out) {
  Batch* bt = (Batch*)h;
  const int B = bt->B, T = bt->T, nx = bt->nx, nu = bt->nu;
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
#pragma omp parallel for schedule(dynamic) num_threads(nthreads)
  for (int b = 0; b < B; ++b) {
    std::vector<Vec> xs = unpack(init_xs ? init_xs + (size_t)b * (T + 1) * nx : nullptr, T + 1, nx);
    std::vector<Vec> us = unpack(init_us ? init_us + (size_t)b * T * nu : nullptr, T, nu);
    bt->converged[b] = bt->solvers[b]->solve(xs, us, maxiter, is_feasible != 0, reginit) ? 1 : 0;
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  if (seconds_out) *seconds_out = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
  int n = 0;
  for (int b = 0; b < B; ++b) n += bt->converged[b];
  return n;
}

int orc_max_threads() { return omp_get_max_threads(); }

// ---- step-wise API (mirrors what unittest/bindings/test_solvers.py drives on the reference) ----
void orc_batch_set_candidate(void* h, const double* xs, const double* us, int is_feasible) {
  Batch* bt = (Batch*)h;
  const int T = bt->T, nx = bt->nx, nu = bt->nu;
  for (int b = 0; b < bt->B; ++b) {
    SolverDDP& s = *bt->solvers[b];
    s.xs_try[0] = bt->problems[b]->x0;
    s.setCandidate(unpack(xs ? xs + (size_t)b * (T + 1) * nx : nullptr, T + 1, nx),
                   unpack(us ? us + (size_t)b * T * nu : nullptr, T, nu), is_feasible != 0);
    s.was_feasible = false;
    s.iter = 0;
  }
}
// problem.calc / calcDiff at the current candidate; costs[b] out (may be NULL)
void orc_batch_calc(void* h, double* costs) {
  Batch* bt = (Batch*)h;
  for (int b = 0; b < bt->B; ++b) {
    double c = bt->problems[b]->calc(bt->solvers[b]->xs, bt->solvers[b]->us);
    if (costs) costs[b] = c;
  }
}
void orc_batch_calcdiff(void* h, double* costs) {  // SolverDDP::calcDiff (calc at iter 0 + calcDiff + gaps)
  Batch* bt = (Batch*)h;
  for (int b = 0; b < bt->B; ++b) {
    double c = bt->solvers[b]->calcDiff();
    if (costs) costs[b] = c;
  }
}
// status[b] = 0 ok, 1 backward_error
void orc_batch_backward(void* h, int* status) {
  Batch* bt = (Batch*)h;
  for (int b = 0; b < bt->B; ++b) {
    int st = 0;
    try {
      bt->solvers[b]->backwardPass();
    } catch (SolveError&) {
      st = 1;
    }
    if (status) status[b] = st;
  }
}
// dV[b] = tryStep(alpha); status[b]=1 on forward_error
void orc_batch_try_step(void* h, double alpha, double* dV, int* status) {
  Batch* bt = (Batch*)h;
  for (int b = 0; b < bt->B; ++b) {
    int st = 0;
    double v = NAN;
    try {
      v = bt->solvers[b]->tryStep(alpha);
    } catch (SolveError&) {
      st = 1;
    }
    if (dV) dV[b] = v;
    if (status) status[b] = st;
  }
}
// d[2*b..] = expectedImprovement() (FDDP: updateExpectedImprovement() first), stop[b] = stoppingCriteria()
void orc_batch_expected_improvement(void* h, double* d, double* stop) {
  Batch* bt = (Batch*)h;
  for (int b = 0; b < bt->B; ++b) {
    SolverDDP& s = *bt->solvers[b];
    SolverFDDP* f = dynamic_cast<SolverFDDP*>(&s);
    if (f) f->updateExpectedImprovement();
    const double* dd = s.expectedImprovement();
    if (d) {
      d[2 * b] = dd[0];
      d[2 * b + 1] = dd[1];
    }
    if (stop) stop[b] = s.stoppingCriteria();
  }
}

// ---- getters ----
// field at node t of problem b, column-major; returns number of doubles written or -1
int orc_get(void* h, const char* name, int b, int t, double* out) {
  Batch* bt = (Batch*)h;
  if (b < 0 || b >= bt->B) return -1;
  SolverDDP& s = *bt->solvers[b];
  ShootingProblem& p = *bt->problems[b];
  const int T = bt->T;
  std::string n(name);
  if (t < 0 || t > T) return -1;
  if (n == "xs") return copy_vec(s.xs[t], out);
  if (n == "xs_try") return copy_vec(s.xs_try[t], out);
  if (n == "fs") return copy_vec(s.fs[t], out);
  if (n == "Vx") return copy_vec(s.Vx[t], out);
  if (n == "Vxx") return copy_mat(s.Vxx[t], out);
  ActionData& d = (t == T) ? *p.terminal_data : *p.running_datas[t];
  if (n == "Fx") return copy_mat(d.Fx, out);
  if (n == "Fu") return copy_mat(d.Fu, out);
  if (n == "Lx") return copy_vec(d.Lx, out);
  if (n == "Lu") return copy_vec(d.Lu, out);
  if (n == "Lxx") return copy_mat(d.Lxx, out);
  if (n == "Lxu") return copy_mat(d.Lxu, out);
  if (n == "Luu") return copy_mat(d.Luu, out);
  if (n == "xnext") return copy_vec(d.xnext, out);
  if (n == "node_cost") {
    out[0] = d.cost;
    return 1;
  }
  if (t == T) return -1;
  if (n == "us") return copy_vec(s.us[t], out);
  if (n == "us_try") return copy_vec(s.us_try[t], out);
  if (n == "K") return copy_mat(s.K[t], out);
  if (n == "k") return copy_vec(s.k[t], out);
  if (n == "Qxx") return copy_mat(s.Qxx[t], out);
  if (n == "Qxu") return copy_mat(s.Qxu[t], out);
  if (n == "Quu") return copy_mat(s.Quu[t], out);
  if (n == "Qx") return copy_vec(s.Qx[t], out);
  if (n == "Qu") return copy_vec(s.Qu[t], out);
  if (n == "Quuk") return copy_vec(s.Quuk[t], out);
  return -1;
}

// out[16*b + i]: 0 cost 1 stop 2 d0 3 d1 4 xreg 5 ureg 6 steplength 7 dV 8 dVexp 9 iter 10 is_feasible
//                11 converged 12 n_records 13 cost_try 14 was_feasible 15 reserved
void orc_batch_scalars(void* h, double* out) {
  Batch* bt = (Batch*)h;
  for (int b = 0; b < bt->B; ++b) {
    SolverDDP& s = *bt->solvers[b];
    double* o = out + 16 * (size_t)b;
    o[0] = s.cost;
    o[1] = s.stop;
    o[2] = s.d[0];
    o[3] = s.d[1];
    o[4] = s.xreg;
    o[5] = s.ureg;
    o[6] = s.steplength;
    o[7] = s.dV;
    o[8] = s.dVexp;
    o[9] = s.iter;
    o[10] = s.is_feasible;
    o[11] = bt->converged[b];
    o[12] = (double)s.trace.size();
    o[13] = s.cost_try;
    o[14] = s.was_feasible;
    o[15] = 0;
  }
}

// per-iteration records of problem b: out[13*i + j], j = iter cost stop d0 d1 xreg ureg step feas dV dVexp
// alpha_idx retries; returns the number of records (call with out=NULL to query)
int orc_get_trace(void* h, int b, double* out, int max_records) {
  Batch* bt = (Batch*)h;
  SolverDDP& s = *bt->solvers[b];
  const int n = (int)s.trace.size();
  if (!out) return n;
  for (int i = 0; i < n && i < max_records; ++i) {
    const IterRecord& r = s.trace[i];
    double* o = out + 13 * (size_t)i;
    o[0] = r.iter;
    o[1] = r.cost;
    o[2] = r.stop;
    o[3] = r.d0;
    o[4] = r.d1;
    o[5] = r.xreg;
    o[6] = r.ureg;
    o[7] = r.steplength;
    o[8] = r.is_feasible;
    o[9] = r.dV;
    o[10] = r.dVexp;
    o[11] = s.trace_alpha_idx[i];
    o[12] = s.trace_retries[i];
  }
  return n;
}

// Independent cross-check (unittest/test_solvers.cpp:65-110): solve problem b with the dense KKT solver
// from the given warm start; writes xs ((T+1)*nx) and us (T*nu). Returns 1 if converged.
int orc_kkt_solve(void* h, int b, const double* init_xs, const double* init_us, int maxiter, double* xs_out,
                  double* us_out) {
  Batch* bt = (Batch*)h;
  const int T = bt->T, nx = bt->nx, nu = bt->nu;
  SolverKKT kkt(bt->problems[b]);
  bool ok = false;
  try {
    ok = kkt.solve(unpack(init_xs, T + 1, nx), unpack(init_us, T, nu), maxiter, false);
  } catch (std::exception& e) {
    std::fprintf(stderr, "orc_kkt_solve: %s\n", e.what());
    return -1;
  }
  for (int t = 0; t <= T; ++t) std::memcpy(xs_out + (size_t)t * nx, kkt.xs[t].data(), nx * sizeof(double));
  for (int t = 0; t < T; ++t) std::memcpy(us_out + (size_t)t * nu, kkt.us[t].data(), nu * sizeof(double));
  return ok ? 1 : 0;
}

// Finite-difference check of the node model of problem b at (x,u): fills the analytic and the
// numerically differentiated [Fx, Fu, Lx, Lu] (method of core/numdiff/action.hxx:48-105, disturbance
// sqrt(2*eps)). Layout of each out buffer: Fx(ndx*ndx) Fu(ndx*nu) Lx(ndx) Lu(nu), column-major.
int orc_numdiff(void* h, int b, int terminal, const double* x, const double* u, double* analytic, double* numeric) {
  Batch* bt = (Batch*)h;
  ShootingProblem& p = *bt->problems[b];
  ActionModel& m = terminal ? *p.terminal_model : *p.running_models[0];
  const int nx = m.nx, nu = m.nu;
  Vec xv(x, x + nx), uv(u, u + nu);
  std::shared_ptr<ActionData> d0 = m.createData();
  m.calc(*d0, xv, uv);
  m.calcDiff(*d0, xv, uv);
  double* a = analytic;
  a += copy_mat(d0->Fx, a);
  a += copy_mat(d0->Fu, a);
  a += copy_vec(d0->Lx, a);
  a += copy_vec(d0->Lu, a);
  const double h_ = std::sqrt(2.0 * std::numeric_limits<double>::epsilon());
  std::shared_ptr<ActionData> d1 = m.createData();
  double* nFx = numeric;
  double* nFu = nFx + nx * nx;
  double* nLx = nFu + nx * nu;
  double* nLu = nLx + nx;
  for (int j = 0; j < nx; ++j) {
    Vec xp = xv;
    xp[j] += h_;
    m.calc(*d1, xp, uv);
    for (int i = 0; i < nx; ++i) nFx[j * nx + i] = (d1->xnext[i] - d0->xnext[i]) / h_;
    nLx[j] = (d1->cost - d0->cost) / h_;
  }
  for (int j = 0; j < nu; ++j) {
    Vec up = uv;
    up[j] += h_;
    m.calc(*d1, xv, up);
    for (int i = 0; i < nx; ++i) nFu[j * nx + i] = (d1->xnext[i] - d0->xnext[i]) / h_;
    nLu[j] = (d1->cost - d0->cost) / h_;
  }
  return nx * nx + nx * nu + nx + nu;
}

}  // extern "C"

// Rigid-body primitive of the arm oracle by name (for the self-consistency tests of the restated
// Pinocchio algorithms); see arm.hpp arm_op.
extern "C" int orc_arm_op(const double* params, long nparams, const char* op, const double* in, double* out) {
  try {
    return oracle::arm_op(params, nparams, op, in, out);
  } catch (std::exception& e) {
    std::fprintf(stderr, "orc_arm_op: %s\n", e.what());
    return -1;
  }
}
