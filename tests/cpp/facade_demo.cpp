// Test program for the C++ facade (include/crocoddyl_b200.hpp): the flow of benchmark/unicycle-optctrl.cpp:25-47 —
// model = make_shared<ActionModelUnicycle>(), runningModels(N, model), ShootingProblem(x0, runningModels, model),
// SolverFDDP(problem).solve(xs, us) — for one problem, printing what the pytest side compares with the oracle; then the
// node plugin interface (createData / calc / calcDiff on one node) and a control-limited solve (SolverBoxFDDP).
// Without a CUDA device the library must refuse (no CPU fallback): prints the exception and exits 3.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "crocoddyl_b200.hpp"

namespace cb = crocoddyl_b200;

int main(int argc, char** argv) {
  const int T = argc > 1 ? atoi(argv[1]) : 20;
  const bool difflqr = argc > 2 && !strcmp(argv[2], "difflqr");
  const bool lqr = difflqr || (argc > 2 && !strcmp(argv[2], "lqr"));
  try {
    std::shared_ptr<cb::ActionModelAbstract> model;
    if (difflqr)  // IntegratedActionModelEuler(DifferentialActionModelLQR(3, 2, drift), dt) on the LQR kernels
      model = cb::IntegratedActionModelEuler(cb::DifferentialActionModelLQR(3, 2, false), 0.05);
    else if (lqr)
      model = std::make_shared<cb::ActionModelLQR>(6, 2, true);
    else
      model = std::make_shared<cb::ActionModelUnicycle>();
    const std::size_t nx = model->get_state()->get_nx(), nu = model->get_nu();
    std::vector<double> x0(nx, 0.);
    if (!lqr) {
      x0[0] = -1.;
      x0[1] = -1.;
      x0[2] = 1.;
    }
    std::vector<std::shared_ptr<cb::ActionModelAbstract> > runningModels(T, model);
    auto problem = std::make_shared<cb::ShootingProblem>(x0, runningModels, model);
    cb::SolverFDDP ddp(problem);
    std::vector<double> xs, us((std::size_t)T * nu, 0.);
    for (int t = 0; t <= T; ++t) xs.insert(xs.end(), x0.begin(), x0.end());
    const bool ok = ddp.solve(xs, us, 100, false, 1e-9);
    printf("converged %d iter %zu cost %.17g stop %.17g feasible %d\n", (int)ok, ddp.get_iter(), ddp.get_cost(), ddp.get_stop(),
           (int)ddp.get_is_feasible());
    const std::vector<double> xo = ddp.get_xs(), uo = ddp.get_us(), K0 = ddp.get("K", 0, 0);
    printf("xs");
    for (double v : xo) printf(" %.17g", v);
    printf("\nus");
    for (double v : uo) printf(" %.17g", v);
    printf("\nK0");
    for (double v : K0) printf(" %.17g", v);
    printf("\n");
    // the fused solve + getters overload must hand back exactly what the getters hold
    {
      std::vector<double> xf, uf;
      const bool ok2 = ddp.solve(xs, us, xf, uf, 100, false, 1e-9);
      if (ok2 != ok || xf != xo || uf != uo || xf != ddp.get_xs()) {
        printf("solve_fetch mismatch\n");
        return 6;
      }
    }
    // argument checking mirrors the reference's throw_pretty messages
    try {
      ddp.solve(std::vector<double>(3, 0.), us);
      return 4;
    } catch (const cb::Exception& e) {
      printf("expected error: %s\n", e.what());
    }
    // node plugin interface: data = model->createData(); model->calc(data, x, u); model->calcDiff(data, x, u)
    {
      std::shared_ptr<cb::ActionDataAbstract> data = model->createData();
      std::vector<double> x(nx), u(nu);
      for (std::size_t i = 0; i < nx; ++i) x[i] = 0.3 + 0.1 * (double)i;
      for (std::size_t i = 0; i < nu; ++i) u[i] = -0.2 + 0.5 * (double)i;
      model->calc(data, x, u);
      model->calcDiff(data, x, u);
      printf("node cost %.17g\nnode_xnext", data->cost);
      for (double v : data->xnext) printf(" %.17g", v);
      printf("\nnode_Fx");
      for (double v : data->Fx) printf(" %.17g", v);
      printf("\nnode_Lu");
      for (double v : data->Lu) printf(" %.17g", v);
      printf("\n");
      model->calc(data, x);  // x-only overload (terminal node): u = unone
      printf("term cost %.17g\n", data->cost);
    }
    // control-limited: set_u_lb / set_u_ub on the model, SolverBoxFDDP
    {
      std::shared_ptr<cb::ActionModelAbstract> bm;
      if (difflqr)
        bm = cb::IntegratedActionModelEuler(cb::DifferentialActionModelLQR(3, 2, false), 0.05);
      else if (lqr)
        bm = std::make_shared<cb::ActionModelLQR>(6, 2, true);
      else
        bm = std::make_shared<cb::ActionModelUnicycle>();
      bm->set_u_lb(std::vector<double>(nu, -0.4));
      bm->set_u_ub(std::vector<double>(nu, 0.4));
      if (!bm->get_has_control_limits()) return 7;
      auto bp = std::make_shared<cb::ShootingProblem>(x0, std::vector<std::shared_ptr<cb::ActionModelAbstract> >(T, bm), bm);
      cb::SolverBoxFDDP box(bp);
      const bool okb = box.solve(xs, us, 100, false, 1e-9);
      printf("box converged %d iter %zu cost %.17g th_stop %.3g\nbox_us", (int)okb, box.get_iter(), box.get_cost(), box.get_th_stop());
      for (double v : box.get_us()) printf(" %.17g", v);
      printf("\n");
    }
    return 0;
  } catch (const cb::Exception& e) {
    printf("exception %d: %s\n", e.code(), e.what());
    return e.code() == CGX_ERR_CUDA ? 3 : 5;
  }
}
