// Test program for the C++ facade (include/crocoddyl_b200.hpp): the flow of benchmark/unicycle-optctrl.cpp:25-47 —
// build the problem, solve from xs = {x0}, us = {0} — for one problem, printing what the pytest side compares with
// the oracle. Without a CUDA device the library must refuse (no CPU fallback): prints the exception and exits 3.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "crocoddyl_b200.hpp"

namespace cb = crocoddyl_b200;

int main(int argc, char** argv) {
  const int T = argc > 1 ? atoi(argv[1]) : 20;
  const bool lqr = argc > 2 && !strcmp(argv[2], "lqr");
  try {
    cb::ActionModel model = lqr ? cb::ActionModelLQR(6, 2, true) : cb::ActionModelUnicycle();
    std::vector<double> x0(model.nx, 0.);
    if (!lqr) {
      x0[0] = -1.;
      x0[1] = -1.;
      x0[2] = 1.;
    }
    auto problem = std::make_shared<cb::ShootingProblem>(x0, (std::size_t)T, model);
    cb::SolverFDDP ddp(problem);
    std::vector<double> xs, us((std::size_t)T * model.nu, 0.);
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
    return 0;
  } catch (const cb::Exception& e) {
    printf("exception %d: %s\n", e.code(), e.what());
    return e.code() == CGX_ERR_CUDA ? 3 : 5;
  }
}
