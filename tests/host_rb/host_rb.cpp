// TEST INFRASTRUCTURE — host instantiation of the product's rigid-body arithmetic (crocoddyl_b200/csrc/rigid_body.cuh
// is __host__ __device__), so that it can be compared with the oracle on the CPU-only box. Not a product path: the
// product library never calls these on the host.
#include "../../crocoddyl_b200/csrc/rigid_body.cuh"

using namespace cgx::rb;

template <int NV>
static int run(const char op, const double* params, int terminal, const double* x, const double* u, double* out) {
  ArmParams P{params};
  constexpr int n = 2 * NV;
  if (op == 'c') {  // calc: out = [cost, xnext(n), qdd(NV)]
    out[0] = arm_calc_node<NV>(P, terminal != 0, x, u, out + 1, out + 1 + n);
    return 1 + n + NV;
  }
  if (op == 'd') {  // calcDiff: out = [Fx(n*n), Fu(n*NV), Lx(n), Lu(NV), Lxx(n*n), Luu(NV*NV)]
    double* Fx = out;
    double* Fu = Fx + n * n;
    double* Lx = Fu + n * NV;
    double* Lu = Lx + n;
    double* Lxx = Lu + NV;
    double* Luu = Lxx + n * n;
    arm_calcdiff_node<NV>(P, terminal != 0, x, u, Fx, Fu, Lx, Lu, Lxx, Luu);
    return 2 * n * n + n * NV + n + NV + NV * NV;
  }
  if (op == 'q') {  // quasiStatic: out = u(NV)
    arm_quasi_static_node<NV>(P, x, out);
    return NV;
  }
  return -1;
}

extern "C" int host_rb_eval(int nv, char op, const double* params, int terminal, const double* x, const double* u,
                            double* out) {
  switch (nv) {
    case 1: return run<1>(op, params, terminal, x, u, out);
    case 2: return run<2>(op, params, terminal, x, u, out);
    case 3: return run<3>(op, params, terminal, x, u, out);
    case 4: return run<4>(op, params, terminal, x, u, out);
    case 6: return run<6>(op, params, terminal, x, u, out);
    case 7: return run<7>(op, params, terminal, x, u, out);
    default: return -2;
  }
}
