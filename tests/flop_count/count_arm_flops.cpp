// Test infrastructure (includes oracle/): FP64 operation counts of one arm node's calc / calcDiff, obtained by running the
// oracle's rigid-body algorithms (oracle/rbd.hpp, templated on the scalar) on a COUNTING scalar — every + - * / sqrt and
// trig call that the published algorithms (Featherstone ABA / RNEA, Carpentier-Mansard derivatives, log6 / Jlog6) execute
// on the synthetic 7-DoF arm at a generic state. SURVEY 8(d) asks for "ABA flops" per node without giving a closed form;
// this is the algorithmic count (the straightforward statement, no sparsity exploited beyond what the algorithm itself
// has), as opposed to tools/count_fp64_ops.py which counts what the compiled kernel issues.
//   g++ -std=c++17 -O1 -I oracle -o tests/flop_count/_build/count tests/flop_count/count_arm_flops.cpp && ./count params.bin
// Output: one JSON object on stdout.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

struct Ops {
  long add = 0, mul = 0, div = 0, sqrt_ = 0, trig = 0, cmp = 0;
  long flops() const { return add + mul + div + sqrt_ + trig; }
};
static Ops g_ops;

struct Cnt {
  double v;
  Cnt() : v(0.) {}
  Cnt(double x) : v(x) {}
  Cnt(int x) : v(x) {}
  explicit operator double() const { return v; }
  Cnt& operator+=(const Cnt& o) { ++g_ops.add; v += o.v; return *this; }
  Cnt& operator-=(const Cnt& o) { ++g_ops.add; v -= o.v; return *this; }
  Cnt& operator*=(const Cnt& o) { ++g_ops.mul; v *= o.v; return *this; }
  Cnt& operator/=(const Cnt& o) { ++g_ops.div; v /= o.v; return *this; }
};
inline Cnt operator+(const Cnt& a, const Cnt& b) { ++g_ops.add; return Cnt(a.v + b.v); }
inline Cnt operator-(const Cnt& a, const Cnt& b) { ++g_ops.add; return Cnt(a.v - b.v); }
inline Cnt operator*(const Cnt& a, const Cnt& b) { ++g_ops.mul; return Cnt(a.v * b.v); }
inline Cnt operator/(const Cnt& a, const Cnt& b) { ++g_ops.div; return Cnt(a.v / b.v); }
inline Cnt operator-(const Cnt& a) { return Cnt(-a.v); }  // sign flip: not counted
inline bool operator<(const Cnt& a, const Cnt& b) { ++g_ops.cmp; return a.v < b.v; }
inline bool operator>(const Cnt& a, const Cnt& b) { ++g_ops.cmp; return a.v > b.v; }
inline bool operator<=(const Cnt& a, const Cnt& b) { ++g_ops.cmp; return a.v <= b.v; }
inline bool operator>=(const Cnt& a, const Cnt& b) { ++g_ops.cmp; return a.v >= b.v; }
inline bool operator==(const Cnt& a, const Cnt& b) { ++g_ops.cmp; return a.v == b.v; }
inline bool operator!=(const Cnt& a, const Cnt& b) { ++g_ops.cmp; return a.v != b.v; }
inline Cnt sqrt(const Cnt& a) { ++g_ops.sqrt_; return Cnt(std::sqrt(a.v)); }
inline Cnt sin(const Cnt& a) { ++g_ops.trig; return Cnt(std::sin(a.v)); }
inline Cnt cos(const Cnt& a) { ++g_ops.trig; return Cnt(std::cos(a.v)); }
inline Cnt acos(const Cnt& a) { ++g_ops.trig; return Cnt(std::acos(a.v)); }
inline Cnt fabs(const Cnt& a) { return Cnt(std::fabs(a.v)); }
inline Cnt abs(const Cnt& a) { return Cnt(std::fabs(a.v)); }

#include "arm.hpp"

using oracle::ArmEval;
using oracle::ArmParams;
namespace rbd = oracle::rbd;

static void print(const char* name, const Ops& o, bool last = false) {
  std::printf("  \"%s\": {\"add\": %ld, \"mul\": %ld, \"div\": %ld, \"sqrt\": %ld, \"trig\": %ld, \"flops\": %ld}%s\n", name, o.add,
              o.mul, o.div, o.sqrt_, o.trig, o.flops(), last ? "" : ",");
}

int main(int argc, char** argv) {
  if (argc < 3) {
    std::fprintf(stderr, "usage: count params.bin nv\n");
    return 2;
  }
  const int nv = std::atoi(argv[2]);
  std::vector<double> params(oracle::arm_param_count(nv));
  FILE* f = std::fopen(argv[1], "rb");
  if (!f || std::fread(params.data(), sizeof(double), params.size(), f) != params.size()) return 3;
  std::fclose(f);
  ArmParams<Cnt> P;
  P.parse(params.data(), (long)params.size(), nv);
  const int n = 2 * nv;
  std::vector<Cnt> x(n), u(nv), a(nv);
  for (int i = 0; i < nv; ++i) {
    x[i] = Cnt(0.3 + 0.17 * i);
    x[nv + i] = Cnt(-0.4 + 0.11 * i);
    u[i] = Cnt(0.9 - 0.23 * i);
  }
  Cnt r6[6];
  rbd::SE3<Cnt> rMf;
  std::printf("{\n  \"nv\": %d,\n", nv);

  // ---- calc: ABA + frame placement + log6 + the three cost terms (free-fwddyn.hxx:44-80), then the Euler step
  rbd::ChainData<Cnt> cd;
  g_ops = Ops();
  Cnt cost = ArmEval<Cnt>::calc(P, cd, x.data(), u.data(), a.data(), rMf, r6);
  (void)cost;
  Ops calc = g_ops;
  calc.mul += 3 * nv + 2;  // euler.hxx:65-69: v dt, a dt^2 (dt^2 once), a dt per joint; dt * cost
  calc.add += 3 * nv;      // dq = v dt + a dt^2 ; q + dq ; v + dv
  print("calc", calc);

  // ---- calcDiff, piece by piece
  g_ops = Ops();
  rbd::aba(P.chain, cd, x.data(), x.data() + nv, u.data(), a.data());
  print("aba_alone", g_ops);
  std::vector<Cnt> dq(nv * nv), dv(nv * nv), Minv(nv * nv);
  g_ops = Ops();
  rbd::abaDerivatives(P.chain, cd, x.data(), x.data() + nv, a.data(), dq.data(), dv.data(), Minv.data());
  Ops abad = g_ops;
  print("aba_derivatives", abad);
  Cnt rJf[36];
  g_ops = Ops();
  rbd::Jlog6(rMf, rJf);
  Ops jl = g_ops;
  print("jlog6", jl);
  rbd::aba(P.chain, cd, x.data(), x.data() + nv, u.data(), a.data());
  std::vector<Cnt> fJf(6 * nv);
  g_ops = Ops();
  rbd::frameJacobianLocal(P.chain, cd, fJf.data());
  Ops fj = g_ops;
  print("frame_jacobian_local", fj);
  // the assembly loops of calcDiff (oracle/arm.hpp ActionModelArm::calcDiff, cost-sum.hxx:140-158, euler.hxx:100-121)
  Ops as;
  as.mul += 36 * nv; as.add += 30 * nv;                  // J = Jlog6 * fJf           (6 x 6 x nv)
  as.mul += 6 * nv + nv; as.add += 5 * nv + nv;          // Lx += w J^T r
  as.mul += 6 * nv * nv + nv * nv; as.add += 5 * nv * nv + nv * nv;  // Lxx += w J^T J
  as.mul += nv + n; as.add += nv + nv + n + n;           // uReg, xReg
  as.mul += 6 * nv * nv + 1;                             // Fx, Fu blocks * dt, dt^2
  as.add += nv + n;                                      // + dt, + 1 on the diagonals
  as.mul += n + nv + n * n + nv * nv;                    // dt * (Lx, Lu, Lxx, Luu)
  print("calcdiff_assembly", as);
  Ops cdiff;
  for (const Ops* o : {&abad, &jl, &fj, &as}) {
    cdiff.add += o->add; cdiff.mul += o->mul; cdiff.div += o->div; cdiff.sqrt_ += o->sqrt_; cdiff.trig += o->trig;
  }
  print("calcdiff", cdiff, true);
  std::printf("}\n");
  return 0;
}
