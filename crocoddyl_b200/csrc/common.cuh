// crocoddyl_b200 — shared device/host definitions of the batched DDP/FDDP path.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stddef.h>
#include <stdint.h>

namespace cgx {

// Thresholds and constants of SolverAbstract / SolverDDP / SolverFDDP
// (reference: src/core/solver-base.cpp:14-25, src/core/solvers/ddp.cpp:17-36, fddp.cpp:15).
struct SolverOpts {
  double th_acceptstep, th_stop, th_grad, th_stepdec, th_stepinc, th_acceptnegstep;
  double regfactor, regmin, regmax;
  double alphas[16];
  int nalphas;
};

// per-problem scalar slots (double), stored as scal[slot * Bp + b]
enum {
  S_COST = 0, S_STOP, S_D0, S_D1, S_XREG, S_UREG, S_STEP, S_DV, S_DVEXP, S_COST_TRY, S_DG, S_DQ, S_DVGAP,
  S_LBXREG, S_LBUREG, S_FLXREG /* thread-per-problem core: regularisation of the last successful / first failed backward pass */,
  S_NSLOTS
};
// per-problem int slots, iscal[slot * Bp + b]
enum {
  I_ITER = 0, I_FEAS, I_WASFEAS, I_CONVERGED, I_CUR, I_NREC, I_BSTATUS, I_FSTATUS, I_NBWD, I_NFWD, I_FSBUF,
  I_LOOP /* resumable solve loop: bit 0 recalcDiff, bit 1 gaps must be recomputed */, I_FINISHED,
  I_PHASE /* phase-split solve (cta.cuh): 0 computeDirection pending, 1 line search pending, 2 solve() returned */,
  I_ALPHA /* phase-split solve: index of the next step length to try */, I_RETRIES, I_NSLOTS
};

// raiseIfNaN (reference: src/core/solver-base.cpp:175-181)
__host__ __device__ inline bool raise_if_nan(double v) { return isnan(v) || isinf(v) || v >= 1e30; }

// increaseRegularization / decreaseRegularization (reference: src/core/solvers/ddp.cpp:312-326)
__host__ __device__ inline void reg_increase(const SolverOpts& o, double& xreg, double& ureg) {
  xreg *= o.regfactor;
  if (xreg > o.regmax) xreg = o.regmax;
  ureg = xreg;
}
__host__ __device__ inline void reg_decrease(const SolverOpts& o, double& xreg, double& ureg) {
  xreg /= o.regfactor;
  if (xreg < o.regmin) xreg = o.regmin;
  ureg = xreg;
}

// x / p in the LLT and the triangular solves of computeGains: through the reciprocal of the pivot (default build;
// <= 1 ulp from the division, 50.6 -> 38.8 ms on config C2) or as the IEEE division the reference performs
// (parity build: -DCGX_PARITY together with -fmad=false, crocoddyl_b200/build.py variant "parity").
__host__ __device__ inline double div_pivot(double x, double p, double pinv) {
#ifdef CGX_PARITY
  (void)pinv;
  return x / p;
#else
  (void)p;
  return x * pinv;
#endif
}

// Pivot of the unblocked LLT: p -> sqrt(p) and 1 / sqrt(p). The serial chain sqrt -> reciprocal (~2 x 10 dependent FP64
// instructions) sits on the critical path of every node of the backward sweeps; the default build takes the reciprocal
// square root directly (rsqrt, <= 1 ulp, the accuracy class of 1 / sqrt(p)) and forms sqrt(p) = p * rsqrt(p), which the
// default build only stores (all divisions by the pivot are multiplications by the reciprocal, div_pivot above). The
// parity build keeps the reference's sqrt and division.
__host__ __device__ inline void llt_pivot(double p, double& sp, double& pinv) {
#if defined(CGX_PARITY) || !defined(__CUDA_ARCH__)
  sp = sqrt(p);
  pinv = 1. / sp;
#else
  pinv = rsqrt(p);
  sp = p * pinv;
#endif
}

constexpr int NTRACE = 13;  // matches CGX_NTRACE

}  // namespace cgx
