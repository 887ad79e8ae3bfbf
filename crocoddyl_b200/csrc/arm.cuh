// crocoddyl_b200 — node model of the CTA-per-problem core for the fixed-base revolute-chain arm:
// IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics(StateMultibody, ActuationModelFull,
// CostModelSum{gripperPose, uReg, xReg}), dt) for the running nodes and the same differential model with dt = 0 for
// the terminal node (reference: benchmark/factory/arm.hpp:60-90, include/crocoddyl/core/integrator/euler.hxx:41-131,
// include/crocoddyl/multibody/actions/free-fwddyn.hxx:44-169). The arithmetic of one node is in rigid_body.cuh; this
// file maps nodes onto the threads of the block that owns the problem:
//   calc_all / calcdiff_all : one node per thread (ShootingProblem::calc/calcDiff, shooting.hxx:133-195 — the
//                             reference's OpenMP loop over t becomes the thread index);
//   calc_block              : the serial rollout step (fddp.cpp:168,200) — ABA is a 3-sweep recursion over the joints.
// Node derivatives vary along the horizon, so Fx, Fu, Lxx, Luu live per node in HBM ([B][T+1][.], column-major);
// Lxu is identically zero for this cost set and is neither stored per node nor read by the backward pass.
#pragma once
#include "cta.cuh"
#include "rigid_body.cuh"
#include "arm_warp.cuh"

// Block shape of the arm kernels. The backward pass and the rollout are one warp's work, so extra warps only wait at
// barriers; measured on B200 (B = 8192, T = 100 FDDP): 128 thr x 4 CTAs/SM 1.50 s, 64 x 8 1.17 s, 32 x 16 1.23 s (128
// registers: spills), 32 x 8 1.13 s <- default.
#ifndef CGX_ARM_MINBLOCKS
#define CGX_ARM_MINBLOCKS 8
#endif
#ifndef CGX_ARM_DIR_MINBLOCKS
#define CGX_ARM_DIR_MINBLOCKS 8
#endif
#ifndef CGX_ARM_FWD_MINBLOCKS
#define CGX_ARM_FWD_MINBLOCKS 8
#endif
#ifndef CGX_ARM_THREADS
#define CGX_ARM_THREADS 32
#endif

namespace cgx {

template <int NV>
__device__ __noinline__ double arm_calc_node_dev(const double* params, bool terminal, const double* x, const double* u,
                                                  double* xnext) {
  return rb::arm_calc_node<NV>(rb::ArmParams{params}, terminal, x, u, xnext, nullptr);
}
template <int NV>
__device__ __noinline__ void arm_calcdiff_node_dev(const double* params, bool terminal, const double* x, const double* u,
                                                    double* Fx, double* Fu, double* Lx, double* Lu, double* Lxx,
                                                    double* Luu) {
  rb::arm_calcdiff_node<NV>(rb::ArmParams{params}, terminal, x, u, Fx, Fu, Lx, Lu, Lxx, Luu);
}

template <int NV>
__global__ void arm_quasi_static_kernel(const double* params, long nparams, int per_problem, int B, const double* xs,
                                        double* us) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double x[2 * NV], u[NV];
  for (int i = 0; i < 2 * NV; ++i) x[i] = xs[(size_t)b * 2 * NV + i];
  rb::arm_quasi_static_node<NV>(rb::ArmParams{params + (per_problem ? (size_t)b * nparams : 0)}, x, u);
  for (int i = 0; i < NV; ++i) us[(size_t)b * NV + i] = u[i];
}

template <int NV>
struct ArmModel {
  static constexpr bool kPerNodeMats = true;
  // solve() as lockstep phase launches instead of the fused per-problem loop (cta.cuh, measured 2x faster: i-cache)
  static constexpr bool kPhaseSplit = true;
  // calc/calcDiff of the phase-split solve run as ONE launch over all (problem, node) pairs of the batch
  static constexpr bool kNodeKernel = true;
  // the solve kernel is latency-bound (serial recursions over joints and nodes): resident blocks per SM matter more
  // than registers per thread, so the register file is split over CGX_ARM_MINBLOCKS blocks of 128 threads
  static constexpr int kMinThreads = CGX_ARM_THREADS, kMaxThreads = CGX_ARM_THREADS, kMinBlocks = CGX_ARM_MINBLOCKS;
  // resident blocks per SM of the phase-split kernels (backward pass only / one line-search trial)
  static constexpr int kDirMinBlocks = CGX_ARM_DIR_MINBLOCKS, kFwdMinBlocks = CGX_ARM_FWD_MINBLOCKS;
  static constexpr int kN = 2 * NV, kM = NV;  // compile-time node sizes for the warp-level backward pass
  const double* params;
  const double* tparams;  // terminal model's block (default: the running block with dt = 0)

  // host side: device address of a derivative matrix of node t of problem b
  template <class PD>
  static const double* node_matrix(const CtaCtx& c, const PD*, const std::string& name, int b, int t) {
    const size_t n = c.n, m = c.m, node = (size_t)b * (c.T + 1) + t;
    if (name == "Fx") return c.d_Fx + node * n * n;
    if (name == "Fu") return c.d_Fu + node * n * m;
    if (name == "Lxx") return c.d_Lxx + node * n * n;
    if (name == "Lxu") return c.d_Lxu;  // one zero block shared by every node
    return c.d_Luu + node * m * m;
  }
  template <class BE>
  static void quasi_static(BE& be, const CtaCtx& c, const double* d_x, double* d_u, cudaStream_t stream) {
    arm_quasi_static_kernel<NV><<<(c.B + 127) / 128, 128, 0, stream>>>(c.params, c.nparams, c.per_problem, c.B, d_x, d_u);
    ++be.last_launches;
  }
  __device__ void load(const CtaCtx& c, int b) {
    const size_t off = c.per_problem ? (size_t)b * c.nparams : 0;
    params = c.params + off;
    tparams = c.tparams + off;
  }
  __device__ NodeMats mats(const CtaCtx& c, int b, int t) const {
    const size_t n = c.n, m = c.m, node = (size_t)b * (c.T + 1) + t;
    return NodeMats{c.d_Fx + node * n * n, c.d_Fu + node * n * m, c.d_Lxx + node * n * n, nullptr,
                    c.d_Luu + node * m * m, false};
  }
  // one node by the calling thread (xs, us: this problem's trajectories)
  __device__ void calc_node(const CtaCtx& c, int b, int t, const double* xs, const double* us) const {
    const int T = c.T;
    const size_t node = (size_t)b * (T + 1) + t;
    double x[2 * NV], u[NV], xn[2 * NV];
#pragma unroll
    for (int i = 0; i < 2 * NV; ++i) x[i] = xs[(size_t)t * 2 * NV + i];
    if (t < T) {
#pragma unroll
      for (int i = 0; i < NV; ++i) u[i] = us[(size_t)t * NV + i];
    }
    c.d_cost[node] = arm_calc_node_dev<NV>(t == T ? tparams : params, t == T, x, t < T ? u : nullptr, xn);
#pragma unroll
    for (int i = 0; i < 2 * NV; ++i) c.d_xnext[node * 2 * NV + i] = xn[i];
  }
  __device__ void calcdiff_node(const CtaCtx& c, int b, int t, const double* xs, const double* us) const {
    const int T = c.T;
    constexpr size_t n = 2 * NV, m = NV;
    const size_t node = (size_t)b * (T + 1) + t;
    double x[2 * NV], u[NV];
#pragma unroll
    for (int i = 0; i < 2 * NV; ++i) x[i] = xs[(size_t)t * 2 * NV + i];
    if (t < T) {
#pragma unroll
      for (int i = 0; i < NV; ++i) u[i] = us[(size_t)t * NV + i];
    }
    arm_calcdiff_node_dev<NV>(t == T ? tparams : params, t == T, x, t < T ? u : nullptr, c.d_Fx + node * n * n, c.d_Fu + node * n * m,
                              c.d_Lx + node * n, c.d_Lu + node * m, c.d_Lxx + node * n * n, c.d_Luu + node * m * m);
  }
  __device__ void calc_all(const CtaCtx& c, int b, const double* xs, const double* us) const {
    for (int t = threadIdx.x; t <= c.T; t += blockDim.x) calc_node(c, b, t, xs, us);
  }
  __device__ void calcdiff_all(const CtaCtx& c, int b, const double* xs, const double* us) const {
    for (int t = threadIdx.x; t <= c.T; t += blockDim.x) calcdiff_node(c, b, t, xs, us);
  }
  // calc of one node by the calling thread (thread-per-problem rollout of large batches, cta_forward_thread)
  __device__ double calc_thread(const double* x, const double* u, double* xnext) const {
    return arm_calc_node_dev<NV>(u == nullptr ? tparams : params, u == nullptr, x, u, xnext);
  }
  // calc of one node for the rollout: x, u (null at the terminal node) and xnext in shared memory; warp 0 evaluates
  // it cooperatively (arm_warp.cuh), the other warps of the block wait
  __device__ void calc_block(const double* x, const double* u, double* xnext, double*, double* cost_out) const {
    __shared__ ArmWarpSmem<NV> W;
    if (threadIdx.x < 32) arm_calc_warp<NV>(u == nullptr ? tparams : params, u == nullptr, x, u, xnext, cost_out, W);
    __syncthreads();
  }
};

}  // namespace cgx
