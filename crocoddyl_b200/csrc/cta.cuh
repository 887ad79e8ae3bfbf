// crocoddyl_b200 — CTA-per-problem DDP/FDDP core for dense node models (LQR n <= ~72, revolute-chain arm).
//
// One thread block owns one OCP and runs the reference's whole solve() loop on the device; the block's
// threads cooperate on the dense algebra of each node (Vxx, Fx^T Vxx, Fx staged in shared memory, register-
// tiled DFMA products) and on the node-parallel parts (calc / calcDiff of all T+1 nodes at once).
// HBM layout = the host layout of the C ABI: per problem contiguous, [B][nt][size], matrices column-major,
// so a block streams its own problem with coalesced accesses and accessors are plain copies.
//
// Reference functions restated here (paths relative to the reference tree):
//   ShootingProblem::calc/calcDiff              include/crocoddyl/core/optctrl/shooting.hxx:133-195
//   SolverDDP::calcDiff/backwardPass/computeGains  src/core/solvers/ddp.cpp:157-253,298-310
//   SolverDDP::forwardPass / SolverFDDP::forwardPass  ddp.cpp:255-296, fddp.cpp:149-225
//   expectedImprovement / updateExpectedImprovement / stoppingCriteria  ddp.cpp:132-155, fddp.cpp:107-147
//   SolverDDP::solve / SolverFDDP::solve        ddp.cpp:41-118, fddp.cpp:19-105
#pragma once
#include <stdexcept>
#include <string>
#include "boxqp.cuh"
#include "common.cuh"

namespace cgx {

struct CtaCtx {
  int B, Bp, T, n, m;
  int fddp, full;
  int model_kind;         // 0 LQR, 2 ARM
  double* xs[2];          // [B][T+1][n] candidate / trial (I_CUR)
  double* us[2];          // [B][T][m]
  double* fs;             // [B][T+1][n]
  double* K;              // [B][T][m*n]
  double* k;              // [B][T][m]
  double* Vx;             // [B][T+1][n]
  double* Vxx;            // [B][T+1][n*n]
  double* w;              // [B][T+1][n]  Vxx fs
  double* Qxx;            // full only: [B][T][n*n]
  double* Qxu;            // [B][T][n*m]
  double* Quu;            // [B][T][m*m]
  double* Qx;             // [B][T][n]
  double* Qu;             // [B][T][m]
  // ActionData of every node (core/action-base.hpp:101-142)
  double* d_cost;         // [B][T+1]
  double* d_xnext;        // [B][T+1][n]
  double* d_Lx;           // [B][T+1][n]
  double* d_Lu;           // [B][T+1][m]
  double* d_Fx;           // per-node matrices, only for models whose derivatives vary along the horizon
  double* d_Fu;
  double* d_Lxx;
  double* d_Lxu;
  double* d_Luu;
  const double* params;   // model parameter block(s)
  const double* tparams;  // terminal model's block(s): same layout / count / per-problem-ness (shooting.hpp:48-63)
  const double* u_lb;     // [B][m] control limits of the running model (action-base.hxx:117-140)
  const double* u_ub;     // [B][m]
  const int* has_limits;  // [B]
  int box;                // 1 = SolverBoxDDP / SolverBoxFDDP (box-ddp.cpp, box-fddp.cpp)
  double* Quu_inv;        // box only: [B][T][m*m]
  double* box_work;       // box only: [B][box_gains_work_doubles(m)] scratch of the per-node box QP
  int* box_iwork;         // box only: [B][box_gains_work_ints(m)]
  long nparams;
  int per_problem;
  const double* x0;       // [B][n]
  double* scal;           // [S_NSLOTS][Bp]
  int* iscal;             // [I_NSLOTS][Bp]
  double* trace;          // [B][cap][NTRACE] or null
  int trace_cap;
  SolverOpts o;
};

// pointers to the derivative matrices of one node (column-major, leading dimension = rows)
struct NodeMats {
  const double* Fx;
  const double* Fu;
  const double* Lxx;
  const double* Lxu;
  const double* Luu;
  bool constant;  // the same for every running node of the problem (LQR): stage once
};

__device__ __forceinline__ int CTA_I_(const CtaCtx& c, int b, int slot) { return c.iscal[(size_t)slot * c.Bp + b]; }

}  // namespace cgx
#include "cta_small.cuh"
namespace cgx {

__device__ __forceinline__ int cta_ld(int rows) { return rows | 1; }  // odd leading dimension: no bank conflicts

// ---------------------------------------------------------------------------------------------
// Register-tiled product in shared memory, all threads of the block:
//   C(i,j) = Add(i,j) + alpha * sum_k opA(i,k) * B(k,j)      i < M, j < N, k ascending (the oracle's order)
// opA = A^T if TA (A is K x M) else A (M x K). A, B, C in shared memory with leading dims lda, ldb, ldc;
// Add (may be null, may be global memory or C itself) with leading dim ldadd.
template <bool TA>
__device__ void cta_gemm(double* C, int ldc, const double* A, int lda, const double* B, int ldb, int M, int N, int K,
                         const double* Add, int ldadd, double alpha) {
  const int tm = (M + 3) >> 2, tn = (N + 3) >> 2;
  for (int tile = threadIdx.x; tile < tm * tn; tile += blockDim.x) {
    const int i0 = (tile % tm) << 2, j0 = (tile / tm) << 2;
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[a][b] = 0.;
    // clamp the operand indices of partial tiles (results of the clamped lanes are discarded)
    int ia[4], jb[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      ia[a] = min(i0 + a, M - 1);
      jb[a] = min(j0 + a, N - 1);
    }
    for (int kk = 0; kk < K; ++kk) {
      double av[4], bv[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        av[a] = TA ? A[kk + lda * ia[a]] : A[ia[a] + lda * kk];
        bv[a] = B[kk + ldb * jb[a]];
      }
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] += av[a] * bv[b];
    }
#pragma unroll
    for (int b = 0; b < 4; ++b)
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        const int i = i0 + a, j = j0 + b;
        if (i < M && j < N) {
          const double base = Add ? Add[i + ldadd * j] : 0.;
          C[i + ldc * j] = base + alpha * acc[a][b];
        }
      }
  }
}

// ---------------------------------------------------------------------------------------------
// The same product on the FP64 tensor cores (DMMA): mma.sync.aligned.m8n8k4.f64, one warp per 8 x 16 output tile
// (two accumulator fragments sharing the A fragment). Used for the dense n >= 24 products of the Riccati step
// (Fx^T Vxx Fx at n = 72 is a real 72^3 GEMM); tcgen05 has no FP64 path. Fragment layout of m8n8k4 (PTX ISA):
//   A (8x4, row): lane l holds A[l/4][l%4]     B (4x8, col): lane l holds B[l%4][l/4]
//   C/D (8x8)   : lane l holds C[l/4][2(l%4)], C[l/4][2(l%4)+1]
// Out-of-range operands are read as zero, so any M, N, K works. The k-sum is accumulated in groups of four inside the
// tensor core instead of one DFMA per k: differences from cta_gemm are at rounding level.
__host__ __device__ inline int cta_mma_tiles(int M, int N) { return ((M + 23) / 24) * ((N + 23) / 24); }
__device__ __forceinline__ void dmma_8x8x4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}
// 8 x 16 warp tiles (two accumulator fragments sharing the A fragment): more, smaller tiles — the better fit below n = 48
template <bool TA>
__device__ void cta_gemm_mma_8x16(double* C, int ldc, const double* A, int lda, const double* B, int ldb, int M, int N, int K,
                             const double* Add, int ldadd, double alpha) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int tm = (M + 7) >> 3, tn = (N + 15) >> 4;
  const int ar = lane >> 2, ac = lane & 3;  // A fragment: row ar, k-offset ac ; B fragment: k-offset ac, column ar
  for (int tile = warp; tile < tm * tn; tile += nwarps) {
    const int i0 = (tile % tm) << 3, j0 = (tile / tm) << 4;
    const int ia = i0 + ar;
    const int jb0 = j0 + ar, jb1 = j0 + 8 + ar;
    const bool oka = ia < M, okb0 = jb0 < N, okb1 = jb1 < N;
    const bool second = j0 + 8 < N;
    const double* pa = TA ? (A + (size_t)lda * (oka ? ia : 0)) : (A + (oka ? ia : 0));
    const double* pb0 = B + (size_t)ldb * (okb0 ? jb0 : 0);
    const double* pb1 = B + (size_t)ldb * (okb1 ? jb1 : 0);
    double c00 = 0., c01 = 0., c10 = 0., c11 = 0.;
    int kk = 0;
    for (; kk + 4 <= K; kk += 4) {
      const int k = kk + ac;
      const double a = oka ? (TA ? pa[k] : pa[(size_t)lda * k]) : 0.;
      const double b0 = okb0 ? pb0[k] : 0.;
      dmma_8x8x4(c00, c01, a, b0);
      if (second) {
        const double b1 = okb1 ? pb1[k] : 0.;
        dmma_8x8x4(c10, c11, a, b1);
      }
    }
    if (kk < K) {  // K not a multiple of 4: zero-padded last step
      const int k = kk + ac;
      const bool okk = k < K;
      const double a = (oka && okk) ? (TA ? pa[k] : pa[(size_t)lda * k]) : 0.;
      const double b0 = (okb0 && okk) ? pb0[k] : 0.;
      dmma_8x8x4(c00, c01, a, b0);
      if (second) {
        const double b1 = (okb1 && okk) ? pb1[k] : 0.;
        dmma_8x8x4(c10, c11, a, b1);
      }
    }
    const int ci = i0 + ar, cj = j0 + 2 * ac;
    if (ci < M) {
      if (cj < N) C[ci + ldc * cj] = (Add ? Add[ci + ldadd * cj] : 0.) + alpha * c00;
      if (cj + 1 < N) C[ci + ldc * (cj + 1)] = (Add ? Add[ci + ldadd * (cj + 1)] : 0.) + alpha * c01;
      if (cj + 8 < N) C[ci + ldc * (cj + 8)] = (Add ? Add[ci + ldadd * (cj + 8)] : 0.) + alpha * c10;
      if (cj + 9 < N) C[ci + ldc * (cj + 9)] = (Add ? Add[ci + ldadd * (cj + 9)] : 0.) + alpha * c11;
    }
  }
}

// Register blocking: one warp owns a 24 x 24 output tile = 3 x 3 fragments, so a k-step of 4 loads 3 A and 3 B
// fragments from shared memory for 9 DMMAs (0.67 LDS.64 per DMMA; the first version loaded 2 - 3 per DMMA and was bound
// by the shared-memory pipe: ncu, DMMA pipe 13 % busy). `woff` rotates the tile -> warp assignment so that several
// products issued between two barriers land on different warps.
template <bool TA>
__device__ void cta_gemm_mma(double* C, int ldc, const double* A, int lda, const double* B, int ldb, int M, int N, int K,
                             const double* Add, int ldadd, double alpha, int woff = 0) {
  constexpr int RT = 3;  // fragments per tile side
  const int lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int warp = ((threadIdx.x >> 5) + nwarps - (woff % nwarps)) % nwarps;
  const int tm = (M + 8 * RT - 1) / (8 * RT), tn = (N + 8 * RT - 1) / (8 * RT);
  const int ar = lane >> 2, ac = lane & 3;  // A fragment: row ar, k-offset ac ; B fragment: k-offset ac, column ar
  for (int tile = warp; tile < tm * tn; tile += nwarps) {
    const int i0 = (tile % tm) * (8 * RT), j0 = (tile / tm) * (8 * RT);
    const double* pa[RT];
    const double* pb[RT];
    bool oka[RT], okb[RT];
#pragma unroll
    for (int r = 0; r < RT; ++r) {
      const int ia = i0 + 8 * r + ar, jb = j0 + 8 * r + ar;
      oka[r] = ia < M;
      okb[r] = jb < N;
      pa[r] = TA ? (A + (size_t)lda * (oka[r] ? ia : 0)) : (A + (oka[r] ? ia : 0));
      pb[r] = B + (size_t)ldb * (okb[r] ? jb : 0);
    }
    double acc[RT][RT][2];
#pragma unroll
    for (int r = 0; r < RT; ++r)
#pragma unroll
      for (int q = 0; q < RT; ++q) acc[r][q][0] = acc[r][q][1] = 0.;
    const int K4 = K & ~3;
    for (int kk = 0; kk < K4; kk += 4) {
      const int k = kk + ac;
      double a[RT], b[RT];
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        a[r] = oka[r] ? (TA ? pa[r][k] : pa[r][(size_t)lda * k]) : 0.;
        b[r] = okb[r] ? pb[r][k] : 0.;
      }
#pragma unroll
      for (int r = 0; r < RT; ++r)
#pragma unroll
        for (int q = 0; q < RT; ++q) dmma_8x8x4(acc[r][q][0], acc[r][q][1], a[r], b[q]);
    }
    if (K4 < K) {  // K not a multiple of 4: zero-padded last step
      const int k = K4 + ac;
      const bool okk = k < K;
      double a[RT], b[RT];
#pragma unroll
      for (int r = 0; r < RT; ++r) {
        a[r] = (oka[r] && okk) ? (TA ? pa[r][k] : pa[r][(size_t)lda * k]) : 0.;
        b[r] = (okb[r] && okk) ? pb[r][k] : 0.;
      }
#pragma unroll
      for (int r = 0; r < RT; ++r)
#pragma unroll
        for (int q = 0; q < RT; ++q) dmma_8x8x4(acc[r][q][0], acc[r][q][1], a[r], b[q]);
    }
#pragma unroll
    for (int r = 0; r < RT; ++r) {
      const int ci = i0 + 8 * r + ar;
      if (ci >= M) continue;
#pragma unroll
      for (int q = 0; q < RT; ++q) {
        const int cj = j0 + 8 * q + 2 * ac;
        if (cj < N) C[ci + ldc * cj] = (Add ? Add[ci + ldadd * cj] : 0.) + alpha * acc[r][q][0];
        if (cj + 1 < N) C[ci + ldc * (cj + 1)] = (Add ? Add[ci + ldadd * (cj + 1)] : 0.) + alpha * acc[r][q][1];
      }
    }
  }
}
// dispatcher: tensor cores for the large products, DFMA register tiles otherwise
template <bool TA>
__device__ __forceinline__ void cta_gemm_auto(double* C, int ldc, const double* A, int lda, const double* B, int ldb, int M,
                                              int N, int K, const double* Add, int ldadd, double alpha, int woff = 0) {
#ifndef CGX_NO_DMMA
  if (K >= 16 && M >= 8 && N >= 8) {
    if (M >= 48 || N >= 48)
      cta_gemm_mma<TA>(C, ldc, A, lda, B, ldb, M, N, K, Add, ldadd, alpha, woff);
    else
      cta_gemm_mma_8x16<TA>(C, ldc, A, lda, B, ldb, M, N, K, Add, ldadd, alpha);
  } else
#endif
    cta_gemm<TA>(C, ldc, A, lda, B, ldb, M, N, K, Add, ldadd, alpha);
}

// ---------------------------------------------------------------------------------------------
// ActionModelLQR (reference: include/crocoddyl/core/actions/lqr.hxx:14-70, lqr.hpp:88-95).
// params = [drift_free, Fx(n*n), Fu(n*m), f0(n), Lxx(n*n), Lxu(n*m), Luu(m*m), lx(n), lu(m)], column-major.
struct LqrModel {
  static constexpr bool kPerNodeMats = false;
  static constexpr bool kPhaseSplit = false;  // the fused solve kernel is small for LQR
  static constexpr bool kNodeKernel = false;
  static constexpr int kMinThreads = 64, kMaxThreads = 384, kMinBlocks = 1, kDirMinBlocks = 1, kFwdMinBlocks = 1;
  static constexpr int kN = 0, kM = 0;  // run-time sizes
  // matrices of one ActionModelLQR inside its parameter block
  struct Blk {
    const double *Fx, *Fu, *f0, *Lxx, *Lxu, *Luu, *lx, *lu;
    bool drift_free;
    __device__ void set(const double* p, int n, int m) {
      drift_free = p[0] != 0.;
      Fx = p + 1;
      Fu = Fx + n * n;
      f0 = Fu + n * m;
      Lxx = f0 + n;
      Lxu = Lxx + n * n;
      Luu = Lxu + n * m;
      lx = Luu + m * m;
      lu = lx + n;
    }
  };
  Blk run, term;  // running nodes / terminal node (u == nullptr selects the terminal model: action-base.hxx:27-31)
  int n, m;
  // host side: device address of a constant derivative matrix (ActionDataLQR holds copies, lqr.hxx:65-69)
  template <class PD>
  static const double* node_matrix(const CtaCtx& c, const PD*, const std::string& name, int b, int t) {
    const size_t n = c.n, m = c.m;
    const double* p = (t == c.T ? c.tparams : c.params) + (c.per_problem ? (size_t)b * c.nparams : 0);
    if (name == "Fx") return p + 1;
    if (name == "Fu") return p + 1 + n * n;
    if (name == "Lxx") return p + 1 + n * n + n * m + n;
    if (name == "Lxu") return p + 1 + 2 * n * n + n * m + n;
    return p + 1 + 2 * n * n + 2 * n * m + n;  // Luu
  }
  template <class BE>
  static void quasi_static(BE&, const CtaCtx&, const double*, double*, cudaStream_t) {
    throw std::runtime_error("quasiStatic is not implemented for ActionModelLQR");
  }
  __device__ void load(const CtaCtx& c, int b) {
    n = c.n;
    m = c.m;
    const size_t off = c.per_problem ? (size_t)b * c.nparams : 0;
    run.set(c.params + off, n, m);
    term.set(c.tparams + off, n, m);
  }
  __device__ NodeMats mats(const CtaCtx& c, int, int t) const {
    const Blk& q = t == c.T ? term : run;
    return NodeMats{q.Fx, q.Fu, q.Lxx, q.Lxu, q.Luu, true};
  }

  // calc of one node by ONE thread, in the oracle's operation order (lqr.hxx:31-49). u == nullptr: zeros.
  __device__ double calc_serial(const double* x, const double* u, double* xnext) const {
    const Blk& q = u ? run : term;
    for (int i = 0; i < n; ++i) {
      double a = 0., bsum = 0.;
      for (int j = 0; j < n; ++j) a += q.Fx[i + n * j] * x[j];
      if (u)
        for (int j = 0; j < m; ++j) bsum += q.Fu[i + n * j] * u[j];
      xnext[i] = q.drift_free ? (a + bsum) : (a + bsum + q.f0[i]);
    }
    double xLx = 0., uLu = 0., xLu = 0., lxx_ = 0., luu_ = 0.;
    for (int i = 0; i < n; ++i) {
      double s = 0., s2 = 0.;
      for (int j = 0; j < n; ++j) s += q.Lxx[i + n * j] * x[j];
      if (u)
        for (int j = 0; j < m; ++j) s2 += q.Lxu[i + n * j] * u[j];
      xLx += x[i] * s;
      xLu += x[i] * s2;
      lxx_ += q.lx[i] * x[i];
    }
    if (u)
      for (int i = 0; i < m; ++i) {
        double s = 0.;
        for (int j = 0; j < m; ++j) s += q.Luu[i + m * j] * u[j];
        uLu += u[i] * s;
        luu_ += q.lu[i] * u[i];
      }
    return 0.5 * xLx + 0.5 * uLu + xLu + lxx_ + luu_;
  }
  // Lx, Lu of one node by one thread (lqr.hxx:63-64); the five matrices are the model's constants
  __device__ void calcdiff_serial(const double* x, const double* u, double* Lx, double* Lu) const {
    const Blk& q = u ? run : term;
    for (int i = 0; i < n; ++i) {
      double a = 0., bsum = 0.;
      for (int j = 0; j < n; ++j) a += q.Lxx[i + n * j] * x[j];
      if (u)
        for (int j = 0; j < m; ++j) bsum += q.Lxu[i + n * j] * u[j];
      Lx[i] = q.lx[i] + a + bsum;
    }
    for (int i = 0; i < m; ++i) {
      double cc = 0., e = 0.;
      for (int j = 0; j < n; ++j) cc += q.Lxu[j + n * i] * x[j];
      if (u)
        for (int j = 0; j < m; ++j) e += q.Luu[i + m * j] * u[j];
      Lu[i] = q.lu[i] + cc + e;
    }
  }
  // all nodes of the problem at once: one node per thread (ShootingProblem::calc, shooting.hxx:133-161)
  __device__ void calc_all(const CtaCtx& c, int b, const double* xs, const double* us) const {
    const int T = c.T;
    for (int t = threadIdx.x; t <= T; t += blockDim.x) {
      const size_t node = (size_t)b * (T + 1) + t;
      c.d_cost[node] = calc_serial(xs + (size_t)t * n, t < T ? us + (size_t)t * m : nullptr, c.d_xnext + node * n);
    }
  }
  __device__ void calcdiff_all(const CtaCtx& c, int b, const double* xs, const double* us) const {
    const int T = c.T;
    for (int t = threadIdx.x; t <= T; t += blockDim.x) {
      const size_t node = (size_t)b * (T + 1) + t;
      calcdiff_serial(xs + (size_t)t * n, t < T ? us + (size_t)t * m : nullptr, c.d_Lx + node * n, c.d_Lu + node * m);
    }
  }
  __device__ double calc_thread(const double* x, const double* u, double* xnext) const { return calc_serial(x, u, xnext); }
  // calc of one node by the whole block (rollout): x, u in shared memory; writes xnext (shared), returns the
  // cost to thread 0 through *cost_out (shared). scratch: >= 2n+m doubles of shared memory.
  __device__ void calc_block(const double* x, const double* u, double* xnext, double* scratch, double* cost_out) const {
    const Blk& q = u ? run : term;
    double* Lxx_x = scratch;
    double* Lxu_u = scratch + n;
    double* Luu_u = scratch + 2 * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      double a = 0., bsum = 0., s = 0., s2 = 0.;
      for (int j = 0; j < n; ++j) {
        a += q.Fx[i + n * j] * x[j];
        s += q.Lxx[i + n * j] * x[j];
      }
      if (u)
        for (int j = 0; j < m; ++j) {
          bsum += q.Fu[i + n * j] * u[j];
          s2 += q.Lxu[i + n * j] * u[j];
        }
      xnext[i] = q.drift_free ? (a + bsum) : (a + bsum + q.f0[i]);
      Lxx_x[i] = s;
      Lxu_u[i] = s2;
    }
    if (u)
      for (int i = threadIdx.x; i < m; i += blockDim.x) {
        double s = 0.;
        for (int j = 0; j < m; ++j) s += q.Luu[i + m * j] * u[j];
        Luu_u[i] = s;
      }
    __syncthreads();
    if (threadIdx.x == 0) {
      double xLx = 0., uLu = 0., xLu = 0., lxx_ = 0., luu_ = 0.;
      for (int i = 0; i < n; ++i) {
        xLx += x[i] * Lxx_x[i];
        xLu += x[i] * Lxu_u[i];
        lxx_ += q.lx[i] * x[i];
      }
      if (u)
        for (int i = 0; i < m; ++i) {
          uLu += u[i] * Luu_u[i];
          luu_ += q.lu[i] * u[i];
        }
      *cost_out = 0.5 * xLx + 0.5 * uLu + xLu + lxx_ + luu_;
    }
    __syncthreads();
  }
};

// ---------------------------------------------------------------------------------------------
// ---------------------------------------------------------------------------------------------
// Shared-memory work area of one block (doubles). Leading dimensions are odd (cta_ld).
struct CtaSmem {
  double *V, *A, *Fx, *Fu, *Qxu, *G, *Kg, *Quu, *Lc;  // matrices
  double *Vx, *Qx, *Qu, *kg, *Quuk, *x, *u, *dx, *xn, *scr;  // vectors
  double* red;   // 8 doubles of reduction / broadcast scratch
  int* flag;
  int ldn, ldm;
  __device__ static size_t doubles(int n, int m) {
    const size_t ldn = n | 1, ldm = (m | 1);
    return 3 * ldn * n + 2 * ldn * m + 2 * ldm * n + 2 * ldm * m + 6 * (size_t)n + 4 * (size_t)m + (2 * n + m) + 16;
  }
  __device__ void carve(double* base, int n, int m) {
    ldn = n | 1;
    ldm = m | 1;
    double* p = base;
    V = p; p += ldn * n;
    A = p; p += ldn * n;
    Fx = p; p += ldn * n;
    Fu = p; p += ldn * m;
    Qxu = p; p += ldn * m;
    G = p; p += ldm * n;
    Kg = p; p += ldm * n;
    Quu = p; p += ldm * m;
    Lc = p; p += ldm * m;
    Vx = p; p += n;
    Qx = p; p += n;
    x = p; p += n;
    dx = p; p += n;
    xn = p; p += n;
    p += n;  // spare
    Qu = p; p += m;
    kg = p; p += m;
    Quuk = p; p += m;
    u = p; p += m;
    scr = p; p += 2 * n + m;
    red = p; p += 8;
    flag = reinterpret_cast<int*>(p);
  }
};
// the block-wide plan alone (CtaSmem::carve): what the rollout / line-search kernels touch
inline size_t cta_blockwide_smem_bytes(int n, int m) {
  const size_t ldn = n | 1, ldm = (m | 1);
  return 8 * (3 * ldn * n + 2 * ldn * m + 2 * ldm * n + 2 * ldm * m + 6 * (size_t)n + 4 * (size_t)m + (2 * n + m) + 16);
}
inline size_t cta_smem_bytes(int n, int m) {
  const size_t blockwide = cta_blockwide_smem_bytes(n, m);
  const size_t warpwide = cta_is_small(n, m) ? cta_small_smem_bytes(n, m) : 0;  // the two plans overlay each other
  return blockwide > warpwide ? blockwide : warpwide;
}

// copy a column-major rows x cols matrix from global memory to shared memory (leading dim ld)
__device__ __forceinline__ void cta_stage(double* dst, int ld, const double* src, int rows, int cols) {
  for (int e = threadIdx.x; e < rows * cols; e += blockDim.x) dst[(e % rows) + ld * (e / rows)] = src[e];
}
__device__ __forceinline__ void cta_unstage(double* dst, const double* src, int ld, int rows, int cols) {
  for (int e = threadIdx.x; e < rows * cols; e += blockDim.x) dst[e] = src[(e % rows) + ld * (e / rows)];
}

// ---------------------------------------------------------------------------------------------
// SolverDDP::calcDiff (ddp.cpp:157-178) for one problem: calc (only when do_calc: iter_ == 0), calcDiff of all
// nodes, total cost in the reference's order (shooting.hxx:188-192) and the gaps fs. Returns the cost to all.
template <class M, bool nodes_done = false>
__device__ double cta_calcdiff(const CtaCtx& c, const M& mod, CtaSmem& sm, int b, int cur, bool do_calc,
                               bool is_feasible, bool was_feasible) {
  const int T = c.T, n = c.n, m = c.m;
  const double* xs = c.xs[cur] + (size_t)b * (T + 1) * n;
  const double* us = c.us[cur] + (size_t)b * T * m;
  if constexpr (!nodes_done) {  // else: evaluated by cta_phase_nodes_kernel in an earlier launch
    if (do_calc) mod.calc_all(c, b, xs, us);
    mod.calcdiff_all(c, b, xs, us);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double cost = 0.;
    for (int t = 0; t <= T; ++t) cost += c.d_cost[(size_t)b * (T + 1) + t];
    sm.red[0] = cost;
  }
  double* fs = c.fs + (size_t)b * (T + 1) * n;
  if (!is_feasible) {
    const double* xnext = c.d_xnext + (size_t)b * (T + 1) * n;
    for (int e = threadIdx.x; e < (T + 1) * n; e += blockDim.x) {
      const int t = e / n, i = e % n;
      fs[e] = (t == 0) ? (c.x0[(size_t)b * n + i] - xs[i]) : (xnext[(size_t)(t - 1) * n + i] - xs[e]);
    }
  } else if (!was_feasible) {
    for (int e = threadIdx.x; e < (T + 1) * n; e += blockDim.x) fs[e] = 0.;
  }
  __syncthreads();
  return sm.red[0];
}

// Unblocked lower LLT of Quu (m <= MR <= 32) by ONE warp with lane i holding row i in registers: left-looking, the
// pivot row's entries reach the other lanes by shuffles, so a step is a chain of multiply-adds without shared-memory
// round trips or warp barriers. Same operation order as the oracle / Eigen's unblocked LLT (sums over j ascending);
// x / p as x * (1 / p) unless CGX_PARITY. Writes the factor to sm.Lc; sets *sm.flag when a pivot is not positive.
// Measured on B200: C1 (LQR 37/12, one problem) 5.03 -> 4.02 ms per call, n = 36 sweep 30.7 -> 24.7 ms, n = 72 sweep
// 158.1 -> 155.7 ms (there the m = 24 instantiation is 1.5 k straight-line instructions run once per node by one warp:
// ncu, no_instruction = 71 % of its stalls; a rolled version with predicated register selects was slower still:
// 229 ms).
template <int MR>
__device__ __forceinline__ void cta_llt_warp_regs(CtaSmem& sm, int m, int ldm) {
  const int lane = threadIdx.x & 31;
  double row[MR];
#pragma unroll
  for (int j = 0; j < MR; ++j) row[j] = (lane < m && j <= lane && j < m) ? sm.Quu[lane + ldm * j] : 0.;
  bool bad = false;
#pragma unroll
  for (int kk = 0; kk < MR; ++kk) {
    if (kk < m) {
      double p = row[kk], s = row[kk];
#pragma unroll
      for (int j = 0; j < kk; ++j) {
        const double lkj = __shfl_sync(0xffffffffu, row[j], kk);
        p -= row[j] * row[j];  // the pivot row's own sum (meaningful on lane kk)
        s -= row[j] * lkj;
      }
      p = __shfl_sync(0xffffffffu, p, kk);
      if (p <= 0.) bad = true;
      double sp, pinv;
      llt_pivot(p, sp, pinv);
      if (lane == kk)
        row[kk] = sp;
      else if (lane > kk)
        row[kk] = div_pivot(s, sp, pinv);
    }
  }
  if (bad && lane == 0) *sm.flag = 1;
  if (lane < m) {
#pragma unroll
    for (int j = 0; j < MR; ++j)
      if (j <= lane && j < m) sm.Lc[lane + ldm * j] = row[j];
  }
}

// K = Quu^-1 Qxu^T, k = Quu^-1 Qu with the right-hand side of a thread in registers (m <= MR): only the factor is read
// from shared memory (broadcast loads). Same operation order as the in-place version in cta_backward.
template <int MR>
__device__ __forceinline__ void cta_solve_rhs_regs(CtaSmem& sm, int n, int m, int ldn, int ldm, int tid, int P) {
  for (int col = tid; col <= n; col += P) {
    double r[MR], rinv[MR];  // the reciprocal pivots once per right-hand side instead of 2 m divisions (common.cuh)
#pragma unroll
    for (int i = 0; i < MR; ++i) {
      r[i] = i < m ? ((col < n) ? sm.Qxu[col + ldn * i] : sm.Qu[i]) : 0.;
      rinv[i] = i < m ? 1. / sm.Lc[i + ldm * i] : 0.;
    }
#pragma unroll
    for (int i = 0; i < MR; ++i) {
      if (i < m) {
        double s = r[i];
#pragma unroll
        for (int j = 0; j < i; ++j) s -= sm.Lc[i + ldm * j] * r[j];
        r[i] = div_pivot(s, sm.Lc[i + ldm * i], rinv[i]);
      }
    }
#pragma unroll
    for (int i = MR - 1; i >= 0; --i) {
      if (i < m) {
        double s = r[i];
#pragma unroll
        for (int j = i + 1; j < MR; ++j)
          if (j < m) s -= sm.Lc[j + ldm * i] * r[j];
        r[i] = div_pivot(s, sm.Lc[i + ldm * i], rinv[i]);
      }
    }
    double* dst = (col < n) ? (sm.Kg + ldm * col) : sm.kg;
#pragma unroll
    for (int i = 0; i < MR; ++i)
      if (i < m) dst[i] = r[i];
  }
}

// the 12 / 4 instantiation out of line (see OUTLINE_SMALL at cta_backward)
template <class M>
__device__ __noinline__ int cta_backward_warp_12_4(const CtaCtx& c, const M& mod, int b, int cur, bool is_feasible, double xreg,
                                                   double ureg, double& dg_out, double& dq_out, double& stop_out) {
  return cta_backward_warp<12, 4>(c, mod, b, cur, is_feasible, xreg, ureg, dg_out, dq_out, stop_out);
}

// ---------------------------------------------------------------------------------------------
// SolverDDP::backwardPass + computeGains (ddp.cpp:180-253,298-310) for one problem by the whole block, with the
// reductions of updateExpectedImprovement (fddp.cpp:126-147) / expectedImprovement (ddp.cpp:144-155) /
// stoppingCriteria (ddp.cpp:132-142). Returns 0 or 1 ("backward_error") to all threads; dg,dq,stop valid on
// every thread.
// OUTLINE_SMALL: whether the compile-time 12 / 4 instantiation below is called out of line. Measured on B200, same
// source otherwise: inline, the fused solve kernel's register allocation changes and the single-OCP case C1 (n = 37)
// goes from 3.7 to 4.6 ms; out of line, the batched sweep kernel slows down (n = 36: 24.2 -> 27.0 ms, n = 72: 148.7 ->
// 157.2 ms). So the fused solve kernel calls it out of line and the step-wise / phase kernels inline it.
template <class M, bool OUTLINE_SMALL = false>
__device__ int cta_backward(const CtaCtx& c, const M& mod, CtaSmem& sm, int b, int cur, bool is_feasible, double xreg,
                            double ureg, double& dg_out, double& dq_out, double& stop_out) {
#ifndef CGX_NO_SMALL_BACKWARD
  if (cta_is_small(c.n, c.m)) {
    // run-time-sized models (LQR): the size of BASELINE config C3 gets the compile-time instantiation (outputs per lane
    // known, LLT and right-hand sides in registers), the others the run-time loops
    if constexpr (M::kN == 0) {
      if (c.n == 12 && c.m == 4) {
        if constexpr (OUTLINE_SMALL)
          return cta_backward_warp_12_4(c, mod, b, cur, is_feasible, xreg, ureg, dg_out, dq_out, stop_out);
        else
          return cta_backward_warp<12, 4>(c, mod, b, cur, is_feasible, xreg, ureg, dg_out, dq_out, stop_out);
      }
    }
    return cta_backward_warp<M::kN, M::kM>(c, mod, b, cur, is_feasible, xreg, ureg, dg_out, dq_out, stop_out);
  }
#endif
  const int T = c.T, n = c.n, m = c.m, ldn = sm.ldn, ldm = sm.ldm;
  const int tid = threadIdx.x, P = blockDim.x;
  const bool fddp = c.fddp != 0, full = c.full != 0;
  const bool has_xreg = !isnan(xreg), has_ureg = !isnan(ureg);
  const double* fs = c.fs + (size_t)b * (T + 1) * n;
  double dg = 0., dq = 0., stop = 0.;  // thread 0 accumulates
  // ---- terminal node (ddp.cpp:181-191)
  {
    const NodeMats nm = mod.mats(c, b, T);
    cta_stage(sm.V, ldn, nm.Lxx, n, n);
    for (int i = tid; i < n; i += P) sm.Vx[i] = c.d_Lx[((size_t)b * (T + 1) + T) * n + i];
    __syncthreads();
    if (has_xreg)
      for (int i = tid; i < n; i += P) sm.V[i + ldn * i] += xreg;
    __syncthreads();
    if (!is_feasible) {
      for (int i = tid; i < n; i += P) {
        double s = 0.;
        for (int j = 0; j < n; ++j) s += sm.V[i + ldn * j] * fs[(size_t)T * n + j];
        sm.scr[i] = s;
        sm.Vx[i] += s;
      }
      __syncthreads();
    }
    for (int i = tid; i < n; i += P) {
      c.Vx[((size_t)b * (T + 1) + T) * n + i] = sm.Vx[i];
      c.w[((size_t)b * (T + 1) + T) * n + i] = is_feasible ? 0. : sm.scr[i];
    }
    cta_unstage(c.Vxx + ((size_t)b * (T + 1) + T) * n * n, sm.V, ldn, n, n);
    if (tid == 0 && fddp && !is_feasible) {  // fddp.cpp:131-135
      double a = 0., q = 0.;
      for (int i = 0; i < n; ++i) {
        a += sm.Vx[i] * fs[(size_t)T * n + i];
        q += fs[(size_t)T * n + i] * sm.scr[i];
      }
      dg -= a;
      dq += q;
    }
    __syncthreads();
  }
  bool staged = false;
  for (int t = T - 1; t >= 0; --t) {
    const NodeMats nm = mod.mats(c, b, t);
    const size_t node = (size_t)b * (T + 1) + t, rnode = (size_t)b * T + t;
    if (!(nm.constant && staged)) {
      cta_stage(sm.Fx, ldn, nm.Fx, n, n);
      cta_stage(sm.Fu, ldn, nm.Fu, n, m);
      staged = true;
    }
    __syncthreads();
    // A = Fx^T Vxx' ; G = Fu^T Vxx' ; Qx = Lx + Fx^T Vx' ; Qu = Lu + Fu^T Vx'   (ddp.cpp:204,206,211,214)
    cta_gemm_auto<true>(sm.A, ldn, sm.Fx, ldn, sm.V, ldn, n, n, n, nullptr, 0, 1.);
    cta_gemm_auto<true>(sm.G, ldm, sm.Fu, ldn, sm.V, ldn, m, n, n, nullptr, 0, 1., cta_mma_tiles(n, n));
    for (int j = tid; j < n + m; j += P) {
      double s = 0.;
      if (j < n) {
        for (int i = 0; i < n; ++i) s += sm.Fx[i + ldn * j] * sm.Vx[i];
        sm.Qx[j] = c.d_Lx[node * n + j] + s;
      } else {
        const int jj = j - n;
        for (int i = 0; i < n; ++i) s += sm.Fu[i + ldn * jj] * sm.Vx[i];
        sm.Qu[jj] = c.d_Lu[node * m + jj] + s;
      }
    }
    __syncthreads();
    // Qxx = Lxx + A Fx (into V: Vxx' is dead) ; Qxu = Lxu + A Fu ; Quu = Luu + G Fu (+ ureg)   (ddp.cpp:202-218)
    cta_gemm_auto<false>(sm.V, ldn, sm.A, ldn, sm.Fx, ldn, n, n, n, nm.Lxx, n, 1.);
    cta_gemm_auto<false>(sm.Qxu, ldn, sm.A, ldn, sm.Fu, ldn, n, m, n, nm.Lxu, n, 1., cta_mma_tiles(n, n));
    cta_gemm_auto<false>(sm.Quu, ldm, sm.G, ldm, sm.Fu, ldn, m, m, n, nm.Luu, m, 1., cta_mma_tiles(n, n) + cta_mma_tiles(n, m));
    __syncthreads();
    if (has_ureg)
      for (int i = tid; i < m; i += P) sm.Quu[i + ldm * i] += ureg;
    if (tid == 0) *sm.flag = 0;
    __syncthreads();
    if (full) {
      cta_unstage(c.Qxx + rnode * n * n, sm.V, ldn, n, n);
      cta_unstage(c.Qxu + rnode * n * m, sm.Qxu, ldn, n, m);
      cta_unstage(c.Quu + rnode * m * m, sm.Quu, ldm, m, m);
      for (int i = tid; i < n; i += P) c.Qx[rnode * n + i] = sm.Qx[i];
      for (int i = tid; i < m; i += P) c.Qu[rnode * m + i] = sm.Qu[i];
    }
    // SolverBoxDDP/BoxFDDP::computeGains (box-ddp.cpp:46-74): box QP on the control step when the model has control
    // limits and the candidate is feasible (one thread; work arrays in the problem's global scratch)
    const bool boxed = c.box && is_feasible && c.has_limits[b] != 0;
    if (boxed) {
      if (tid == 0) {
        const double* u_t = c.us[cur] + rnode * m;  // us_[t] of the candidate
        for (int i = 0; i < m; ++i) sm.kg[i] = c.k[rnode * m + i];  // k_[t] of the previous backward pass (warm start)
        *sm.flag = box_compute_gains(n, m, sm.Quu, ldm, sm.Qu, sm.Qxu, ldn, u_t, c.u_lb + (size_t)b * m,
                                     c.u_ub + (size_t)b * m, sm.Kg, ldm, sm.kg, c.Quu_inv + rnode * m * m, c.o,
                                     c.box_work + (size_t)b * box_gains_work_doubles(m),
                                     c.box_iwork + (size_t)b * box_gains_work_ints(m));
      }
      __syncthreads();
      if (*sm.flag) return 1;
      if (full)
        for (int i = tid; i < m; i += P) c.Qu[rnode * m + i] = sm.Qu[i];  // get_Qu returns the clamped gradient
    } else {
    // computeGains (ddp.cpp:298-310): unblocked lower LLT by warp 0
    if (tid < 32 && m <= 8) {
      cta_llt_warp_regs<8>(sm, m, ldm);
    } else if (tid < 32 && m <= 12) {
      cta_llt_warp_regs<12>(sm, m, ldm);
    } else if (tid < 32 && m <= 16) {
      cta_llt_warp_regs<16>(sm, m, ldm);
    } else if (tid < 32 && m <= 24) {
      cta_llt_warp_regs<24>(sm, m, ldm);
    } else if (tid < 32) {
      for (int e = tid; e < m * m; e += 32) sm.Lc[(e % m) + ldm * (e / m)] = sm.Quu[(e % m) + ldm * (e / m)];
      __syncwarp();
      // lane i owns rows i, i+32, ...: left-looking through shared memory. (A register-resident right-looking variant
      // with rows/right-hand sides in registers was measured on B200 and is slower: it spills under the 168-register
      // budget of 384-thread blocks — n = 72: 69.8 vs 64.2 us per sweep.)
      for (int kk = 0; kk < m; ++kk) {
        double p = sm.Lc[kk + ldm * kk];
        for (int j = 0; j < kk; ++j) p -= sm.Lc[kk + ldm * j] * sm.Lc[kk + ldm * j];
        if (p <= 0.) *sm.flag = 1;
        p = sqrt(p);
        __syncwarp();
        if (tid == 0) sm.Lc[kk + ldm * kk] = p;
        for (int i = kk + 1 + tid; i < m; i += 32) {
          double s = sm.Lc[i + ldm * kk];
          for (int j = 0; j < kk; ++j) s -= sm.Lc[i + ldm * j] * sm.Lc[kk + ldm * j];
          sm.Lc[i + ldm * kk] = s / p;
        }
        __syncwarp();
      }
    }
    __syncthreads();
    if (*sm.flag) return 1;
    // K = Quu^-1 Qxu^T, k = Quu^-1 Qu: one right-hand side per thread (forward / back substitution in the oracle's
    // order). m <= 24: the right-hand side lives in registers and only the factor is read from shared memory
    // (broadcast loads); the in-place version below pays a shared-memory round trip per multiply-add.
    if (m <= 8)
      cta_solve_rhs_regs<8>(sm, n, m, ldn, ldm, tid, P);
    else if (m <= 12)
      cta_solve_rhs_regs<12>(sm, n, m, ldn, ldm, tid, P);
    else if (m <= 16)
      cta_solve_rhs_regs<16>(sm, n, m, ldn, ldm, tid, P);
    else if (m <= 24)
      cta_solve_rhs_regs<24>(sm, n, m, ldn, ldm, tid, P);
    else
    for (int col = tid; col <= n; col += P) {
      double* r = (col < n) ? (sm.Kg + ldm * col) : sm.kg;
      for (int i = 0; i < m; ++i) r[i] = (col < n) ? sm.Qxu[col + ldn * i] : sm.Qu[i];
      for (int i = 0; i < m; ++i) {
        double s = r[i];
        for (int j = 0; j < i; ++j) s -= sm.Lc[i + ldm * j] * r[j];
        r[i] = s / sm.Lc[i + ldm * i];
      }
      for (int i = m - 1; i >= 0; --i) {
        double s = r[i];
        for (int j = i + 1; j < m; ++j) s -= sm.Lc[j + ldm * i] * r[j];
        r[i] = s / sm.Lc[i + ldm * i];
      }
    }
    }  // !boxed
    __syncthreads();
    for (int i = tid; i < m; i += P) {
      double s = 0.;
      for (int j = 0; j < m; ++j) s += sm.Quu[i + ldm * j] * sm.kg[j];
      sm.Quuk[i] = s;  // ddp.cpp:229
    }
    __syncthreads();
    // Vx = Qx + K^T Quuk - 2 K^T Qu ; Vxx = Qxx - Qxu K   (ddp.cpp:223-233)
    for (int j = tid; j < n; j += P) {
      double s1 = 0., s2 = 0.;
      for (int i = 0; i < m; ++i) {
        s1 += sm.Kg[i + ldm * j] * sm.Quuk[i];
        s2 += sm.Kg[i + ldm * j] * sm.Qu[i];
      }
      double v = sm.Qx[j];
      if (has_ureg) {
        v += s1;
        v += -2. * s2;
      } else {
        v += -1. * s2;
      }
      sm.Vx[j] = v;
    }
    cta_gemm_auto<false>(sm.V, ldn, sm.Qxu, ldn, sm.Kg, ldm, n, n, m, sm.V, ldn, -1.);
    if (tid == 0) {  // ddp.cpp:132-155 / fddp.cpp:137-140
      double a = 0., q = 0., st = 0.;
      for (int i = 0; i < m; ++i) {
        a += sm.Qu[i] * sm.kg[i];
        q += sm.kg[i] * sm.Quuk[i];
        st += sm.Qu[i] * sm.Qu[i];
      }
      dg += a;
      dq -= q;
      stop += st;
    }
    __syncthreads();
    // Vxx = .5 (Vxx + Vxx^T) ; diag += xreg   (ddp.cpp:235-239)
    for (int e = tid; e < n * n; e += P) {
      const int i = e % n, j = e / n;
      if (i > j) {
        const double s = 0.5 * (sm.V[i + ldn * j] + sm.V[j + ldn * i]);
        sm.V[i + ldn * j] = s;
        sm.V[j + ldn * i] = s;
      } else if (i == j) {
        double s = 0.5 * (sm.V[i + ldn * i] + sm.V[i + ldn * i]);
        if (has_xreg) s += xreg;
        sm.V[i + ldn * i] = s;
      }
    }
    __syncthreads();
    if (!is_feasible) {  // Vx += Vxx fs[t]   (ddp.cpp:242-244)
      for (int i = tid; i < n; i += P) {
        double s = 0.;
        for (int j = 0; j < n; ++j) s += sm.V[i + ldn * j] * fs[(size_t)t * n + j];
        sm.scr[i] = s;
        sm.Vx[i] += s;
      }
      __syncthreads();
    }
    // raiseIfNaN(|Vx|_inf), raiseIfNaN(|Vxx|_inf)   (ddp.cpp:246-251)
    int bad = 0;
    for (int i = tid; i < n; i += P) bad |= raise_if_nan(fabs(sm.Vx[i]));
    for (int e = tid; e < n * n; e += P) bad |= raise_if_nan(fabs(sm.V[(e % n) + ldn * (e / n)]));
    // stores
    for (int i = tid; i < n; i += P) {
      c.Vx[node * n + i] = sm.Vx[i];
      c.w[node * n + i] = is_feasible ? 0. : sm.scr[i];
    }
    for (int i = tid; i < m; i += P) c.k[rnode * m + i] = sm.kg[i];
    cta_unstage(c.Vxx + node * n * n, sm.V, ldn, n, n);
    cta_unstage(c.K + rnode * m * n, sm.Kg, ldm, m, n);
    if (tid == 0 && fddp && !is_feasible) {  // fddp.cpp:141-145
      double a = 0., q = 0.;
      for (int i = 0; i < n; ++i) {
        a += sm.Vx[i] * fs[(size_t)t * n + i];
        q += fs[(size_t)t * n + i] * sm.scr[i];
      }
      dg -= a;
      dq += q;
    }
    if (__syncthreads_or(bad)) return 1;
  }
  if (tid == 0) {
    sm.red[0] = dg;
    sm.red[1] = dq;
    sm.red[2] = stop;
  }
  __syncthreads();
  dg_out = sm.red[0];
  dq_out = sm.red[1];
  stop_out = sm.red[2];
  __syncthreads();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// SolverFDDP::forwardPass (fddp.cpp:149-225) / SolverDDP::forwardPass (ddp.cpp:255-296) for one problem by the
// whole block, fused with SolverFDDP::expectedImprovement's gap term dv = sum_t w_t^T (xs_try_t - xs_t)
// (fddp.cpp:107-124). Node data (xnext, cost) are overwritten by the trial exactly like the reference's datas.
// Returns 0 or 1 ("forward_error") to all threads; cost_try and dv valid on every thread.
// where a trial goes when it is not the problem's own trial buffers (concurrent line search): whole-batch arrays
struct TrialBufs {
  double* xs;     // [B][T+1][n]
  double* us;     // [B][T][m]
  double* xnext;  // [B][T+1][n]
  double* cost;   // [B][T+1]
};
template <class M>
__device__ int cta_forward(const CtaCtx& c, const M& mod, CtaSmem& sm, int b, int cur, double alpha, bool is_feasible,
                           double& cost_try_out, double& dv_out, const TrialBufs* tb = nullptr) {
  const int T = c.T, n = c.n, m = c.m;
  const int tid = threadIdx.x, P = blockDim.x;
  const bool fddp = c.fddp != 0;
  const bool closed = !fddp || is_feasible || (alpha == 1.);
  const bool use_w = fddp && !is_feasible;
  const double am1 = alpha - 1.;
  // SolverBoxDDP::forwardPass keeps the gap of the TERMINAL node only (box-ddp.cpp:121-125); both box solvers clamp
  // the control to the model's limits (box-ddp.cpp:95-97, box-fddp.cpp:96-98)
  const bool term_open = c.box && !fddp && !(is_feasible || (alpha == 1.));
  const bool clamp = c.box && c.has_limits[b] != 0;
  const double* u_lb = c.u_lb + (size_t)b * m;
  const double* u_ub = c.u_ub + (size_t)b * m;
  const double* xs = c.xs[cur] + (size_t)b * (T + 1) * n;
  const double* us = c.us[cur] + (size_t)b * T * m;
  double* xt_ = (tb ? tb->xs : c.xs[cur ^ 1]) + (size_t)b * (T + 1) * n;
  double* ut_ = (tb ? tb->us : c.us[cur ^ 1]) + (size_t)b * T * m;
  double* xnext_ = tb ? tb->xnext : c.d_xnext;
  double* cost_ = tb ? tb->cost : c.d_cost;
  const double* fs = c.fs + (size_t)b * (T + 1) * n;
  const double* w = c.w + (size_t)b * (T + 1) * n;
  // Models with compile-time node sizes (the arm): the next node's xs, fs, w, us, k, K are copied into shared memory
  // with cp.async while the current node's dynamics are evaluated. The rollout is serial in t and only a few warps
  // are resident per SM, so every global load in the loop is an exposed L2/HBM round trip otherwise (two per node).
  constexpr bool PF = M::kN > 0;
  constexpr int KN = M::kN, KM = M::kM;
  constexpr int PFS = PF ? 3 * KN + 2 * KM + KM * KN : 1;
  __shared__ double pf[2][PFS];
  auto prefetch = [&](int t, int buf) {
    double* d = pf[buf];
    for (int i = tid; i < KN; i += P) {
      cp_async8(d + i, xs + (size_t)t * KN + i);
      cp_async8(d + KN + i, fs + (size_t)t * KN + i);
      cp_async8(d + 2 * KN + i, w + (size_t)t * KN + i);
    }
    if (t < T) {
      const double* Kt = c.K + ((size_t)b * T + t) * KM * KN;
      for (int i = tid; i < KM; i += P) {
        cp_async8(d + 3 * KN + i, us + (size_t)t * KM + i);
        cp_async8(d + 3 * KN + KM + i, c.k + ((size_t)b * T + t) * KM + i);
      }
      for (int e = tid; e < KM * KN; e += P) cp_async8(d + 3 * KN + 2 * KM + e, Kt + e);
    }
  };
  double cost_try = 0., dv = 0.;  // thread 0
  for (int i = tid; i < n; i += P) sm.xn[i] = c.x0[(size_t)b * n + i];
  if constexpr (PF) prefetch(0, 0);
  __syncthreads();
  for (int t = 0; t <= T; ++t) {
    const double *xs_t = xs + (size_t)t * n, *fs_t = fs + (size_t)t * n, *w_t = w + (size_t)t * n;
    const double *us_t = us + (size_t)t * m, *Kt = c.K + ((size_t)b * T + t) * m * n, *kt = c.k + ((size_t)b * T + t) * m;
    if constexpr (PF) {
      cp_async_wait_all();
      __syncthreads();  // node t has landed for every thread; node t - 1's buffer is free again
      const double* d = pf[t & 1];
      xs_t = d;
      fs_t = d + KN;
      w_t = d + 2 * KN;
      us_t = d + 3 * KN;
      kt = d + 3 * KN + KM;
      Kt = d + 3 * KN + 2 * KM;
      if (t < T) prefetch(t + 1, (t + 1) & 1);
    }
    for (int i = tid; i < n; i += P) {
      const double xt = (closed && !(term_open && t == T)) ? sm.xn[i] : sm.xn[i] + fs_t[i] * am1;
      sm.x[i] = xt;
      sm.dx[i] = xt - xs_t[i];
      xt_[(size_t)t * n + i] = xt;
    }
    __syncthreads();
    if (t < T) {
      for (int i = tid; i < m; i += P) {
        double s = 0.;
        for (int j = 0; j < n; ++j) s += Kt[i + m * j] * sm.dx[j];
        double u = us_t[i] - kt[i] * alpha - s;
        if (clamp) u = fmin(fmax(u, u_lb[i]), u_ub[i]);
        sm.u[i] = u;
        ut_[(size_t)t * m + i] = u;
      }
    }
    if (tid == 0 && use_w) {
      double s = 0.;
      for (int i = 0; i < n; ++i) s += w_t[i] * sm.dx[i];
      dv += s;
    }
    __syncthreads();
    mod.calc_block(sm.x, t < T ? sm.u : nullptr, sm.xn, sm.scr, &sm.red[3]);
    // datas[t]->xnext / cost (overwritten by every trial)
    const size_t node = (size_t)b * (T + 1) + t;
    for (int i = tid; i < n; i += P) xnext_[node * n + i] = sm.xn[i];
    int bad = 0;
    if (tid == 0) {
      cost_[node] = sm.red[3];
      cost_try += sm.red[3];
      bad = raise_if_nan(cost_try);
    }
    if (t < T)
      for (int i = tid; i < n; i += P) bad |= raise_if_nan(fabs(sm.xn[i]));
    if (__syncthreads_or(bad)) {
      if constexpr (PF) cp_async_wait_all();  // nothing in flight into shared memory when the caller moves on
      return 1;
    }
  }
  if (tid == 0) {
    sm.red[0] = cost_try;
    sm.red[1] = dv;
  }
  __syncthreads();
  cost_try_out = sm.red[0];
  dv_out = sm.red[1];
  __syncthreads();
  return 0;
}

// The same rollout by ONE THREAD per problem, for models with compile-time node sizes (M::kN > 0, the arm). One warp
// per problem spends most of its lanes waiting on the serial joint recursions of ABA (ncu, B = 8192: 11 of 32 lanes
// active per instruction, 3980 warp instructions per node, 7.8 ms per trial of the whole batch); with one thread per
// problem a warp instruction advances 32 rollouts, so a large batch needs ~20x fewer of them. The latency of a lone
// rollout is about the same, so the host uses this version for large batches only (CtaBackend::solve_phases).
// Same operation order per problem as cta_forward; the node function is the single-thread M::calc_thread, whose 6x6
// products are summed in a different order than the warp version's (rounding level, arm_warp.cuh).
template <class M>
__device__ int cta_forward_thread(const CtaCtx& c, const M& mod, int b, int cur, double alpha, bool is_feasible,
                                  double& cost_try_out, double& dv_out, const TrialBufs* tb = nullptr) {
  constexpr int n = M::kN > 0 ? M::kN : 1, m = M::kM > 0 ? M::kM : 1;
  const int T = c.T;
  const bool fddp = c.fddp != 0;
  const bool closed = !fddp || is_feasible || (alpha == 1.);
  const bool use_w = fddp && !is_feasible;
  const double am1 = alpha - 1.;
  const bool term_open = c.box && !fddp && !(is_feasible || (alpha == 1.));
  const bool clamp = c.box && c.has_limits[b] != 0;
  const double* u_lb = c.u_lb + (size_t)b * m;
  const double* u_ub = c.u_ub + (size_t)b * m;
  const double* xs = c.xs[cur] + (size_t)b * (T + 1) * n;
  const double* us = c.us[cur] + (size_t)b * T * m;
  double* xt_ = (tb ? tb->xs : c.xs[cur ^ 1]) + (size_t)b * (T + 1) * n;
  double* ut_ = (tb ? tb->us : c.us[cur ^ 1]) + (size_t)b * T * m;
  double* xnext_ = (tb ? tb->xnext : c.d_xnext) + (size_t)b * (T + 1) * n;
  double* cost_ = (tb ? tb->cost : c.d_cost) + (size_t)b * (T + 1);
  const double* fs = c.fs + (size_t)b * (T + 1) * n;
  const double* w = c.w + (size_t)b * (T + 1) * n;
  const double* Kb = c.K + (size_t)b * T * m * n;
  const double* kb = c.k + (size_t)b * T * m;
  double cost_try = 0., dv = 0.;
  double xn[n], x[n], dx[n], u[m];
#pragma unroll
  for (int i = 0; i < n; ++i) xn[i] = c.x0[(size_t)b * n + i];
  for (int t = 0; t <= T; ++t) {
    const bool open = !(closed && !(term_open && t == T));
#pragma unroll
    for (int i = 0; i < n; ++i) {
      const double xt = open ? xn[i] + fs[(size_t)t * n + i] * am1 : xn[i];
      x[i] = xt;
      dx[i] = xt - xs[(size_t)t * n + i];
      xt_[(size_t)t * n + i] = xt;
    }
    if (t < T) {
      const double* Kt = Kb + (size_t)t * m * n;
#pragma unroll
      for (int i = 0; i < m; ++i) {
        double s = 0.;
#pragma unroll
        for (int j = 0; j < n; ++j) s += Kt[i + m * j] * dx[j];
        double uu = us[(size_t)t * m + i] - kb[(size_t)t * m + i] * alpha - s;
        if (clamp) uu = fmin(fmax(uu, u_lb[i]), u_ub[i]);
        u[i] = uu;
        ut_[(size_t)t * m + i] = uu;
      }
    }
    if (use_w) {
      double s = 0.;
#pragma unroll
      for (int i = 0; i < n; ++i) s += w[(size_t)t * n + i] * dx[i];
      dv += s;
    }
    const double cst = mod.calc_thread(x, t < T ? u : nullptr, xn);
#pragma unroll
    for (int i = 0; i < n; ++i) xnext_[(size_t)t * n + i] = xn[i];
    cost_[t] = cst;
    cost_try += cst;
    bool bad = raise_if_nan(cost_try);
    if (t < T) {
#pragma unroll
      for (int i = 0; i < n; ++i) bad = bad || raise_if_nan(fabs(xn[i]));
    }
    if (bad) return 1;
  }
  cost_try_out = cost_try;
  dv_out = dv;
  return 0;
}

__device__ inline bool cta_accept(const SolverOpts& o, bool fddp, bool is_feasible, double d0, double dV, double dVexp) {
  if (fddp) {
    if (dVexp >= 0) return d0 < o.th_grad || dV > o.th_acceptstep * dVexp;
    return dV > o.th_acceptnegstep * dVexp;
  }
  if (dVexp >= 0) return d0 < o.th_grad || !is_feasible || dV > o.th_acceptstep * dVexp;
  return false;
}

#define CTA_S(slot) c.scal[(size_t)(slot) * c.Bp + b]
#define CTA_I(slot) c.iscal[(size_t)(slot) * c.Bp + b]

// ---------------------------------------------------------------------------------------------
// Whole SolverDDP::solve / SolverFDDP::solve loop for one problem per block (ddp.cpp:41-118, fddp.cpp:19-105).
// All control-flow variables are block-uniform (every thread computes the same scalars from broadcast values).
template <class M>
__global__ void __launch_bounds__(M::kMaxThreads, M::kMinBlocks) cta_solve_kernel(const __grid_constant__ CtaCtx c, int maxiter, int is_feasible0, double reginit) {
  extern __shared__ double cta_smem_raw[];
  const int b = blockIdx.x;
  CtaSmem sm;
  sm.carve(cta_smem_raw, c.n, c.m);
  M mod;
  mod.load(c, b);
  const bool fddp = c.fddp != 0;
  int cur = CTA_I(I_CUR);
  bool feas = is_feasible0 != 0, wasfeas = false;
  double xreg = isnan(reginit) ? c.o.regmin : reginit, ureg = xreg;
  double cost = CTA_S(S_COST), stop = CTA_S(S_STOP), d0 = CTA_S(S_D0), d1 = CTA_S(S_D1);
  double steplength = CTA_S(S_STEP), dV = CTA_S(S_DV), dVexp = CTA_S(S_DVEXP), cost_try = CTA_S(S_COST_TRY);
  double dg = 0., dq = 0.;
  int nbwd = 0, nfwd = 0, nrec = 0, converged = 0, iter = 0;
  bool recalc = true, done = false;
  __syncthreads();
  for (iter = 0; iter < maxiter && !done; ++iter) {
    int retries = 0;
    while (true) {  // computeDirection(recalcDiff) with regularisation retry (ddp.cpp:55-68)
      if (recalc) cost = cta_calcdiff(c, mod, sm, b, cur, iter == 0, feas, wasfeas);
      const int st = cta_backward<M, true>(c, mod, sm, b, cur, feas, xreg, ureg, dg, dq, stop);
      ++nbwd;
      if (st == 0) break;
      recalc = false;
      reg_increase(c.o, xreg, ureg);
      ++retries;
      if (xreg == c.o.regmax) {
        done = true;
        break;
      }
    }
    if (done) break;
    if (!fddp) {
      d0 = dg;
      d1 = dq;
    }
    recalc = false;
    int accepted = -1;
    for (int a = 0; a < c.o.nalphas; ++a) {
      steplength = c.o.alphas[a];
      double dv = 0.;
      const int st = cta_forward(c, mod, sm, b, cur, steplength, feas, cost_try, dv);
      ++nfwd;
      if (st) continue;
      dV = cost - cost_try;
      if (fddp) {
        d0 = dg + dv;
        d1 = dq - 2 * dv;
      }
      dVexp = steplength * (d0 + 0.5 * steplength * d1);
      if (cta_accept(c.o, fddp, feas, d0, dV, dVexp)) {
        wasfeas = feas;
        feas = fddp ? (wasfeas || (steplength == 1.)) : true;
        cost = cost_try;
        cur ^= 1;
        recalc = true;
        accepted = a;
        break;
      }
    }
    if (steplength > c.o.th_stepdec) reg_decrease(c.o, xreg, ureg);
    if (steplength <= c.o.th_stepinc) {
      reg_increase(c.o, xreg, ureg);
      if (xreg == c.o.regmax) break;
    }
    if (c.trace && nrec < c.trace_cap && threadIdx.x == 0) {
      double* tr = c.trace + ((size_t)b * c.trace_cap + nrec) * NTRACE;
      tr[0] = iter;
      tr[1] = cost;
      tr[2] = stop;
      tr[3] = d0;
      tr[4] = d1;
      tr[5] = xreg;
      tr[6] = ureg;
      tr[7] = steplength;
      tr[8] = feas ? 1. : 0.;
      tr[9] = dV;
      tr[10] = dVexp;
      tr[11] = accepted;
      tr[12] = retries;
    }
    ++nrec;
    if (wasfeas && stop < c.o.th_stop) {
      converged = 1;
      break;
    }
  }
  if (threadIdx.x == 0) {
    CTA_S(S_COST) = cost;
    CTA_S(S_STOP) = stop;
    CTA_S(S_D0) = d0;
    CTA_S(S_D1) = d1;
    CTA_S(S_XREG) = xreg;
    CTA_S(S_UREG) = ureg;
    CTA_S(S_STEP) = steplength;
    CTA_S(S_DV) = dV;
    CTA_S(S_DVEXP) = dVexp;
    CTA_S(S_COST_TRY) = cost_try;
    CTA_S(S_DG) = dg;
    CTA_S(S_DQ) = dq;
    CTA_I(I_ITER) = iter;
    CTA_I(I_FEAS) = feas;
    CTA_I(I_WASFEAS) = wasfeas;
    CTA_I(I_CONVERGED) = converged;
    CTA_I(I_CUR) = cur;
    CTA_I(I_NREC) = nrec;
    CTA_I(I_NBWD) = nbwd;
    CTA_I(I_NFWD) = nfwd;
  }
}

// ---------------------------------------------------------------------------------------------
// Phase-split solve: the same loop as cta_solve_kernel cut at its two heavy calls, one launch per phase for the whole
// batch, driven by the host (CtaBackend::solve_phases) until every problem has returned. Why: the fused arm kernel is
// ~50k SASS instructions; with every warp of an SM at a different place of the loop the instruction cache thrashes
// (ncu: stall_no_instruction = half of all stall cycles) and the fused kernel measured 2.3x slower than the sum of
// its phases run in lockstep. The loop variables live in the per-problem slots between launches; the arithmetic is
// the same device functions in the same order, so results are identical to the fused kernel.
//   cta_phase_direction_kernel : problems in phase 0 — computeDirection incl. ONE regularisation retry per launch
//   cta_phase_forward_kernel   : problems in phase 1 — ONE trial of the line search + the end-of-iteration bookkeeping
//   cta_phase_forward_wide_kernel + cta_phase_select_kernel : the same with CTA_LS_WIDTH trials at once (tail of a batch)
static __global__ void cta_phase_init_kernel(const __grid_constant__ CtaCtx c, int maxiter, int is_feasible0, double reginit) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const double xreg = isnan(reginit) ? c.o.regmin : reginit;
  CTA_S(S_XREG) = xreg;
  CTA_S(S_UREG) = xreg;
  CTA_S(S_DG) = 0.;
  CTA_S(S_DQ) = 0.;
  CTA_I(I_FEAS) = is_feasible0 != 0;
  CTA_I(I_WASFEAS) = 0;
  CTA_I(I_ITER) = 0;
  CTA_I(I_NREC) = 0;
  CTA_I(I_NBWD) = 0;
  CTA_I(I_NFWD) = 0;
  CTA_I(I_CONVERGED) = 0;
  CTA_I(I_LOOP) = 1;  // recalcDiff
  CTA_I(I_ALPHA) = 0;
  CTA_I(I_RETRIES) = 0;
  CTA_I(I_PHASE) = maxiter > 0 ? 0 : 2;
}

#ifndef CGX_NODES_MINBLOCKS
#define CGX_NODES_MINBLOCKS 4
#endif
// calc (first iteration only) + calcDiff of every node of every problem whose next step is computeDirection(recalc =
// true): one thread per (problem, node) pair over the whole batch, so a lone problem gets T + 1 threads instead of one
// warp's four passes and every warp is full (T + 1 = 101 = 3 x 32 + 5 otherwise)
template <class M>
__global__ void __launch_bounds__(64, CGX_NODES_MINBLOCKS) cta_phase_nodes_kernel(const __grid_constant__ CtaCtx c) {
  const int nt = c.T + 1;
  const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (long)c.B * nt) return;
  const int b = (int)(g / nt), t = (int)(g % nt);
  if (CTA_I(I_PHASE) != 0 || !(CTA_I(I_LOOP) & 1)) return;
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR);
  const double* xs = c.xs[cur] + (size_t)b * nt * c.n;
  const double* us = c.us[cur] + (size_t)b * c.T * c.m;
  if (CTA_I(I_ITER) == 0) mod.calc_node(c, b, t, xs, us);
  mod.calcdiff_node(c, b, t, xs, us);
}

template <class M>
__global__ void __launch_bounds__(M::kMaxThreads, M::kDirMinBlocks) cta_phase_direction_kernel(const __grid_constant__ CtaCtx c, int* active) {
  extern __shared__ double cta_smem_raw[];
  const int b = blockIdx.x;
  const int phase = CTA_I(I_PHASE);
  if (phase != 2 && threadIdx.x == 0) atomicAdd(active, 1);  // problems whose solve() has not returned yet
  if (phase != 0) return;
  CtaSmem sm;
  sm.carve(cta_smem_raw, c.n, c.m);
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR), iter = CTA_I(I_ITER);
  const bool feas = CTA_I(I_FEAS) != 0, wasfeas = CTA_I(I_WASFEAS) != 0, recalc = (CTA_I(I_LOOP) & 1) != 0;
  double xreg = CTA_S(S_XREG), ureg = CTA_S(S_UREG);
  double dg = CTA_S(S_DG), dq = CTA_S(S_DQ), stop = CTA_S(S_STOP);  // kept when the backward pass fails
  __syncthreads();
  double cost = 0.;
  if (recalc) cost = cta_calcdiff<M, M::kNodeKernel>(c, mod, sm, b, cur, iter == 0, feas, wasfeas);
  const int st = cta_backward(c, mod, sm, b, cur, feas, xreg, ureg, dg, dq, stop);
  if (threadIdx.x != 0) return;
  if (recalc) CTA_S(S_COST) = cost;
  ++CTA_I(I_NBWD);
  CTA_I(I_LOOP) = 0;
  CTA_S(S_DG) = dg;
  CTA_S(S_DQ) = dq;
  CTA_S(S_STOP) = stop;
  if (st != 0) {  // ddp.cpp:55-68: increase the regularisation and try again (next launch)
    reg_increase(c.o, xreg, ureg);
    CTA_S(S_XREG) = xreg;
    CTA_S(S_UREG) = ureg;
    ++CTA_I(I_RETRIES);
    if (xreg == c.o.regmax) CTA_I(I_PHASE) = 2;
    return;
  }
  if (!c.fddp) {
    CTA_S(S_D0) = dg;
    CTA_S(S_D1) = dq;
  }
  CTA_I(I_ALPHA) = 0;
  CTA_I(I_PHASE) = 1;
}

// What SolverDDP/FDDP::solve does with the result of ONE trial of the line search (ddp.cpp:72-116, fddp.cpp:56-103),
// by one thread: the acceptance test, and — when the step is accepted or the step lengths are exhausted — the end of
// the iteration (regularisation update, trace record, stopping test). st / cost_try / dv: what cta_forward returned for
// step length index I_ALPHA. Leaves I_PHASE = 1 with I_ALPHA advanced when the next step length has to be tried.
__device__ inline void cta_phase_trial_done(const CtaCtx& c, int b, int maxiter, int st, double cost_try_new, double dv) {
  const bool fddp = c.fddp != 0;
  int cur = CTA_I(I_CUR);
  bool feas = CTA_I(I_FEAS) != 0, wasfeas = CTA_I(I_WASFEAS) != 0;
  const int a = CTA_I(I_ALPHA);
  const double steplength = c.o.alphas[a];
  double cost = CTA_S(S_COST), d0 = CTA_S(S_D0), d1 = CTA_S(S_D1);
  double dV = CTA_S(S_DV), dVexp = CTA_S(S_DVEXP);
  const double dg = CTA_S(S_DG), dq = CTA_S(S_DQ), stop = CTA_S(S_STOP);
  double xreg = CTA_S(S_XREG), ureg = CTA_S(S_UREG);
  ++CTA_I(I_NFWD);
  int accepted = -1;
  bool recalc = false;
  if (!st) {
    const double cost_try = cost_try_new;
    CTA_S(S_COST_TRY) = cost_try;
    dV = cost - cost_try;
    if (fddp) {
      d0 = dg + dv;
      d1 = dq - 2 * dv;
    }
    dVexp = steplength * (d0 + 0.5 * steplength * d1);
    if (cta_accept(c.o, fddp, feas, d0, dV, dVexp)) {
      wasfeas = feas;
      feas = fddp ? (wasfeas || (steplength == 1.)) : true;
      cost = cost_try;
      cur ^= 1;
      recalc = true;
      accepted = a;
    }
  }
  CTA_S(S_STEP) = steplength;
  CTA_S(S_DV) = dV;
  CTA_S(S_DVEXP) = dVexp;
  CTA_S(S_D0) = d0;
  CTA_S(S_D1) = d1;
  if (accepted < 0 && a + 1 < c.o.nalphas) {  // next step length
    CTA_I(I_ALPHA) = a + 1;
    return;
  }
  CTA_S(S_COST) = cost;
  CTA_I(I_CUR) = cur;
  CTA_I(I_FEAS) = feas;
  CTA_I(I_WASFEAS) = wasfeas;
  CTA_I(I_LOOP) = recalc ? 1 : 0;
  // end of the iteration: ddp.cpp:93-116 / fddp.cpp:80-103
  if (steplength > c.o.th_stepdec) reg_decrease(c.o, xreg, ureg);
  bool done = false;
  if (steplength <= c.o.th_stepinc) {
    reg_increase(c.o, xreg, ureg);
    if (xreg == c.o.regmax) done = true;
  }
  CTA_S(S_XREG) = xreg;
  CTA_S(S_UREG) = ureg;
  int iter = CTA_I(I_ITER);
  if (!done) {
    const int nrec = CTA_I(I_NREC);
    if (c.trace && nrec < c.trace_cap) {
      double* tr = c.trace + ((size_t)b * c.trace_cap + nrec) * NTRACE;
      tr[0] = iter;
      tr[1] = cost;
      tr[2] = stop;
      tr[3] = d0;
      tr[4] = d1;
      tr[5] = xreg;
      tr[6] = ureg;
      tr[7] = steplength;
      tr[8] = feas ? 1. : 0.;
      tr[9] = dV;
      tr[10] = dVexp;
      tr[11] = accepted;
      tr[12] = CTA_I(I_RETRIES);
    }
    CTA_I(I_NREC) = nrec + 1;
    if (wasfeas && stop < c.o.th_stop) {
      CTA_I(I_CONVERGED) = 1;
      done = true;
    } else {
      ++iter;
      if (iter >= maxiter) done = true;
    }
  }
  CTA_I(I_ITER) = iter;
  CTA_I(I_RETRIES) = 0;
  CTA_I(I_PHASE) = done ? 2 : 0;
}

template <class M>
__global__ void __launch_bounds__(M::kMaxThreads, M::kFwdMinBlocks) cta_phase_forward_kernel(const __grid_constant__ CtaCtx c, int maxiter) {
  extern __shared__ double cta_smem_raw[];
  const int b = blockIdx.x;
  if (CTA_I(I_PHASE) != 1) return;
  CtaSmem sm;
  sm.carve(cta_smem_raw, c.n, c.m);
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR);
  const bool feas = CTA_I(I_FEAS) != 0;
  const double steplength = c.o.alphas[CTA_I(I_ALPHA)];
  __syncthreads();
  double cost_try = 0., dv = 0.;
  const int st = cta_forward(c, mod, sm, b, cur, steplength, feas, cost_try, dv);
  if (threadIdx.x == 0) cta_phase_trial_done(c, b, maxiter, st, cost_try, dv);
}

// the same phase with one thread per problem (cta_forward_thread): large batches of compile-time-sized models
template <class M>
__global__ void __launch_bounds__(32) cta_phase_forward_tpp_kernel(const __grid_constant__ CtaCtx c, int maxiter) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  if (CTA_I(I_PHASE) != 1) return;
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR);
  const bool feas = CTA_I(I_FEAS) != 0;
  const double steplength = c.o.alphas[CTA_I(I_ALPHA)];
  double cost_try = 0., dv = 0.;
  const int st = cta_forward_thread(c, mod, b, cur, steplength, feas, cost_try, dv);
  cta_phase_trial_done(c, b, maxiter, st, cost_try, dv);
}

// Concurrent line search (the trials of one iteration are independent given xs, us, K, k): block (b, g)
// rolls out step length index I_ALPHA + g of problem b. Slot 0 writes the problem's own trial buffers, slot g > 0 the
// g-th spare set; nothing of the solver state changes here. cta_phase_select_kernel then walks the slots in the
// reference's order, so the accepted step, the counters and the trace are those of the sequential search; the
// trajectories and node data of the slot the search ended on become the problem's trial. Used by the host while the
// unfinished problems x width fit the GPU in one wave (the tail of a batch), where it shortens the critical path of
// the slowest problems from (trials) to (1) rollouts per iteration; with the machine full it would only add work.
constexpr int CTA_LS_WIDTH = 4;
struct TrialSlots {
  TrialBufs spare[CTA_LS_WIDTH - 1];
  double* res;  // [CTA_LS_WIDTH][3][Bp]: status, cost_try, dv of every slot
};
template <class M>
__global__ void __launch_bounds__(M::kMaxThreads, M::kFwdMinBlocks) cta_phase_forward_wide_kernel(const __grid_constant__ CtaCtx c, const __grid_constant__ TrialSlots ts) {
  extern __shared__ double cta_smem_raw[];
  const int b = blockIdx.x, g = blockIdx.y;
  if (CTA_I(I_PHASE) != 1) return;
  const int a = CTA_I(I_ALPHA) + g;
  if (a >= c.o.nalphas) return;
  CtaSmem sm;
  sm.carve(cta_smem_raw, c.n, c.m);
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR);
  const bool feas = CTA_I(I_FEAS) != 0;
  const double steplength = c.o.alphas[a];
  __syncthreads();
  double cost_try = 0., dv = 0.;
  const int st = cta_forward(c, mod, sm, b, cur, steplength, feas, cost_try, dv, g ? &ts.spare[g - 1] : nullptr);
  if (threadIdx.x == 0) {
    double* r = ts.res + (size_t)g * 3 * c.Bp;
    r[b] = st;
    r[(size_t)c.Bp + b] = cost_try;
    r[(size_t)2 * c.Bp + b] = dv;
  }
}
template <class M>
__global__ void __launch_bounds__(32) cta_phase_forward_wide_tpp_kernel(const __grid_constant__ CtaCtx c, const __grid_constant__ TrialSlots ts) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x, g = blockIdx.y;
  if (b >= c.B) return;
  if (CTA_I(I_PHASE) != 1) return;
  const int a = CTA_I(I_ALPHA) + g;
  if (a >= c.o.nalphas) return;
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR);
  const bool feas = CTA_I(I_FEAS) != 0;
  double cost_try = 0., dv = 0.;
  const int st = cta_forward_thread(c, mod, b, cur, c.o.alphas[a], feas, cost_try, dv, g ? &ts.spare[g - 1] : nullptr);
  double* r = ts.res + (size_t)g * 3 * c.Bp;
  r[b] = st;
  r[(size_t)c.Bp + b] = cost_try;
  r[(size_t)2 * c.Bp + b] = dv;
}
static __global__ void cta_phase_select_kernel(const __grid_constant__ CtaCtx c, const __grid_constant__ TrialSlots ts, int maxiter) {
  const int b = blockIdx.x;
  __shared__ int s_slot, s_cur;
  if (CTA_I(I_PHASE) != 1) return;
  __syncthreads();  // every thread has seen phase 1 before thread 0 changes it
  if (threadIdx.x == 0) {
    s_cur = CTA_I(I_CUR);
    int slot = -1;
    for (int g = 0; g < CTA_LS_WIDTH; ++g) {
      if (CTA_I(I_ALPHA) >= c.o.nalphas) break;
      const double* r = ts.res + (size_t)g * 3 * c.Bp;
      cta_phase_trial_done(c, b, maxiter, (int)r[b], r[(size_t)c.Bp + b], r[(size_t)2 * c.Bp + b]);
      if (CTA_I(I_PHASE) != 1) {  // the iteration ended on this slot
        slot = g;
        break;
      }
    }
    s_slot = slot;
  }
  __syncthreads();
  const int slot = s_slot;
  if (slot <= 0) return;  // slot 0 is the problem's own trial buffers; -1: the search goes on
  const size_t T = c.T, n = c.n, m = c.m;
  const TrialBufs& sp = ts.spare[slot - 1];
  double* xt = c.xs[s_cur ^ 1] + (size_t)b * (T + 1) * n;
  double* ut = c.us[s_cur ^ 1] + (size_t)b * T * m;
  for (size_t e = threadIdx.x; e < (T + 1) * n; e += blockDim.x) {
    xt[e] = sp.xs[(size_t)b * (T + 1) * n + e];
    c.d_xnext[(size_t)b * (T + 1) * n + e] = sp.xnext[(size_t)b * (T + 1) * n + e];
  }
  for (size_t e = threadIdx.x; e < T * m; e += blockDim.x) ut[e] = sp.us[(size_t)b * T * m + e];
  for (size_t e = threadIdx.x; e <= T; e += blockDim.x) c.d_cost[(size_t)b * (T + 1) + e] = sp.cost[(size_t)b * (T + 1) + e];
}

// ---------------------------------------------------------------------------------------------
// Step-wise kernels. what: 0 = calcDiff only, 1 = backwardPass only, 2 = computeDirection (calcDiff if recalc, then
// backwardPass) — ddp.cpp:120-125,157-253. The step-wise API runs with iter_ == 0, so problem_->calc is evaluated.
template <class M>
__global__ void __launch_bounds__(M::kMaxThreads, M::kMinBlocks) cta_direction_kernel(const __grid_constant__ CtaCtx c, int what, int recalc) {
  extern __shared__ double cta_smem_raw[];
  const int b = blockIdx.x;
  CtaSmem sm;
  sm.carve(cta_smem_raw, c.n, c.m);
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR);
  const bool feas = CTA_I(I_FEAS) != 0, wasfeas = CTA_I(I_WASFEAS) != 0;
  __syncthreads();
  if (what == 0 || (what == 2 && recalc)) {
    const double cost = cta_calcdiff(c, mod, sm, b, cur, true, feas, wasfeas);
    if (threadIdx.x == 0) CTA_S(S_COST) = cost;
  }
  if (what == 0) return;
  double dg = 0., dq = 0., stop = 0.;
  const int st = cta_backward(c, mod, sm, b, cur, feas, CTA_S(S_XREG), CTA_S(S_UREG), dg, dq, stop);
  if (threadIdx.x == 0) {
    CTA_I(I_BSTATUS) = st;
    ++CTA_I(I_NBWD);
    if (st == 0) {
      CTA_S(S_DG) = dg;
      CTA_S(S_DQ) = dq;
      CTA_S(S_STOP) = stop;
      if (!c.fddp) {
        CTA_S(S_D0) = dg;
        CTA_S(S_D1) = dq;
      }
    }
  }
}

template <class M>
__global__ void __launch_bounds__(M::kMaxThreads, M::kMinBlocks) cta_trystep_kernel(const __grid_constant__ CtaCtx c, double alpha) {
  extern __shared__ double cta_smem_raw[];
  const int b = blockIdx.x;
  CtaSmem sm;
  sm.carve(cta_smem_raw, c.n, c.m);
  M mod;
  mod.load(c, b);
  const int cur = CTA_I(I_CUR);
  const bool feas = CTA_I(I_FEAS) != 0;
  __syncthreads();
  double cost_try = 0., dv = 0.;
  const int st = cta_forward(c, mod, sm, b, cur, alpha, feas, cost_try, dv);
  if (threadIdx.x == 0) {
    CTA_I(I_FSTATUS) = st;
    if (st == 0) {
      CTA_S(S_COST_TRY) = cost_try;
      CTA_S(S_DV) = CTA_S(S_COST) - cost_try;
      CTA_S(S_DVGAP) = dv;
      CTA_S(S_STEP) = alpha;
      if (c.fddp) {
        CTA_S(S_D0) = CTA_S(S_DG) + dv;
        CTA_S(S_D1) = CTA_S(S_DQ) - 2 * dv;
      }
    }
  }
}

// ShootingProblem::calc / calcDiff on caller-provided trajectories (user layout): what 0 = calc, 1 = calcDiff
template <class M>
__global__ void cta_problem_kernel(const __grid_constant__ CtaCtx c, const double* xs, const double* us, int diff,
                                   double* costs) {
  const int b = blockIdx.x;
  M mod;
  mod.load(c, b);
  const double* x = xs + (size_t)b * (c.T + 1) * c.n;
  const double* u = us + (size_t)b * c.T * c.m;
  if (diff)
    mod.calcdiff_all(c, b, x, u);
  else
    mod.calc_all(c, b, x, u);
  __syncthreads();
  if (threadIdx.x == 0) {
    double cost = 0.;
    for (int t = 0; t <= c.T; ++t) cost += c.d_cost[(size_t)b * (c.T + 1) + t];
    costs[b] = cost;
  }
}

// ShootingProblem::rollout (shooting.hxx:198-223) by the block
template <class M>
__global__ void __launch_bounds__(M::kMaxThreads, M::kMinBlocks) cta_rollout_kernel(const __grid_constant__ CtaCtx c, const double* us, double* xs) {
  extern __shared__ double cta_smem_raw[];
  const int b = blockIdx.x, T = c.T, n = c.n, m = c.m;
  CtaSmem sm;
  sm.carve(cta_smem_raw, n, m);
  M mod;
  mod.load(c, b);
  for (int i = threadIdx.x; i < n; i += blockDim.x) sm.x[i] = c.x0[(size_t)b * n + i];
  __syncthreads();
  for (int t = 0; t <= T; ++t) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) xs[((size_t)b * (T + 1) + t) * n + i] = sm.x[i];
    if (t < T)
      for (int i = threadIdx.x; i < m; i += blockDim.x) sm.u[i] = us[((size_t)b * T + t) * m + i];
    __syncthreads();
    mod.calc_block(sm.x, t < T ? sm.u : nullptr, sm.xn, sm.scr, &sm.red[3]);
    const size_t node = (size_t)b * (T + 1) + t;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      c.d_xnext[node * n + i] = sm.xn[i];
      if (t < T) sm.x[i] = sm.xn[i];
    }
    if (threadIdx.x == 0) c.d_cost[node] = sm.red[3];
    __syncthreads();
  }
}

// per-problem bookkeeping of setCandidate + scalar export
static __global__ void cta_reset_kernel(CtaCtx c, int is_feasible) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  CTA_I(I_ITER) = 0;
  CTA_I(I_FEAS) = is_feasible;
  CTA_I(I_WASFEAS) = 0;
  CTA_I(I_CONVERGED) = 0;
  CTA_I(I_CUR) = 0;
  CTA_I(I_NREC) = 0;
  CTA_I(I_BSTATUS) = 0;
  CTA_I(I_FSTATUS) = 0;
  CTA_I(I_NBWD) = 0;
  CTA_I(I_NFWD) = 0;
  CTA_S(S_DVGAP) = 0.;
}
static __global__ void cta_scalars_kernel(CtaCtx c, double* out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  double* o = out + (size_t)b * 17;
  o[0] = CTA_S(S_COST);
  o[1] = CTA_S(S_STOP);
  o[2] = CTA_S(S_D0);
  o[3] = CTA_S(S_D1);
  o[4] = CTA_S(S_XREG);
  o[5] = CTA_S(S_UREG);
  o[6] = CTA_S(S_STEP);
  o[7] = CTA_S(S_DV);
  o[8] = CTA_S(S_DVEXP);
  o[9] = CTA_I(I_ITER);
  o[10] = CTA_I(I_FEAS);
  o[11] = CTA_I(I_CONVERGED);
  o[12] = CTA_I(I_NREC);
  o[13] = CTA_S(S_COST_TRY);
  o[14] = CTA_I(I_WASFEAS);
  o[15] = CTA_I(I_NBWD);
  o[16] = CTA_I(I_NFWD);
}
static __global__ void cta_setreg_kernel(CtaCtx c, double xreg, double ureg, int set_x, int set_u) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  if (set_x) CTA_S(S_XREG) = xreg;
  if (set_u) CTA_S(S_UREG) = ureg;
}
// out[b][...] = field of problems b0.. : per-problem contiguous copy, selecting the candidate/trial buffer
static __global__ void cta_copy_sel_kernel(double* out, const double* a0, const double* a1, const int* sel, int flip,
                                    size_t per_problem, int b0, int to_field) {
  const int b = b0 + blockIdx.y;
  const double* src = (sel && ((sel[b] ^ flip) & 1)) ? a1 : a0;
  double* fld = const_cast<double*>(src) + (size_t)b * per_problem;
  double* usr = out + (size_t)blockIdx.y * per_problem;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < per_problem; e += (size_t)gridDim.x * blockDim.x) {
    if (to_field)
      fld[e] = usr[e];
    else
      usr[e] = fld[e];
  }
}

}  // namespace cgx
