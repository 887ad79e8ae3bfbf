// crocoddyl_b200 — thread-per-problem (TPP) DDP/FDDP core for tiny node models (nx <= ~4).
//
// One CUDA thread owns one OCP and runs the reference's whole solve() loop on the device
// (per-problem state machine, no host round trip per iteration).
//
// HBM layout ("lane-sector" records): every per-node quantity of problem b lives in a small record made of
// whole 32-byte sectors, records of consecutive problems are adjacent:
//     rec(t, b) = base + (t*Bp + b) * R        (R = record size in doubles, a multiple of 4)
//   XU[2] : { x[N], u[U], aux | fs[N] } candidate / trial (double-buffered, picked by I_CUR); aux = what calc leaves
//           for calcDiff (unicycle: sin/cos of the heading), fs in its own sector(s)
//   KK    : { K[U*N] column-major, k[U] }
//   V     : { w[N] = Vxx fs, Vx[N], lower triangle of Vxx (column-major) }
//   Q     : { Qxx, Qxu, Quu, Qx, Qu }  only with full storage
// A lane moves only whole sectors with 256-bit LDG/STG, so (i) a full warp reads/writes 32 adjacent sectors
// (1 KB, coalesced) and (ii) a warp that diverged — each OCP has its own iteration count — still uses every
// byte of every sector it touches. (ncu on the first batch-fastest SoA version: 2.6x DRAM traffic over the
// algorithmic bytes at 9.9 active lanes per warp instruction; profiles/.)
//
// Node derivatives (Fx,Fu,L**) and the gaps are generated INSIDE the backward sweep from (xs,us);
// they never touch HBM. The accepted trial becomes the candidate by flipping I_CUR (no copy).
//
// Reference functions restated here (paths relative to the reference tree):
//   SolverDDP::calcDiff/backwardPass/computeGains  src/core/solvers/ddp.cpp:157-253,298-310
//   SolverDDP::forwardPass/expectedImprovement/stoppingCriteria  ddp.cpp:132-155,255-296
//   SolverFDDP::forwardPass/(update)expectedImprovement  src/core/solvers/fddp.cpp:107-225
//   SolverDDP::solve / SolverFDDP::solve  ddp.cpp:41-118, fddp.cpp:19-105
#pragma once
#include "boxqp.cuh"
#include "common.cuh"

namespace cgx {

struct TppCtx {
  int B, Bp, T;
  int fddp;       // 1 = SolverFDDP, 0 = SolverDDP
  int full;       // keep Qxx,Qxu,Quu,Qx,Qu per node
  double* xu[2];  // [T+1][Bp][R_XU] candidate / trial: x, u, aux
  double* fs;     // [T+1][Bp][R_FS] gaps of the candidate the last calcDiff ran on
  double* w;      // [T+1][Bp][R_FS] w = Vxx fs of the last backward pass (infeasible candidates only)
  double* kk;     // [T][Bp][R_KK]
  double* vx;     // [T+1][Bp][R_FS] Vx
  double* vl;     // [T+1][Bp][R_VL] lower triangle of Vxx
  double* q;      // [T][Bp][R_Q] (full only)
  int store_v;    // 1: every backward pass stores Vx / Vxx; 0: the solve loop materialises them once, at its end
  double* x0;     // [NX][Bp]
  double* scal;   // [S_NSLOTS][Bp]
  int* iscal;     // [I_NSLOTS][Bp]
  double* trace;  // [cap][NTRACE][Bp] or null
  int trace_cap;
  const double* mparams;  // shared model parameters
  const double* tparams;  // terminal model's parameters (shooting.hpp:48-63; default = the running block)
  const double* u_lb;     // [B][U] control limits of the running model (action-base.hxx:117-140)
  const double* u_ub;
  const int* has_limits;  // [B]
  int box;                // 1 = SolverBoxDDP / SolverBoxFDDP
  double* quu_inv;        // box only: [T][Bp][U*U] (get_Quu_inv, box-ddp.hpp)
  SolverOpts o;
};

__host__ __device__ constexpr int rup4(int n) { return (n + 3) / 4 * 4; }
__host__ __device__ constexpr int rup8(int n) { return (n + 7) / 8 * 8; }

template <class M>
struct TppLayout {
  static constexpr int N = M::NX, U = M::NU;
  static constexpr int NL = N * (N + 1) / 2;
  // Every record is a whole number of 64-byte DRAM atoms or a single 32-byte sector in an array of its own (adjacent
  // lanes then complete each other's atoms): ncu of the 96-byte records of the first layouts showed 1.6x the algorithmic
  // DRAM bytes (reads of 64 bytes straddling two atoms, partial-atom writes).
  //   XU : { x[N], u[U], aux[NA] }  candidate / trial; aux = per-node values calc leaves for calcDiff (unicycle: sin,
  //        cos of the heading, so that the backward pass needs no sincos of its own)
  //   FS : { fs[N] }   W : { w[N] = Vxx fs }   VX : { Vx[N] }   VL : { lower triangle of Vxx, column-major }
  //   KK : { K[U*N] column-major, k[U] }
  static constexpr int NA = M::NAUX;
  static constexpr int XU_X = 0, XU_U = N, XU_A = N + U, R_XU = rup8(N + U + NA);
  static constexpr int R_FS = rup4(N), R_VL = rup8(NL);
  static constexpr int KK_K = 0, KK_k = U * N, R_KK = rup8(U * N + U);
  static constexpr int Q_XX = 0, Q_XU = N * N, Q_UU = Q_XU + N * U, Q_X = Q_UU + U * U, Q_U = Q_X + N,
                       R_Q = rup4(Q_U + U);
  // index of (i,j), i >= j, in the column-major packed lower triangle
  __host__ __device__ static constexpr int low(int i, int j) { return j * N - j * (j - 1) / 2 + (i - j); }
  __host__ __device__ static constexpr int sym(int i, int j) { return i >= j ? low(i, j) : low(j, i); }
};

// 256-bit sector moves (LDG.E.256 / STG.E.256 on sm_100a) and L2 prefetch
__device__ __forceinline__ void ld_sector(const double* p, double* o) {
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(o[0]), "=d"(o[1]), "=d"(o[2]), "=d"(o[3]) : "l"(p));
}
__device__ __forceinline__ void st_sector(double* p, const double* o) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(o[0]), "d"(o[1]), "d"(o[2]), "d"(o[3]) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const double* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// sectors [s0, s1) of a record
template <int S0, int S1>
__device__ __forceinline__ void ld_rec(const double* p, double* o) {
#pragma unroll
  for (int s = S0; s < S1; ++s) ld_sector(p + 4 * s, o + 4 * s);
}
template <int S0, int S1>
__device__ __forceinline__ void st_rec(double* p, const double* o) {
#pragma unroll
  for (int s = S0; s < S1; ++s) st_sector(p + 4 * s, o + 4 * s);
}
template <int S0, int S1>
__device__ __forceinline__ void pf_rec(const double* p) {
#pragma unroll
  for (int s = S0; s < S1; ++s) prefetch_l2(p + 4 * s);
}

// Per-thread staging area in shared memory for the NEXT node's records: cp.async (LDGSTS, 16 bytes per request) moves
// them from L2 while the current node is computed, without holding them in registers (under the 128-register cap of
// the one-wave solve kernel a register prefetch is spilled as soon as it lands — ncu: 30 % of all stall samples on the
// spill store and the first move of the prefetched record). Layout [buffer][16-byte chunk][thread]: the lanes of a warp
// touch consecutive 16-byte words (no bank conflicts); groups are per thread, so diverged lanes stage independently.
#define TPP_BLOCK 64
#define TPP_STG_CHUNKS 12  // x,u,aux (4) + K,k (4) + fs (2) + w (2) for the unicycle; checked per model below
struct TppStage {
  double2 s[2][TPP_STG_CHUNKS][TPP_BLOCK];
};
// sectors [0, NSEC) of the record at g -> chunks [C0, C0 + 2 NSEC) of buffer buf
template <int C0, int NSEC>
__device__ __forceinline__ void stg_issue(TppStage& S, int buf, const double* g) {
#pragma unroll
  for (int k = 0; k < 2 * NSEC; ++k) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(&S.s[buf][C0 + k][threadIdx.x]);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(g + 2 * k) : "memory");
  }
}
__device__ __forceinline__ void stg_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void stg_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
template <int C0, int NSEC>
__device__ __forceinline__ void stg_read(const TppStage& S, int buf, double* o) {
#pragma unroll
  for (int k = 0; k < 2 * NSEC; ++k) {
    const double2 v = S.s[buf][C0 + k][threadIdx.x];
    o[2 * k] = v.x;
    o[2 * k + 1] = v.y;
  }
}

// ---------------------------------------------------------------------------------------------
// ActionModelUnicycle (reference: include/crocoddyl/core/actions/unicycle.hxx:13-73, unicycle.hpp:78-80)
struct UnicycleModel {
  static constexpr int NX = 3;
  static constexpr int NU = 2;
  static constexpr int NAUX = 2;  // sin, cos of x[2]: computed by calc, reused by calcDiff (unicycle.hxx:30-31,57-58)
  static constexpr int NR = 5;    // residual data->r = [w0 x; w1 u] (unicycle.hxx:39-40)
  // Structure of the derivatives (unicycle.hxx:45-73, data ctor unicycle.hpp:78-80): 0 = identically zero, 1 = one,
  // 2 = general. The Riccati products skip the structural zeros (x + 0*y = x for finite y) and do not multiply by the
  // ones, in the reference's summation order: same values, a third of the multiply-adds.
  __host__ __device__ static constexpr int fx_kind(int i, int j) { return i == j ? 1 : ((j == 2 && i < 2) ? 2 : 0); }
  __host__ __device__ static constexpr int fu_kind(int i, int j) { return ((j == 0 && i < 2) || (j == 1 && i == 2)) ? 2 : 0; }
  __host__ __device__ static constexpr int lxx_kind(int i, int j) { return i == j ? 2 : 0; }
  __host__ __device__ static constexpr int lxu_kind(int, int) { return 0; }
  __host__ __device__ static constexpr int luu_kind(int i, int j) { return i == j ? 2 : 0; }
  double dt, w0, w1;
  __device__ void load(const double* p) {
    dt = p[0];
    w0 = p[1];
    w1 = p[2];
  }
  // cost (unicycle.hxx:38-41)
  __device__ double cost(const double* x, const double* u) const {
    double rr = 0.;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double r = w0 * x[i];
      rr += r * r;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const double r = w1 * u[i];
      rr += r * r;
    }
    return 0.5 * rr;
  }
  // calc: xnext, cost (unicycle.hxx:22-42); aux = {sin, cos}
  __device__ double calc(const double* x, const double* u, double* xnext, double* aux) const {
    double s, c;
    sincos(x[2], &s, &c);
    aux[0] = s;
    aux[1] = c;
    xnext[0] = x[0] + c * u[0] * dt;
    xnext[1] = x[1] + s * u[0] * dt;
    xnext[2] = x[2] + u[1] * dt;
    return cost(x, u);
  }
  __device__ double calc(const double* x, const double* u, double* xnext) const {
    double aux[2];
    return calc(x, u, xnext, aux);
  }
  // ActionDataAbstract::r as calc leaves it (unicycle.hxx:39-40)
  __device__ void residual(const double* x, const double* u, double* r) const {
#pragma unroll
    for (int i = 0; i < 3; ++i) r[i] = w0 * x[i];
#pragma unroll
    for (int i = 0; i < 2; ++i) r[3 + i] = w1 * u[i];
  }
  // calc + calcDiff in one go from the sin / cos calc left in aux. Matrices column-major. (unicycle.hxx:45-73)
  __device__ double calcBothAux(const double* x, const double* u, const double* aux, double* xnext, double* Fx,
                                double* Fu, double* Lx, double* Lu, double* Lxx, double* Lxu, double* Luu) const {
    const double s = aux[0], c = aux[1];
    xnext[0] = x[0] + c * u[0] * dt;
    xnext[1] = x[1] + s * u[0] * dt;
    xnext[2] = x[2] + u[1] * dt;
    const double wx = w0 * w0, wu = w1 * w1;
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      Fx[i] = 0.;
      Lxx[i] = 0.;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      Fu[i] = 0.;
      Lxu[i] = 0.;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) Luu[i] = 0.;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      Lx[i] = x[i] * wx;
      Lxx[i + 3 * i] = wx;
      Fx[i + 3 * i] = 1.;  // data ctor (unicycle.hpp:78-80)
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      Lu[i] = u[i] * wu;
      Luu[i + 2 * i] = wu;
    }
    Fx[0 + 3 * 2] = -s * u[0] * dt;
    Fx[1 + 3 * 2] = c * u[0] * dt;
    Fu[0 + 3 * 0] = c * dt;
    Fu[1 + 3 * 0] = s * dt;
    Fu[2 + 3 * 1] = dt;
    return cost(x, u);
  }
  __device__ double calcBoth(const double* x, const double* u, double* xnext, double* Fx, double* Fu, double* Lx,
                             double* Lu, double* Lxx, double* Lxu, double* Luu) const {
    double aux[2];
    sincos(x[2], &aux[0], &aux[1]);
    return calcBothAux(x, u, aux, xnext, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu);
  }
};

#ifdef TPP_NOINLINE
#define TPP_SWEEP_ATTR __device__ __noinline__
#else
#define TPP_SWEEP_ATTR __device__ __forceinline__
#endif
#ifndef TPP_PF_DIST
#define TPP_PF_DIST 3  // L2 prefetch distance in nodes (the register prefetch is one node ahead)
#endif
#define TPP_REC(base, t, R) ((base) + ((size_t)(t) * (size_t)Bp + b) * (R))

// ---------------------------------------------------------------------------------------------
// problem_->calc(xs_, us_) in the reference's summation order (shooting.hxx:133-161): returns total cost and leaves
// aux (what calcDiff reuses from calc) in every running node's record.
template <class M>
TPP_SWEEP_ATTR double tpp_calc_pass(const TppCtx& c, const M& m, int b, double* __restrict__ xu) {
  typedef TppLayout<M> L;
  constexpr int N = M::NX, U = M::NU, SXA = L::R_XU / 4, SA0 = L::XU_A / 4;
  const int Bp = c.Bp, T = c.T;
  const size_t stride = (size_t)Bp * L::R_XU;
  double* p = xu + (size_t)b * L::R_XU;
  double cost = 0.;
  double rq[4 * SXA], r[4 * SXA], xn[N];
  ld_rec<0, SXA>(p, rq);
  for (int t = 0; t < T; ++t) {
#pragma unroll
    for (int i = 0; i < 4 * SXA; ++i) r[i] = rq[i];
    ld_rec<0, SXA>(p + stride, rq);
    if (t + TPP_PF_DIST <= T) pf_rec<0, SXA>(p + TPP_PF_DIST * stride);
    cost += m.calc(r + L::XU_X, r + L::XU_U, xn, r + L::XU_A);
    st_rec<SA0, SXA>(p, r);
    p += stride;
  }
  {
    double u0[U];
#pragma unroll
    for (int i = 0; i < U; ++i) u0[i] = 0.;
    M mt;  // terminal model (loaded here: its parameters are live only for this node)
    mt.load(c.tparams);
    cost += mt.cost(rq + L::XU_X, u0);
  }
  return cost;
}

// ---------------------------------------------------------------------------------------------
// SolverDDP::calcDiff (gaps) + backwardPass + computeGains, fused with the reductions of
// updateExpectedImprovement / expectedImprovement(DDP) / stoppingCriteria.
// Returns 0, or 1 = "backward_error" (LLT pivot <= 0, ddp.cpp:302-304; NaN/inf/>=1e30 in Vx/Vxx, :246-251).
// The gap term of node t+1 (Vx += Vxx fs, ddp.cpp:189-191,242-244) is applied when node t is visited,
// because fs[t+1] = xnext(xs[t],us[t]) - xs[t+1] (ddp.cpp:160-171) needs node t's calc; the operations
// and their order per node are unchanged. Vxx is exactly symmetric after ddp.cpp:235, so only its lower
// triangle is carried and stored.
// structured accumulation: s (+)= f * v where the factor f is structurally 0 (skip), 1 (add v) or general; `have`
// says whether s already holds a term — both are compile-time constants once the loops are unrolled
#define TPP_ACC(s, have, kind, f, v)             \
  do {                                           \
    if ((kind) == 1) {                           \
      s = (have) ? s + (v) : (v);                \
      have = true;                               \
    } else if ((kind) == 2) {                    \
      s = (have) ? s + (f) * (v) : (f) * (v);    \
      have = true;                               \
    }                                            \
  } while (0)

// fszero: bit 0 set = the gap array holds zeros (kept across the iterations of one solve so that a feasible candidate
// zero-fills it once, ddp.cpp:172-176).
// GAPS = !is_feasible: the feasible instantiation carries no gap / w arithmetic and no data of node t + 1 at all
// (most sweeps of a solve run on a feasible candidate).
// STORE: 0 = gains only (K, k; Vx / Vxx stay in registers: the solve loop materialises them once, after its last
// backward pass), 1 = everything (step-wise API, full storage), 2 = Vx / Vxx only (the materialising pass: recomputes
// the last backward pass on the same candidate, gaps and regularisation and stores nothing else).
template <class M, bool BOX, bool GAPS, int STORE>
TPP_SWEEP_ATTR int tpp_backward_impl(const TppCtx& c, const M& m, TppStage& stg, int b, int cur, bool recalc, double xreg,
                                     double ureg, double& dg_out, double& dq_out, double& stop_out, int& fszero) {
  constexpr bool is_feasible = !GAPS;
  constexpr bool ST_V = STORE != 0, ST_K = STORE != 2;
  typedef TppLayout<M> L;
  constexpr int N = M::NX, U = M::NU, NL = L::NL;
  constexpr int SXA = L::R_XU / 4, SFS = L::R_FS / 4;
  const int Bp = c.Bp, T = c.T;
  const bool fddp = c.fddp != 0;
  const bool full = c.full != 0 && ST_K;
  const size_t sxu = (size_t)Bp * L::R_XU, skk = (size_t)Bp * L::R_KK, sfs = (size_t)Bp * L::R_FS, svl = (size_t)Bp * L::R_VL;
  const double* pxu = c.xu[cur] + ((size_t)T * Bp + b) * L::R_XU;  // record of node t + 1 while node t is visited
  double* pfs = c.fs + ((size_t)T * Bp + b) * L::R_FS;             // gaps / w / Vx of node t + 1
  double* pw = c.w + ((size_t)T * Bp + b) * L::R_FS;
  double* pvx = c.vx + ((size_t)T * Bp + b) * L::R_FS;
  double* pvl = c.vl + ((size_t)T * Bp + b) * L::R_VL;
  double* pkk = c.kk + ((size_t)(T - 1) * Bp + b) * L::R_KK;       // record of node t
  double dg = 0., dq = 0., stop = 0.;
  const bool has_xreg = !isnan(xreg), has_ureg = !isnan(ureg);  // ddp.cpp:185,216,226,237
  // gaps: read back when they are not recomputed (only an infeasible candidate uses them)
  const bool load_fs = GAPS && !recalc;
  const bool zero_fs = !GAPS && recalc && !(fszero & 1);
  double Vl[NL], Vx[N];
  double Fx[N * N], Fu[N * U], Lx[N], Lu[U], Lxx[N * N], Lxu[N * U], Luu[U * U], xn[N];
  // node t's record comes out of the shared-memory staging buffer; node t - 1 is on its way into the other buffer while
  // node t is computed. GAPS only: xp / fp = x and gaps of node t + 1, fr = gaps of node t (read back when not recomputed)
  constexpr int CF = 2 * SXA;  // staging chunks of the gap record
  static_assert(2 * (SXA + SFS) <= TPP_STG_CHUNKS, "staging area too small for this model");
  double r[L::R_XU], xp[N], fp[L::R_FS], fr[L::R_FS];
#pragma unroll
  for (int i = 0; i < L::R_FS; ++i) fp[i] = fr[i] = 0.;
  int buf = 0;
  stg_wait();  // nothing of an earlier sweep (one that returned on an error) may still be landing in the staging area
  // node T - 1 into the staging buffer, then the terminal node (ddp.cpp:181-187): u = unone_ = 0 (action-base.hxx:27-37)
  stg_issue<0, SXA>(stg, buf, pxu - sxu);
  if constexpr (GAPS) {
    if (load_fs) stg_issue<CF, SFS>(stg, buf, pfs - sfs);
  }
  stg_commit();
  ld_rec<0, SXA>(pxu, r);
  if constexpr (GAPS) {
    if (load_fs) ld_rec<0, SFS>(pfs, fp);
#pragma unroll
    for (int i = 0; i < N; ++i) xp[i] = r[L::XU_X + i];
  }
  {
    double u0[U];
#pragma unroll
    for (int i = 0; i < U; ++i) u0[i] = 0.;
    M mt;
    mt.load(c.tparams);
    mt.calcBothAux(r + L::XU_X, u0, r + L::XU_A, xn, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu);  // only Lx, Lxx are used
  }
#pragma unroll
  for (int j = 0; j < N; ++j)
#pragma unroll
    for (int i = j; i < N; ++i) Vl[L::low(i, j)] = Lxx[i + N * j] + ((i == j && has_xreg) ? xreg : 0.);
#pragma unroll
  for (int i = 0; i < N; ++i) Vx[i] = Lx[i];

  for (int t = T - 1; t >= -1; --t) {
    if (t >= 0) {
      stg_wait();
      stg_read<0, SXA>(stg, buf, r);
      if constexpr (GAPS) {
        if (load_fs) stg_read<CF, SFS>(stg, buf, fr);
      }
      buf ^= 1;
      // node t-1's record is requested (cp.async into the other staging buffer) before node t's dependent Riccati
      // chain starts, node t-1-TPP_PF_DIST's record is pulled into L2
      if (t > 0) {
        stg_issue<0, SXA>(stg, buf, pxu - 2 * sxu);
        if constexpr (GAPS) {
          if (load_fs) stg_issue<CF, SFS>(stg, buf, pfs - 2 * sfs);
        }
        stg_commit();
      }
      if (t > TPP_PF_DIST) {
        pf_rec<0, SXA>(pxu - (2 + TPP_PF_DIST) * sxu);
        if constexpr (GAPS) {
          if (load_fs) pf_rec<0, SFS>(pfs - (2 + TPP_PF_DIST) * sfs);
        }
      }
      m.calcBothAux(r + L::XU_X, r + L::XU_U, r + L::XU_A, xn, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu);
    } else {
      // gap of node 0: fs[0] = diff(xs[0], x0) = x0 - xs[0]   (ddp.cpp:162)
#pragma unroll
      for (int i = 0; i < N; ++i) xn[i] = c.x0[(size_t)i * Bp + b];
    }
    // ---- finish node t+1: gap term, w, Vx / Vxx records ----
    if constexpr (GAPS) {
      if (recalc) {
#pragma unroll
        for (int i = 0; i < N; ++i) fp[i] = xn[i] - xp[i];
        if (ST_K) st_rec<0, SFS>(pfs, fp);
      }
      double wv[L::R_FS];
#pragma unroll
      for (int i = 0; i < L::R_FS; ++i) wv[i] = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
#pragma unroll
        for (int j = 0; j < N; ++j) s += Vl[L::sym(i, j)] * fp[j];
        wv[i] = s;
        Vx[i] += s;
      }
      if (ST_K) st_rec<0, SFS>(pw, wv);
      if (fddp) {  // updateExpectedImprovement gap terms (fddp.cpp:131-135,141-145)
        double a = 0., q = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) {
          a += Vx[i] * fp[i];
          q += fp[i] * wv[i];
        }
        dg -= a;
        dq += q;
      }
    } else if (zero_fs) {
      st_rec<0, SFS>(pfs, fp);  // zeros: ddp.cpp:172-176
    }
    {
      double nrm = 0.;
#pragma unroll
      for (int i = 0; i < N; ++i) {
        const double v = fabs(Vx[i]);
        if (!(v <= nrm)) nrm = v;
      }
      if (ST_V) {
        double vrec[L::R_FS], lrec[L::R_VL];
#pragma unroll
        for (int i = 0; i < L::R_FS; ++i) vrec[i] = i < N ? Vx[i] : 0.;
#pragma unroll
        for (int i = 0; i < L::R_VL; ++i) lrec[i] = i < NL ? Vl[i] : 0.;
        st_rec<0, SFS>(pvx, vrec);
        st_rec<0, L::R_VL / 4>(pvl, lrec);
      }
      if (t + 1 < T && raise_if_nan(nrm)) return 1;  // ddp.cpp:246-248
    }
    if (t < 0) break;

    // ---- node t (ddp.cpp:202-244) ----
    double A[N * N];  // FxTVxx_p = Fx^T Vxx'
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
        bool have = false;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) TPP_ACC(s, have, M::fx_kind(kk, i), Fx[kk + N * i], Vl[L::sym(kk, j)]);
        A[i + N * j] = s;
      }
    double Qxx[N * N], Qx[N], Qxu[N * U], Quu[U * U], Qu[U], G[U * N];  // G = FuTVxx_p = Fu^T Vxx'
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
        bool have = false;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) TPP_ACC(s, have, M::fx_kind(kk, j), Fx[kk + N * j], A[i + N * kk]);
        Qxx[i + N * j] = M::lxx_kind(i, j) ? Lxx[i + N * j] + s : s;
      }
#pragma unroll
    for (int j = 0; j < N; ++j) {
      double s = 0.;
      bool have = false;
#pragma unroll
      for (int i = 0; i < N; ++i) TPP_ACC(s, have, M::fx_kind(i, j), Fx[i + N * j], Vx[i]);
      Qx[j] = Lx[j] + s;
    }
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < U; ++i) {
        double s = 0.;
        bool have = false;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) TPP_ACC(s, have, M::fu_kind(kk, i), Fu[kk + N * i], Vl[L::sym(kk, j)]);
        G[i + U * j] = s;
      }
#pragma unroll
    for (int j = 0; j < U; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
        bool have = false;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) TPP_ACC(s, have, M::fu_kind(kk, j), Fu[kk + N * j], A[i + N * kk]);
        Qxu[i + N * j] = M::lxu_kind(i, j) ? Lxu[i + N * j] + s : s;
      }
#pragma unroll
    for (int j = 0; j < U; ++j)
#pragma unroll
      for (int i = 0; i < U; ++i) {
        double s = 0.;
        bool have = false;
#pragma unroll
        for (int kk = 0; kk < N; ++kk) TPP_ACC(s, have, M::fu_kind(kk, j), Fu[kk + N * j], G[i + U * kk]);
        Quu[i + U * j] = M::luu_kind(i, j) ? Luu[i + U * j] + s : s;
      }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      double s = 0.;
      bool have = false;
#pragma unroll
      for (int i = 0; i < N; ++i) TPP_ACC(s, have, M::fu_kind(i, j), Fu[i + N * j], Vx[i]);
      Qu[j] = Lu[j] + s;
      if (has_ureg) Quu[j + U * j] += ureg;
    }
    double kkrec[L::R_KK];
#pragma unroll
    for (int i = 0; i < L::R_KK; ++i) kkrec[i] = 0.;
    double* Kg = kkrec + L::KK_K;
    double* kg = kkrec + L::KK_k;
    bool boxed = false;
    if constexpr (BOX) {
      // SolverBoxDDP/BoxFDDP::computeGains (box-ddp.cpp:46-74): box QP on the control step when the model has control
      // limits and the candidate is feasible; vanilla gains otherwise
      boxed = is_feasible && c.has_limits[b] != 0;
      if (boxed) {
        double lb[U], ub[U], work[box_gains_work_doubles(U)], qinv[U * U];
        int iwork[box_gains_work_ints(U)];
        double kold[L::R_KK];
        ld_rec<0, L::R_KK / 4>(pkk, kold);  // k_[t] of the previous backward pass: the warm start
#pragma unroll
        for (int i = 0; i < U; ++i) {
          lb[i] = c.u_lb[(size_t)b * U + i];
          ub[i] = c.u_ub[(size_t)b * U + i];
          kg[i] = kold[L::KK_k + i];
        }
        if (box_compute_gains(N, U, Quu, U, Qu, Qxu, N, r + L::XU_U, lb, ub, Kg, U, kg, qinv, c.o, work, iwork)) return 1;
        if (c.quu_inv) {
          double* qo = c.quu_inv + ((size_t)t * Bp + b) * (U * U);  // box solvers only
#pragma unroll
          for (int i = 0; i < U * U; ++i) qo[i] = qinv[i];
        }
      }
    }
    if (!boxed) {
    // computeGains (ddp.cpp:298-310): unblocked lower LLT, failure iff pivot <= 0
    double Lc[U * U];
#pragma unroll
    for (int i = 0; i < U * U; ++i) Lc[i] = Quu[i];
    // Divisions by a pivot are multiplications by its reciprocal, here and in the two triangular solves below: an
    // FP64 division is ~12 dependent instructions on this path and the 2 U (N+1) + U(U-1)/2 of them were the largest
    // single item of the node (ncu: 750 instructions per node visit; B200: 50.6 -> 38.8 ms for config C2).
    // s * (1/d) differs from s / d by at most one ulp, the same class of difference as the FMA contraction around it.
    bool fail = false;
    double Linv[U];
#pragma unroll
    for (int kk = 0; kk < U; ++kk) {
      double p = Lc[kk + U * kk];
#pragma unroll
      for (int j = 0; j < kk; ++j) p -= Lc[kk + U * j] * Lc[kk + U * j];
      if (p <= 0.) fail = true;
      llt_pivot(p, p, Linv[kk]);
      Lc[kk + U * kk] = p;
#pragma unroll
      for (int i = kk + 1; i < U; ++i) {
        double s = Lc[i + U * kk];
#pragma unroll
        for (int j = 0; j < kk; ++j) s -= Lc[i + U * j] * Lc[kk + U * j];
        Lc[i + U * kk] = div_pivot(s, p, Linv[kk]);
      }
    }
    if (fail) return 1;
    // K = Quu^-1 Qxu^T  (columns of K), k = Quu^-1 Qu
#pragma unroll
    for (int col = 0; col <= N; ++col) {
      double rr[U];
#pragma unroll
      for (int i = 0; i < U; ++i) rr[i] = (col < N) ? Qxu[(col < N ? col : 0) + N * i] : Qu[i];
#pragma unroll
      for (int i = 0; i < U; ++i) {
        double s = rr[i];
#pragma unroll
        for (int j = 0; j < i; ++j) s -= Lc[i + U * j] * rr[j];
        rr[i] = div_pivot(s, Lc[i + U * i], Linv[i]);
      }
#pragma unroll
      for (int i = U - 1; i >= 0; --i) {
        double s = rr[i];
#pragma unroll
        for (int j = i + 1; j < U; ++j) s -= Lc[j + U * i] * rr[j];
        rr[i] = div_pivot(s, Lc[i + U * i], Linv[i]);
      }
#pragma unroll
      for (int i = 0; i < U; ++i) {
        if (col < N)
          Kg[i + U * (col < N ? col : 0)] = rr[i];
        else
          kg[i] = rr[i];
      }
    }
    }  // !boxed
    double Quuk[U];
#pragma unroll
    for (int i = 0; i < U; ++i) {
      double s = 0.;
#pragma unroll
      for (int j = 0; j < U; ++j) s += Quu[i + U * j] * kg[j];
      Quuk[i] = s;
    }
    // reductions: ddp.cpp:132-155 / fddp.cpp:137-140
    {
      double a = 0., q = 0., st = 0.;
#pragma unroll
      for (int i = 0; i < U; ++i) {
        a += Qu[i] * kg[i];
        q += kg[i] * Quuk[i];
        st += Qu[i] * Qu[i];
      }
      dg += a;
      dq -= q;
      stop += st;
    }
    if (ST_K) st_rec<0, L::R_KK / 4>(pkk, kkrec);
    if (full) {
      double qrec[L::R_Q];
#pragma unroll
      for (int i = 0; i < L::R_Q; ++i) qrec[i] = 0.;
#pragma unroll
      for (int i = 0; i < N * N; ++i) qrec[L::Q_XX + i] = Qxx[i];
#pragma unroll
      for (int i = 0; i < N * U; ++i) qrec[L::Q_XU + i] = Qxu[i];
#pragma unroll
      for (int i = 0; i < U * U; ++i) qrec[L::Q_UU + i] = Quu[i];
#pragma unroll
      for (int i = 0; i < N; ++i) qrec[L::Q_X + i] = Qx[i];
#pragma unroll
      for (int i = 0; i < U; ++i) qrec[L::Q_U + i] = Qu[i];
      st_rec<0, L::R_Q / 4>(c.q + ((size_t)t * Bp + b) * L::R_Q, qrec);
    }
    // Vx = Qx + K^T Quuk - 2 K^T Qu ; Vxx = Qxx - Qxu K   (ddp.cpp:223-233)
#pragma unroll
    for (int j = 0; j < N; ++j) {
      double s1 = 0., s2 = 0.;
#pragma unroll
      for (int i = 0; i < U; ++i) {
        s1 += Kg[i + U * j] * Quuk[i];
        s2 += Kg[i + U * j] * Qu[i];
      }
      double v = Qx[j];
      if (has_ureg) {
        v += s1;
        v += -2. * s2;
      } else {
        v += -1. * s2;  // ddp.cpp:226-227
      }
      Vx[j] = v;
    }
    double Vr[N * N];
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = 0; i < N; ++i) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < U; ++kk) s += Qxu[i + N * kk] * Kg[kk + U * j];
        Vr[i + N * j] = Qxx[i + N * j] + (-1.) * s;
      }
    double nrm = 0.;
#pragma unroll
    for (int j = 0; j < N; ++j)
#pragma unroll
      for (int i = j; i < N; ++i) {
        double v = 0.5 * (Vr[i + N * j] + Vr[j + N * i]);  // ddp.cpp:235
        if (i == j && has_xreg) v += xreg;                 // ddp.cpp:237-239
        Vl[L::low(i, j)] = v;
        const double av = fabs(v);
        if (!(av <= nrm)) nrm = av;
      }
    if (raise_if_nan(nrm)) return 1;  // ddp.cpp:249-251
    if constexpr (GAPS) {
#pragma unroll
      for (int i = 0; i < N; ++i) xp[i] = r[L::XU_X + i];
      if (load_fs) {
#pragma unroll
        for (int i = 0; i < L::R_FS; ++i) fp[i] = fr[i];
      }
    }
    pxu -= sxu;
    pfs -= sfs;
    pw -= sfs;
    pvx -= sfs;
    pvl -= svl;
    pkk -= skk;
  }
  if (recalc && ST_K) fszero = is_feasible ? 1 : 0;  // what the gap array holds now
  dg_out = dg;
  dq_out = dq;
  stop_out = stop;
  return 0;
}
// store: 0 = gains only, 1 = everything, 2 = Vx / Vxx only (see tpp_backward_impl)
template <class M, bool BOX = false>
TPP_SWEEP_ATTR int tpp_backward(const TppCtx& c, const M& m, TppStage& stg, int b, int cur, bool recalc, bool is_feasible,
                                int store, double xreg, double ureg, double& dg_out, double& dq_out, double& stop_out,
                                int& fszero) {
#define TPP_BWD(G, S) tpp_backward_impl<M, BOX, G, S>(c, m, stg, b, cur, recalc, xreg, ureg, dg_out, dq_out, stop_out, fszero)
  if (store == 0) return is_feasible ? TPP_BWD(false, 0) : TPP_BWD(true, 0);
  if (store == 2) return is_feasible ? TPP_BWD(false, 2) : TPP_BWD(true, 2);
  return is_feasible ? TPP_BWD(false, 1) : TPP_BWD(true, 1);
#undef TPP_BWD
}

// ---------------------------------------------------------------------------------------------
// SolverFDDP::forwardPass (fddp.cpp:149-225) / SolverDDP::forwardPass (ddp.cpp:255-296, == the closed
// branch with xs_try[0] = x0) fused with SolverFDDP::expectedImprovement's gap reduction
// dv = -sum_t fs_t^T Vxx_t (xs_t - xs_try_t) = sum_t w_t^T dx_t (fddp.cpp:107-124).
// Reads the candidate records XU[cur], writes the trial records XU[cur^1]. Returns 0, or 1 = "forward_error".
// GAPS = the rollout may be open (FDDP / box-DDP from an infeasible candidate): only that instantiation reads the gaps
// and w = Vxx fs; a rollout from a feasible candidate moves x, u, K, k only
template <class M, bool BOX, bool GAPS>
TPP_SWEEP_ATTR int tpp_forward_impl(const TppCtx& c, const M& m, TppStage& stg, int b, int cur, double alpha,
                                    bool is_feasible, double& cost_try_out, double& dv_out) {
  typedef TppLayout<M> L;
  constexpr int N = M::NX, U = M::NU;
  constexpr int SXA = L::R_XU / 4, SK = L::R_KK / 4, SFS = L::R_FS / 4;
  const int Bp = c.Bp, T = c.T;
  const bool fddp = c.fddp != 0;
  const bool closed = !GAPS || !fddp || is_feasible || (alpha == 1.);
  const bool use_w = GAPS && fddp && !is_feasible;
  const double am1 = alpha - 1.;
  // SolverBoxDDP::forwardPass keeps the gap of the TERMINAL node only (box-ddp.cpp:121-125)
  const bool term_open = GAPS && BOX && !fddp && !(is_feasible || (alpha == 1.));
  bool clamp = false;
  double ulb[U], uub[U];
  if constexpr (BOX) {
    clamp = c.has_limits[b] != 0;  // box-ddp.cpp:95-97 / box-fddp.cpp:96-98
#pragma unroll
    for (int i = 0; i < U; ++i) {
      ulb[i] = c.u_lb[(size_t)b * U + i];
      uub[i] = c.u_ub[(size_t)b * U + i];
    }
  }
  const bool need_fs = !closed || term_open;  // the gaps are read only by an open rollout
  const size_t sxu = (size_t)Bp * L::R_XU, skk = (size_t)Bp * L::R_KK, sfs = (size_t)Bp * L::R_FS;
  const double* __restrict__ pxu = c.xu[cur] + (size_t)b * L::R_XU;  // candidate record of node t + 1 (prefetched)
  double* __restrict__ pxt = c.xu[cur ^ 1] + (size_t)b * L::R_XU;    // trial record of node t
  const double* __restrict__ pkk = c.kk + (size_t)b * L::R_KK;
  const double* __restrict__ pfs = c.fs + (size_t)b * L::R_FS;
  const double* __restrict__ pw = c.w + (size_t)b * L::R_FS;
  double cost_try = 0., dv = 0.;
  double xn[N], xt[N], dx[N];
  // node t+1's inputs (none of them depends on the rollout recursion) travel into the shared-memory staging buffer
  // while node t is computed: chunks [0, CK) = x,u,aux  [CK, CFS) = K,k  then the gaps and w (GAPS only)
  constexpr int CK = 2 * SXA, CFS = CK + 2 * SK, CW = CFS + 2 * SFS;
  static_assert(CW + 2 * SFS <= TPP_STG_CHUNKS, "staging area too small for this model");
  int buf = 0;
  stg_wait();  // nothing of an earlier sweep may still be landing in the staging area
  auto fetch = [&](int t) {  // node t's inputs; the pointers stand on node t
    stg_issue<0, SXA>(stg, buf, pxu);
    if constexpr (GAPS) {
      if (need_fs) stg_issue<CFS, SFS>(stg, buf, pfs);
      if (use_w) stg_issue<CW, SFS>(stg, buf, pw);
    }
    if (t < T) stg_issue<CK, SK>(stg, buf, pkk);
    stg_commit();
  };
#pragma unroll
  for (int i = 0; i < N; ++i) xn[i] = c.x0[(size_t)i * Bp + b];
  fetch(0);
  for (int t = 0; t <= T; ++t) {
    double ca[L::R_XU], ck[L::R_KK], cf[L::R_FS], cw[L::R_FS], o[L::R_XU];
#pragma unroll
    for (int i = 0; i < L::R_FS; ++i) cf[i] = cw[i] = 0.;
    stg_wait();
    stg_read<0, SXA>(stg, buf, ca);
    if (t < T) stg_read<CK, SK>(stg, buf, ck);
    if constexpr (GAPS) {
      if (need_fs) stg_read<CFS, SFS>(stg, buf, cf);
      if (use_w) stg_read<CW, SFS>(stg, buf, cw);
    }
    buf ^= 1;
    pxu += sxu;
    pkk += skk;
    pfs += sfs;
    pw += sfs;
    if (t < T) fetch(t + 1);
    if (t + TPP_PF_DIST <= T) {  // node t + TPP_PF_DIST into L2
      pf_rec<0, SXA>(pxu + (TPP_PF_DIST - 1) * sxu);
      if constexpr (GAPS) {
        if (need_fs) pf_rec<0, SFS>(pfs + (TPP_PF_DIST - 1) * sfs);
        if (use_w) pf_rec<0, SFS>(pw + (TPP_PF_DIST - 1) * sfs);
      }
      if (t + TPP_PF_DIST < T) pf_rec<0, SK>(pkk + (TPP_PF_DIST - 1) * skk);
    }
    if (closed && !(term_open && t == T)) {
#pragma unroll
      for (int i = 0; i < N; ++i) xt[i] = xn[i];
    } else if constexpr (GAPS) {
#pragma unroll
      for (int i = 0; i < N; ++i) xt[i] = xn[i] + cf[i] * am1;  // integrate(xnext, fs*(a-1))
    }
#pragma unroll
    for (int i = 0; i < 4 * SXA; ++i) o[i] = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      dx[i] = xt[i] - ca[L::XU_X + i];
      o[L::XU_X + i] = xt[i];
    }
    if constexpr (GAPS) {
      if (use_w) {
        double s = 0.;
#pragma unroll
        for (int i = 0; i < N; ++i) s += cw[i] * dx[i];
        dv += s;
      }
    }
    if (t == T) {
      st_rec<0, SXA>(pxt, o);
      break;
    }
    double* u = o + L::XU_U;
#pragma unroll
    for (int i = 0; i < U; ++i) {
      double s = 0.;
#pragma unroll
      for (int j = 0; j < N; ++j) s += ck[L::KK_K + i + U * j] * dx[j];
      u[i] = ca[L::XU_U + i] - ck[L::KK_k + i] * alpha - s;
      if (BOX && clamp) u[i] = fmin(fmax(u[i], ulb[i]), uub[i]);  // cwiseMax(lb).cwiseMin(ub)
    }
    // calc leaves aux (what the next backward pass reuses) in the trial record
    cost_try += m.calc(xt, u, xn, o + L::XU_A);
    st_rec<0, SXA>(pxt, o);
    pxt += sxu;
    double nrm = 0.;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const double v = fabs(xn[i]);
      if (!(v <= nrm)) nrm = v;
    }
    if (raise_if_nan(cost_try) || raise_if_nan(nrm)) return 1;
  }
  {
    double u0[U];
#pragma unroll
    for (int i = 0; i < U; ++i) u0[i] = 0.;
    M mt;
    mt.load(c.tparams);
    cost_try += mt.cost(xt, u0);
  }
  if (raise_if_nan(cost_try)) return 1;
  cost_try_out = cost_try;
  dv_out = dv;
  return 0;
}
template <class M, bool BOX = false>
TPP_SWEEP_ATTR int tpp_forward(const TppCtx& c, const M& m, TppStage& stg, int b, int cur, double alpha, bool is_feasible,
                               double& cost_try_out, double& dv_out) {
  if (!is_feasible && (c.fddp != 0 || BOX)) return tpp_forward_impl<M, BOX, true>(c, m, stg, b, cur, alpha, is_feasible, cost_try_out, dv_out);
  return tpp_forward_impl<M, BOX, false>(c, m, stg, b, cur, alpha, is_feasible, cost_try_out, dv_out);
}

// acceptance tests of the line search (ddp.cpp:84-93 / fddp.cpp:64-80)
__device__ inline bool accept_step(const SolverOpts& o, bool fddp, bool is_feasible, double d0, double dV,
                                   double dVexp) {
  if (fddp) {
    if (dVexp >= 0) return d0 < o.th_grad || dV > o.th_acceptstep * dVexp;
    return dV > o.th_acceptnegstep * dVexp;
  }
  if (dVexp >= 0) return d0 < o.th_grad || !is_feasible || dV > o.th_acceptstep * dVexp;
  return false;
}

#define TPP_S(slot) c.scal[(size_t)(slot) * Bp + b]
#define TPP_I(slot) c.iscal[(size_t)(slot) * Bp + b]

// ---------------------------------------------------------------------------------------------
// Whole SolverDDP::solve / SolverFDDP::solve loop for one problem per thread (ddp.cpp:41-118, fddp.cpp:19-105).
// The loop is RESUMABLE: a launch runs at most `chunk` iterations of every problem it owns (chunk <= 0: all of them),
// saves the loop state in the per-problem slots and sets I_FINISHED when solve() has returned. Between launches the
// host compacts the unfinished problems into `list` (tpp_compact_*), so that a launch only spends lanes on OCPs that
// still iterate: each OCP has its own iteration count, and without compaction one slow lane keeps its whole warp
// alive (ncu, profiles/README.md: 19 % of the warp slots active at B = 65,536).
template <class M, bool BOX>
__device__ __forceinline__ void tpp_solve_body(const TppCtx& c, TppStage& stg, int maxiter, int is_feasible0, double reginit,
                                               int chunk, int first, const int* __restrict__ list,
                                               const int* __restrict__ nactive, int b_lo, int b_hi) {
  // a launch owns the problems [b_lo, b_hi) (slices of the batch on different streams, TppBackend::solve_pipelined)
  const int gid = b_lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (list ? *nactive : b_hi)) return;
  const int b = list ? list[gid] : gid;
  const int Bp = c.Bp;
  const bool fddp = c.fddp != 0;
  M m;
  m.load(c.mparams);
  int cur = TPP_I(I_CUR), fsbuf = TPP_I(I_FSBUF);
  bool feas = is_feasible0 != 0, wasfeas = false;
  double xreg = isnan(reginit) ? c.o.regmin : reginit, ureg = xreg;
  double cost = TPP_S(S_COST), stop = TPP_S(S_STOP), d0 = TPP_S(S_D0), d1 = TPP_S(S_D1);
  double steplength = TPP_S(S_STEP), dV = TPP_S(S_DV), dVexp = TPP_S(S_DVEXP), cost_try = TPP_S(S_COST_TRY);
  double dg = 0., dq = 0.;
  int nbwd = 0, nfwd = 0, nrec = 0, converged = 0, iter = 0;
  int fszero = 0;  // does the gap array hold zeros (tpp_backward)
  bool recalc = true, need_gaps = true, done = false;
  // Vx / Vxx are not stored by the backward passes of the loop (store == 0): they are materialised once, when solve()
  // returns, by repeating the last successful backward pass (lb_*) — and the failed attempts after it, if the loop ended
  // on them (fl_*: the reference's arrays hold their partial overwrites) — with stores of Vx / Vxx only
  const int store = c.store_v ? 1 : 0;
  bool has_lb = false, lb_feas = false;
  int lb_cur = 0, nfail = 0;
  double lb_xreg = 0., lb_ureg = 0., fl_xreg = 0.;
  if (!first) {  // resume
    feas = TPP_I(I_FEAS) != 0;
    wasfeas = TPP_I(I_WASFEAS) != 0;
    xreg = TPP_S(S_XREG);
    ureg = TPP_S(S_UREG);
    dg = TPP_S(S_DG);
    dq = TPP_S(S_DQ);
    nbwd = TPP_I(I_NBWD);
    nfwd = TPP_I(I_NFWD);
    nrec = TPP_I(I_NREC);
    iter = TPP_I(I_ITER);
    const int lp = TPP_I(I_LOOP);
    recalc = (lp & 1) != 0;
    need_gaps = (lp & 2) != 0;
    fszero = (lp >> 2) & 1;
    has_lb = (lp >> 3) & 1;
    lb_cur = (lp >> 4) & 1;
    lb_feas = (lp >> 5) & 1;
    nfail = lp >> 8;
    lb_xreg = TPP_S(S_LBXREG);
    lb_ureg = TPP_S(S_LBUREG);
    fl_xreg = TPP_S(S_FLXREG);
  }
  int budget = chunk > 0 ? chunk : maxiter + 1;
  bool paused = false;
  for (; iter < maxiter && !done; ++iter) {
    if (budget-- == 0) {
      paused = true;
      break;
    }
    int retries = 0;
    if (recalc && iter == 0) {  // ddp.cpp:158
      cost = tpp_calc_pass(c, m, b, c.xu[cur]);
      ++nfwd;
    }
    if (recalc) need_gaps = true;
    while (true) {  // computeDirection with regularisation retry (ddp.cpp:55-68)
      if (need_gaps) fsbuf = cur;
      const int st = tpp_backward<M, BOX>(c, m, stg, b, cur, need_gaps, feas, store, xreg, ureg, dg, dq, stop, fszero);
      ++nbwd;
      if (st == 0) {
        need_gaps = false;
        has_lb = true;
        lb_cur = cur;
        lb_feas = feas;
        lb_xreg = xreg;
        lb_ureg = ureg;
        nfail = 0;
        break;
      }
      if (nfail++ == 0) fl_xreg = xreg;
      recalc = false;
      reg_increase(c.o, xreg, ureg);
      ++retries;
      if (xreg == c.o.regmax) {
        done = true;
        break;
      }
    }
    if (done) break;  // return false (iter_ keeps its value)
    if (!fddp) {      // SolverDDP::expectedImprovement before the line search (ddp.cpp:71)
      d0 = dg;
      d1 = dq;
    }
    recalc = false;
    int accepted = -1;
    for (int a = 0; a < c.o.nalphas; ++a) {
      steplength = c.o.alphas[a];
      double dv = 0.;
      const int st = tpp_forward<M, BOX>(c, m, stg, b, cur, steplength, feas, cost_try, dv);
      ++nfwd;
      if (st) continue;
      dV = cost - cost_try;
      if (fddp) {
        d0 = dg + dv;
        d1 = dq - 2 * dv;
      }
      dVexp = steplength * (d0 + 0.5 * steplength * d1);
      if (accept_step(c.o, fddp, feas, d0, dV, dVexp)) {
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
      if (xreg == c.o.regmax) break;  // return false
    }
    if (c.trace && nrec < c.trace_cap) {
      double* tr = c.trace + (size_t)nrec * NTRACE * Bp + b;
      tr[0 * (size_t)Bp] = iter;
      tr[1 * (size_t)Bp] = cost;
      tr[2 * (size_t)Bp] = stop;
      tr[3 * (size_t)Bp] = d0;
      tr[4 * (size_t)Bp] = d1;
      tr[5 * (size_t)Bp] = xreg;
      tr[6 * (size_t)Bp] = ureg;
      tr[7 * (size_t)Bp] = steplength;
      tr[8 * (size_t)Bp] = feas ? 1. : 0.;
      tr[9 * (size_t)Bp] = dV;
      tr[10 * (size_t)Bp] = dVexp;
      tr[11 * (size_t)Bp] = accepted;
      tr[12 * (size_t)Bp] = retries;
    }
    ++nrec;
    if (wasfeas && stop < c.o.th_stop) {
      converged = 1;
      break;  // return true with iter_ = iter
    }
  }
  if (!paused && store == 0) {  // solve() has returned: materialise Vx / Vxx (see above)
    double g0, g1, g2;
    int fz = 0;
    if (has_lb) tpp_backward<M, BOX>(c, m, stg, b, lb_cur, false, lb_feas, 2, lb_xreg, lb_ureg, g0, g1, g2, fz);
    double fx = fl_xreg, fu = fl_xreg;
    for (int i = 0; i < nfail; ++i) {  // the failed attempts the loop ended on, in their order
      tpp_backward<M, BOX>(c, m, stg, b, cur, true, feas, 2, fx, fu, g0, g1, g2, fz);
      reg_increase(c.o, fx, fu);
    }
  }
  TPP_S(S_LBXREG) = lb_xreg;
  TPP_S(S_LBUREG) = lb_ureg;
  TPP_S(S_FLXREG) = fl_xreg;
  TPP_S(S_COST) = cost;
  TPP_S(S_STOP) = stop;
  TPP_S(S_D0) = d0;
  TPP_S(S_D1) = d1;
  TPP_S(S_XREG) = xreg;
  TPP_S(S_UREG) = ureg;
  TPP_S(S_STEP) = steplength;
  TPP_S(S_DV) = dV;
  TPP_S(S_DVEXP) = dVexp;
  TPP_S(S_COST_TRY) = cost_try;
  TPP_S(S_DG) = dg;
  TPP_S(S_DQ) = dq;
  TPP_I(I_ITER) = iter;
  TPP_I(I_FEAS) = feas;
  TPP_I(I_WASFEAS) = wasfeas;
  TPP_I(I_CONVERGED) = converged;
  TPP_I(I_CUR) = cur;
  TPP_I(I_FSBUF) = fsbuf;
  TPP_I(I_NREC) = nrec;
  TPP_I(I_NBWD) = nbwd;
  TPP_I(I_NFWD) = nfwd;
  TPP_I(I_LOOP) = (recalc ? 1 : 0) | (need_gaps ? 2 : 0) | ((fszero & 1) << 2) | ((has_lb ? 1 : 0) << 3) | (lb_cur << 4) |
                  ((lb_feas ? 1 : 0) << 5) | (nfail << 8);
  TPP_I(I_FINISHED) = paused ? 0 : 1;
}
// The whole batch in ONE wave: 8 blocks of 64 threads per SM = 128 registers per thread (measured: with 144 registers
// the last blocks of a 65,536-problem batch wait for a second wave behind the slowest problems).
template <class M, bool BOX = false>
__global__ void __launch_bounds__(TPP_BLOCK, 8)
    tpp_solve_kernel(const __grid_constant__ TppCtx c, int maxiter, int is_feasible0, double reginit, int chunk, int first,
                     const int* __restrict__ list, const int* __restrict__ nactive, int b_lo, int b_hi) {
  __shared__ TppStage stg;
  tpp_solve_body<M, BOX>(c, stg, maxiter, is_feasible0, reginit, chunk, first, list, nactive, b_lo, b_hi);
}
// The same loop for the TAIL of a batch (the few problems that are still iterating when most have returned, compacted
// into `list`): occupancy is irrelevant there — each warp runs alone on its scheduler at the speed of its own dependent
// chain — so this instantiation takes all the registers it wants (no spills on the chain) and one warp per block.
template <class M, bool BOX = false>
__global__ void __launch_bounds__(TPP_BLOCK, 2)
    tpp_solve_tail_kernel(const __grid_constant__ TppCtx c, int maxiter, int is_feasible0, double reginit, int chunk,
                          int first, const int* __restrict__ list, const int* __restrict__ nactive, int b_lo, int b_hi) {
  __shared__ TppStage stg;
  tpp_solve_body<M, BOX>(c, stg, maxiter, is_feasible0, reginit, chunk, first, list, nactive, b_lo, b_hi);
}

// Ordered compaction of the unfinished problems (keeps neighbouring OCPs in neighbouring lanes):
//   pass 1: counts[blk] = unfinished problems among the 1024 of block blk
//   pass 2: list[prefix(blk) + rank in block] = b ; nactive = total
__global__ void __launch_bounds__(1024) tpp_compact_count_kernel(const int* __restrict__ finished, int B, int* counts) {
  const int b = blockIdx.x * 1024 + threadIdx.x;
  const int act = (b < B && finished[b] == 0) ? 1 : 0;
  const int n = __syncthreads_count(act);
  if (threadIdx.x == 0) counts[blockIdx.x] = n;
}
__global__ void __launch_bounds__(1024) tpp_compact_scatter_kernel(const int* __restrict__ finished, int B,
                                                                   const int* __restrict__ counts, int* list, int* nactive,
                                                                   int base) {
  __shared__ int warp_off[32];
  __shared__ int base_off;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int b = blockIdx.x * 1024 + tid;
  const int act = (b < B && finished[b] == 0) ? 1 : 0;
  const unsigned bal = __ballot_sync(0xffffffffu, act);
  if (lane == 0) warp_off[wid] = __popc(bal);
  if (tid == 0) {
    int s = 0;
    for (int i = 0; i < (int)blockIdx.x; ++i) s += counts[i];
    base_off = s;
    if (blockIdx.x == gridDim.x - 1) *nactive = s + counts[blockIdx.x];
  }
  __syncthreads();
  if (wid == 0) {  // exclusive scan of the 32 warp counts
    int v = warp_off[lane], x = v;
    for (int d = 1; d < 32; d <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    warp_off[lane] = x - v;
  }
  __syncthreads();
  if (act) list[base_off + warp_off[wid] + __popc(bal & ((1u << lane) - 1u))] = base + b;  // base: first problem of the slice
}

// ---------------------------------------------------------------------------------------------
// Step-wise kernels (same device functions, state kept in scal/iscal between calls).
// computeDirection(recalc) (ddp.cpp:120-125): calcDiff (calc + gaps, ddp.cpp:157-178; the step-wise API
// always runs with iter_ == 0, so problem_->calc is evaluated) + backwardPass.
template <class M, bool BOX = false>
__global__ void __launch_bounds__(64) tpp_direction_kernel(const __grid_constant__ TppCtx c, int recalc) {
  __shared__ TppStage stg;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const int Bp = c.Bp;
  M m;
  m.load(c.mparams);
  const int cur = TPP_I(I_CUR);
  const bool feas = TPP_I(I_FEAS) != 0, wasfeas = TPP_I(I_WASFEAS) != 0;
  if (recalc) {
    TPP_S(S_COST) = tpp_calc_pass(c, m, b, c.xu[cur]);
    TPP_I(I_FSBUF) = cur;
  }
  double dg = 0., dq = 0., stop = 0.;
  int fszero = 0;  // step-wise call: a feasible candidate always zero-fills the gap array
  (void)wasfeas;
  const int st = tpp_backward<M, BOX>(c, m, stg, b, cur, recalc != 0, feas, 1, TPP_S(S_XREG), TPP_S(S_UREG), dg, dq, stop,
                                      fszero);
  TPP_I(I_BSTATUS) = st;
  if (st == 0) {
    TPP_S(S_DG) = dg;
    TPP_S(S_DQ) = dq;
    TPP_S(S_STOP) = stop;
    if (!c.fddp) {
      TPP_S(S_D0) = dg;
      TPP_S(S_D1) = dq;
    }
  }
}

// tryStep(alpha) (ddp.cpp:127-130)
template <class M, bool BOX = false>
__global__ void __launch_bounds__(64) tpp_trystep_kernel(const __grid_constant__ TppCtx c, double alpha) {
  __shared__ TppStage stg;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  const int Bp = c.Bp;
  M m;
  m.load(c.mparams);
  const int cur = TPP_I(I_CUR);
  const bool feas = TPP_I(I_FEAS) != 0;
  double cost_try = 0., dv = 0.;
  const int st = tpp_forward<M, BOX>(c, m, stg, b, cur, alpha, feas, cost_try, dv);
  TPP_I(I_FSTATUS) = st;
  if (st == 0) {
    TPP_S(S_COST_TRY) = cost_try;
    TPP_S(S_DV) = TPP_S(S_COST) - cost_try;
    TPP_S(S_DVGAP) = dv;
    TPP_S(S_STEP) = alpha;
    if (c.fddp) {
      TPP_S(S_D0) = TPP_S(S_DG) + dv;
      TPP_S(S_D1) = TPP_S(S_DQ) - 2 * dv;
    }
  }
}

// ShootingProblem::calc / calcDiff for one (problem, node) per thread; inputs and results in the host ("user")
// layout [B][T+1][size] (shooting.hxx:133-195).
template <class M>
__global__ void tpp_node_data_kernel(TppCtx c, const double* __restrict__ xs, const double* __restrict__ us,
                                     int diff, double* cost, double* xnext, double* Fx, double* Fu, double* Lx,
                                     double* Lu, double* Lxx, double* Lxu, double* Luu, double* R) {
  constexpr int N = M::NX, U = M::NU;
  const int T = c.T;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  const int t = blockIdx.y;
  if (b >= c.B) return;
  M m;
  m.load(t < T ? c.mparams : c.tparams);
  double x[N], u[U], xn[N], fx[N * N], fu[N * U], lx[N], lu[U], lxx[N * N], lxu[N * U], luu[U * U];
  const size_t node = (size_t)b * (T + 1) + t;
#pragma unroll
  for (int i = 0; i < N; ++i) x[i] = xs[node * N + i];
#pragma unroll
  for (int i = 0; i < U; ++i) u[i] = (t < T) ? us[((size_t)b * T + (t < T ? t : 0)) * U + i] : 0.;
  if (!diff) {
    cost[node] = m.calc(x, u, xn);
    for (int i = 0; i < N; ++i) xnext[node * N + i] = xn[i];
    double r[M::NR];
    m.residual(x, u, r);
    for (int i = 0; i < M::NR; ++i) R[node * M::NR + i] = r[i];
  } else {
    m.calcBoth(x, u, xn, fx, fu, lx, lu, lxx, lxu, luu);
    for (int i = 0; i < N * N; ++i) Fx[node * N * N + i] = fx[i];
    for (int i = 0; i < N * U; ++i) Fu[node * N * U + i] = fu[i];
    for (int i = 0; i < N; ++i) Lx[node * N + i] = lx[i];
    for (int i = 0; i < U; ++i) Lu[node * U + i] = lu[i];
    for (int i = 0; i < N * N; ++i) Lxx[node * N * N + i] = lxx[i];
    for (int i = 0; i < N * U; ++i) Lxu[node * N * U + i] = lxu[i];
    for (int i = 0; i < U * U; ++i) Luu[node * U * U + i] = luu[i];
  }
}

// ShootingProblem::rollout (shooting.hxx:198-223); us/xs in the user layout
template <class M>
__global__ void tpp_rollout_kernel(TppCtx c, const double* __restrict__ us, double* __restrict__ xs) {
  constexpr int N = M::NX, U = M::NU;
  const int Bp = c.Bp, T = c.T;
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.B) return;
  M m;
  m.load(c.mparams);
  double x[N], u[U], xn[N];
#pragma unroll
  for (int i = 0; i < N; ++i) x[i] = c.x0[(size_t)i * Bp + b];
  for (int t = 0; t < T; ++t) {
#pragma unroll
    for (int i = 0; i < N; ++i) xs[((size_t)b * (T + 1) + t) * N + i] = x[i];
#pragma unroll
    for (int i = 0; i < U; ++i) u[i] = us[((size_t)b * T + t) * U + i];
    m.calc(x, u, xn);
#pragma unroll
    for (int i = 0; i < N; ++i) x[i] = xn[i];
  }
#pragma unroll
  for (int i = 0; i < N; ++i) xs[((size_t)b * (T + 1) + T) * N + i] = x[i];
}

// ---------------------------------------------------------------------------------------------
// Layout conversion between the host layout [nb][nt][size] and the record arrays. One thread per
// (problem, node): record side coalesced across the warp, host side strided (one pass per solve; small next to
// the solve itself). sel (may be null) picks base0/base1 per problem: buffer = (sel[b] ^ flip) & 1.
struct TppFieldMap {
  int R, size, nt;
  int off[16];
};
__global__ void tpp_gather_kernel(double* user, double* base0, double* base1, const int* sel, int flip,
                                  TppFieldMap fm, int Bp, int b0, int nb, int to_rec) {
  const int bl = blockIdx.x * blockDim.x + threadIdx.x;
  const int t = blockIdx.y;
  if (bl >= nb) return;
  const int b = b0 + bl;
  double* base = (sel && ((sel[b] ^ flip) & 1)) ? base1 : base0;
  double* rec = base + ((size_t)t * Bp + b) * fm.R;
  double* up = user + ((size_t)bl * fm.nt + t) * fm.size;
  if (to_rec) {
    for (int e = 0; e < fm.size; ++e) rec[fm.off[e]] = up[e];
  } else {
    for (int e = 0; e < fm.size; ++e) up[e] = rec[fm.off[e]];
  }
}
// Records of the LISTED problems -> the caller's array, one block per listed problem, consecutive threads writing
// consecutive doubles of that problem's [nt][size] block (coalesced: the destination may be page-locked HOST memory
// written through the unified address space, TppBackend::solve_pipelined). list holds global problem indices.
__global__ void tpp_gather_list_kernel(double* user, const double* base0, const double* base1, const int* sel, int flip,
                                       TppFieldMap fm, int Bp, const int* __restrict__ list, const int* __restrict__ nlist) {
  const int n = *nlist;
  const int per = fm.nt * fm.size;
  for (int j = blockIdx.x; j < n; j += gridDim.x) {
    const int b = list[j];
    const double* base = (sel && ((sel[b] ^ flip) & 1)) ? base1 : base0;
    double* up = user + (size_t)b * per;
    for (int i = threadIdx.x; i < per; i += blockDim.x) {
      const int t = i / fm.size, e = i - t * fm.size;
      up[i] = base[((size_t)t * Bp + b) * fm.R + fm.off[e]];
    }
  }
}
// x0: host [B][N] -> [N][Bp]
__global__ void tpp_x0_kernel(const double* user, double* x0, int N, int B, int Bp) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  for (int i = 0; i < N; ++i) x0[(size_t)i * Bp + b] = user[(size_t)b * N + i];
}

// setCandidate bookkeeping (solver-base.cpp:42-67) + reset of the step-wise state
__global__ void tpp_reset_kernel(TppCtx c, int is_feasible) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= c.Bp) return;
  const int Bp = c.Bp;
  TPP_I(I_ITER) = 0;
  TPP_I(I_FEAS) = is_feasible;
  TPP_I(I_WASFEAS) = 0;
  TPP_I(I_CONVERGED) = 0;
  TPP_I(I_CUR) = 0;
  TPP_I(I_FSBUF) = 0;
  TPP_I(I_NREC) = 0;
  TPP_I(I_BSTATUS) = 0;
  TPP_I(I_FSTATUS) = 0;
  TPP_I(I_NBWD) = 0;
  TPP_I(I_NFWD) = 0;
  TPP_S(S_DVGAP) = 0.;
}

}  // namespace cgx
