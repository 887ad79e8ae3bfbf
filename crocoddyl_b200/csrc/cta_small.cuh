// crocoddyl_b200 — SolverDDP::backwardPass for SMALL dense nodes (n <= 16, m <= 8: the 7-DoF arm, LQR-12): ONE WARP
// per OCP. The block-wide version in cta.cuh pays ~25 block barriers, several exposed HBM round trips and serial
// single-thread sections per node (measured on B200: 20 us per node at n = 14, m = 7 for a lone CTA); here the whole
// node is one warp's work between __syncwarp()s, the next node's derivatives are copied into shared memory with
// cp.async while the current node is factorised, and every phase is a strided sweep of dot products over the lanes.
// Operation order inside every dot product is the reference's (k ascending), so results are those of cta_backward.
//
// Reference: src/core/solvers/ddp.cpp:180-253 (backwardPass), 298-310 (computeGains), 132-155 (stoppingCriteria,
// expectedImprovement), src/core/solvers/fddp.cpp:126-147 (updateExpectedImprovement).
#pragma once
#include "boxqp.cuh"
#include "common.cuh"

namespace cgx {

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gsrc) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// shared-memory plan of the warp (doubles)
struct SmallPlan {
  int ldn, ldm;
  int stage_sz;                                      // one staging buffer: Fx, Fu, Lxx, Lxu, Luu, Lx, Lu, fs
  int oFx, oFu, oLxx, oLxu, oLuu, oLx, oLu, ofs;     // offsets inside a staging buffer
  int oV, oVx, oA, oG, oQx, oQu, oQxu, oQuu, oLc, oKg, okg, oQuuk, ow, ored, total;
  __host__ __device__ void make(int n, int m) {
    ldn = n | 1;
    ldm = m | 1;
    int p = 0;
    oFx = p; p += ldn * n;
    oFu = p; p += ldn * m;
    oLxx = p; p += ldn * n;
    oLxu = p; p += ldn * m;
    oLuu = p; p += ldm * m;
    oLx = p; p += n;
    oLu = p; p += m;
    ofs = p; p += n;
    stage_sz = p;
    p = 2 * stage_sz;
    oV = p; p += ldn * n;
    oVx = p; p += n;
    oA = p; p += ldn * n;
    oG = p; p += ldm * n;
    oQx = p; p += n;
    oQu = p; p += m;
    oQxu = p; p += ldn * m;
    oQuu = p; p += ldm * m;
    oLc = p; p += ldm * m;
    oKg = p; p += ldm * n;
    okg = p; p += m;
    oQuuk = p; p += m;
    ow = p; p += n;
    ored = p; p += 8;
    total = p + 2;
  }
};
inline size_t cta_small_smem_bytes(int n, int m) {
  SmallPlan pl;
  pl.make(n, m);
  return 8 * (size_t)pl.total;
}
__host__ __device__ inline bool cta_is_small(int n, int m) { return n <= 16 && m <= 8 && m >= 1; }

// strided sweep helper: element index e = lane, lane+32, ... of a rows x cols column-major matrix, tracked as (r, c)
// without integer division in the loop
struct ColMajorIter {
  int r, c, dr, dc, rows;
  __device__ ColMajorIter(int lane, int rows_) : rows(rows_) {
    r = lane % rows_;
    c = lane / rows_;
    dr = 32 % rows_;
    dc = 32 / rows_;
  }
  __device__ void next() {
    r += dr;
    c += dc;
    if (r >= rows) {
      r -= rows;
      ++c;
    }
  }
};

// asynchronous copy of a rows x cols column-major matrix from global memory (unpadded) into shared memory (ld)
__device__ __forceinline__ void warp_cp_matrix(double* dst, int ld, const double* src, int rows, int cols, int lane) {
  ColMajorIter it(lane, rows);
  for (int e = lane; e < rows * cols; e += 32, it.next()) cp_async8(dst + it.r + ld * it.c, src + e);
}
__device__ __forceinline__ void warp_st_matrix(double* gdst, const double* src, int ld, int rows, int cols, int lane) {
  ColMajorIter it(lane, rows);
  for (int e = lane; e < rows * cols; e += 32, it.next()) gdst[e] = src[it.r + ld * it.c];
}

// KN / KM: compile-time n / m (0 = run-time sizes bounded by n <= 16, m <= 8). Every lane keeps its outputs of a
// phase in registers and advances all of them together along k, so the shared-memory loads of one k-step are
// independent and pipelined (a single accumulator per lane serialises LDS -> DFMA: measured 20 us per node).
template <int KN, int KM, class M>
__device__ int cta_backward_warp(const CtaCtx& c, const M& mod, int b, int cur, bool is_feasible, double xreg, double ureg,
                                 double& dg_out, double& dq_out, double& stop_out) {
  extern __shared__ double cta_smem_raw[];
  double* S = cta_smem_raw;
  const int T = c.T, n = KN ? KN : c.n, m = KM ? KM : c.m;
  constexpr int O1 = KN ? ((KN + KM) * (KN + 1) + 31) / 32 : 13;   // outputs per lane of P1: (n+m) x (n+1)
  constexpr int O2 = KN ? (KN * (KN + KM) + 31) / 32 : 12;         // P2a: n x (n+m)
  constexpr int O3 = KN ? (KM * KM + 31) / 32 : 2;                 // P2b: m x m
  constexpr int O5 = KN ? (KN * KN + 31) / 32 : 8;                 // P5: n x n
  SmallPlan pl;
  pl.make(n, m);
  const int ldn = pl.ldn, ldm = pl.ldm;
  int* flag = reinterpret_cast<int*>(S + pl.total - 2);
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    const bool fddp = c.fddp != 0, full = c.full != 0;
    const bool has_xreg = !isnan(xreg), has_ureg = !isnan(ureg);
    const double* fsg = c.fs + (size_t)b * (T + 1) * n;
    double *V = S + pl.oV, *Vx = S + pl.oVx, *A = S + pl.oA, *G = S + pl.oG, *Qx = S + pl.oQx, *Qu = S + pl.oQu;
    double *Qxu = S + pl.oQxu, *Quu = S + pl.oQuu, *Lc = S + pl.oLc, *Kg = S + pl.oKg, *kg = S + pl.okg;
    double *Quuk = S + pl.oQuuk, *w = S + pl.ow;
    double dg = 0., dq = 0., stop = 0.;  // lane 0 accumulates
    int status = 0;
    bool mats_staged = false;
    // issue the asynchronous loads of node t into staging buffer st
    auto load_node = [&](int t, double* st) {
      const NodeMats nm = mod.mats(c, b, t);
      const size_t node = (size_t)b * (T + 1) + t;
      if (!(nm.constant && mats_staged)) {
        double* mb = nm.constant ? S : st;  // constant matrices live in buffer 0 for the whole sweep
        warp_cp_matrix(mb + pl.oFx, ldn, nm.Fx, n, n, lane);
        warp_cp_matrix(mb + pl.oFu, ldn, nm.Fu, n, m, lane);
        warp_cp_matrix(mb + pl.oLxx, ldn, nm.Lxx, n, n, lane);
        if (nm.Lxu) warp_cp_matrix(mb + pl.oLxu, ldn, nm.Lxu, n, m, lane);
        warp_cp_matrix(mb + pl.oLuu, ldm, nm.Luu, m, m, lane);
        mats_staged = true;
      }
      if (lane < n) cp_async8(st + pl.oLx + lane, c.d_Lx + node * n + lane);
      if (lane < m && t < T) cp_async8(st + pl.oLu + lane, c.d_Lu + node * m + lane);
      if (!is_feasible && lane < n) cp_async8(st + pl.ofs + lane, fsg + (size_t)t * n + lane);
    };
    const bool has_Lxu = mod.mats(c, b, 0).Lxu != nullptr;
    const bool constant = mod.mats(c, b, 0).constant;
    // ---- terminal node (ddp.cpp:181-191)
    {
      double* st = S;  // buffer 0
      load_node(T, st);
      cp_async_wait_all();
      __syncwarp();
      const double* mb = S;  // terminal matrices are in buffer 0 in both cases
      {
        ColMajorIter it(lane, n);
        for (int e = lane; e < n * n; e += 32, it.next()) {
          double v = mb[pl.oLxx + it.r + ldn * it.c];
          if (has_xreg && it.r == it.c) v += xreg;
          V[it.r + ldn * it.c] = v;
        }
      }
      if (lane < n) Vx[lane] = st[pl.oLx + lane];
      __syncwarp();
      if (!is_feasible && lane < n) {
        double s = 0.;
        for (int j = 0; j < n; ++j) s += V[lane + ldn * j] * st[pl.ofs + j];
        w[lane] = s;
        Vx[lane] += s;
      }
      __syncwarp();
      const size_t node = (size_t)b * (T + 1) + T;
      if (lane < n) {
        c.Vx[node * n + lane] = Vx[lane];
        c.w[node * n + lane] = is_feasible ? 0. : w[lane];
      }
      warp_st_matrix(c.Vxx + node * n * n, V, ldn, n, n, lane);
      if (lane == 0 && fddp && !is_feasible) {  // fddp.cpp:131-135
        double a = 0., q = 0.;
        for (int i = 0; i < n; ++i) {
          a += Vx[i] * st[pl.ofs + i];
          q += st[pl.ofs + i] * w[i];
        }
        dg -= a;
        dq += q;
      }
      __syncwarp();
      mats_staged = false;  // the terminal model may differ from the running one (shooting.hpp:48-63): restage
      if (T > 0) load_node(T - 1, S + pl.stage_sz);  // buffer 1
    }
    for (int t = T - 1; t >= 0; --t) {
      double* st = S + (((T - t) & 1) ? pl.stage_sz : 0);
      const double* mb = constant ? S : st;
      const size_t node = (size_t)b * (T + 1) + t, rnode = (size_t)b * T + t;
      cp_async_wait_all();
      __syncwarp();
      if (t > 0) load_node(t - 1, S + (((T - t + 1) & 1) ? pl.stage_sz : 0));
      const double* F = mb + pl.oFx;  // [Fx Fu], n x (n+m), leading dimension ldn
      // P1: [A; G | Qx; Qu] = [Fx Fu]^T [Vxx' Vx']   (ddp.cpp:204,206,211,214)
      if constexpr (KN > 0 && KM > 0 && KN == 2 * KM) {
        // register tiles: lane (rg, cg) owns rows {3 rg .. 3 rg + 2} x columns {TC cg ..} of the (n + m) x (n + 1)
        // product, so a k-step loads 3 + TC operands for 3 TC multiply-adds (the strided version below loads 2 per
        // multiply-add and was bound by the shared-memory pipe: ncu, FP64 pipe 10 %). Each output still sums over k
        // ascending: same values.
        constexpr int RG = KM, CG = 32 / RG, TR = 3, NC = KN + 1, TC = (NC + CG - 1) / CG;
        const int rg = lane % RG, cg = lane / RG;
        if (cg < CG) {
          double acc[TR][TC];
          int fo[TR], vo[TC];
#pragma unroll
          for (int a = 0; a < TR; ++a) fo[a] = ldn * (TR * rg + a);
#pragma unroll
          for (int bb = 0; bb < TC; ++bb) vo[bb] = ldn * min(TC * cg + bb, NC - 1);
#pragma unroll
          for (int a = 0; a < TR; ++a)
#pragma unroll
            for (int bb = 0; bb < TC; ++bb) acc[a][bb] = 0.;
#pragma unroll
          for (int k = 0; k < KN; ++k) {
            double fv[TR], vv[TC];
#pragma unroll
            for (int a = 0; a < TR; ++a) fv[a] = F[fo[a] + k];
#pragma unroll
            for (int bb = 0; bb < TC; ++bb) vv[bb] = V[vo[bb] + k];
#pragma unroll
            for (int a = 0; a < TR; ++a)
#pragma unroll
              for (int bb = 0; bb < TC; ++bb) acc[a][bb] += fv[a] * vv[bb];
          }
#pragma unroll
          for (int a = 0; a < TR; ++a)
#pragma unroll
            for (int bb = 0; bb < TC; ++bb) {
              const int i = TR * rg + a, j = TC * cg + bb;
              if (j >= NC) continue;
              if (j < n) {
                if (i < n)
                  A[i + ldn * j] = acc[a][bb];
                else
                  G[(i - n) + ldm * j] = acc[a][bb];
              } else if (i < n) {
                Qx[i] = st[pl.oLx + i] + acc[a][bb];
              } else {
                Qu[i - n] = st[pl.oLu + i - n] + acc[a][bb];
              }
            }
        }
      } else
      {
        const int R1 = n + m, NO = R1 * (n + 1);
        double acc[O1];
        int fo[O1], vo[O1];
        {
          ColMajorIter it(lane, R1);
#pragma unroll
          for (int q = 0; q < O1; ++q, it.next()) {
            const bool ok = lane + 32 * q < NO;
            fo[q] = ok ? ldn * it.r : 0;
            vo[q] = ok ? ldn * it.c : 0;  // column n of [V | Vx] is Vx (contiguous after V)
            acc[q] = 0.;
          }
        }
#pragma unroll
        for (int k = 0; k < (KN ? KN : 16); ++k) {
          if (!KN && k >= n) break;
#pragma unroll
          for (int q = 0; q < O1; ++q) acc[q] += F[fo[q] + k] * V[vo[q] + k];
        }
        ColMajorIter it(lane, R1);
#pragma unroll
        for (int q = 0; q < O1; ++q, it.next()) {
          if (lane + 32 * q >= NO) break;
          const int i = it.r, j = it.c;
          if (j < n) {
            if (i < n)
              A[i + ldn * j] = acc[q];
            else
              G[(i - n) + ldm * j] = acc[q];
          } else if (i < n) {
            Qx[i] = st[pl.oLx + i] + acc[q];
          } else {
            Qu[i - n] = st[pl.oLu + i - n] + acc[q];
          }
        }
      }
      __syncwarp();
      // P2: Qxx = Lxx + A Fx (into V: Vxx' is dead), Qxu = Lxu + A Fu, Quu = Luu + G Fu (+ ureg)   (ddp.cpp:202-218)
      {
        const int NO = n * (n + m);
        double acc[O2];
        int ao[O2], fo[O2];
        {
          ColMajorIter it(lane, n);
#pragma unroll
          for (int q = 0; q < O2; ++q, it.next()) {
            const bool ok = lane + 32 * q < NO;
            ao[q] = ok ? it.r : 0;
            fo[q] = ok ? ldn * it.c : 0;
            acc[q] = 0.;
          }
        }
        double acu[O3];
        int go[O3], fu[O3];
        {
          ColMajorIter iu(lane, m);
#pragma unroll
          for (int q = 0; q < O3; ++q, iu.next()) {
            const bool ok = lane + 32 * q < m * m;
            go[q] = ok ? iu.r : 0;
            fu[q] = ok ? ldn * (n + iu.c) : 0;
            acu[q] = 0.;
          }
        }
        constexpr bool TILED = KN > 0 && KM > 0 && KN == 2 * KM;
#pragma unroll
        for (int k = 0; k < (KN ? KN : 16); ++k) {
          if (!KN && k >= n) break;
          if constexpr (!TILED) {
#pragma unroll
            for (int q = 0; q < O2; ++q) acc[q] += A[ao[q] + ldn * k] * F[fo[q] + k];
          }
#pragma unroll
          for (int q = 0; q < O3; ++q) acu[q] += G[go[q] + ldm * k] * F[fu[q] + k];
        }
        if constexpr (TILED) {
          // [Qxx Qxu] = [Lxx Lxu] + A [Fx Fu]: lane (rg, cg) owns rows {2 rg, 2 rg + 1} x TC columns of n x (n + m)
          constexpr int RG = KM, CG = 32 / RG, TR = 2, NC = KN + KM, TC = (NC + CG - 1) / CG;
          const int rg = lane % RG, cg = lane / RG;
          if (cg < CG) {
            double t[TR][TC];
            int fo2[TC];
#pragma unroll
            for (int bb = 0; bb < TC; ++bb) fo2[bb] = ldn * min(TC * cg + bb, NC - 1);
#pragma unroll
            for (int a = 0; a < TR; ++a)
#pragma unroll
              for (int bb = 0; bb < TC; ++bb) t[a][bb] = 0.;
#pragma unroll
            for (int k = 0; k < KN; ++k) {
              double av[TR], fv[TC];
#pragma unroll
              for (int a = 0; a < TR; ++a) av[a] = A[TR * rg + a + ldn * k];
#pragma unroll
              for (int bb = 0; bb < TC; ++bb) fv[bb] = F[fo2[bb] + k];
#pragma unroll
              for (int a = 0; a < TR; ++a)
#pragma unroll
                for (int bb = 0; bb < TC; ++bb) t[a][bb] += av[a] * fv[bb];
            }
#pragma unroll
            for (int a = 0; a < TR; ++a)
#pragma unroll
              for (int bb = 0; bb < TC; ++bb) {
                const int i = TR * rg + a, j = TC * cg + bb;
                if (j >= NC) continue;
                if (j < n)
                  V[i + ldn * j] = mb[pl.oLxx + i + ldn * j] + t[a][bb];
                else
                  Qxu[i + ldn * (j - n)] = (has_Lxu ? mb[pl.oLxu + i + ldn * (j - n)] : 0.) + t[a][bb];
              }
          }
        }
        ColMajorIter it(lane, n);
#pragma unroll
        for (int q = 0; q < O2; ++q, it.next()) {
          if (TILED || lane + 32 * q >= NO) break;
          const int i = it.r, j = it.c;
          if (j < n)
            V[i + ldn * j] = mb[pl.oLxx + i + ldn * j] + acc[q];
          else
            Qxu[i + ldn * (j - n)] = (has_Lxu ? mb[pl.oLxu + i + ldn * (j - n)] : 0.) + acc[q];
        }
        ColMajorIter iu(lane, m);
#pragma unroll
        for (int q = 0; q < O3; ++q, iu.next()) {
          if (lane + 32 * q >= m * m) break;
          const int i = iu.r, j = iu.c;
          double v = mb[pl.oLuu + i + ldm * j] + acu[q];
          if (has_ureg && i == j) v += ureg;
          Quu[i + ldm * j] = v;
          Lc[i + ldm * j] = v;
        }
        if (lane == 0) *flag = 0;
      }
      __syncwarp();
      if (full) {
        warp_st_matrix(c.Qxx + rnode * n * n, V, ldn, n, n, lane);
        warp_st_matrix(c.Qxu + rnode * n * m, Qxu, ldn, n, m, lane);
        warp_st_matrix(c.Quu + rnode * m * m, Quu, ldm, m, m, lane);
        if (lane < n) c.Qx[rnode * n + lane] = Qx[lane];
        if (lane < m) c.Qu[rnode * m + lane] = Qu[lane];
      }
      // P3 + P4: computeGains (ddp.cpp:298-310): unblocked lower LLT of Quu, then K = Quu^-1 Qxu^T, k = Quu^-1 Qu.
      // SolverBoxDDP/BoxFDDP::computeGains (box-ddp.cpp:46-74): box QP on the control step when the model has control
      // limits and the candidate is feasible (lane 0; work arrays in the problem's global scratch)
      const bool boxed = c.box && is_feasible && c.has_limits[b] != 0;
      if (boxed) {
        if (lane == 0) {
          const double* u_t = c.us[cur] + rnode * m;  // us_[t] of the candidate
          for (int i = 0; i < m; ++i) kg[i] = c.k[rnode * m + i];  // k_[t] of the previous backward pass (warm start)
          *flag = box_compute_gains(n, m, Quu, ldm, Qu, Qxu, ldn, u_t, c.u_lb + (size_t)b * m, c.u_ub + (size_t)b * m, Kg,
                                    ldm, kg, c.Quu_inv + rnode * m * m, c.o,
                                    c.box_work + (size_t)b * box_gains_work_doubles(m),
                                    c.box_iwork + (size_t)b * box_gains_work_ints(m));
        }
        __syncwarp();
        if (*flag) {
          status = 1;
          break;
        }
        if (full && lane < m) c.Qu[rnode * m + lane] = Qu[lane];  // get_Qu returns the clamped gradient
      } else if constexpr (KM > 0) {
        // Compile-time m: EVERY lane factorises Quu redundantly in registers — no barriers and no shared-memory round
        // trips inside the serial chain — keeps the reciprocal pivots, and solves its own right-hand side in registers.
        // (ncu source page of the shared-memory version below: 36 % of the backward kernel's samples in these lines.)
        // Same operation order as Eigen's unblocked LLT / the oracle; x / p becomes x * (1 / p) (<= 1 ulp).
        constexpr int NLR = KM * (KM + 1) / 2;
        double Lr[NLR], rinv[KM];
#pragma unroll
        for (int i = 0; i < KM; ++i)
#pragma unroll
          for (int j = 0; j <= i; ++j) Lr[i * (i + 1) / 2 + j] = Lc[i + ldm * j];
        bool notpd = false;
#pragma unroll
        for (int kk = 0; kk < KM; ++kk) {
          double p = Lr[kk * (kk + 1) / 2 + kk];
#pragma unroll
          for (int j = 0; j < kk; ++j) p -= Lr[kk * (kk + 1) / 2 + j] * Lr[kk * (kk + 1) / 2 + j];
          if (p <= 0.) notpd = true;  // a NaN pivot does not fail here; raiseIfNaN catches it below
          double pinv;
          llt_pivot(p, p, pinv);
          Lr[kk * (kk + 1) / 2 + kk] = p;
          rinv[kk] = pinv;
#pragma unroll
          for (int i = kk + 1; i < KM; ++i) {
            double sv = Lr[i * (i + 1) / 2 + kk];
#pragma unroll
            for (int j = 0; j < kk; ++j) sv -= Lr[i * (i + 1) / 2 + j] * Lr[kk * (kk + 1) / 2 + j];
            Lr[i * (i + 1) / 2 + kk] = div_pivot(sv, p, pinv);
          }
        }
        if (notpd) {
          status = 1;
          break;
        }
        if (lane <= n) {
          double rr[KM];
#pragma unroll
          for (int i = 0; i < KM; ++i) rr[i] = (lane < n) ? Qxu[lane + ldn * i] : Qu[i];
#pragma unroll
          for (int i = 0; i < KM; ++i) {
            double sv = rr[i];
#pragma unroll
            for (int j = 0; j < i; ++j) sv -= Lr[i * (i + 1) / 2 + j] * rr[j];
            rr[i] = div_pivot(sv, Lr[i * (i + 1) / 2 + i], rinv[i]);
          }
#pragma unroll
          for (int i = KM - 1; i >= 0; --i) {
            double sv = rr[i];
#pragma unroll
            for (int j = i + 1; j < KM; ++j) sv -= Lr[j * (j + 1) / 2 + i] * rr[j];
            rr[i] = div_pivot(sv, Lr[i * (i + 1) / 2 + i], rinv[i]);
          }
          double* r = (lane < n) ? (Kg + ldm * lane) : kg;
#pragma unroll
          for (int i = 0; i < KM; ++i) r[i] = rr[i];
        }
      } else {
      // run-time m: unblocked lower LLT in shared memory, lane i owns row i
      for (int kk = 0; kk < m; ++kk) {
        double p = Lc[kk + ldm * kk];
        for (int j = 0; j < kk; ++j) p -= Lc[kk + ldm * j] * Lc[kk + ldm * j];
        if (p <= 0.) *flag = 1;
        p = sqrt(p);
        __syncwarp();
        if (lane == 0) Lc[kk + ldm * kk] = p;
        if (lane > kk && lane < m) {
          double s = Lc[lane + ldm * kk];
          for (int j = 0; j < kk; ++j) s -= Lc[lane + ldm * j] * Lc[kk + ldm * j];
          Lc[lane + ldm * kk] = s / p;
        }
        __syncwarp();
      }
      if (*flag) {
        status = 1;
        break;
      }
      // P4: K = Quu^-1 Qxu^T, k = Quu^-1 Qu: one right-hand side per lane
      if (lane <= n) {
        double* r = (lane < n) ? (Kg + ldm * lane) : kg;
        for (int i = 0; i < m; ++i) r[i] = (lane < n) ? Qxu[lane + ldn * i] : Qu[i];
        for (int i = 0; i < m; ++i) {
          double s = r[i];
          for (int j = 0; j < i; ++j) s -= Lc[i + ldm * j] * r[j];
          r[i] = s / Lc[i + ldm * i];
        }
        for (int i = m - 1; i >= 0; --i) {
          double s = r[i];
          for (int j = i + 1; j < m; ++j) s -= Lc[j + ldm * i] * r[j];
          r[i] = s / Lc[i + ldm * i];
        }
      }
      }
      __syncwarp();
      if (lane < m) {  // ddp.cpp:229
        double s = 0.;
        for (int j = 0; j < m; ++j) s += Quu[lane + ldm * j] * kg[j];
        Quuk[lane] = s;
      }
      __syncwarp();
      // P5: Vx = Qx + K^T Quuk - 2 K^T Qu ; Vxx = Qxx - Qxu K   (ddp.cpp:223-233)
      if (lane < n) {
        double s1 = 0., s2 = 0.;
        for (int i = 0; i < m; ++i) {
          s1 += Kg[i + ldm * lane] * Quuk[i];
          s2 += Kg[i + ldm * lane] * Qu[i];
        }
        double v = Qx[lane];
        if (has_ureg) {
          v += s1;
          v += -2. * s2;
        } else {
          v += -1. * s2;
        }
        Vx[lane] = v;
      }
      if constexpr (KN > 0 && KM > 0 && KN == 2 * KM) {
        // Vxx = Qxx - Qxu K: lane (rg, cg) owns rows {2 rg, 2 rg + 1} x TC columns of n x n
        constexpr int RG = KM, CG = 32 / RG, TR = 2, TC = (KN + CG - 1) / CG;
        const int rg = lane % RG, cg = lane / RG;
        if (cg < CG) {
          double t[TR][TC];
          int ko2[TC];
#pragma unroll
          for (int bb = 0; bb < TC; ++bb) ko2[bb] = ldm * min(TC * cg + bb, KN - 1);
#pragma unroll
          for (int a = 0; a < TR; ++a)
#pragma unroll
            for (int bb = 0; bb < TC; ++bb) t[a][bb] = 0.;
#pragma unroll
          for (int k = 0; k < KM; ++k) {
            double qv[TR], kv[TC];
#pragma unroll
            for (int a = 0; a < TR; ++a) qv[a] = Qxu[TR * rg + a + ldn * k];
#pragma unroll
            for (int bb = 0; bb < TC; ++bb) kv[bb] = Kg[ko2[bb] + k];
#pragma unroll
            for (int a = 0; a < TR; ++a)
#pragma unroll
              for (int bb = 0; bb < TC; ++bb) t[a][bb] += qv[a] * kv[bb];
          }
#pragma unroll
          for (int a = 0; a < TR; ++a)
#pragma unroll
            for (int bb = 0; bb < TC; ++bb) {
              const int i = TR * rg + a, j = TC * cg + bb;
              if (j < KN) V[i + ldn * j] = V[i + ldn * j] + -1. * t[a][bb];
            }
        }
      } else {
        double acc[O5];
        int qo[O5], ko[O5];
        {
          ColMajorIter it(lane, n);
#pragma unroll
          for (int q = 0; q < O5; ++q, it.next()) {
            const bool ok = lane + 32 * q < n * n;
            qo[q] = ok ? it.r : 0;
            ko[q] = ok ? ldm * it.c : 0;
            acc[q] = 0.;
          }
        }
#pragma unroll
        for (int k = 0; k < (KM ? KM : 8); ++k) {
          if (!KM && k >= m) break;
#pragma unroll
          for (int q = 0; q < O5; ++q) acc[q] += Qxu[qo[q] + ldn * k] * Kg[ko[q] + k];
        }
        ColMajorIter it(lane, n);
#pragma unroll
        for (int q = 0; q < O5; ++q, it.next()) {
          if (lane + 32 * q >= n * n) break;
          V[it.r + ldn * it.c] = V[it.r + ldn * it.c] + -1. * acc[q];
        }
      }
      if (lane == 0) {  // ddp.cpp:132-155 / fddp.cpp:137-140
        double a = 0., q = 0., sq = 0.;
        for (int i = 0; i < m; ++i) {
          a += Qu[i] * kg[i];
          q += kg[i] * Quuk[i];
          sq += Qu[i] * Qu[i];
        }
        dg += a;
        dq -= q;
        stop += sq;
      }
      __syncwarp();
      // P6: Vxx = .5 (Vxx + Vxx^T) ; diag += xreg   (ddp.cpp:235-239)
      {
        ColMajorIter it(lane, n);
        for (int o = lane; o < n * n; o += 32, it.next()) {
          const int i = it.r, j = it.c;
          if (i > j) {
            const double s = 0.5 * (V[i + ldn * j] + V[j + ldn * i]);
            V[i + ldn * j] = s;
            V[j + ldn * i] = s;
          } else if (i == j) {
            double s = 0.5 * (V[i + ldn * i] + V[i + ldn * i]);
            if (has_xreg) s += xreg;
            V[i + ldn * i] = s;
          }
        }
      }
      __syncwarp();
      if (!is_feasible && lane < n) {  // Vx += Vxx fs[t]   (ddp.cpp:242-244)
        double s = 0.;
        for (int j = 0; j < n; ++j) s += V[lane + ldn * j] * st[pl.ofs + j];
        w[lane] = s;
        Vx[lane] += s;
      }
      __syncwarp();
      // raiseIfNaN(|Vx|_inf), raiseIfNaN(|Vxx|_inf)   (ddp.cpp:246-251) + stores
      int bad = 0;
      if (lane < n) {
        bad |= raise_if_nan(fabs(Vx[lane]));
        c.Vx[node * n + lane] = Vx[lane];
        c.w[node * n + lane] = is_feasible ? 0. : w[lane];
      }
      if (lane < m) c.k[rnode * m + lane] = kg[lane];
      {
        ColMajorIter it(lane, n);
        double* gV = c.Vxx + node * n * n;
        for (int e = lane; e < n * n; e += 32, it.next()) {
          const double v = V[it.r + ldn * it.c];
          bad |= raise_if_nan(fabs(v));
          gV[e] = v;
        }
      }
      warp_st_matrix(c.K + rnode * m * n, Kg, ldm, m, n, lane);
      if (lane == 0 && fddp && !is_feasible) {  // fddp.cpp:141-145
        double a = 0., q = 0.;
        for (int i = 0; i < n; ++i) {
          a += Vx[i] * st[pl.ofs + i];
          q += st[pl.ofs + i] * w[i];
        }
        dg -= a;
        dq += q;
      }
      if (__any_sync(0xffffffffu, bad)) {
        status = 1;
        break;
      }
    }
    cp_async_wait_all();
    __syncwarp();
    if (lane == 0) {
      S[pl.ored + 0] = dg;
      S[pl.ored + 1] = dq;
      S[pl.ored + 2] = stop;
      *flag = status;
    }
  }
  __syncthreads();
  dg_out = S[pl.ored + 0];
  dq_out = S[pl.ored + 1];
  stop_out = S[pl.ored + 2];
  const int st = *flag;
  __syncthreads();
  return st;
}

}  // namespace cgx
