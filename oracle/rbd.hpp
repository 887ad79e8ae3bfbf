// TEST INFRASTRUCTURE — CPU oracle (see linalg.hpp header). Not part of the product.
//
// Rigid-body algorithms for a fixed-base serial chain of 1-DoF revolute joints, restating the PUBLISHED
// algorithms of the third-party dependency the reference calls and that is NOT vendored under /root/reference:
// Pinocchio (>= 2.4.7, CMakeLists.txt:63). Call sites in the reference:
//   pinocchio::aba / updateGlobalPlacements      include/crocoddyl/multibody/actions/free-fwddyn.hxx:64-65
//   pinocchio::computeABADerivatives             free-fwddyn.hxx:104-105
//   pinocchio::rnea (quasiStatic)                free-fwddyn.hxx:158-159
//   updateFramePlacement, log6, Jlog6, getFrameJacobian(LOCAL)   multibody/costs/frame-placement.hxx:57-76
// PARITY UNPINNED for this arithmetic: the reference holds no golden values of ABA / ABA derivatives and Pinocchio
// is not installed here, so these functions are pinned only by self-consistency (tests/test_oracle_arm_cpu.py):
// finite differences, M(q) qdd + b(q,v) = tau with M from CRBA and b from RNEA, d(qdd)/d(tau) = M^-1 symmetric,
// closed-form pendulum, Jlog6 against the derivative of log6.
// Conventions (Pinocchio's): spatial motion = [linear; angular], force = [force; torque]; SE3 (R,p) maps child
// coordinates to parent coordinates; joint i: placement in parent at q = 0 (Rp, pp), unit axis a_i, S_i = [0; a_i];
// body inertia = (mass, COM lever c, rotational inertia about the COM); gravity enters as base acceleration -g.
// Algorithms: Featherstone's ABA, RNEA, CRBA (local frames); ABA derivatives as
//   d(qdd)/d(tau) = M^-1,  d(qdd)/dq = -M^-1 dRNEA/dq |_qdd,  d(qdd)/dv = -M^-1 dRNEA/dv
// with the RNEA partial derivatives from the recursive chain rule (Carpentier & Mansard, RSS 2018, local form).
#pragma once
#include <cmath>
#include <limits>
#include <stdexcept>
#include <vector>

namespace oracle {
namespace rbd {

constexpr int MAXNV = 16;

template <typename S>
struct V3 {
  S x, y, z;
};
template <typename S>
inline V3<S> add(const V3<S>& a, const V3<S>& b) { return V3<S>{a.x + b.x, a.y + b.y, a.z + b.z}; }
template <typename S>
inline V3<S> sub(const V3<S>& a, const V3<S>& b) { return V3<S>{a.x - b.x, a.y - b.y, a.z - b.z}; }
template <typename S>
inline V3<S> scale(const S& s, const V3<S>& a) { return V3<S>{s * a.x, s * a.y, s * a.z}; }
template <typename S>
inline V3<S> cross(const V3<S>& a, const V3<S>& b) {
  return V3<S>{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <typename S>
inline S dot(const V3<S>& a, const V3<S>& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// 3x3 row-major
template <typename S>
struct M3 {
  S m[9];
  S& operator()(int i, int j) { return m[3 * i + j]; }
  const S& operator()(int i, int j) const { return m[3 * i + j]; }
};
template <typename S>
inline V3<S> mul(const M3<S>& A, const V3<S>& v) {
  return V3<S>{A.m[0] * v.x + A.m[1] * v.y + A.m[2] * v.z, A.m[3] * v.x + A.m[4] * v.y + A.m[5] * v.z,
               A.m[6] * v.x + A.m[7] * v.y + A.m[8] * v.z};
}
template <typename S>
inline V3<S> mulT(const M3<S>& A, const V3<S>& v) {
  return V3<S>{A.m[0] * v.x + A.m[3] * v.y + A.m[6] * v.z, A.m[1] * v.x + A.m[4] * v.y + A.m[7] * v.z,
               A.m[2] * v.x + A.m[5] * v.y + A.m[8] * v.z};
}
template <typename S>
inline M3<S> mul(const M3<S>& A, const M3<S>& B) {
  M3<S> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C(i, j) = A(i, 0) * B(0, j) + A(i, 1) * B(1, j) + A(i, 2) * B(2, j);
  return C;
}
template <typename S>
inline M3<S> mulTA(const M3<S>& A, const M3<S>& B) {  // A^T B
  M3<S> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C(i, j) = A(0, i) * B(0, j) + A(1, i) * B(1, j) + A(2, i) * B(2, j);
  return C;
}

template <typename S>
struct SE3 {
  M3<S> R;
  V3<S> p;
};
template <typename S>
inline SE3<S> compose(const SE3<S>& a, const SE3<S>& b) {  // a * b
  return SE3<S>{mul(a.R, b.R), add(a.p, mul(a.R, b.p))};
}
template <typename S>
inline SE3<S> inverse(const SE3<S>& a) {
  SE3<S> r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.R(i, j) = a.R(j, i);
  V3<S> t = mulT(a.R, a.p);
  r.p = V3<S>{-t.x, -t.y, -t.z};
  return r;
}
// a^-1 * b  (Pinocchio's actInv on a placement)
template <typename S>
inline SE3<S> actInv(const SE3<S>& a, const SE3<S>& b) {
  return SE3<S>{mulTA(a.R, b.R), mulT(a.R, sub(b.p, a.p))};
}

// spatial motion / force: lin = linear (force), ang = angular (torque)
template <typename S>
struct Sp {
  V3<S> lin, ang;
};
template <typename S>
inline Sp<S> add(const Sp<S>& a, const Sp<S>& b) { return Sp<S>{add(a.lin, b.lin), add(a.ang, b.ang)}; }
template <typename S>
inline Sp<S> scale(const S& s, const Sp<S>& a) { return Sp<S>{scale(s, a.lin), scale(s, a.ang)}; }
// motion in child coordinates -> parent coordinates, M = (R,p) child->parent
template <typename S>
inline Sp<S> actM(const SE3<S>& M, const Sp<S>& v) {
  V3<S> w = mul(M.R, v.ang);
  return Sp<S>{add(mul(M.R, v.lin), cross(M.p, w)), w};
}
template <typename S>
inline Sp<S> actInvM(const SE3<S>& M, const Sp<S>& v) {
  return Sp<S>{mulT(M.R, sub(v.lin, cross(M.p, v.ang))), mulT(M.R, v.ang)};
}
template <typename S>
inline Sp<S> actF(const SE3<S>& M, const Sp<S>& f) {
  V3<S> l = mul(M.R, f.lin);
  return Sp<S>{l, add(mul(M.R, f.ang), cross(M.p, l))};
}
template <typename S>
inline Sp<S> actInvF(const SE3<S>& M, const Sp<S>& f) {
  return Sp<S>{mulT(M.R, f.lin), mulT(M.R, sub(f.ang, cross(M.p, f.lin)))};
}
// motion x motion, motion x* force
template <typename S>
inline Sp<S> crossM(const Sp<S>& a, const Sp<S>& b) {
  return Sp<S>{add(cross(a.ang, b.lin), cross(a.lin, b.ang)), cross(a.ang, b.ang)};
}
template <typename S>
inline Sp<S> crossF(const Sp<S>& v, const Sp<S>& f) {
  return Sp<S>{cross(v.ang, f.lin), add(cross(v.ang, f.ang), cross(v.lin, f.lin))};
}
template <typename S>
inline S dotS(const Sp<S>& a, const Sp<S>& b) { return dot(a.lin, b.lin) + dot(a.ang, b.ang); }

// rigid-body inertia (mass, COM lever, rotational inertia about the COM)
template <typename S>
struct Inertia {
  S mass;
  V3<S> c;
  M3<S> I;
};
template <typename S>
inline Sp<S> mulI(const Inertia<S>& Y, const Sp<S>& v) {
  V3<S> l = scale(Y.mass, sub(v.lin, cross(Y.c, v.ang)));
  return Sp<S>{l, add(mul(Y.I, v.ang), cross(Y.c, l))};
}

// 6x6 column-major matrix acting on [lin; ang]
template <typename S>
struct M6 {
  S m[36];
  S& operator()(int i, int j) { return m[i + 6 * j]; }
  const S& operator()(int i, int j) const { return m[i + 6 * j]; }
};
template <typename S>
inline void toArr(const Sp<S>& v, S* a) {
  a[0] = v.lin.x; a[1] = v.lin.y; a[2] = v.lin.z; a[3] = v.ang.x; a[4] = v.ang.y; a[5] = v.ang.z;
}
template <typename S>
inline Sp<S> fromArr(const S* a) { return Sp<S>{V3<S>{a[0], a[1], a[2]}, V3<S>{a[3], a[4], a[5]}}; }
template <typename S>
inline Sp<S> mulM6(const M6<S>& A, const Sp<S>& v) {
  S a[6], r[6];
  toArr(v, a);
  for (int i = 0; i < 6; ++i) {
    S s = S(0);
    for (int j = 0; j < 6; ++j) s += A(i, j) * a[j];
    r[i] = s;
  }
  return fromArr(r);
}
template <typename S>
inline M6<S> inertiaMatrix(const Inertia<S>& Y) {
  M6<S> A;
  for (int j = 0; j < 6; ++j) {
    S e[6] = {S(0), S(0), S(0), S(0), S(0), S(0)};
    e[j] = S(1);
    S col[6];
    toArr(mulI(Y, fromArr(e)), col);
    for (int i = 0; i < 6; ++i) A(i, j) = col[i];
  }
  return A;
}

// ---------------------------------------------------------------------------------------------
template <typename S>
struct Chain {
  int nv;
  V3<S> axis[MAXNV];
  SE3<S> placement[MAXNV];  // joint i in its parent joint frame at q_i = 0 (parent of joint 0 = world)
  Inertia<S> inertia[MAXNV];
  V3<S> gravity;            // world linear acceleration of gravity, e.g. (0,0,-9.81)
  int frame_parent;         // joint the operational frame is attached to
  SE3<S> frame_placement;   // frame in its parent joint frame
};

// rotation about a unit axis (Rodrigues); joint transform liMi = placement * (R(axis,q), 0)
template <typename S>
inline M3<S> rotAxis(const V3<S>& a, const S& q) {
  using std::cos;
  using std::sin;
  const S c = cos(q), s = sin(q), v = S(1) - c;
  M3<S> R;
  R(0, 0) = a.x * a.x * v + c;
  R(0, 1) = a.x * a.y * v - a.z * s;
  R(0, 2) = a.x * a.z * v + a.y * s;
  R(1, 0) = a.x * a.y * v + a.z * s;
  R(1, 1) = a.y * a.y * v + c;
  R(1, 2) = a.y * a.z * v - a.x * s;
  R(2, 0) = a.x * a.z * v - a.y * s;
  R(2, 1) = a.y * a.z * v + a.x * s;
  R(2, 2) = a.z * a.z * v + c;
  return R;
}
template <typename S>
inline SE3<S> jointTransform(const Chain<S>& ch, int i, const S& q) {
  SE3<S> J{rotAxis(ch.axis[i], q), V3<S>{S(0), S(0), S(0)}};
  return compose(ch.placement[i], J);
}
template <typename S>
inline Sp<S> jointS(const Chain<S>& ch, int i) { return Sp<S>{V3<S>{S(0), S(0), S(0)}, ch.axis[i]}; }

// data shared by the algorithms (what pinocchio::Data holds for this path)
template <typename S>
struct ChainData {
  SE3<S> liMi[MAXNV], oMi[MAXNV];
  Sp<S> v[MAXNV], a[MAXNV], f[MAXNV];
  S ddq[MAXNV];
};

// Articulated-body algorithm: qdd = aba(q, v, tau); also fills liMi, oMi (updateGlobalPlacements).
template <typename S>
inline void aba(const Chain<S>& ch, ChainData<S>& d, const S* q, const S* qd, const S* tau, S* qdd) {
  const int n = ch.nv;
  M6<S> Ia[MAXNV];
  Sp<S> pa[MAXNV], c[MAXNV], U[MAXNV];
  S Dinv[MAXNV], u[MAXNV];
  for (int i = 0; i < n; ++i) {
    d.liMi[i] = jointTransform(ch, i, q[i]);
    d.oMi[i] = (i == 0) ? d.liMi[i] : compose(d.oMi[i - 1], d.liMi[i]);
    const Sp<S> vj = scale(qd[i], jointS(ch, i));
    d.v[i] = (i == 0) ? vj : add(actInvM(d.liMi[i], d.v[i - 1]), vj);
    c[i] = crossM(d.v[i], vj);
    Ia[i] = inertiaMatrix(ch.inertia[i]);
    pa[i] = crossF(d.v[i], mulI(ch.inertia[i], d.v[i]));
  }
  for (int i = n - 1; i >= 0; --i) {
    const Sp<S> Si = jointS(ch, i);
    U[i] = mulM6(Ia[i], Si);
    const S D = dotS(Si, U[i]);
    Dinv[i] = S(1) / D;
    u[i] = tau[i] - dotS(Si, pa[i]);
    if (i > 0) {
      S Ua[6];
      toArr(U[i], Ua);
      M6<S> Iaa = Ia[i];
      for (int r = 0; r < 6; ++r)
        for (int cc = 0; cc < 6; ++cc) Iaa(r, cc) -= Ua[r] * Dinv[i] * Ua[cc];
      Sp<S> p = add(add(pa[i], mulM6(Iaa, c[i])), scale(Dinv[i] * u[i], U[i]));
      // propagate to the parent: Ia_parent += X* Iaa X^-1, pa_parent += X* p   (X = liMi)
      for (int cc = 0; cc < 6; ++cc) {
        S e[6] = {S(0), S(0), S(0), S(0), S(0), S(0)};
        e[cc] = S(1);
        // column of X* Iaa X^-1: act on a parent-frame motion e: bring to child, apply, bring the force back
        Sp<S> col = actF(d.liMi[i], mulM6(Iaa, actInvM(d.liMi[i], fromArr(e))));
        S ca[6];
        toArr(col, ca);
        for (int r = 0; r < 6; ++r) Ia[i - 1](r, cc) += ca[r];
      }
      pa[i - 1] = add(pa[i - 1], actF(d.liMi[i], p));
    }
  }
  const Sp<S> a0{V3<S>{-ch.gravity.x, -ch.gravity.y, -ch.gravity.z}, V3<S>{S(0), S(0), S(0)}};
  for (int i = 0; i < n; ++i) {
    const Sp<S> ap = actInvM(d.liMi[i], i == 0 ? a0 : d.a[i - 1]);
    const Sp<S> ai = add(ap, c[i]);
    qdd[i] = Dinv[i] * (u[i] - dotS(U[i], ai));
    d.a[i] = add(ai, scale(qdd[i], jointS(ch, i)));
    d.ddq[i] = qdd[i];
  }
}

// Recursive Newton-Euler: tau = rnea(q, v, a)
template <typename S>
inline void rnea(const Chain<S>& ch, ChainData<S>& d, const S* q, const S* qd, const S* qdd, S* tau) {
  const int n = ch.nv;
  const Sp<S> a0{V3<S>{-ch.gravity.x, -ch.gravity.y, -ch.gravity.z}, V3<S>{S(0), S(0), S(0)}};
  for (int i = 0; i < n; ++i) {
    d.liMi[i] = jointTransform(ch, i, q[i]);
    d.oMi[i] = (i == 0) ? d.liMi[i] : compose(d.oMi[i - 1], d.liMi[i]);
    const Sp<S> Si = jointS(ch, i);
    const Sp<S> vj = scale(qd[i], Si);
    d.v[i] = (i == 0) ? vj : add(actInvM(d.liMi[i], d.v[i - 1]), vj);
    d.a[i] = add(add(actInvM(d.liMi[i], i == 0 ? a0 : d.a[i - 1]), scale(qdd[i], Si)), crossM(d.v[i], vj));
    d.f[i] = add(mulI(ch.inertia[i], d.a[i]), crossF(d.v[i], mulI(ch.inertia[i], d.v[i])));
  }
  for (int i = n - 1; i >= 0; --i) {
    tau[i] = dotS(jointS(ch, i), d.f[i]);
    if (i > 0) d.f[i - 1] = add(d.f[i - 1], actF(d.liMi[i], d.f[i]));
  }
}

// Composite-rigid-body algorithm: joint-space inertia M (column-major nv x nv)
template <typename S>
inline void crba(const Chain<S>& ch, ChainData<S>& d, const S* q, S* M) {
  const int n = ch.nv;
  M6<S> Yc[MAXNV];
  for (int i = 0; i < n; ++i) {
    d.liMi[i] = jointTransform(ch, i, q[i]);
    Yc[i] = inertiaMatrix(ch.inertia[i]);
  }
  for (int i = n - 1; i >= 0; --i) {
    Sp<S> F = mulM6(Yc[i], jointS(ch, i));
    M[i + n * i] = dotS(jointS(ch, i), F);
    for (int j = i; j > 0; --j) {
      F = actF(d.liMi[j], F);
      M[(j - 1) + n * i] = dotS(jointS(ch, j - 1), F);
      M[i + n * (j - 1)] = M[(j - 1) + n * i];
    }
    if (i > 0)
      for (int cc = 0; cc < 6; ++cc) {
        S e[6] = {S(0), S(0), S(0), S(0), S(0), S(0)};
        e[cc] = S(1);
        Sp<S> col = actF(d.liMi[i], mulM6(Yc[i], actInvM(d.liMi[i], fromArr(e))));
        S ca[6];
        toArr(col, ca);
        for (int r = 0; r < 6; ++r) Yc[i - 1](r, cc) += ca[r];
      }
  }
}

// Partial derivatives of RNEA: dtau_dq, dtau_dv (column-major nv x nv) at (q, v, a), by the recursive chain rule.
template <typename S>
inline void rneaDerivatives(const Chain<S>& ch, ChainData<S>& d, const S* q, const S* qd, const S* qdd, S* dtau_dq,
                            S* dtau_dv) {
  const int n = ch.nv;
  S tau[MAXNV];
  rnea(ch, d, q, qd, qdd, tau);  // fills liMi, v, a and the ACCUMULATED forces f
  const Sp<S> a0{V3<S>{-ch.gravity.x, -ch.gravity.y, -ch.gravity.z}, V3<S>{S(0), S(0), S(0)}};
  // dv[i][k], da[i][k] (k <= i) w.r.t. q_k and qd_k; df accumulated backward
  static thread_local std::vector<Sp<S> > dvq, daq, dvv, dav, dfq, dfv;
  dvq.assign(n * n, Sp<S>());
  daq.assign(n * n, Sp<S>());
  dvv.assign(n * n, Sp<S>());
  dav.assign(n * n, Sp<S>());
  dfq.assign(n * n, Sp<S>());
  dfv.assign(n * n, Sp<S>());
  const Sp<S> zero{V3<S>{S(0), S(0), S(0)}, V3<S>{S(0), S(0), S(0)}};
  for (int i = 0; i < n; ++i) {
    const Sp<S> Si = jointS(ch, i);
    const Sp<S> vj = scale(qd[i], Si);
    const Sp<S> vpar = (i == 0) ? zero : actInvM(d.liMi[i], d.v[i - 1]);  // X_i v_parent
    const Sp<S> apar = actInvM(d.liMi[i], i == 0 ? a0 : d.a[i - 1]);      // X_i a_parent
    for (int k = 0; k <= i; ++k) {
      Sp<S> dv_q, da_q, dv_v, da_v;
      if (k < i) {
        dv_q = actInvM(d.liMi[i], dvq[(i - 1) * n + k]);
        da_q = add(actInvM(d.liMi[i], daq[(i - 1) * n + k]), crossM(dv_q, vj));
        dv_v = actInvM(d.liMi[i], dvv[(i - 1) * n + k]);
        da_v = add(actInvM(d.liMi[i], dav[(i - 1) * n + k]), crossM(dv_v, vj));
      } else {
        dv_q = crossM(vpar, Si);                       // d(X_i v_parent)/dq_i = (X_i v_parent) x S_i
        da_q = add(crossM(apar, Si), crossM(dv_q, vj));
        dv_v = Si;
        da_v = add(crossM(d.v[i], Si), crossM(Si, vj));  // d(v_i x S_i qd_i)/dqd_i
      }
      dvq[i * n + k] = dv_q;
      daq[i * n + k] = da_q;
      dvv[i * n + k] = dv_v;
      dav[i * n + k] = da_v;
      const Sp<S> Iv = mulI(ch.inertia[i], d.v[i]);
      dfq[i * n + k] = add(mulI(ch.inertia[i], da_q), add(crossF(dv_q, Iv), crossF(d.v[i], mulI(ch.inertia[i], dv_q))));
      dfv[i * n + k] = add(mulI(ch.inertia[i], da_v), add(crossF(dv_v, Iv), crossF(d.v[i], mulI(ch.inertia[i], dv_v))));
    }
  }
  for (int i = n - 1; i >= 0; --i) {
    const Sp<S> Si = jointS(ch, i);
    for (int k = 0; k < n; ++k) {
      dtau_dq[i + n * k] = dotS(Si, dfq[i * n + k]);
      dtau_dv[i + n * k] = dotS(Si, dfv[i * n + k]);
    }
    if (i > 0) {
      for (int k = 0; k < n; ++k) {
        Sp<S> t = actF(d.liMi[i], dfq[i * n + k]);
        // d(X_i^* f_i)/dq_i = X_i^* (S_i x* f_i), f_i = accumulated force of body i
        if (k == i) t = add(t, actF(d.liMi[i], crossF(Si, d.f[i])));
        dfq[(i - 1) * n + k] = add(dfq[(i - 1) * n + k], t);
        dfv[(i - 1) * n + k] = add(dfv[(i - 1) * n + k], actF(d.liMi[i], dfv[i * n + k]));
      }
    }
  }
}

// symmetric positive definite inverse by Cholesky (column-major n x n); returns false if not SPD
template <typename S>
inline bool spdInverse(const S* A, int n, S* Ainv) {
  using std::sqrt;
  std::vector<S> L(n * n, S(0)), Li(n * n, S(0));
  for (int k = 0; k < n; ++k) {
    S x = A[k + n * k];
    for (int j = 0; j < k; ++j) x -= L[k + n * j] * L[k + n * j];
    if (!(x > S(0))) return false;
    x = sqrt(x);
    L[k + n * k] = x;
    for (int i = k + 1; i < n; ++i) {
      S s = A[i + n * k];
      for (int j = 0; j < k; ++j) s -= L[i + n * j] * L[k + n * j];
      L[i + n * k] = s / x;
    }
  }
  for (int c = 0; c < n; ++c) {  // Li = L^-1 (lower)
    for (int i = c; i < n; ++i) {
      S s = (i == c) ? S(1) : S(0);
      for (int j = c; j < i; ++j) s -= L[i + n * j] * Li[j + n * c];
      Li[i + n * c] = s / L[i + n * i];
    }
  }
  for (int i = 0; i < n; ++i)
    for (int j = 0; j <= i; ++j) {
      S s = S(0);
      for (int k = i; k < n; ++k) s += Li[k + n * i] * Li[k + n * j];
      Ainv[i + n * j] = s;
      Ainv[j + n * i] = s;
    }
  return true;
}

// computeABADerivatives semantics (free-fwddyn.hxx:104-105): ddq_dq, ddq_dv, Minv (column-major nv x nv) at
// (q, v, tau); qdd must be aba(q, v, tau).
template <typename S>
inline void abaDerivatives(const Chain<S>& ch, ChainData<S>& d, const S* q, const S* qd, const S* qdd, S* ddq_dq,
                           S* ddq_dv, S* Minv) {
  const int n = ch.nv;
  std::vector<S> M(n * n), tq(n * n), tv(n * n);
  crba(ch, d, q, M.data());
  if (!spdInverse(M.data(), n, Minv)) throw std::runtime_error("joint-space inertia is not positive definite");
  rneaDerivatives(ch, d, q, qd, qdd, tq.data(), tv.data());
  for (int j = 0; j < n; ++j)
    for (int i = 0; i < n; ++i) {
      S s1 = S(0), s2 = S(0);
      for (int k = 0; k < n; ++k) {
        s1 += Minv[i + n * k] * tq[k + n * j];
        s2 += Minv[i + n * k] * tv[k + n * j];
      }
      ddq_dq[i + n * j] = -s1;
      ddq_dv[i + n * j] = -s2;
    }
}

// frame placement in the world (updateFramePlacement) and its LOCAL Jacobian (getFrameJacobian(LOCAL)): 6 x nv,
// column-major, rows [linear; angular]; needs d.liMi / d.oMi
template <typename S>
inline SE3<S> framePlacement(const Chain<S>& ch, const ChainData<S>& d) {
  return compose(d.oMi[ch.frame_parent], ch.frame_placement);
}
template <typename S>
inline void frameJacobianLocal(const Chain<S>& ch, const ChainData<S>& d, S* J) {
  const int n = ch.nv;
  for (int e = 0; e < 6 * n; ++e) J[e] = S(0);
  SE3<S> fMk = inverse(ch.frame_placement);  // joint frame_parent -> frame coordinates
  for (int k = ch.frame_parent; k >= 0; --k) {
    S col[6];
    toArr(actM(fMk, jointS(ch, k)), col);
    for (int r = 0; r < 6; ++r) J[r + 6 * k] = col[r];
    if (k > 0) fMk = compose(fMk, inverse(d.liMi[k]));
  }
}

// ---------------------------------------------------------------------------------------------
// log3 / log6 / Jlog3 / Jlog6 (Pinocchio's spatial/explog.hpp formulas; frame-placement.hxx:59,75)
template <typename S>
inline S taylorPrec() {
  using std::pow;
  return pow(std::numeric_limits<double>::epsilon(), 0.25);  // TaylorSeriesExpansion<double>::precision<3>()
}
template <typename S>
inline V3<S> log3(const M3<S>& R, S& theta) {
  using std::acos;
  using std::cos;
  using std::sin;
  using std::sqrt;
  const double PI = 3.14159265358979323846;
  const S tr = R(0, 0) + R(1, 1) + R(2, 2);
  if (tr > S(3))
    theta = S(0);
  else if (tr < S(-1))
    theta = S(PI);
  else
    theta = acos((tr - S(1)) / S(2));
  V3<S> res;
  if (theta >= S(PI - 1e-2)) {
    const S cphi = cos(theta - S(PI));
    const S beta = theta * theta / (S(1) + cphi);
    const S t0 = (R(0, 0) + cphi) * beta, t1 = (R(1, 1) + cphi) * beta, t2 = (R(2, 2) + cphi) * beta;
    res.x = (R(2, 1) > R(1, 2) ? S(1) : S(-1)) * (t0 > S(0) ? sqrt(t0) : S(0));
    res.y = (R(0, 2) > R(2, 0) ? S(1) : S(-1)) * (t1 > S(0) ? sqrt(t1) : S(0));
    res.z = (R(1, 0) > R(0, 1) ? S(1) : S(-1)) * (t2 > S(0) ? sqrt(t2) : S(0));
  } else {
    const S t = ((theta > taylorPrec<S>()) ? theta / sin(theta) : S(1)) / S(2);
    res.x = t * (R(2, 1) - R(1, 2));
    res.y = t * (R(0, 2) - R(2, 0));
    res.z = t * (R(1, 0) - R(0, 1));
  }
  return res;
}
template <typename S>
inline Sp<S> log6(const SE3<S>& M) {
  using std::cos;
  using std::sin;
  S t;
  const V3<S> w = log3(M.R, t);
  const S t2 = t * t;
  S alpha, beta;
  if (t < taylorPrec<S>()) {
    alpha = S(1) - t2 / S(12) - t2 * t2 / S(720);
    beta = S(1) / S(12) + t2 / S(720);
  } else {
    const S st = sin(t), ct = cos(t);
    alpha = t * st / (S(2) * (S(1) - ct));
    beta = S(1) / t2 - st / (S(2) * t * (S(1) - ct));
  }
  Sp<S> r;
  r.lin = add(sub(scale(alpha, M.p), scale(S(0.5), cross(w, M.p))), scale(beta * dot(w, M.p), w));
  r.ang = w;
  return r;
}
template <typename S>
inline M3<S> Jlog3(const S& theta, const V3<S>& lg) {
  using std::cos;
  using std::sin;
  S alpha, diag;
  if (theta < taylorPrec<S>()) {
    const S t2 = theta * theta;
    alpha = S(1) / S(12) + t2 / S(720);
    diag = S(0.5) * (S(2) - t2 / S(6));
  } else {
    const S ct = cos(theta), st = sin(theta);
    const S st_1mct = st / (S(1) - ct);
    alpha = S(1) / (theta * theta) - st_1mct / (S(2) * theta);
    diag = S(0.5) * (theta * st_1mct);
  }
  const S l[3] = {lg.x, lg.y, lg.z};
  M3<S> J;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) J(i, j) = alpha * l[i] * l[j];
  for (int i = 0; i < 3; ++i) J(i, i) += diag;
  // addSkew(0.5 * log, J)
  J(0, 1) -= S(0.5) * lg.z;
  J(0, 2) += S(0.5) * lg.y;
  J(1, 0) += S(0.5) * lg.z;
  J(1, 2) -= S(0.5) * lg.x;
  J(2, 0) -= S(0.5) * lg.y;
  J(2, 1) += S(0.5) * lg.x;
  return J;
}
// Jlog6: 6x6 column-major, rows/cols [linear; angular]: [[A, B], [0, A]]
template <typename S>
inline void Jlog6(const SE3<S>& M, S* J) {
  using std::cos;
  using std::sin;
  S t;
  const V3<S> w = log3(M.R, t);
  const M3<S> A = Jlog3(t, w);
  const S t2 = t * t;
  S beta, beta_dot_over_theta;
  if (t < taylorPrec<S>()) {
    beta = S(1) / S(12) + t2 / S(720);
    beta_dot_over_theta = S(1) / S(360);
  } else {
    const S tinv = S(1) / t, t2inv = tinv * tinv;
    const S st = sin(t), ct = cos(t);
    const S inv_2_2ct = S(1) / (S(2) * (S(1) - ct));
    beta = t2inv - st * tinv * inv_2_2ct;
    beta_dot_over_theta = -S(2) * t2inv * t2inv + (S(1) + st * tinv) * t2inv * inv_2_2ct;
  }
  const V3<S>& p = M.p;
  const S wTp = dot(w, p);
  const V3<S> v3 = sub(scale(beta_dot_over_theta * wTp, w), scale(t2 * beta_dot_over_theta + S(2) * beta, p));
  const S wv[3] = {w.x, w.y, w.z}, pv[3] = {p.x, p.y, p.z}, vv[3] = {v3.x, v3.y, v3.z};
  M3<S> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C(i, j) = vv[i] * wv[j] + beta * wv[i] * pv[j];
  for (int i = 0; i < 3; ++i) C(i, i) += wTp * beta;
  C(0, 1) -= S(0.5) * p.z;
  C(0, 2) += S(0.5) * p.y;
  C(1, 0) += S(0.5) * p.z;
  C(1, 2) -= S(0.5) * p.x;
  C(2, 0) -= S(0.5) * p.y;
  C(2, 1) += S(0.5) * p.x;
  const M3<S> Bm = mul(C, A);
  for (int e = 0; e < 36; ++e) J[e] = S(0);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      J[i + 6 * j] = A(i, j);
      J[i + 6 * (j + 3)] = Bm(i, j);
      J[(i + 3) + 6 * (j + 3)] = A(i, j);
    }
}

}  // namespace rbd
}  // namespace oracle
