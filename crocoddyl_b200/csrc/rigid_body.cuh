// crocoddyl_b200 — rigid-body arithmetic for fixed-base serial chains of 1-DoF revolute joints (FP64), the part of
// the path the reference delegates to Pinocchio (call sites, paths relative to the reference tree):
//   pinocchio::aba + updateGlobalPlacements   include/crocoddyl/multibody/actions/free-fwddyn.hxx:64-65
//   pinocchio::computeABADerivatives          free-fwddyn.hxx:104-105     (ddq_dq, ddq_dv, Minv)
//   pinocchio::rnea (quasiStatic)             free-fwddyn.hxx:158-159
//   updateFramePlacement, log6, Jlog6, getFrameJacobian(LOCAL)   include/crocoddyl/multibody/costs/frame-placement.hxx:57-76
// Conventions (Pinocchio's): spatial motion = [linear; angular], force = [force; torque]; SE3 (R,p) maps child
// coordinates to parent coordinates; joint i has a placement in its parent joint frame at q_i = 0 and a unit axis a_i
// (S_i = [0; a_i]); body inertia = (mass, COM lever, rotational inertia about the COM); gravity enters as the base
// acceleration -g.
//
// Everything here is __host__ __device__ and works on ONE node (one (x,u) pair) held by one thread: the callers
// (arm.cuh) spread nodes / Jacobian columns / step lengths over threads. NV is a compile-time joint count so every
// per-joint array is fully unrolled into registers/local memory.
//
// Articulated inertias are kept as three 3x3 blocks {A,B,C}: force = A lin + B ang, torque = B^T lin + C ang, and are
// moved to the parent frame blockwise (A' = R A R^T, B' = R B R^T - A'[p]x, C' = R C R^T - (R B R^T)^T [p]x + [p]x B').
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define CGX_HD __host__ __device__ __forceinline__
#else
#define CGX_HD inline
#endif

namespace cgx {
namespace rb {

// Wide stores of a thread's own contiguous outputs (one node per thread: a scalar store per element touches 32
// different sectors per warp instruction and uses 8 of their 32 bytes; ncu: 30 % of the node kernel's samples sat on
// those stores). p must be 16-byte aligned. Host build: plain stores. 256-bit st.global.v4.f64 was tried and dropped:
// ptxas emitted it only in the kernel without launch bounds (cuobjdump: no STG.256 in the other clones of the node
// function, wrong Lxx on the GPU), so pairs are the widest store used here.
CGX_HD void st2(double* p, double a, double b) {
#if defined(__CUDA_ARCH__)
  asm volatile("st.global.v2.f64 [%0], {%1,%2};" ::"l"(p), "d"(a), "d"(b) : "memory");
#else
  p[0] = a;
  p[1] = b;
#endif
}
struct V3 {
  double x, y, z;
};
CGX_HD V3 v3(double x, double y, double z) { return V3{x, y, z}; }
CGX_HD V3 operator+(const V3& a, const V3& b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
CGX_HD V3 operator-(const V3& a, const V3& b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
CGX_HD V3 operator*(double s, const V3& a) { return V3{s * a.x, s * a.y, s * a.z}; }
CGX_HD V3 cross(const V3& a, const V3& b) {
  return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
CGX_HD double dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// 3x3, row-major
struct M3 {
  double m[9];
  CGX_HD double& operator()(int i, int j) { return m[3 * i + j]; }
  CGX_HD const double& operator()(int i, int j) const { return m[3 * i + j]; }
};
CGX_HD V3 mul(const M3& A, const V3& v) {
  return V3{A.m[0] * v.x + A.m[1] * v.y + A.m[2] * v.z, A.m[3] * v.x + A.m[4] * v.y + A.m[5] * v.z,
            A.m[6] * v.x + A.m[7] * v.y + A.m[8] * v.z};
}
CGX_HD V3 mulT(const M3& A, const V3& v) {
  return V3{A.m[0] * v.x + A.m[3] * v.y + A.m[6] * v.z, A.m[1] * v.x + A.m[4] * v.y + A.m[7] * v.z,
            A.m[2] * v.x + A.m[5] * v.y + A.m[8] * v.z};
}
CGX_HD M3 mul(const M3& A, const M3& B) {
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C(i, j) = A(i, 0) * B(0, j) + A(i, 1) * B(1, j) + A(i, 2) * B(2, j);
  return C;
}
CGX_HD M3 mulABt(const M3& A, const M3& B) {  // A B^T
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C(i, j) = A(i, 0) * B(j, 0) + A(i, 1) * B(j, 1) + A(i, 2) * B(j, 2);
  return C;
}
CGX_HD M3 mulAtB(const M3& A, const M3& B) {  // A^T B
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C(i, j) = A(0, i) * B(0, j) + A(1, i) * B(1, j) + A(2, i) * B(2, j);
  return C;
}
CGX_HD M3 transpose(const M3& A) {
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C(i, j) = A(j, i);
  return C;
}
CGX_HD M3 operator+(const M3& A, const M3& B) {
  M3 C;
#pragma unroll
  for (int e = 0; e < 9; ++e) C.m[e] = A.m[e] + B.m[e];
  return C;
}
CGX_HD M3 operator-(const M3& A, const M3& B) {
  M3 C;
#pragma unroll
  for (int e = 0; e < 9; ++e) C.m[e] = A.m[e] - B.m[e];
  return C;
}
CGX_HD M3 zero3() {
  M3 C;
#pragma unroll
  for (int e = 0; e < 9; ++e) C.m[e] = 0.;
  return C;
}
// M [p]x : column j = M (p x e_j)
CGX_HD M3 rightSkew(const M3& M, const V3& p) {
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    C(i, 0) = M(i, 1) * p.z - M(i, 2) * p.y;
    C(i, 1) = M(i, 2) * p.x - M(i, 0) * p.z;
    C(i, 2) = M(i, 0) * p.y - M(i, 1) * p.x;
  }
  return C;
}
// [p]x M : column j = p x (M column j)
CGX_HD M3 leftSkew(const V3& p, const M3& M) {
  M3 C;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    C(0, j) = p.y * M(2, j) - p.z * M(1, j);
    C(1, j) = p.z * M(0, j) - p.x * M(2, j);
    C(2, j) = p.x * M(1, j) - p.y * M(0, j);
  }
  return C;
}

struct SE3 {
  M3 R;
  V3 p;
};
CGX_HD SE3 compose(const SE3& a, const SE3& b) { return SE3{mul(a.R, b.R), a.p + mul(a.R, b.p)}; }
CGX_HD SE3 inverse(const SE3& a) {
  const V3 t = mulT(a.R, a.p);
  return SE3{transpose(a.R), V3{-t.x, -t.y, -t.z}};
}
CGX_HD SE3 actInv(const SE3& a, const SE3& b) { return SE3{mulAtB(a.R, b.R), mulT(a.R, b.p - a.p)}; }  // a^-1 b

// spatial motion / force
struct Sp {
  V3 lin, ang;
};
CGX_HD Sp operator+(const Sp& a, const Sp& b) { return Sp{a.lin + b.lin, a.ang + b.ang}; }
CGX_HD Sp operator*(double s, const Sp& a) { return Sp{s * a.lin, s * a.ang}; }
CGX_HD Sp sp_zero() { return Sp{V3{0., 0., 0.}, V3{0., 0., 0.}}; }
CGX_HD Sp actM(const SE3& M, const Sp& v) {  // motion child -> parent
  const V3 w = mul(M.R, v.ang);
  return Sp{mul(M.R, v.lin) + cross(M.p, w), w};
}
CGX_HD Sp actInvM(const SE3& M, const Sp& v) {  // motion parent -> child
  return Sp{mulT(M.R, v.lin - cross(M.p, v.ang)), mulT(M.R, v.ang)};
}
CGX_HD Sp actF(const SE3& M, const Sp& f) {  // force child -> parent
  const V3 l = mul(M.R, f.lin);
  return Sp{l, mul(M.R, f.ang) + cross(M.p, l)};
}
CGX_HD Sp crossM(const Sp& a, const Sp& b) {  // motion x motion
  return Sp{cross(a.ang, b.lin) + cross(a.lin, b.ang), cross(a.ang, b.ang)};
}
CGX_HD Sp crossF(const Sp& v, const Sp& f) {  // motion x* force
  return Sp{cross(v.ang, f.lin), cross(v.ang, f.ang) + cross(v.lin, f.lin)};
}
CGX_HD double dotS(const Sp& a, const Sp& b) { return dot(a.lin, b.lin) + dot(a.ang, b.ang); }
// specialisations for a revolute joint S = [0; a]
CGX_HD Sp crossM_S(const Sp& v, const V3& a, double s) {  // v x (s S)
  return Sp{s * cross(v.lin, a), s * cross(v.ang, a)};
}

// rigid-body inertia
struct Inertia {
  double mass;
  V3 c;
  M3 I;
};
CGX_HD Sp mulI(const Inertia& Y, const Sp& v) {
  const V3 l = Y.mass * (v.lin - cross(Y.c, v.ang));
  return Sp{l, mul(Y.I, v.ang) + cross(Y.c, l)};
}

// articulated / composite inertia as 3x3 blocks
struct I6 {
  M3 A, B, C;
};
CGX_HD I6 i6_zero() { return I6{zero3(), zero3(), zero3()}; }
CGX_HD I6 i6_from(const Inertia& Y) {
  I6 r;
  const double m = Y.mass;
  const V3& c = Y.c;
  r.A = zero3();
  r.A(0, 0) = r.A(1, 1) = r.A(2, 2) = m;
  r.B = zero3();  // -m [c]x
  r.B(0, 1) = m * c.z;
  r.B(0, 2) = -m * c.y;
  r.B(1, 0) = -m * c.z;
  r.B(1, 2) = m * c.x;
  r.B(2, 0) = m * c.y;
  r.B(2, 1) = -m * c.x;
  const double cc = dot(c, c);
  r.C = Y.I;  // I_c + m (|c|^2 1 - c c^T)
  r.C(0, 0) += m * (cc - c.x * c.x);
  r.C(1, 1) += m * (cc - c.y * c.y);
  r.C(2, 2) += m * (cc - c.z * c.z);
  r.C(0, 1) -= m * c.x * c.y;
  r.C(1, 0) -= m * c.x * c.y;
  r.C(0, 2) -= m * c.x * c.z;
  r.C(2, 0) -= m * c.x * c.z;
  r.C(1, 2) -= m * c.y * c.z;
  r.C(2, 1) -= m * c.y * c.z;
  return r;
}
CGX_HD I6 operator+(const I6& a, const I6& b) { return I6{a.A + b.A, a.B + b.B, a.C + b.C}; }
CGX_HD Sp mulI6(const I6& Y, const Sp& v) {
  return Sp{mul(Y.A, v.lin) + mul(Y.B, v.ang), mulT(Y.B, v.lin) + mul(Y.C, v.ang)};
}
// X* Y X^-1 : inertia expressed in the child frame -> parent frame, X = (R,p) child -> parent
CGX_HD I6 i6_to_parent(const SE3& X, const I6& Y) {
  I6 r;
  r.A = mulABt(mul(X.R, Y.A), X.R);
  const M3 Bt = mulABt(mul(X.R, Y.B), X.R);
  r.B = Bt - rightSkew(r.A, X.p);
  r.C = mulABt(mul(X.R, Y.C), X.R) - rightSkew(transpose(Bt), X.p) + leftSkew(X.p, r.B);
  return r;
}

// rotation about a unit axis (Rodrigues)
CGX_HD M3 rotAxis(const V3& a, double q) {
  double s, c;
  sincos(q, &s, &c);
  const double v = 1. - c;
  M3 R;
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

// ---------------------------------------------------------------------------------------------
// Parameter block of the arm model (include/crocoddyl_b200.h "ARM"): 36 header doubles + 28 per joint.
constexpr int ARM_HEADER = 36, ARM_PER_JOINT = 28;
// goal cost of the frame (header [32]): CostModelFramePlacement / FrameTranslation / FrameRotation
constexpr int GOAL_PLACEMENT = 0, GOAL_TRANSLATION = 1, GOAL_ROTATION = 2;
CGX_HD SE3 read_se3(const double* p) {
  SE3 M;
#pragma unroll
  for (int i = 0; i < 9; ++i) M.R.m[i] = p[i];
  M.p = V3{p[9], p[10], p[11]};
  return M;
}
struct ArmParams {
  const double* p;
  CGX_HD double dt() const { return p[0]; }
  CGX_HD V3 gravity() const { return V3{p[1], p[2], p[3]}; }
  CGX_HD double w_gp() const { return p[4]; }
  CGX_HD double w_x() const { return p[5]; }
  CGX_HD double w_u() const { return p[6]; }
  CGX_HD int frame_parent() const { return (int)p[7]; }
  CGX_HD SE3 frame_placement() const { return read_se3(p + 8); }
  CGX_HD SE3 Mref() const { return read_se3(p + 20); }
  CGX_HD int goal_kind() const { return (int)p[32]; }
  CGX_HD const double* joint(int i) const { return p + ARM_HEADER + ARM_PER_JOINT * i; }
  CGX_HD V3 axis(int i) const {
    const double* j = joint(i);
    return V3{j[0], j[1], j[2]};
  }
  CGX_HD SE3 placement(int i) const { return read_se3(joint(i) + 3); }
  // rotor inertia seen at joint i (free-fwddyn.hxx:66-74: M.diagonal() += armature); 0 = none
  CGX_HD double armature(int i) const { return joint(i)[25]; }
  // ActivationModelWeightedQuad weights of the state cost for q_i and v_i (weighted-quadratic.hpp:39-49); 1 = plain quad
  CGX_HD double wq(int i) const { return joint(i)[26]; }
  CGX_HD double wv(int i) const { return joint(i)[27]; }
  CGX_HD Inertia inertia(int i) const {
    const double* j = joint(i);
    Inertia Y;
    Y.mass = j[15];
    Y.c = V3{j[16], j[17], j[18]};
    Y.I(0, 0) = j[19];
    Y.I(0, 1) = Y.I(1, 0) = j[20];
    Y.I(0, 2) = Y.I(2, 0) = j[21];
    Y.I(1, 1) = j[22];
    Y.I(1, 2) = Y.I(2, 1) = j[23];
    Y.I(2, 2) = j[24];
    return Y;
  }
  // liMi = placement * (R(axis, q), 0)
  CGX_HD SE3 joint_transform(int i, double q) const {
    const SE3 pl = placement(i);
    return SE3{mul(pl.R, rotAxis(axis(i), q)), pl.p};
  }
};

// ---------------------------------------------------------------------------------------------
// what pinocchio::Data keeps for this path: joint placements and body velocities / accelerations / forces
template <int NV>
struct Kin {
  SE3 liMi[NV];
  Sp v[NV];
};
template <int NV>
struct Dyn {
  Sp a[NV], f[NV];  // f = ACCUMULATED force of the subtree rooted at body i (after the backward RNEA sweep)
};

// Articulated-body algorithm (Featherstone): qdd = aba(q, qd, tau); tau == nullptr means zero torques.
// Fills k.liMi (and hence updateGlobalPlacements' oMi = product of liMi) and k.v.
template <int NV>
CGX_HD void aba(const ArmParams& P, const double* q, const double* qd, const double* tau, double* qdd, Kin<NV>& k) {
  Sp c[NV], U[NV];
  double Dinv[NV], u[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    k.liMi[i] = P.joint_transform(i, q[i]);
    const V3 ax = P.axis(i);
    const Sp vj{V3{0., 0., 0.}, qd[i] * ax};
    k.v[i] = (i == 0) ? vj : actInvM(k.liMi[i], k.v[i - 1]) + vj;
    c[i] = crossM_S(k.v[i], ax, qd[i]);
  }
  I6 Iacc = i6_zero();
  Sp pacc = sp_zero();
#pragma unroll
  for (int i = NV - 1; i >= 0; --i) {
    const Inertia Y = P.inertia(i);
    const V3 ax = P.axis(i);
    I6 Ia = i6_from(Y) + Iacc;
    const Sp pa = crossF(k.v[i], mulI(Y, k.v[i])) + pacc;
    U[i] = Sp{mul(Ia.B, ax), mul(Ia.C, ax)};
    // joint-space pivot; the armature (rotor inertia) adds to it, which makes the sweep the solution of
    // (M + diag(armature)) qdd = tau - nle, the system free-fwddyn.hxx:66-74 solves through Minv
    const double D = dot(ax, U[i].ang) + P.armature(i);
    Dinv[i] = 1. / D;
    u[i] = (tau ? tau[i] : 0.) - dot(ax, pa.ang);
    if (i > 0) {
      const double ul[3] = {U[i].lin.x, U[i].lin.y, U[i].lin.z}, ua[3] = {U[i].ang.x, U[i].ang.y, U[i].ang.z};
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          Ia.A(r, s) -= ul[r] * Dinv[i] * ul[s];
          Ia.B(r, s) -= ul[r] * Dinv[i] * ua[s];
          Ia.C(r, s) -= ua[r] * Dinv[i] * ua[s];
        }
      const Sp p = pa + mulI6(Ia, c[i]) + (Dinv[i] * u[i]) * U[i];
      Iacc = i6_to_parent(k.liMi[i], Ia);
      pacc = actF(k.liMi[i], p);
    }
  }
  const V3 g = P.gravity();
  Sp a{V3{-g.x, -g.y, -g.z}, V3{0., 0., 0.}};
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const Sp ai = actInvM(k.liMi[i], a) + c[i];
    qdd[i] = Dinv[i] * (u[i] - dotS(U[i], ai));
    a = ai + Sp{V3{0., 0., 0.}, qdd[i] * P.axis(i)};
  }
}

// Recursive Newton-Euler: tau = rnea(q, qd, qdd) (qd / qdd == nullptr mean zeros); fills k and d.
template <int NV>
CGX_HD void rnea(const ArmParams& P, const double* q, const double* qd, const double* qdd, double* tau, Kin<NV>& k,
                 Dyn<NV>& d, bool kin_ready) {
  const V3 g = P.gravity();
  const Sp a0{V3{-g.x, -g.y, -g.z}, V3{0., 0., 0.}};
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    if (!kin_ready) k.liMi[i] = P.joint_transform(i, q[i]);
    const V3 ax = P.axis(i);
    const double qdi = qd ? qd[i] : 0., qddi = qdd ? qdd[i] : 0.;
    const Sp vj{V3{0., 0., 0.}, qdi * ax};
    k.v[i] = (i == 0) ? vj : actInvM(k.liMi[i], k.v[i - 1]) + vj;
    d.a[i] = actInvM(k.liMi[i], i == 0 ? a0 : d.a[i - 1]) + Sp{V3{0., 0., 0.}, qddi * ax} + crossM_S(k.v[i], ax, qdi);
    const Inertia Y = P.inertia(i);
    d.f[i] = mulI(Y, d.a[i]) + crossF(k.v[i], mulI(Y, k.v[i]));
  }
#pragma unroll
  for (int i = NV - 1; i >= 0; --i) {
    tau[i] = dot(P.axis(i), d.f[i].ang);
    if (i > 0) d.f[i - 1] = d.f[i - 1] + actF(k.liMi[i], d.f[i]);
  }
}

// Composite-rigid-body algorithm: joint-space inertia M (column-major NV x NV); needs k.liMi
template <int NV>
CGX_HD void crba(const ArmParams& P, const Kin<NV>& k, double* M) {
  I6 Yacc = i6_zero();
#pragma unroll
  for (int i = NV - 1; i >= 0; --i) {
    const I6 Yc = i6_from(P.inertia(i)) + Yacc;
    const V3 ax = P.axis(i);
    Sp F{mul(Yc.B, ax), mul(Yc.C, ax)};
    M[i + NV * i] = dot(ax, F.ang);
#pragma unroll
    for (int j = i; j > 0; --j) {
      F = actF(k.liMi[j], F);
      const double mji = dot(P.axis(j - 1), F.ang);
      M[(j - 1) + NV * i] = mji;
      M[i + NV * (j - 1)] = mji;
    }
    if (i > 0) Yacc = i6_to_parent(k.liMi[i], Yc);
  }
}

// inverse of a symmetric positive definite matrix through its Cholesky factor (column-major, in registers/local).
// Returns false if a pivot is not positive.
template <int N>
CGX_HD bool spd_inverse(const double* A, double* Ainv) {
  double L[N * N], Li[N * N];
#pragma unroll
  for (int e = 0; e < N * N; ++e) L[e] = Li[e] = 0.;
  bool ok = true;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    double x = A[k + N * k];
#pragma unroll
    for (int j = 0; j < k; ++j) x -= L[k + N * j] * L[k + N * j];
    if (!(x > 0.)) ok = false;
    x = sqrt(x);
    L[k + N * k] = x;
#pragma unroll
    for (int i = k + 1; i < N; ++i) {
      double s = A[i + N * k];
#pragma unroll
      for (int j = 0; j < k; ++j) s -= L[i + N * j] * L[k + N * j];
      L[i + N * k] = s / x;
    }
  }
#pragma unroll
  for (int c = 0; c < N; ++c)
#pragma unroll
    for (int i = c; i < N; ++i) {
      double s = (i == c) ? 1. : 0.;
#pragma unroll
      for (int j = c; j < i; ++j) s -= L[i + N * j] * Li[j + N * c];
      Li[i + N * c] = s / L[i + N * i];
    }
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double s = 0.;
#pragma unroll
      for (int kk = i; kk < N; ++kk) s += Li[kk + N * i] * Li[kk + N * j];
      Ainv[i + N * j] = s;
      Ainv[j + N * i] = s;
    }
  return ok;
}

// One column (joint kcol) of the partial derivatives of RNEA, dtau/dq_kcol and dtau/dqd_kcol (NV entries each), by
// the recursive chain rule in local frames (Carpentier & Mansard 2018). Needs k (liMi, v) and d (a, accumulated f)
// from rnea() at the same (q, qd, qdd).
template <int NV>
CGX_HD void rnea_derivative_column(const ArmParams& P, const Kin<NV>& k, const Dyn<NV>& d, const double* qd, int kcol,
                                   double* tq, double* tv) {
  Sp ownq[NV], ownv[NV];
  Sp dv_q = sp_zero(), da_q = sp_zero(), dv_v = sp_zero(), da_v = sp_zero();
  const V3 g = P.gravity();
  const Sp a0{V3{-g.x, -g.y, -g.z}, V3{0., 0., 0.}};
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    if (i < kcol) continue;
    const V3 ax = P.axis(i);
    const Sp Si{V3{0., 0., 0.}, ax};
    if (i == kcol) {
      const Sp vpar = (i == 0) ? sp_zero() : actInvM(k.liMi[i], k.v[i - 1]);
      const Sp apar = actInvM(k.liMi[i], i == 0 ? a0 : d.a[i - 1]);
      dv_q = crossM(vpar, Si);
      da_q = crossM(apar, Si) + crossM_S(dv_q, ax, qd[i]);
      dv_v = Si;
      da_v = crossM(k.v[i], Si);
    } else {
      dv_q = actInvM(k.liMi[i], dv_q);
      da_q = actInvM(k.liMi[i], da_q) + crossM_S(dv_q, ax, qd[i]);
      dv_v = actInvM(k.liMi[i], dv_v);
      da_v = actInvM(k.liMi[i], da_v) + crossM_S(dv_v, ax, qd[i]);
    }
    const Inertia Y = P.inertia(i);
    const Sp Iv = mulI(Y, k.v[i]);
    ownq[i] = mulI(Y, da_q) + (crossF(dv_q, Iv) + crossF(k.v[i], mulI(Y, dv_q)));
    ownv[i] = mulI(Y, da_v) + (crossF(dv_v, Iv) + crossF(k.v[i], mulI(Y, dv_v)));
  }
  Sp accq = sp_zero(), accv = sp_zero();
#pragma unroll
  for (int i = NV - 1; i >= 0; --i) {
    if (i >= kcol) {
      accq = ownq[i] + accq;
      accv = ownv[i] + accv;
    }
    const V3 ax = P.axis(i);
    tq[i] = dot(ax, accq.ang);
    tv[i] = dot(ax, accv.ang);
    if (i > 0) {
      Sp t = actF(k.liMi[i], accq);
      if (kcol == i) t = t + actF(k.liMi[i], crossF(Sp{V3{0., 0., 0.}, ax}, d.f[i]));
      accq = t;
      accv = actF(k.liMi[i], accv);
    }
  }
}

// frame placement in the world (updateFramePlacement): oMf = oMi[parent] * placement; needs k.liMi
template <int NV>
CGX_HD SE3 frame_placement(const ArmParams& P, const Kin<NV>& k) {
  const int fp = P.frame_parent();
  SE3 oM = k.liMi[0];
#pragma unroll
  for (int i = 1; i < NV; ++i)
    if (i <= fp) oM = compose(oM, k.liMi[i]);
  return compose(oM, P.frame_placement());
}
// getFrameJacobian(LOCAL): 6 x NV column-major, rows [linear; angular]
template <int NV>
CGX_HD void frame_jacobian_local(const ArmParams& P, const Kin<NV>& k, double* J) {
  const int fp = P.frame_parent();
  SE3 fMk = inverse(P.frame_placement());
#pragma unroll
  for (int kk = NV - 1; kk >= 0; --kk) {
    if (kk > fp) {
#pragma unroll
      for (int r = 0; r < 6; ++r) J[r + 6 * kk] = 0.;
      continue;
    }
    const Sp col = actM(fMk, Sp{V3{0., 0., 0.}, P.axis(kk)});
    J[0 + 6 * kk] = col.lin.x;
    J[1 + 6 * kk] = col.lin.y;
    J[2 + 6 * kk] = col.lin.z;
    J[3 + 6 * kk] = col.ang.x;
    J[4 + 6 * kk] = col.ang.y;
    J[5 + 6 * kk] = col.ang.z;
    if (kk > 0) fMk = compose(fMk, inverse(k.liMi[kk]));
  }
}

// ---------------------------------------------------------------------------------------------
// log3 / log6 / Jlog3 / Jlog6 (the formulas of Pinocchio's spatial/explog.hpp; frame-placement.hxx:59,75).
// Small-angle switch at eps^(1/4) = 2^-13 (TaylorSeriesExpansion<double>::precision<3>()).
#define CGX_TAYLOR_PREC 0.0001220703125
#define CGX_PI 3.14159265358979323846
CGX_HD V3 log3(const M3& R, double& theta) {
  const double tr = R(0, 0) + R(1, 1) + R(2, 2);
  if (tr > 3.)
    theta = 0.;
  else if (tr < -1.)
    theta = CGX_PI;
  else
    theta = acos((tr - 1.) / 2.);
  V3 res;
  if (theta >= CGX_PI - 1e-2) {
    const double cphi = cos(theta - CGX_PI);
    const double beta = theta * theta / (1. + cphi);
    const double t0 = (R(0, 0) + cphi) * beta, t1 = (R(1, 1) + cphi) * beta, t2 = (R(2, 2) + cphi) * beta;
    res.x = (R(2, 1) > R(1, 2) ? 1. : -1.) * (t0 > 0. ? sqrt(t0) : 0.);
    res.y = (R(0, 2) > R(2, 0) ? 1. : -1.) * (t1 > 0. ? sqrt(t1) : 0.);
    res.z = (R(1, 0) > R(0, 1) ? 1. : -1.) * (t2 > 0. ? sqrt(t2) : 0.);
  } else {
    const double t = ((theta > CGX_TAYLOR_PREC) ? theta / sin(theta) : 1.) / 2.;
    res.x = t * (R(2, 1) - R(1, 2));
    res.y = t * (R(0, 2) - R(2, 0));
    res.z = t * (R(1, 0) - R(0, 1));
  }
  return res;
}
// r = log6(M) as [linear; angular]; also returns theta and w = log3(R) for Jlog6
CGX_HD Sp log6(const SE3& M, double& t, V3& w) {
  w = log3(M.R, t);
  const double t2 = t * t;
  double alpha, beta;
  if (t < CGX_TAYLOR_PREC) {
    alpha = 1. - t2 / 12. - t2 * t2 / 720.;
    beta = 1. / 12. + t2 / 720.;
  } else {
    double st, ct;
    sincos(t, &st, &ct);
    alpha = t * st / (2. * (1. - ct));
    beta = 1. / t2 - st / (2. * t * (1. - ct));
  }
  Sp r;
  r.lin = (alpha * M.p - 0.5 * cross(w, M.p)) + (beta * dot(w, M.p)) * w;
  r.ang = w;
  return r;
}
CGX_HD M3 Jlog3(double theta, const V3& lg) {
  double alpha, diag;
  if (theta < CGX_TAYLOR_PREC) {
    const double t2 = theta * theta;
    alpha = 1. / 12. + t2 / 720.;
    diag = 0.5 * (2. - t2 / 6.);
  } else {
    double st, ct;
    sincos(theta, &st, &ct);
    const double st_1mct = st / (1. - ct);
    alpha = 1. / (theta * theta) - st_1mct / (2. * theta);
    diag = 0.5 * (theta * st_1mct);
  }
  const double l[3] = {lg.x, lg.y, lg.z};
  M3 J;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) J(i, j) = alpha * l[i] * l[j];
#pragma unroll
  for (int i = 0; i < 3; ++i) J(i, i) += diag;
  J(0, 1) -= 0.5 * lg.z;
  J(0, 2) += 0.5 * lg.y;
  J(1, 0) += 0.5 * lg.z;
  J(1, 2) -= 0.5 * lg.x;
  J(2, 0) -= 0.5 * lg.y;
  J(2, 1) += 0.5 * lg.x;
  return J;
}
// Jlog6(M) = [[A, B], [0, A]] (rows/cols [linear; angular]); returns the blocks A and B
CGX_HD void Jlog6(const SE3& M, double t, const V3& w, M3& A, M3& Bm) {
  A = Jlog3(t, w);
  const double t2 = t * t;
  double beta, beta_dot_over_theta;
  if (t < CGX_TAYLOR_PREC) {
    beta = 1. / 12. + t2 / 720.;
    beta_dot_over_theta = 1. / 360.;
  } else {
    const double tinv = 1. / t, t2inv = tinv * tinv;
    double st, ct;
    sincos(t, &st, &ct);
    const double inv_2_2ct = 1. / (2. * (1. - ct));
    beta = t2inv - st * tinv * inv_2_2ct;
    beta_dot_over_theta = -2. * t2inv * t2inv + (1. + st * tinv) * t2inv * inv_2_2ct;
  }
  const V3& p = M.p;
  const double wTp = dot(w, p);
  const V3 v3_ = (beta_dot_over_theta * wTp) * w - (t2 * beta_dot_over_theta + 2. * beta) * p;
  const double wv[3] = {w.x, w.y, w.z}, pv[3] = {p.x, p.y, p.z}, vv[3] = {v3_.x, v3_.y, v3_.z};
  M3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C(i, j) = vv[i] * wv[j] + beta * wv[i] * pv[j];
#pragma unroll
  for (int i = 0; i < 3; ++i) C(i, i) += wTp * beta;
  C(0, 1) -= 0.5 * p.z;
  C(0, 2) += 0.5 * p.y;
  C(1, 0) += 0.5 * p.z;
  C(1, 2) -= 0.5 * p.x;
  C(2, 0) -= 0.5 * p.y;
  C(2, 1) += 0.5 * p.x;
  Bm = mul(C, A);
}

// ---------------------------------------------------------------------------------------------
// Node level: IntegratedActionModelEuler(DifferentialActionModelFreeFwdDynamics) with ActuationModelFull and
// CostModelSum{gripperPose, uReg, xReg} (reference: core/integrator/euler.hxx:41-131, multibody/actions/
// free-fwddyn.hxx:44-118, multibody/actuations/full.hpp:38-63, multibody/costs/cost-sum.hxx:89-160 — std::map name
// order gripperPose < uReg < xReg —, costs/state.hxx:130-169, control.hxx:56-87, frame-placement.hxx:52-86,
// core/activations/quadratic.hpp:34-60).

// residual of the frame-placement cost and what its derivative needs
struct FrameRes {
  SE3 rMf;  // placement: Mref^-1 oMf; rotation: R = Rref^T oRf; translation: R = oRf
  Sp r;     // residual as [linear; angular], the unused half zero
  double theta;
  V3 w;
};
// residual of the goal cost of the frame placed at oMf: CostModelFramePlacement (frame-placement.hxx:52-64),
// CostModelFrameTranslation (frame-translation.hxx:51-61) or CostModelFrameRotation (frame-rotation.hxx:51-65)
CGX_HD FrameRes frame_residual_at(const ArmParams& P, const SE3& oMf) {
  FrameRes fr;
  const int kind = P.goal_kind();
  if (kind == GOAL_TRANSLATION) {
    const SE3 ref = P.Mref();
    fr.rMf = SE3{oMf.R, V3{0., 0., 0.}};
    fr.r = Sp{oMf.p - ref.p, V3{0., 0., 0.}};
    fr.theta = 0.;
    fr.w = V3{0., 0., 0.};
  } else if (kind == GOAL_ROTATION) {
    fr.rMf = SE3{mulAtB(P.Mref().R, oMf.R), V3{0., 0., 0.}};  // oRf_inv_ * oRf
    fr.w = log3(fr.rMf.R, fr.theta);
    fr.r = Sp{V3{0., 0., 0.}, fr.w};
  } else {
    fr.rMf = actInv(P.Mref(), oMf);  // oMf_inv_ * oMf (frame-placement.hxx:58)
    fr.r = log6(fr.rMf, fr.theta, fr.w);
  }
  return fr;
}
template <int NV>
CGX_HD FrameRes frame_residual(const ArmParams& P, const Kin<NV>& k) {
  return frame_residual_at(P, frame_placement<NV>(P, k));
}
// d r / d q = [[A, Bm], [0, C]] fJf with fJf = getFrameJacobian(LOCAL) as [linear; angular] rows:
// placement: Jlog6(rMf) (frame-placement.hxx:75-76); translation: oRf * fJf.topRows(3) (frame-translation.hxx:71-72);
// rotation: Jlog3(rRf) * fJf.bottomRows(3) (frame-rotation.hxx:76-78)
CGX_HD void frame_residual_jacobian(const ArmParams& P, const FrameRes& fr, M3& A, M3& Bm, M3& C) {
  const int kind = P.goal_kind();
  if (kind == GOAL_TRANSLATION) {
    A = fr.rMf.R;
    Bm = zero3();
    C = zero3();
  } else if (kind == GOAL_ROTATION) {
    A = zero3();
    Bm = zero3();
    C = Jlog3(fr.theta, fr.w);
  } else {
    Jlog6(fr.rMf, fr.theta, fr.w, A, Bm);
    C = A;
  }
}
// differential cost (cost-sum.hxx:103-118, in name order); u == nullptr: zeros
template <int NV>
CGX_HD double diff_cost(const ArmParams& P, const FrameRes& fr, const double* x, const double* u) {
  double c_gp = 0., c_u = 0., c_x = 0.;
  const double r6[6] = {fr.r.lin.x, fr.r.lin.y, fr.r.lin.z, fr.r.ang.x, fr.r.ang.y, fr.r.ang.z};
#pragma unroll
  for (int i = 0; i < 6; ++i) c_gp += r6[i] * r6[i];
  if (u) {
#pragma unroll
    for (int i = 0; i < NV; ++i) c_u += u[i] * u[i];
  }
#pragma unroll
  for (int i = 0; i < 2 * NV; ++i) c_x += x[i] * ((i < NV ? P.wq(i) : P.wv(i - NV)) * x[i]);  // r . (w o r)
  double cost = 0.;
  cost += P.w_gp() * (0.5 * c_gp);
  cost += P.w_u() * (0.5 * c_u);
  cost += P.w_x() * (0.5 * c_x);
  return cost;
}

// calc of one node (euler.hxx:41-77 over free-fwddyn.hxx:44-80). terminal: dt = 0 (benchmark/factory/arm.hpp:89-90),
// u == nullptr means the zero control of the x-only overload (action-base.hxx:28-31). qdd_out (NV, may be null)
// receives DifferentialActionData::xout. Returns the node cost.
template <int NV>
CGX_HD double arm_calc_node(const ArmParams& P, bool terminal, const double* x, const double* u, double* xnext,
                            double* qdd_out) {
  Kin<NV> k;
  double a[NV];
  aba<NV>(P, x, x + NV, u, a, k);
  const FrameRes fr = frame_residual<NV>(P, k);
  const double cost = diff_cost<NV>(P, fr, x, u);
  double dt = P.dt();  // the terminal node has its own block (dt = 0 by default, benchmark/factory/arm.hpp:89-90)
  if (dt < 0.) dt = 1e-3;  // euler.hxx:27-31
  if (qdd_out) {
#pragma unroll
    for (int i = 0; i < NV; ++i) qdd_out[i] = a[i];
  }
  if (dt != 0.) {  // euler.hxx:65-69
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const double dq = x[NV + i] * dt + a[i] * (dt * dt);
      const double dv = a[i] * dt;
      xnext[i] = x[i] + dq;
      xnext[NV + i] = x[NV + i] + dv;
    }
    return dt * cost;
  }
#pragma unroll
  for (int i = 0; i < 2 * NV; ++i) xnext[i] = x[i];  // euler.hxx:70-74
  return cost;
}

// quasiStatic (free-fwddyn.hxx:146-169): u = pinv(dtau_du = I) rnea(q, 0, 0)
template <int NV>
CGX_HD void arm_quasi_static_node(const ArmParams& P, const double* x, double* u) {
  Kin<NV> k;
  Dyn<NV> d;
  rnea<NV>(P, x, nullptr, nullptr, u, k, d, false);
}

// calcDiff of one node (euler.hxx:80-131 over free-fwddyn.hxx:83-118 and the cost derivatives), self-contained:
// recomputes what calc left in the reference's data (ABA, frame residual) from (x, u). Outputs column-major:
// Fx (n x n), Fu (n x NV), Lx (n), Lu (NV), Lxx (n x n), Luu (NV x NV), n = 2 NV. Lxu is identically zero for this
// cost set and is never written. terminal / u as in arm_calc_node.
template <int NV>
CGX_HD void arm_calcdiff_node(const ArmParams& P, bool terminal, const double* x, const double* u, double* Fx, double* Fu,
                              double* Lx, double* Lu, double* Lxx, double* Luu) {
  constexpr int n = 2 * NV;
  double dt = P.dt();  // the terminal node has its own block (dt = 0 by default, benchmark/factory/arm.hpp:89-90)
  if (dt < 0.) dt = 1e-3;
  const bool integ = dt != 0.;
  const double sc = integ ? dt : 1.;  // L* = dt * L*_diff when integrating (euler.hxx:117-121), unscaled otherwise
  Kin<NV> k;
  double a[NV];
  aba<NV>(P, x, x + NV, u, a, k);
  // ---- cost derivatives (frame-placement.hxx:68-86: J = Jlog6(rMf) fJf, Lx = J^T r, Lxx = J^T J)
  {
    const FrameRes fr = frame_residual<NV>(P, k);
    M3 A, Bm, Cb;
    frame_residual_jacobian(P, fr, A, Bm, Cb);
    double fJf[6 * NV], J[6 * NV];
    frame_jacobian_local<NV>(P, k, fJf);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const V3 l{fJf[0 + 6 * j], fJf[1 + 6 * j], fJf[2 + 6 * j]}, w{fJf[3 + 6 * j], fJf[4 + 6 * j], fJf[5 + 6 * j]};
      const V3 top = mul(A, l) + mul(Bm, w), bot = mul(Cb, w);
      J[0 + 6 * j] = top.x;
      J[1 + 6 * j] = top.y;
      J[2 + 6 * j] = top.z;
      J[3 + 6 * j] = bot.x;
      J[4 + 6 * j] = bot.y;
      J[5 + 6 * j] = bot.z;
    }
    const double r6[6] = {fr.r.lin.x, fr.r.lin.y, fr.r.lin.z, fr.r.ang.x, fr.r.ang.y, fr.r.ang.z};
    const double wgp = P.w_gp(), wx = P.w_x(), wu = P.w_u();
    double q2[2];  // n * n = 4 NV^2 elements, written as 16-byte pairs in flat (column-major) order
#pragma unroll
    for (int j = 0; j < n; ++j) {
      double lx = 0.;
      if (j < NV) {
        double s = 0.;
#pragma unroll
        for (int kk = 0; kk < 6; ++kk) s += J[kk + 6 * j] * r6[kk];
        lx += wgp * s;
      }
      const double wj = (j < NV) ? P.wq(j) : P.wv(j - NV);
      lx += wx * (wj * x[j]);
      Lx[j] = sc * lx;
#pragma unroll
      for (int i = 0; i < n; ++i) {
        double lxx = 0.;
        if (i < NV && j < NV) {
          double s = 0.;
#pragma unroll
          for (int kk = 0; kk < 6; ++kk) s += J[kk + 6 * i] * J[kk + 6 * j];
          lxx += wgp * s;
        }
        if (i == j) lxx += wx * wj;
        const int e = i + n * j;
        q2[e & 1] = sc * lxx;
        if ((e & 1) == 1) st2(Lxx + (e - 1), q2[0], q2[1]);
      }
    }
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      Lu[j] = sc * (0. + wu * (u ? u[j] : 0.));
#pragma unroll
      for (int i = 0; i < NV; ++i) Luu[i + NV * j] = sc * (i == j ? 0. + wu * 1. : 0.);
    }
  }
  if (!integ) {  // euler.hxx:122-130: Fx = I, Fu = 0
#pragma unroll
    for (int j = 0; j < n; ++j)
#pragma unroll
      for (int i = 0; i < n; ++i) Fx[i + n * j] = (i == j) ? 1. : 0.;
#pragma unroll
    for (int e = 0; e < n * NV; ++e) Fu[e] = 0.;
    return;
  }
  // ---- dynamics derivatives (free-fwddyn.hxx:103-107): ddq_dq = -Minv dtau_dq, ddq_dv = -Minv dtau_dv, da_du = Minv
  const double dt2 = dt * dt;
  double Minv[NV * NV];
  {
    double M[NV * NV];
    crba<NV>(P, k, M);
#pragma unroll
    for (int i = 0; i < NV; ++i) M[i + NV * i] += P.armature(i);  // free-fwddyn.hxx:69
    spd_inverse<NV>(M, Minv);
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) {  // column j of Fu = [dt^2 Minv(:,j); dt Minv(:,j)], 16-byte stores
    double col[n];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      col[i] = Minv[i + NV * j] * dt2;
      col[NV + i] = Minv[i + NV * j] * dt;
    }
#pragma unroll
    for (int i = 0; i < n; i += 2) st2(Fu + n * j + i, col[i], col[i + 1]);
  }
  Dyn<NV> d;
  {
    double tau[NV];
    rnea<NV>(P, x, x + NV, a, tau, k, d, true);
  }
#if defined(CGX_RB_COL_NOUNROLL)
#pragma unroll 1
#endif
  for (int kc = 0; kc < NV; ++kc) {
    double tq[NV], tv[NV];
    rnea_derivative_column<NV>(P, k, d, x + NV, kc, tq, tv);
    double cq[n], cv[n];  // columns kc and NV + kc of Fx
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      double s1 = 0., s2 = 0.;
#pragma unroll
      for (int kk = 0; kk < NV; ++kk) {
        s1 += Minv[i + NV * kk] * tq[kk];
        s2 += Minv[i + NV * kk] * tv[kk];
      }
      const double dq = -s1, dv = -s2;
      double f00 = dq * dt2, f01 = dv * dt2, f11 = dv * dt;
      if (i == kc) {
        f01 += dt;
        f00 += 1.;
        f11 += 1.;
      }
      cq[i] = f00;
      cv[i] = f01;
      cq[NV + i] = dq * dt;
      cv[NV + i] = f11;
    }
#pragma unroll
    for (int i = 0; i < n; i += 2) {
      st2(Fx + n * kc + i, cq[i], cq[i + 1]);
      st2(Fx + n * (NV + kc) + i, cv[i], cv[i + 1]);
    }
  }
}

}  // namespace rb
}  // namespace cgx
