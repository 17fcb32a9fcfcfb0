// pg_dynamics.cuh -- one Bullet stepSimulation() of the 7-DoF Panda arm, one environment per thread, fp32.
//
// Replaces pybullet.stepSimulation behind panda_gym/pybullet.py:68-71 (reference) for the robot the fork
// loads (franka_panda_custom_0, panda.py:54).  Same mathematics as the oracle (oracle/pgo_dyn.c), different
// algorithm: instead of Bullet's articulated-body pass plus seven impulse-response passes, the thread builds
//   tau_b = RNEA(q, qd, qdd = 0)  (gravity + Coriolis + btMultiBody per-link damping), link-local frames
//   M     = CRBA(q)               (7x7, 28 unique entries, registers)
//   M^-1  = (L L^T)^-1            (Cholesky, explicit inverse: the PGS rows need its columns)
// and then runs Bullet's projected Gauss-Seidel over joint-limit + motor rows exactly as
// btMultiBodyConstraintSolver does (50 sweeps, alternating direction, residual early-out 1e-7), followed by
// the semi-implicit Euler position update.  All 7 joints are revolute about their local z axis and every
// joint origin is Rx(0 / +-90 deg), which the templates below fold into sign flips.
#pragma once
#include "pg_model.cuh"
#include <cuda_runtime.h>

namespace pg {

template <typename T> struct V3T { T x, y, z; };
using V3 = V3T<float>;
using D3 = V3T<double>;
template <typename T> struct ident { using type = T; };
template <typename T> __host__ __device__ __forceinline__ V3T<T> v3t(T x, T y, T z) { V3T<T> r; r.x = x; r.y = y; r.z = z; return r; }
__host__ __device__ __forceinline__ V3 v3(float x, float y, float z) { return v3t<float>(x, y, z); }
template <typename T> __host__ __device__ __forceinline__ V3T<T> operator+(V3T<T> a, V3T<T> b) { return v3t<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> __host__ __device__ __forceinline__ V3T<T> operator-(V3T<T> a, V3T<T> b) { return v3t<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> __host__ __device__ __forceinline__ V3T<T> operator*(typename ident<T>::type s, V3T<T> a) { return v3t<T>(s * a.x, s * a.y, s * a.z); }
template <typename T> __host__ __device__ __forceinline__ V3T<T> cross(V3T<T> a, V3T<T> b) { return v3t<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
template <typename T> __host__ __device__ __forceinline__ T dot(V3T<T> a, V3T<T> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// y + s * x, one fused multiply-add per component
template <typename T> __host__ __device__ __forceinline__ V3T<T> axpy(typename ident<T>::type s, V3T<T> x, V3T<T> y) { return v3t<T>(s * x.x + y.x, s * x.y + y.y, s * x.z + y.z); }
// a compile-time vector (URDF constant) against a run-time one: the constant stays a scalar float operand
__device__ __forceinline__ V3 cross_c(V3 a, V3 b) { return cross(a, b); }
__device__ __forceinline__ float norm(V3 a) { return sqrtf(dot(a, a)); }
__host__ __device__ __forceinline__ double norm(D3 a) { return sqrt(dot(a, a)); }

// child -> parent:  E = Rx(ROT*90deg) * Rz(q)
template <int ROT, typename T>
__device__ __forceinline__ V3T<T> to_parent(V3T<T> v, T c, T s) {
  T x = c * v.x - s * v.y, y = s * v.x + c * v.y, z = v.z;
  if (ROT == 1) return v3t<T>(x, -z, y);
  if (ROT == -1) return v3t<T>(x, z, -y);
  return v3t<T>(x, y, z);
}
// parent -> child:  E^T
template <int ROT, typename T>
__device__ __forceinline__ V3T<T> to_child(V3T<T> v, T c, T s) {
  V3T<T> u = v;
  if (ROT == 1) u = v3t<T>(v.x, v.z, -v.y);
  if (ROT == -1) u = v3t<T>(v.x, -v.z, v.y);
  return v3t<T>(c * u.x + s * u.y, c * u.y - s * u.x, u.z);
}

constexpr float kDt = 1.0f / 500.0f;      // pybullet.py:50
constexpr float kGravity = 9.81f;         // pybullet.py:54
constexpr float kDamp = 0.04f;            // btMultiBody linear/angular damping (pinned by the KAT, see oracle)
constexpr float kMaxCoordVel = 100.0f;    // btMultiBody m_maxCoordinateVelocity
constexpr float kLimitErp = 0.2f;         // solverInfo.m_erp on joint-limit rows
constexpr float kLimitMaxImpulse = 100.f; // btMultiBodyConstraint m_maxAppliedImpulse
constexpr int kSolverIters = 50;          // PyBullet numSolverIterations
constexpr float kResidualThr = 1e-7f;     // PyBullet solverResidualThreshold

// kinematic state of a link frame, expressed in that frame
template <typename T> struct KinT { V3T<T> w, al, a, v; };

// wrench (about the frame origin) of one rigid body riding on a frame with state k:
// Newton-Euler with qdd = 0 plus btMultiBody's velocity damping  m v (k + k|v|),  I w (k + k|w|).
// m, c, inertia are URDF constants (scalar float operands for both T).
template <typename T>
__device__ __forceinline__ void body_wrench(const KinT<T>& k, float m, V3 c, float ixx, float iyy, float izz, float ixy, float ixz,
                                            float iyz, V3T<T>& f, V3T<T>& n) {
  V3T<T> wc = cross_c(k.w, c);
  V3T<T> vc = k.v + wc;
  V3T<T> ac = k.a + cross_c(k.al, c) + cross(k.w, wc);
  T dl = kDamp + kDamp * norm(vc);
  V3T<T> F = m * axpy<T>(dl, vc, ac);
  V3T<T> Iw, Ia;
  if (ixy == 0.f && ixz == 0.f && iyz == 0.f) {  // folded at compile time (arm links have diagonal tensors)
    Iw = v3t<T>(ixx * k.w.x, iyy * k.w.y, izz * k.w.z);
    Ia = v3t<T>(ixx * k.al.x, iyy * k.al.y, izz * k.al.z);
  } else {
    Iw = v3t<T>(ixx * k.w.x + ixy * k.w.y + ixz * k.w.z, ixy * k.w.x + iyy * k.w.y + iyz * k.w.z, ixz * k.w.x + iyz * k.w.y + izz * k.w.z);
    Ia = v3t<T>(ixx * k.al.x + ixy * k.al.y + ixz * k.al.z, ixy * k.al.x + iyy * k.al.y + iyz * k.al.z, ixz * k.al.x + iyz * k.al.y + izz * k.al.z);
  }
  T da = kDamp + kDamp * norm(k.w);
  V3T<T> N = axpy<T>(da, Iw, Ia) + cross(k.w, Iw);
  f = f + F;
  n = n + N + cross_c(c, F);
}

template <int I, typename T>
__device__ __forceinline__ KinT<T> kin_step(const KinT<T>& p, T c, T s, T qd) {
  using L = pgm::Link<I>;
  const V3 pp = v3(L::px, L::py, L::pz);
  KinT<T> k;
  V3T<T> wc = to_child<L::rot>(p.w, c, s);
  k.w = v3t<T>(wc.x, wc.y, wc.z + qd);
  V3T<T> alc = to_child<L::rot>(p.al, c, s);
  k.al = v3t<T>(alc.x + wc.y * qd, alc.y - wc.x * qd, alc.z);
  V3T<T> wxp = cross_c(p.w, pp);
  k.a = to_child<L::rot>(p.a + cross_c(p.al, pp) + cross(p.w, wxp), c, s);
  k.v = to_child<L::rot>(p.v + wxp, c, s);
  return k;
}

template <int I, typename T>
__device__ __forceinline__ void link_wrench(const KinT<T>& k, V3T<T>& f, V3T<T>& n) {
  using L = pgm::Link<I>;
  f = v3t<T>(0.f, 0.f, 0.f); n = v3t<T>(0.f, 0.f, 0.f);
  body_wrench<T>(k, L::m, v3(L::cx, L::cy, L::cz), L::ixx, L::iyy, L::izz, 0.f, 0.f, 0.f, f, n);
}
template <int K, typename T>
__device__ __forceinline__ void sub_wrench(const KinT<T>& k, V3T<T>& f, V3T<T>& n) {
  using S = pgm::Sub<K>;
  body_wrench<T>(k, S::m, v3(S::cx, S::cy, S::cz), S::ixx, S::iyy, S::izz, S::ixy, S::ixz, S::iyz, f, n);
}

// child wrench -> parent frame, accumulated
template <int I, typename T>
__device__ __forceinline__ void wrench_up(V3T<T> fc, V3T<T> nc, T c, T s, V3T<T>& f, V3T<T>& n) {
  using L = pgm::Link<I>;
  V3T<T> fp = to_parent<L::rot>(fc, c, s);
  V3T<T> np = to_parent<L::rot>(nc, c, s);
  f = f + fp;
  n = n + np + cross_c(v3(L::px, L::py, L::pz), fp);
}

// tau_b[i] = joint torque needed to hold qdd = 0  (== C(q,qd) qd + g(q) + damping)
template <typename T>
__device__ __forceinline__ void rnea_bias(const T* cq, const T* sq, const T* qd, T* tau) {
  KinT<T> k0;
  k0.w = v3t<T>(0.f, 0.f, 0.f); k0.al = v3t<T>(0.f, 0.f, 0.f); k0.v = v3t<T>(0.f, 0.f, 0.f);
  k0.a = v3t<T>(0.f, 0.f, kGravity);  // gravity as base acceleration
  V3T<T> f1, n1, f2, n2, f3, n3, f4, n4, f5, n5, f6, n6, f7, n7;
  KinT<T> k1 = kin_step<1>(k0, cq[0], sq[0], qd[0]); link_wrench<1>(k1, f1, n1);
  KinT<T> k2 = kin_step<2>(k1, cq[1], sq[1], qd[1]); link_wrench<2>(k2, f2, n2);
  KinT<T> k3 = kin_step<3>(k2, cq[2], sq[2], qd[2]); link_wrench<3>(k3, f3, n3);
  KinT<T> k4 = kin_step<4>(k3, cq[3], sq[3], qd[3]); link_wrench<4>(k4, f4, n4);
  KinT<T> k5 = kin_step<5>(k4, cq[4], sq[4], qd[4]); link_wrench<5>(k5, f5, n5);
  KinT<T> k6 = kin_step<6>(k5, cq[5], sq[5], qd[5]); link_wrench<6>(k6, f6, n6);
  KinT<T> k7 = kin_step<7>(k6, cq[6], sq[6], qd[6]); link_wrench<7>(k7, f7, n7);
  sub_wrench<0>(k7, f7, n7); sub_wrench<1>(k7, f7, n7); sub_wrench<2>(k7, f7, n7);
  tau[6] = n7.z; wrench_up<7>(f7, n7, cq[6], sq[6], f6, n6);
  tau[5] = n6.z; wrench_up<6>(f6, n6, cq[5], sq[5], f5, n5);
  tau[4] = n5.z; wrench_up<5>(f5, n5, cq[4], sq[4], f4, n4);
  tau[3] = n4.z; wrench_up<4>(f4, n4, cq[3], sq[3], f3, n3);
  tau[2] = n3.z; wrench_up<3>(f3, n3, cq[2], sq[2], f2, n2);
  tau[1] = n2.z; wrench_up<2>(f2, n2, cq[1], sq[1], f1, n1);
  tau[0] = n1.z;
}

// ---------------------------------------------------------------------------------------------- CRBA
// spatial inertia about a frame origin: mass (a sum of URDF constants: always a scalar), first moment h = m c,
// rotational inertia about the origin
template <typename T> struct SIT { float m; V3T<T> h; T xx, yy, zz, xy, xz, yz; };
using SI = SIT<float>;

__device__ __forceinline__ void si_add_body(SI& S, float m, V3 c, float ixx, float iyy, float izz, float ixy, float ixz, float iyz) {
  S.m += m;
  S.h = S.h + m * c;
  float cc = dot(c, c);
  S.xx += ixx + m * (cc - c.x * c.x); S.yy += iyy + m * (cc - c.y * c.y); S.zz += izz + m * (cc - c.z * c.z);
  S.xy += ixy - m * c.x * c.y; S.xz += ixz - m * c.x * c.z; S.yz += iyz - m * c.y * c.z;
}
template <int I>
__device__ __forceinline__ SI si_link() {
  using L = pgm::Link<I>;
  SI S; S.m = 0; S.h = v3(0, 0, 0); S.xx = S.yy = S.zz = S.xy = S.xz = S.yz = 0;
  si_add_body(S, L::m, v3(L::cx, L::cy, L::cz), L::ixx, L::iyy, L::izz, 0.f, 0.f, 0.f);
  return S;
}
template <int K>
__device__ __forceinline__ void si_add_sub(SI& S) {
  using B = pgm::Sub<K>;
  si_add_body(S, B::m, v3(B::cx, B::cy, B::cz), B::ixx, B::iyy, B::izz, B::ixy, B::ixz, B::iyz);
}
template <typename T>
__device__ __forceinline__ SIT<T> si_cast(const SI& S) {
  SIT<T> R; R.m = S.m; R.h = v3t<T>(S.h.x, S.h.y, S.h.z);
  R.xx = S.xx; R.yy = S.yy; R.zz = S.zz; R.xy = S.xy; R.xz = S.xz; R.yz = S.yz;
  return R;
}
// composite inertia of frame I-1: its own link (P0, constants) + the child composite C (frame I) re-expressed
template <int I, typename T>
__device__ __forceinline__ SIT<T> si_up(const SIT<T>& C, T c, T s, const SI& P0) {
  using L = pgm::Link<I>;
  // rotate about z by q:  T = Rz I Rz^T
  T cc = c * c, ss = s * s, cs = c * s;
  T cs2 = cs + cs, c2 = cc - ss;
  T xx = cc * C.xx - cs2 * C.xy + ss * C.yy;
  T yy = ss * C.xx + cs2 * C.xy + cc * C.yy;
  T xy = cs * (C.xx - C.yy) + c2 * C.xy;
  T xz = c * C.xz - s * C.yz;
  T yz = s * C.xz + c * C.yz;
  T zz = C.zz;
  // then Rx(+-90): (x,y,z) -> (x,-z,y) / (x,z,-y)
  T Xx = xx, Yy = yy, Zz = zz, Xy = xy, Xz = xz, Yz = yz;
  if (L::rot == 1) { Yy = zz; Zz = yy; Xy = -xz; Xz = xy; Yz = -yz; }
  if (L::rot == -1) { Yy = zz; Zz = yy; Xy = xz; Xz = -xy; Yz = -yz; }
  V3T<T> g = to_parent<L::rot>(C.h, c, s);
  constexpr float px = L::px, py = L::py, pz = L::pz, pp = px * px + py * py + pz * pz;
  const float m = C.m;  // compile-time after inlining
  T pg2 = (2.f * px) * g.x + (2.f * py) * g.y + (2.f * pz) * g.z;
  SIT<T> R;
  R.m = P0.m + m;
  R.h = v3t<T>(g.x + (P0.h.x + m * px), g.y + (P0.h.y + m * py), g.z + (P0.h.z + m * pz));
  R.xx = (Xx - (2.f * px) * g.x) + pg2 + (P0.xx + m * (pp - px * px));
  R.yy = (Yy - (2.f * py) * g.y) + pg2 + (P0.yy + m * (pp - py * py));
  R.zz = (Zz - (2.f * pz) * g.z) + pg2 + (P0.zz + m * (pp - pz * pz));
  R.xy = (Xy - (py * g.x + px * g.y)) + (P0.xy - m * px * py);
  R.xz = (Xz - (pz * g.x + px * g.z)) + (P0.xz - m * px * pz);
  R.yz = (Yz - (pz * g.y + py * g.z)) + (P0.yz - m * py * pz);
  return R;
}
// force produced by unit acceleration of a z-axis joint on composite inertia C: n = I[:,z], f = z x h
template <typename T>
__device__ __forceinline__ void si_col(const SIT<T>& C, V3T<T>& f, V3T<T>& n) { f = v3t<T>(-C.h.y, C.h.x, 0.f); n = v3t<T>(C.xz, C.yz, C.zz); }
template <int I, typename T>
__device__ __forceinline__ void col_up(V3T<T>& f, V3T<T>& n, T c, T s) {
  using L = pgm::Link<I>;
  V3T<T> fp = to_parent<L::rot>(f, c, s);
  V3T<T> np = to_parent<L::rot>(n, c, s);
  n = np + cross_c(v3(L::px, L::py, L::pz), fp);
  f = fp;
}

// packed symmetric index, i >= j
__host__ __device__ constexpr int sym(int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; }

// joint-space mass matrix, 28 packed entries
template <typename T>
__device__ __forceinline__ void crba(const T* cq, const T* sq, T* M) {
  V3T<T> f7, n7, f6, n6, f5, n5, f4, n4, f3, n3, f2, n2, f1, n1;
  SI C7 = si_link<7>();
  si_add_sub<0>(C7); si_add_sub<1>(C7); si_add_sub<2>(C7);
  SIT<T> C = si_cast<T>(C7);
  si_col(C, f7, n7); M[sym(6, 6)] = n7.z;
  C = si_up<7>(C, cq[6], sq[6], si_link<6>());
  col_up<7>(f7, n7, cq[6], sq[6]); M[sym(6, 5)] = n7.z;
  si_col(C, f6, n6); M[sym(5, 5)] = n6.z;
  C = si_up<6>(C, cq[5], sq[5], si_link<5>());
  col_up<6>(f7, n7, cq[5], sq[5]); M[sym(6, 4)] = n7.z;
  col_up<6>(f6, n6, cq[5], sq[5]); M[sym(5, 4)] = n6.z;
  si_col(C, f5, n5); M[sym(4, 4)] = n5.z;
  C = si_up<5>(C, cq[4], sq[4], si_link<4>());
  col_up<5>(f7, n7, cq[4], sq[4]); M[sym(6, 3)] = n7.z;
  col_up<5>(f6, n6, cq[4], sq[4]); M[sym(5, 3)] = n6.z;
  col_up<5>(f5, n5, cq[4], sq[4]); M[sym(4, 3)] = n5.z;
  si_col(C, f4, n4); M[sym(3, 3)] = n4.z;
  C = si_up<4>(C, cq[3], sq[3], si_link<3>());
  col_up<4>(f7, n7, cq[3], sq[3]); M[sym(6, 2)] = n7.z;
  col_up<4>(f6, n6, cq[3], sq[3]); M[sym(5, 2)] = n6.z;
  col_up<4>(f5, n5, cq[3], sq[3]); M[sym(4, 2)] = n5.z;
  col_up<4>(f4, n4, cq[3], sq[3]); M[sym(3, 2)] = n4.z;
  si_col(C, f3, n3); M[sym(2, 2)] = n3.z;
  C = si_up<3>(C, cq[2], sq[2], si_link<2>());
  col_up<3>(f7, n7, cq[2], sq[2]); M[sym(6, 1)] = n7.z;
  col_up<3>(f6, n6, cq[2], sq[2]); M[sym(5, 1)] = n6.z;
  col_up<3>(f5, n5, cq[2], sq[2]); M[sym(4, 1)] = n5.z;
  col_up<3>(f4, n4, cq[2], sq[2]); M[sym(3, 1)] = n4.z;
  col_up<3>(f3, n3, cq[2], sq[2]); M[sym(2, 1)] = n3.z;
  si_col(C, f2, n2); M[sym(1, 1)] = n2.z;
  C = si_up<2>(C, cq[1], sq[1], si_link<1>());
  col_up<2>(f7, n7, cq[1], sq[1]); M[sym(6, 0)] = n7.z;
  col_up<2>(f6, n6, cq[1], sq[1]); M[sym(5, 0)] = n6.z;
  col_up<2>(f5, n5, cq[1], sq[1]); M[sym(4, 0)] = n5.z;
  col_up<2>(f4, n4, cq[1], sq[1]); M[sym(3, 0)] = n4.z;
  col_up<2>(f3, n3, cq[1], sq[1]); M[sym(2, 0)] = n3.z;
  col_up<2>(f2, n2, cq[1], sq[1]); M[sym(1, 0)] = n2.z;
  si_col(C, f1, n1); M[sym(0, 0)] = n1.z;
}

// M (packed, SPD) -> M^-1 (packed), in place, via Cholesky
template <typename T>
__device__ __forceinline__ void spd_inverse7(T* A) {
  // L L^T = A  (lower, packed)
#pragma unroll
  for (int j = 0; j < 7; j++) {
    T d = A[sym(j, j)];
#pragma unroll
    for (int k = 0; k < j; k++) d -= A[sym(j, k)] * A[sym(j, k)];
    T inv = rsqrtf(d);
    A[sym(j, j)] = inv;  // store 1 / L_jj
#pragma unroll
    for (int i = j + 1; i < 7; i++) {
      T v = A[sym(i, j)];
#pragma unroll
      for (int k = 0; k < j; k++) v -= A[sym(i, k)] * A[sym(j, k)];
      A[sym(i, j)] = v * inv;
    }
  }
  // W = L^-1 (lower), in place; diagonal already holds 1/L_jj
#pragma unroll
  for (int j = 0; j < 7; j++) {
#pragma unroll
    for (int i = j + 1; i < 7; i++) {
      // entries A[sym(k, j)] for j < k < i already hold W_kj (computed earlier in this i-loop)
      T v = -(A[sym(i, j)] * A[sym(j, j)]);
#pragma unroll
      for (int k = j + 1; k < i; k++) v -= A[sym(i, k)] * A[sym(k, j)];
      A[sym(i, j)] = v * A[sym(i, i)];
    }
  }
  // A^-1 = W^T W:  out_ij = sum_{k >= max(i,j)} W_ki W_kj ; ascending (i, j<=i) order only reads W rows >= i
#pragma unroll
  for (int i = 0; i < 7; i++) {
#pragma unroll
    for (int j = 0; j <= i; j++) {
      T v = A[sym(i, i)] * A[sym(i, j)];
#pragma unroll
      for (int k = i + 1; k < 7; k++) v += A[sym(k, i)] * A[sym(k, j)];
      A[sym(i, j)] = v;
    }
  }
}

// Generic projected Gauss-Seidel (cold path): joint-limit rows present and / or non-standard motor impulse
// limits.  Bullet order: rows = [limits (dof order), motors (dof order)]; even iterations sweep backward, odd
// forward; stop when max |residual| <= sqrt(1e-7).  Not inlined (its register working set is larger than the hot
// loop's); inputs arrive through local memory and are re-loaded into registers, the sweeps are fully unrolled.
static __device__ __noinline__ void pgs_generic(const float* Mi_in, const float* q, const float* qd, const float* rhs_in, const float* dinv_in,
                                         const float* max_imp_in, float* dv_out) {
  const float lo[7] = {pgm::Link<1>::lo, pgm::Link<2>::lo, pgm::Link<3>::lo, pgm::Link<4>::lo, pgm::Link<5>::lo, pgm::Link<6>::lo, pgm::Link<7>::lo};
  const float hi[7] = {pgm::Link<1>::hi, pgm::Link<2>::hi, pgm::Link<3>::hi, pgm::Link<4>::hi, pgm::Link<5>::hi, pgm::Link<6>::hi, pgm::Link<7>::hi};
  float Mi[28], rhs[7], dinv[7], max_imp[7], app[7], dv[7], lrhs[7], lsg[7], lapp[7];
#pragma unroll
  for (int k = 0; k < 28; k++) Mi[k] = Mi_in[k];
#pragma unroll
  for (int d = 0; d < 7; d++) {
    rhs[d] = rhs_in[d]; dinv[d] = dinv_in[d]; max_imp[d] = max_imp_in[d];
    app[d] = 0.f; dv[d] = 0.f; lsg[d] = 0.f; lrhs[d] = 0.f; lapp[d] = 0.f;
    float pl = q[d] - lo[d], pu = hi[d] - q[d];  // btMultiBodyJointLimitConstraint: row only when at / over the limit
    if (pl <= 0.f) { lsg[d] = 1.f; lrhs[d] = (-pl * kLimitErp / kDt - qd[d]) * dinv[d]; }
    else if (pu <= 0.f) { lsg[d] = -1.f; lrhs[d] = (-pu * kLimitErp / kDt + qd[d]) * dinv[d]; }
  }
#define PG_GM_ROW(d)                                                        \
  if (max_imp[d] != 0.f) {                                                  \
    float t = fmaf(-dv[d], dinv[d], app[d] + rhs[d]);                       \
    float cl = fminf(fmaxf(t, -max_imp[d]), max_imp[d]);                    \
    float delta = cl - app[d];                                              \
    app[d] = cl;                                                            \
    _Pragma("unroll") for (int i = 0; i < 7; i++) dv[i] = fmaf(Mi[sym(i, d)], delta, dv[i]); \
    res = fmaxf(res, fabsf(delta * Mi[sym(d, d)]));                         \
  }
#define PG_GL_ROW(d)                                                        \
  if (lsg[d] != 0.f) {                                                      \
    float delta = lrhs[d] - lsg[d] * dv[d] * dinv[d];                       \
    float sum = lapp[d] + delta;                                            \
    if (sum < 0.f) { delta = -lapp[d]; lapp[d] = 0.f; }                     \
    else if (sum > kLimitMaxImpulse) { delta = kLimitMaxImpulse - lapp[d]; lapp[d] = kLimitMaxImpulse; } \
    else lapp[d] = sum;                                                     \
    float sd = lsg[d] * delta;                                              \
    _Pragma("unroll") for (int i = 0; i < 7; i++) dv[i] = fmaf(Mi[sym(i, d)], sd, dv[i]); \
    res = fmaxf(res, fabsf(delta * Mi[sym(d, d)]));                         \
  }
  for (int it = 0; it < kSolverIters; it += 2) {
    float res = 0.f;  // even iteration: backward (motors 6..0, then limits 6..0)
    PG_GM_ROW(6) PG_GM_ROW(5) PG_GM_ROW(4) PG_GM_ROW(3) PG_GM_ROW(2) PG_GM_ROW(1) PG_GM_ROW(0)
    PG_GL_ROW(6) PG_GL_ROW(5) PG_GL_ROW(4) PG_GL_ROW(3) PG_GL_ROW(2) PG_GL_ROW(1) PG_GL_ROW(0)
    if (res <= 3.1622777e-4f) break;
    res = 0.f;        // odd iteration: forward (limits 0..6, then motors 0..6)
    PG_GL_ROW(0) PG_GL_ROW(1) PG_GL_ROW(2) PG_GL_ROW(3) PG_GL_ROW(4) PG_GL_ROW(5) PG_GL_ROW(6)
    PG_GM_ROW(0) PG_GM_ROW(1) PG_GM_ROW(2) PG_GM_ROW(3) PG_GM_ROW(4) PG_GM_ROW(5) PG_GM_ROW(6)
    if (res <= 3.1622777e-4f) break;
  }
#undef PG_GM_ROW
#undef PG_GL_ROW
#pragma unroll
  for (int d = 0; d < 7; d++) dv_out[d] = dv[d];
}

// One stepSimulation().  q, qd updated in place; cq/sq hold cos/sin of the CURRENT q on entry and of the new q
// on exit (incrementally rotated).  tgt = POSITION_CONTROL targets (velocity targets when `velocity`), read
// through tgt[d * tgt_stride] (shared memory in the step kernel); unit_imp: every motor's impulse limit is 1
// (load-time default motors, "pcc" control) instead of panda.py:63 joint_forces * dt.
//
// SYNC: every thread of the block calls this (inactive ones with active = false) and the block re-converges at
// a barrier after each phase.  The step kernel is ~40k SASS instructions of mostly straight-line code, far beyond
// the 32 KB L1.5 instruction cache: un-synchronised warps each stream the code from L2 on their own and the
// kernel becomes instruction-fetch bound (ncu: 43 % stall_no_inst).  Keeping the warps of an SM within one cache
// window of each other turns those fetches into hits.
__device__ __forceinline__ constexpr float motor_imp(int d) { return (d < 4 ? 87.0f : (d == 4 ? 12.0f : 120.0f)) * kDt; }  // panda.py:63

#ifndef PG_SYNC_MASK
#define PG_SYNC_MASK 8   // measured best on B200 (0.795 ms vs 0.843 ms with all five); which of the five per-substep barriers are compiled in (bit 0..3: A..D, bit 4: after the collision pass)
#endif
#define PG_SYNC_A ((PG_SYNC_MASK & 1) != 0)
#define PG_SYNC_B ((PG_SYNC_MASK & 2) != 0)
#define PG_SYNC_C ((PG_SYNC_MASK & 4) != 0)
#define PG_SYNC_D ((PG_SYNC_MASK & 8) != 0)
#define PG_SYNC_E ((PG_SYNC_MASK & 16) != 0)
#ifdef PG_SWEEP_PROBE
static __device__ int g_sweeps[20 * 148 * 512];  // [substep of the step][thread]
static __device__ int g_calls[148 * 512];
#endif
template <bool SYNC>
__device__ __forceinline__ void substep(bool active, bool velocity, bool unit_imp, float* q, float* qd, float* cq, float* sq,
                                        const float* tgt, int stride) {
  // q and tgt are touched twice per substep: the step kernel keeps them in shared memory (element d at [d * stride])
  float tau[7], Mi[28];
  float rhs[7], dinv[7], dv[7];
  if (active) rnea_bias(cq, sq, qd, tau);
  if (SYNC && PG_SYNC_A) __syncthreads();
  if (active) crba(cq, sq, Mi);
  if (SYNC && PG_SYNC_B) __syncthreads();
  bool slow = unit_imp;
  if (active) {
    spd_inverse7(Mi);
    // 1. unconstrained velocity update  qd += dt * M^-1 (-tau_b)
#pragma unroll
    for (int i = 0; i < 7; i++) {
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 7; j++) acc -= Mi[sym(i, j)] * tau[j];
      qd[i] = fminf(fmaxf(qd[i] + acc * kDt, -kMaxCoordVel), kMaxCoordVel);
    }
    // 2. motor rows: unit Jacobian on dof d, rhs = (v* - v) / (M^-1)_dd
    const float lo[7] = {pgm::Link<1>::lo, pgm::Link<2>::lo, pgm::Link<3>::lo, pgm::Link<4>::lo, pgm::Link<5>::lo, pgm::Link<6>::lo, pgm::Link<7>::lo};
    const float hi[7] = {pgm::Link<1>::hi, pgm::Link<2>::hi, pgm::Link<3>::hi, pgm::Link<4>::hi, pgm::Link<5>::hi, pgm::Link<6>::hi, pgm::Link<7>::hi};
#pragma unroll
    for (int d = 0; d < 7; d++) {
      float denom = Mi[sym(d, d)];
      dinv[d] = denom > 1.1920929e-07f ? 1.0f / denom : 0.f;
      float t = tgt[d * stride];
      float qv = q[d * stride];
      float want = velocity ? t : 50.0f * (t - qv);  // kp (q* - q)/dt + v + kd (0 - v): kp .1, kd 1, dt 1/500
      rhs[d] = (want - qd[d]) * dinv[d];
      dv[d] = 0.f;
      slow |= (qv - lo[d] <= 0.f) || (hi[d] - qv <= 0.f);
    }
  }
  if (SYNC && PG_SYNC_C) __syncthreads();
  if (active) {
    // 3. projected Gauss-Seidel exactly as btMultiBodyConstraintSolver: 50 sweeps, even backward / odd forward,
    //    residual early-out.  A lane whose residual test fired stays masked off by divergence; the warp leaves
    //    when all lanes are done.
    if (slow) {
      float mi[7], ql[7];
#pragma unroll
      for (int d = 0; d < 7; d++) { mi[d] = unit_imp ? 1.0f : motor_imp(d); ql[d] = q[d * stride]; }
      pgs_generic(Mi, ql, qd, rhs, dinv, mi, dv);
    } else {
      float app[7];
#pragma unroll
      for (int d = 0; d < 7; d++) app[d] = 0.f;
#define PG_MOTOR_ROW(d)                                                     \
  {                                                                         \
    float t = fmaf(-dv[d], dinv[d], app[d] + rhs[d]);                       \
    float cl = fminf(fmaxf(t, -motor_imp(d)), motor_imp(d));                \
    float delta = cl - app[d];                                              \
    app[d] = cl;                                                            \
    _Pragma("unroll") for (int i = 0; i < 7; i++) dv[i] = fmaf(Mi[sym(i, d)], delta, dv[i]); \
    res = fmaxf(res, fabsf(delta * Mi[sym(d, d)]));                         \
  }
      const float kResAbs = 3.1622777e-4f;  // sqrt(1e-7): compare |residual| instead of residual^2
      // Consecutive sweeps run in opposite directions, so every sweep after the first would start on the row the
      // previous one just finished.  Re-solving a row whose residual was zeroed one instruction ago yields delta = 0
      // in exact arithmetic (clamped or not) and ~1e-8 in fp32 (Bullet, in fp64: ~1e-17); those 49 of 350 row updates
      // are skipped.  The residual test is unaffected (the skipped row would contribute ~0 to the maximum).
      bool live = true;
      float carry = 0.f;
      { float res = 0.f; PG_MOTOR_ROW(6) carry = res; }
      for (int it = 0; it < kSolverIters; it += 2) {
#ifdef PG_SWEEP_PROBE  // variant build only (tools/sweep_probe.py): sweep pairs this environment ran, summed over the step
        if (live) g_sweeps[(size_t)(g_calls[blockIdx.x * blockDim.x + threadIdx.x] % 20) * (148 * 512) + blockIdx.x * blockDim.x + threadIdx.x]++;
#endif
        if (live) {  // even iteration: backward sweep (row 6 done above / by the forward sweep before)
          float res = carry;
          PG_MOTOR_ROW(5) PG_MOTOR_ROW(4) PG_MOTOR_ROW(3) PG_MOTOR_ROW(2) PG_MOTOR_ROW(1) PG_MOTOR_ROW(0)
          if (res <= kResAbs) live = false;
        }
        carry = 0.f;
        if (live) {  // odd iteration: forward sweep (row 0 just solved)
          float res = 0.f;
          PG_MOTOR_ROW(1) PG_MOTOR_ROW(2) PG_MOTOR_ROW(3) PG_MOTOR_ROW(4) PG_MOTOR_ROW(5) PG_MOTOR_ROW(6)
          if (res <= kResAbs) live = false;
        }
        if (!__any_sync(__activemask(), live)) break;
      }
#undef PG_MOTOR_ROW
    }
#ifdef PG_SWEEP_PROBE
    g_calls[blockIdx.x * blockDim.x + threadIdx.x]++;
#endif
    // 4. apply, clamp, integrate
#pragma unroll
    for (int d = 0; d < 7; d++) {
      qd[d] = fminf(fmaxf(qd[d] + dv[d], -kMaxCoordVel), kMaxCoordVel);
      float dq = qd[d] * kDt;  // |dq| <= 0.2 (velocity clamp): rotate (cos, sin) by dq with a 5th-order series,
      q[d * stride] += dq;     // error < 1e-7; the caller re-evaluates sincosf(q) at every env step
      float d2 = dq * dq;
      float cd = fmaf(d2, fmaf(d2, 1.0f / 24.0f, -0.5f), 1.0f);
      float sd = dq * fmaf(d2, fmaf(d2, 1.0f / 120.0f, -1.0f / 6.0f), 1.0f);
      float c0 = cq[d], s0 = sq[d];
      cq[d] = c0 * cd - s0 * sd;
      sq[d] = s0 * cd + c0 * sd;
    }
  }
  if (SYNC && PG_SYNC_D) __syncthreads();
}


}  // namespace pg
