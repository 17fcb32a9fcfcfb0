// pg_geom.cuh -- forward kinematics of the collider frames + closest-distance queries between the robot's
// collision primitives (spheres / cylinders of franka_panda_custom_0) and sphere / axis-aligned-box bodies.
//
// Stands in for pybullet.getClosestPoints as the reference uses it through the pyb_utils-fork CollisionDetector
// (reach_ao.py:896-900 check_collided, :902-959 get_obs) and through check_collision (:1169-1181).
// Shape semantics = oracle/pgo_geom.c (Bullet: URDF cylinder = hull + 1 mm margin; box = core shrunk by its
// margin, re-inflated; sphere exact).  fp32, one environment per thread.
#pragma once
#include "pg_dynamics.cuh"

// PG_CHECK: index / invariant assertions of the debug build (-DPG_BOUNDS_CHECK, see tools/bounds_check_run.sh): every
// run-time-indexed shared / global write site asserts its range and traps with the site's line.  Compiled out otherwise.
#if defined(PG_BOUNDS_CHECK) && defined(__CUDA_ARCH__)
#include <cstdio>
#define PG_CHECK(cond) do { if (!(cond)) { printf("PG_CHECK failed: %s  (%s:%d, block %d thread %d)\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, (int)threadIdx.x); __trap(); } } while (0)
#else
#define PG_CHECK(cond) do { } while (0)
#endif

namespace pg {

constexpr float kHullMargin = 0.001f;

struct Frame { V3 x, y, z, o; };  // columns of R (link -> world) and origin

__device__ __forceinline__ Frame base_frame(V3 base) {
  Frame F; F.x = v3(1, 0, 0); F.y = v3(0, 1, 0); F.z = v3(0, 0, 1); F.o = base; return F;
}
template <int I>
__device__ __forceinline__ Frame frame_step(const Frame& P, float c, float s) {
  using L = pgm::Link<I>;
  Frame F;
  F.o = P.o + L::px * P.x + L::py * P.y + L::pz * P.z;
  V3 qy = P.y, qz = P.z;
  if (L::rot == 1) { qy = P.z; qz = v3(-P.y.x, -P.y.y, -P.y.z); }
  if (L::rot == -1) { qy = v3(-P.z.x, -P.z.y, -P.z.z); qz = P.y; }
  F.x = c * P.x + s * qy;
  F.y = c * qy - s * P.x;
  F.z = qz;
  return F;
}
__device__ __forceinline__ V3 frame_point(const Frame& F, float x, float y, float z) { return F.o + x * F.x + y * F.y + z * F.z; }
__device__ __forceinline__ V3 frame_dir(const Frame& F, float x, float y, float z) { return x * F.x + y * F.y + z * F.z; }

// world frame of link7 for joint cosines / sines (cq, sq)
__device__ __forceinline__ Frame link7_frame(V3 base, const float* cq, const float* sq) {
  Frame F = base_frame(base);
  F = frame_step<1>(F, cq[0], sq[0]); F = frame_step<2>(F, cq[1], sq[1]); F = frame_step<3>(F, cq[2], sq[2]);
  F = frame_step<4>(F, cq[3], sq[3]); F = frame_step<5>(F, cq[4], sq[4]); F = frame_step<6>(F, cq[5], sq[5]);
  F = frame_step<7>(F, cq[6], sq[6]);
  return F;
}

// ------------------------------------------------------------------------------------ obstacle description
struct Obst {        // one obstacle slot of the scenario (shape is per scenario, position per environment)
  int type;          // 0 sphere, 1 box
  float sx, sy, sz;  // sphere: sx = radius; box: half extents
  float margin;      // box collision margin  min(0.04, 0.1 * min half extent)
};
constexpr int kMaxObst = 8;

// d = signed closest distance; dir = unit vector robot -> obstacle such that (pointOnB - pointOnA) = d * dir
template <typename T> struct ClosestT { T d; V3T<T> dir; };
using Closest = ClosestT<float>;
template <typename T> __host__ __device__ __forceinline__ T tmin(T a, T b) { return a < b ? a : b; }
template <typename T> __host__ __device__ __forceinline__ T tmax(T a, T b) { return a > b ? a : b; }
template <typename T> __host__ __device__ __forceinline__ T tabs(T a) { return a < T(0) ? -a : a; }
__host__ __device__ __forceinline__ float tsqrt(float a) { return sqrtf(a); }
__host__ __device__ __forceinline__ double tsqrt(double a) { return sqrt(a); }
template <typename T> __host__ __device__ __forceinline__ T tnorm(V3T<T> a) { return tsqrt(dot(a, a)); }

// ------------------------------------------------------------------------------------ analytic pairs
template <typename T> __host__ __device__ __forceinline__ ClosestT<T> sphere_sphere(V3T<T> c, T r, V3T<T> o, T R) {
  V3T<T> d = o - c;
  T L = tnorm(d);
  V3T<T> n = L > T(1.1920929e-07) ? (T(1.0) / L) * d : v3t<T>(T(1), T(0), T(0));
  ClosestT<T> k; k.d = L - r - R; k.dir = n; return k;
}
// Bullet btSphereBoxCollisionAlgorithm::getSphereDistance on the margin-shrunk core e, margin m
template <typename T> __host__ __device__ __forceinline__ ClosestT<T> sphere_box(V3T<T> c, T r, V3T<T> b, V3T<T> e, T m) {
  V3T<T> rel = c - b;
  V3T<T> cl = v3t<T>(tmin<T>(tmax<T>(rel.x, -e.x), e.x), tmin<T>(tmax<T>(rel.y, -e.y), e.y), tmin<T>(tmax<T>(rel.z, -e.z), e.z));
  V3T<T> v = rel - cl;
  T d2 = dot(v, v);
  ClosestT<T> k; V3T<T> n; T d; V3T<T> onb;
  if (d2 <= T(1.1920929e-07)) {  // centre inside the core: face of least penetration
    T best = e.x - rel.x; n = v3t<T>(T(1), T(0), T(0)); onb = v3t<T>(e.x, rel.y, rel.z);
    if (e.x + rel.x < best) { best = e.x + rel.x; n = v3t<T>(T(-1), T(0), T(0)); onb = v3t<T>(-e.x, rel.y, rel.z); }
    if (e.y - rel.y < best) { best = e.y - rel.y; n = v3t<T>(T(0), T(1), T(0)); onb = v3t<T>(rel.x, e.y, rel.z); }
    if (e.y + rel.y < best) { best = e.y + rel.y; n = v3t<T>(T(0), T(-1), T(0)); onb = v3t<T>(rel.x, -e.y, rel.z); }
    if (e.z - rel.z < best) { best = e.z - rel.z; n = v3t<T>(T(0), T(0), T(1)); onb = v3t<T>(rel.x, rel.y, e.z); }
    if (e.z + rel.z < best) { best = e.z + rel.z; n = v3t<T>(T(0), T(0), T(-1)); onb = v3t<T>(rel.x, rel.y, -e.z); }
    d = -best;
  } else {
    d = tsqrt(d2); n = (T(1.0) / d) * v; onb = cl;
  }
  k.d = d - m - r; k.dir = v3t<T>(-n.x, -n.y, -n.z); return k;
}
// point vs solid cylinder (centre c, unit axis a, radius r, half length h): signed distance + closest point
template <typename T> __host__ __device__ __forceinline__ T point_cylinder(V3T<T> p, V3T<T> c, V3T<T> a, T r, T h, V3T<T>& on) {
  V3T<T> d = p - c;
  T t = dot(d, a);
  V3T<T> rho = d - t * a;
  T rn = tnorm(rho);
  T tc = tmin<T>(tmax<T>(t, -h), h);
  T dt = tabs<T>(t) - h, dr = rn - r;
  V3T<T> rdir = rn > T(1e-30) ? (T(1.0) / rn) * rho : v3t<T>(T(0), T(0), T(0));
  if (dt <= T(0.) && dr <= T(0.)) {
    if (-dt < -dr) { on = c + (t >= T(0.) ? h : -h) * a + rho; return dt; }
    on = c + t * a + r * rdir; return dr;
  }
  on = c + tc * a + tmin<T>(rn, r) * rdir;
  T ddt = tmax<T>(dt, T(0.)), ddr = tmax<T>(dr, T(0.));
  return tsqrt(ddt * ddt + ddr * ddr);
}
template <typename T> __host__ __device__ __forceinline__ ClosestT<T> cyl_sphere(V3T<T> c, V3T<T> a, T r, T h, V3T<T> o, T R) {
  V3T<T> on;
  T d = point_cylinder(o, c, a, r, h, on);
  V3T<T> v = o - on;
  T L = tnorm(v);
  if (d < T(0.)) v = v3t<T>(-v.x, -v.y, -v.z);
  V3T<T> n = L > T(1e-30) ? (T(1.0) / L) * v : v3t<T>(T(1), T(0), T(0));
  ClosestT<T> k; k.d = d - T(kHullMargin) - R; k.dir = n; return k;
}

// ------------------------------------------------------------------------------------ segment vs axis-aligned box, exact
// Closest points between the segment c + t a (|t| <= h, a unit) and the box |x_k| <= e_k centred at the origin.
// f(t) = dist^2(c + t a, box) is convex, C1 and piecewise quadratic, so g = f'/2 = a . (p - clamp(p)) is monotone and
// piecewise LINEAR with breakpoints where a coordinate of p(t) crosses +-e_k.  Shrink a bracket [tl, tr] (g(tl) <= 0 < g(tr))
// over the six breakpoints; no breakpoint is left inside it, so the root is one linear interpolation away.  No iteration,
// no tolerance: this is what lets the collision pass decide cylinder-box pairs without GJK (see cyl_box_bound).
template <typename T> struct SegBoxT { T s, t; V3T<T> pa, pb; };  // distance, parameter, point on the segment, point on the box
template <typename T> __host__ __device__ __forceinline__ V3T<T> box_clamp(V3T<T> p, V3T<T> e) {
  return v3t<T>(tmin<T>(tmax<T>(p.x, -e.x), e.x), tmin<T>(tmax<T>(p.y, -e.y), e.y), tmin<T>(tmax<T>(p.z, -e.z), e.z));
}
template <typename T> __host__ __device__ __forceinline__ T segbox_slope(V3T<T> c, V3T<T> a, V3T<T> e, T t) {
  V3T<T> p = c + t * a;
  return dot(a, p - box_clamp(p, e));
}
template <typename T> __host__ __device__ __forceinline__ SegBoxT<T> seg_box(V3T<T> c, V3T<T> a, T h, V3T<T> e) {
  T tl = -h, tr = h, gl = segbox_slope(c, a, e, tl), gr = segbox_slope(c, a, e, tr), t;
  if (gl >= T(0)) t = tl;
  else if (gr <= T(0)) t = tr;
  else {
    const T ck[3] = {c.x, c.y, c.z}, ak[3] = {a.x, a.y, a.z}, ek[3] = {e.x, e.y, e.z};
#pragma unroll
    for (int k = 0; k < 3; k++) {
      if (tabs<T>(ak[k]) > T(1e-12)) {
        const T inv = T(1) / ak[k];
#pragma unroll
        for (int sg = 0; sg < 2; sg++) {
          const T tb = ((sg ? ek[k] : -ek[k]) - ck[k]) * inv;
          if (tb > tl && tb < tr) {
            const T gb = segbox_slope(c, a, e, tb);
            if (gb <= T(0)) { tl = tb; gl = gb; } else { tr = tb; gr = gb; }
          }
        }
      }
    }
    const T den = gr - gl;
    t = den > T(0) ? tl - gl * (tr - tl) / den : tl;
  }
  SegBoxT<T> r;
  r.t = t; r.pa = c + t * a; r.pb = box_clamp(r.pa, e);
  r.s = tnorm(r.pa - r.pb);
  return r;
}

// Solid cylinder (centre c relative to the box centre, unit axis a, radius r, half length h) vs box core e: what can be said
// without iterating.  Every cylinder point is within r of its axis segment, so with s = dist(segment, box), u = unit(box
// point - segment point) and the plane through the box point normal to u separating the two bodies:
//     dist(cylinder, box) >= s - r sin(theta),  sin(theta) = sqrt(1 - (u.a)^2)                      (lb)
// with EQUALITY when u is perpendicular to the axis (the closest feature is the lateral surface: always the case when the
// segment's closest point is interior): `exact`.  Otherwise the closest feature is an end cap / rim; the distance from the
// box point to the solid cylinder bounds the pair from above (ub).  The link colliders are capsules (cylinder + two
// spheres of the same radius at its ends), so a slanted end contact reaches the sphere first and the band lb < thr <= ub
// is a sliver: only there does the caller fall back to GJK.
template <typename T> struct CylBoxBoundT { T lb, s; V3T<T> u, pb; bool exact; };
template <typename T> __host__ __device__ __forceinline__ CylBoxBoundT<T> cyl_box_bound(V3T<T> c, V3T<T> a, T r, T h, V3T<T> e) {
  SegBoxT<T> sb = seg_box<T>(c, a, h, e);
  CylBoxBoundT<T> k;
  k.s = sb.s; k.pb = sb.pb;
  if (sb.s > T(1e-9)) {
    k.u = (T(1) / sb.s) * (sb.pb - sb.pa);
    const T ua = dot(k.u, a);
    k.exact = tabs<T>(ua) < (sizeof(T) == 8 ? T(1e-9) : T(1e-4));  // r (1 - sin) = r ua^2 / 2: far below the arithmetic's resolution
    k.lb = tmax<T>(sb.s - r * tsqrt(tmax<T>(T(1) - ua * ua, T(0))), T(0));
  } else {  // the axis itself touches the core: cores overlap
    k.u = v3t<T>(T(0), T(0), T(0)); k.exact = true; k.lb = T(0);
  }
  return k;
}

// ------------------------------------------------------------------------------------ GJK: cylinder core vs box core
template <typename T> struct GjkIn { V3T<T> c, a, b, e; T r, h; };
template <typename T> __host__ __device__ __forceinline__ void gjk_support(const GjkIn<T>& g, V3T<T> d, V3T<T>& sa, V3T<T>& sb) {
  T da = dot(d, g.a);
  V3T<T> rad = d - da * g.a;
  T rn = tnorm(rad);
  sa = g.c + (da >= T(0.) ? g.h : -g.h) * g.a + (rn > T(1e-12) ? (g.r / rn) * rad : v3t<T>(T(0), T(0), T(0)));
  sb = g.b + v3t<T>(-d.x >= T(0.) ? g.e.x : -g.e.x, -d.y >= T(0.) ? g.e.y : -g.e.y, -d.z >= T(0.) ? g.e.z : -g.e.z);
}
// closest point to the origin on triangle (a, b, c): barycentric weights l, kept-vertex mask
template <typename T> __host__ __device__ __forceinline__ void tri_closest(V3T<T> a, V3T<T> b, V3T<T> c, T* l, int& mask) {
  V3T<T> ab = b - a, ac = c - a;
  T d1 = -dot(ab, a), d2 = -dot(ac, a);
  if (d1 <= T(0.) && d2 <= T(0.)) { l[0] = 1; l[1] = 0; l[2] = 0; mask = 1; return; }
  T d3 = -dot(ab, b), d4 = -dot(ac, b);
  if (d3 >= T(0.) && d4 <= d3) { l[0] = 0; l[1] = 1; l[2] = 0; mask = 2; return; }
  T vc = d1 * d4 - d3 * d2;
  if (vc <= T(0.) && d1 >= T(0.) && d3 <= T(0.)) { T v = d1 / (d1 - d3); l[0] = T(1) - v; l[1] = v; l[2] = 0; mask = 3; return; }
  T d5 = -dot(ab, c), d6 = -dot(ac, c);
  if (d6 >= T(0.) && d5 <= d6) { l[0] = 0; l[1] = 0; l[2] = 1; mask = 4; return; }
  T vb = d5 * d2 - d1 * d6;
  if (vb <= T(0.) && d2 >= T(0.) && d6 <= T(0.)) { T w = d2 / (d2 - d6); l[0] = T(1) - w; l[1] = 0; l[2] = w; mask = 5; return; }
  T va = d3 * d6 - d5 * d4;
  if (va <= T(0.) && (d4 - d3) >= T(0.) && (d5 - d6) >= T(0.)) { T w = (d4 - d3) / ((d4 - d3) + (d5 - d6)); l[0] = 0; l[1] = T(1) - w; l[2] = w; mask = 6; return; }
  T den = T(1.0) / (va + vb + vc);
  l[1] = vb * den; l[2] = vc * den; l[0] = T(1.) - l[1] - l[2]; mask = 7;
}

// distance between the two CORES (0 when they touch / overlap); qa / qb = closest points on the cores
// DECIDE: the caller only wants to know on which side of `thr` the distance lies.  GJK carries both bounds at every iteration --
// |v| from above (v is a point of the difference body), v.w / |v| from below (w supports it in direction -v) -- so it stops as
// soon as one of them settles the question and returns that bound (<= thr, or > thr; qa / qb are then not set).  Against a
// cylinder's curved surface fp64 GJK converges linearly and runs to its iteration cap for 1e-14; the reset's accept / reject
// tests need a handful of iterations.
template <typename T, bool DECIDE> __host__ __device__ __forceinline__ T gjk_cyl_box_impl(const GjkIn<T> g, V3T<T>* qa, V3T<T>* qb, T thr) {
  V3T<T> W[4], A[4], B[4];
  T lam[4] = {T(1), T(0), T(0), T(0)};
  int n = 0;
  V3T<T> v = g.c - g.b;
  if (dot(v, v) < (sizeof(T) == 8 ? T(1e-24) : T(1e-12))) v = v3t<T>(T(1), T(0), T(0));
  for (int it = 0; it < (sizeof(T) == 8 ? 100 : 20); it++) {
    V3T<T> sa, sb;
    gjk_support(g, v3t<T>(-v.x, -v.y, -v.z), sa, sb);
    V3T<T> w = sa - sb;
    T vv = dot(v, v);
    if (DECIDE && n > 0) {
      if (vv <= thr * thr) return tsqrt(vv);
      const T vw = dot(v, w);
      if (vw > T(0) && vw * vw > thr * thr * vv) return vw / tsqrt(vv);
    }
    if (n > 0 && vv - dot(v, w) <= (sizeof(T) == 8 ? T(1e-14) : T(1e-5)) * vv + (sizeof(T) == 8 ? T(1e-22) : T(1e-12))) break;
    W[n] = w; A[n] = sa; B[n] = sb; n++;
    if (n == 1) { lam[0] = 1; }
    else if (n == 2) {
      V3T<T> ab = W[1] - W[0];
      T t = -dot(W[0], ab), den = dot(ab, ab);
      if (t <= T(0.)) { n = 1; lam[0] = 1; }
      else if (t >= den) { W[0] = W[1]; A[0] = A[1]; B[0] = B[1]; n = 1; lam[0] = 1; }
      else { lam[1] = t / den; lam[0] = T(1.) - lam[1]; }
    } else if (n == 3) {
      T l[3]; int mask;
      tri_closest(W[0], W[1], W[2], l, mask);
      int m = 0;
      for (int k = 0; k < 3; k++) if (mask & (1 << k)) { W[m] = W[k]; A[m] = A[k]; B[m] = B[k]; lam[m] = l[k]; m++; }
      n = m;
    } else {
      T best = T(3.0e38), bl[3] = {T(1), T(0), T(0)}; int bmask = 0, bf = -1; bool inside = true;
      for (int f = 0; f < 4; f++) {
        const int i0 = f == 3 ? 1 : 0, i1 = f == 0 ? 1 : (f == 3 ? 2 : (f == 1 ? 1 : 2)), i2 = f == 0 ? 2 : 3, io = 3 - f;
        // faces: f0 = (0,1,2) opp 3; f1 = (0,1,3) opp 2; f2 = (0,2,3) opp 1; f3 = (1,2,3) opp 0
        V3T<T> p0 = W[i0], p1 = W[i1], p2 = W[i2], po = W[io];
        V3T<T> nn = cross(p1 - p0, p2 - p0);
        V3T<T> t0 = po - p0;
        T so = dot(nn, t0), sp = -dot(nn, p0);
        // inside only if strictly on the opposite-vertex side of every face of a non-degenerate tetrahedron
        if (!(so * sp > T(0)) || tabs<T>(so) <= (sizeof(T) == 8 ? T(1e-10) : T(1e-5)) * tsqrt(dot(nn, nn) * dot(t0, t0))) inside = false;
        T l[3]; int mask;
        tri_closest(p0, p1, p2, l, mask);
        V3T<T> q = l[0] * p0 + l[1] * p1 + l[2] * p2;
        T d2 = dot(q, q);
        if (d2 < best) { best = d2; bf = f; bmask = mask; bl[0] = l[0]; bl[1] = l[1]; bl[2] = l[2]; }
      }
      if (inside) { *qa = g.c; *qb = g.c; return T(0); }
      const int i0 = bf == 3 ? 1 : 0, i1 = bf == 0 ? 1 : (bf == 3 ? 2 : (bf == 1 ? 1 : 2)), i2 = bf == 0 ? 2 : 3;
      V3T<T> TW[3] = {W[i0], W[i1], W[i2]}, TA[3] = {A[i0], A[i1], A[i2]}, TB[3] = {B[i0], B[i1], B[i2]};
      int m = 0;
      for (int k = 0; k < 3; k++) if (bmask & (1 << k)) { W[m] = TW[k]; A[m] = TA[k]; B[m] = TB[k]; lam[m] = bl[k]; m++; }
      n = m;
    }
    v = v3t<T>(T(0), T(0), T(0));
    for (int k = 0; k < n; k++) v = v + lam[k] * W[k];
    if (dot(v, v) < (sizeof(T) == 8 ? T(1e-24) : T(1e-14))) {
      V3T<T> p = v3t<T>(T(0), T(0), T(0));
      for (int k = 0; k < n; k++) p = p + lam[k] * A[k];
      *qa = p; *qb = p; return T(0.);
    }
  }
  V3T<T> pa = v3t<T>(T(0), T(0), T(0)), pb = v3t<T>(T(0), T(0), T(0));
  for (int k = 0; k < n; k++) { pa = pa + lam[k] * A[k]; pb = pb + lam[k] * B[k]; }
  *qa = pa; *qb = pb;
  return tnorm(v);
}
template <typename T> __host__ __device__ __noinline__ T gjk_cyl_box(const GjkIn<T> g, V3T<T>* qa, V3T<T>* qb) {
  return gjk_cyl_box_impl<T, false>(g, qa, qb, T(0));
}
template <typename T> __host__ __device__ __noinline__ T gjk_cyl_box_decide(const GjkIn<T> g, T thr) {
  V3T<T> qa, qb;
  return gjk_cyl_box_impl<T, true>(g, &qa, &qb, thr);
}
// cyl_box(...).d for a caller that only compares it with `cutoff`: a value on the same side of `cutoff` as the distance
template <typename T> __host__ __device__ __forceinline__ T cyl_box_decide(V3T<T> c, V3T<T> a, T r, T h, V3T<T> b, V3T<T> e, T m, T cutoff) {
  GjkIn<T> g; g.c = c; g.a = a; g.b = b; g.e = e; g.r = r; g.h = h;
  return gjk_cyl_box_decide<T>(g, cutoff + T(kHullMargin) + m) - T(kHullMargin) - m;
}

template <typename T> __host__ __device__ __forceinline__ ClosestT<T> cyl_box(V3T<T> c, V3T<T> a, T r, T h, V3T<T> b, V3T<T> e, T m) {
  GjkIn<T> g; g.c = c; g.a = a; g.b = b; g.e = e; g.r = r; g.h = h;
  V3T<T> qa, qb;
  T d = gjk_cyl_box(g, &qa, &qb);
  if (sizeof(T) == 4 && d > T(0)) {
    // GJK's stop rule bounds the DISTANCE error by eps but the closest-point / direction error only by
    // sqrt(eps) (~1e-3 in fp32): polish with alternating exact projections (box clamp <-> solid cylinder).
#pragma unroll
    for (int it = 0; it < 4; it++) {
      V3T<T> rel = qa - b;
      qb = b + v3t<T>(tmin<T>(tmax<T>(rel.x, -e.x), e.x), tmin<T>(tmax<T>(rel.y, -e.y), e.y), tmin<T>(tmax<T>(rel.z, -e.z), e.z));
      T dd = point_cylinder<T>(qb, c, a, r, h, qa);
      d = dd > T(0) ? dd : T(0);
    }
  }
  V3T<T> v = qb - qa;
  T L = tnorm(v);
  V3T<T> n = L > T(1e-30) ? (T(1.0) / L) * v : v3t<T>(T(0), T(0), T(0));
  ClosestT<T> k; k.d = d - T(kHullMargin) - m; k.dir = n; return k;
}

// cylinder vs a LARGE box whose top face is the only reachable feature (the table): exact when every point of the
// cylinder projects inside the core's top rectangle; otherwise falls back to GJK.
__device__ __forceinline__ float cyl_table_distance(V3 c, V3 a, float r, float h, V3 b, V3 e, float m) {
  float bound = sqrtf(r * r + h * h);
  float top = b.z + e.z;
  // only the SIGN of a table distance is ever consumed (collision flag): a cylinder whose bounding sphere clears the
  // plane of the table top cannot touch the box, wherever it is in x / y
  float clear = c.z - bound - top - m - kHullMargin;
  if (clear > 0.f) return clear;
  if (fabsf(c.x - b.x) + bound <= e.x && fabsf(c.y - b.y) + bound <= e.y && c.z >= top) {
    float ext = fabsf(a.z) * h + r * sqrtf(fmaxf(1.f - a.z * a.z, 0.f));
    return (c.z - ext) - top - m - kHullMargin;
  }
  return cyl_box<float>(c, a, r, h, b, e, m).d;
}

}  // namespace pg
