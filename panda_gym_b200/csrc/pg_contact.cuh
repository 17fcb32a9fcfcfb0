// pg_contact.cuh -- one Bullet stepSimulation() of the Push / PickAndPlace world: the 7-DoF arm, one free rigid box
// ("object", push.py:33-39, pick_and_place.py:38-44) and the static table / plane (pybullet.py:759-815), coupled by
// frictional contacts.  One environment per thread, fp32.  Replaces pybullet.stepSimulation behind
// panda_gym/pybullet.py:68-71 for these tasks; same formulation as oracle/pgo_world.c:
//   free body   = single-body ABA (gravity, btMultiBody damping 0.04, gyroscopic term), world-frame velocities,
//                 semi-implicit Euler, exponential-map orientation update (btMultiBody::stepPositionsMultiDof);
//   contacts    = rebuilt every substep from closed-form features, <= 4 closest per body pair (object-table,
//                 robot-object, robot-table), alive below the contact breaking threshold 0.02;
//   rows        = btMultiBodyConstraintSolver::setupMultiBodyContactConstraint: normal row (speculative when
//                 d > 0, erp2 = 0.2 when penetrating) + two btPlaneSpace1 friction rows bounded by mu * normal impulse;
//   solver      = [joint limits, motors] alternating sweep, then normals, then frictions; 50 iterations, 1e-7 early-out.
//
// Where the state lives (this is what makes the kernel fit the machine).  Every robot collider that can touch the object or
// the table hangs off link 7, so ALL arm-side contact rows act on one rigid body: a row along direction t at point p is the
// 6-vector w = [t ; (p - o7) x t] in the spatial velocity space of link 7, V7 = S qd with S (6 x 7) the Jacobian of the link-7
// origin o7.  Its joint-space Jacobian is S^T w and its response M^-1 S^T w = X w (X = M^-1 S^T, 7 x 6, built once per substep),
// and the solver tracks the arm in BOTH spaces: motor rows update qd (a column of M^-1) and hand their impulses to the task
// space once per sweep (V7 += X^T dsum); contact rows read and update V7 through G = S X (6 x 6, registers) and collect their
// impulses in a wrench F7 that is applied to qd once per sweep (qd += X F7).  Same mathematics as a row of 7 + 7 joint-space numbers per contact direction, but a contact keeps only its
// normal, its two points and, per direction, four scalars and the 6-vector G w -- 23 / 41 floats, in SHARED memory
// (structure-of-arrays over the block's threads) -- and the 36 x 26-float row table in thread-local memory of the first version is gone (it spilled past
// L2: 611 MB of DRAM writes per launch, profiles/r01_k_step_obj_summary.md).
#pragma once
#include "pg_geom.cuh"

namespace pg {

constexpr int kPairContacts = 4;               // size of a btPersistentManifold
constexpr int kMaxContacts = 3 * kPairContacts;
constexpr float kBreaking = 0.02f;             // gContactBreakingThreshold
constexpr float kErp2 = 0.2f;                  // contactERP

struct WorldParams {
  float half[3], mass;                         // the object box
  float table_pos[3], table_half[3], plane_top;
  float mu_object, mu_static, mu_link, mu_ee;  // lateral friction (Bullet default 0.5; panda.py:69-70: link 9 = 1.0)
};

struct ObjState { V3 x; float qx, qy, qz, qw; V3 v, w; };  // quaternion xyzw body -> world, world-frame velocities

struct Rot { V3 r0, r1, r2; };  // rows of the body -> world rotation
__device__ __forceinline__ Rot quat_rot(float x, float y, float z, float w) {
  Rot R;
  R.r0 = v3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - z * w), 2.f * (x * z + y * w));
  R.r1 = v3(2.f * (x * y + z * w), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - x * w));
  R.r2 = v3(2.f * (x * z - y * w), 2.f * (y * z + x * w), 1.f - 2.f * (x * x + y * y));
  return R;
}
__device__ __forceinline__ V3 rot_apply(const Rot& R, V3 l) { return v3(dot(R.r0, l), dot(R.r1, l), dot(R.r2, l)); }          // body -> world
__device__ __forceinline__ V3 rot_applyT(const Rot& R, V3 g) { return g.x * R.r0 + g.y * R.r1 + g.z * R.r2; }                   // world -> body


// ---------------------------------------------------------------------------------------------- contact store
// Contacts live in shared memory, structure-of-arrays over the threads of the block: field k of contact slot c of this
// thread is base[(c * kCF + k) * stride] (consecutive threads -> consecutive banks).  Slots: pair 0 (object on table /
// plane) = 0..3, pair 1 (robot on object) = 4..7, pair 2 (robot on table top) = 8..11; each pair keeps its 4 closest,
// sorted by distance.  n = unit normal from B towards A; pr = contact point on the robot, po = on the object.
// fields of a slot: n 0-2 | pr 3-5 | po 6-8 | d 9 | mu 10 | per direction (normal, t1, t2): rhs 11-13 | dinv 14-16 | applied impulse
// 17-19 | denom 20-22 | and, for the pairs with an arm side (slots 4..11), G w of each direction 23-40
constexpr int kCF0 = 23, kCF1 = 41;
constexpr int kCGeom = 11;
constexpr int kCStoreFloats = kPairContacts * kCF0 + 2 * kPairContacts * kCF1;
constexpr int kXG = 42;  // X = M^-1 S^T (7 x 6, row-major) in shared memory; G = S X lives in registers during the row setup
struct CStore {
  float* base; int stride;
  int cf = 0;  // a view(c) (one slot seen as slot 0): that slot's field count, for PG_CHECK only
  __device__ __forceinline__ float& at(int c, int k) const {
    PG_CHECK(cf ? (c == 0 && k >= 0 && k < cf) : (c >= 0 && c < kMaxContacts && k >= 0 && k < (c < kPairContacts ? kCF0 : kCF1)));
    const int off = c < kPairContacts ? c * kCF0 : kPairContacts * kCF0 + (c - kPairContacts) * kCF1;
    return base[(off + k) * stride];
  }
  // slot c as slot 0 of a shifted store: one address computation per contact, immediate offsets after it
  __device__ __forceinline__ CStore view(int c) const {
    PG_CHECK(c >= 0 && c < kMaxContacts);
    CStore V; V.stride = stride; V.cf = c < kPairContacts ? kCF0 : kCF1;
    V.base = base + (c < kPairContacts ? c * kCF0 : kPairContacts * kCF0 + (c - kPairContacts) * kCF1) * stride;
    return V;
  }
  __device__ __forceinline__ V3 vec(int c, int k) const { return v3(at(c, k), at(c, k + 1), at(c, k + 2)); }
  __device__ __forceinline__ void put(int c, int k, V3 v) const { at(c, k) = v.x; at(c, k + 1) = v.y; at(c, k + 2) = v.z; }
};

__device__ __forceinline__ void keep_closest(const CStore& S, int pair, int& n, float d, float mu, V3 nrm, V3 pr, V3 po) {  // sorted insert, stable, capacity 4
  const int c0 = pair * kPairContacts;
  int pos = n;
  while (pos > 0 && d < S.at(c0 + pos - 1, 9)) pos--;
  if (pos >= kPairContacts) return;
  const int last = n < kPairContacts ? n : kPairContacts - 1;
  for (int i = last; i > pos; i--)
    for (int k = 0; k < kCGeom; k++) S.at(c0 + i, k) = S.at(c0 + i - 1, k);
  S.put(c0 + pos, 0, nrm); S.put(c0 + pos, 3, pr); S.put(c0 + pos, 6, po);
  S.at(c0 + pos, 9) = d; S.at(c0 + pos, 10) = mu;
  if (n < kPairContacts) n++;
}

__device__ __forceinline__ bool over_rect(V3 p, const float* pos, const float* half) {
  return fabsf(p.x - pos[0]) <= half[0] && fabsf(p.y - pos[1]) <= half[1];
}

// btPlaneSpace1
__device__ __forceinline__ void plane_space(V3 n, V3& p, V3& q) {
  if (fabsf(n.z) > 0.70710678f) {
    float a = n.y * n.y + n.z * n.z, k = rsqrtf(a);
    p = v3(0.f, -n.z * k, n.y * k);
    q = v3(a * k, -n.x * p.z, n.x * p.y);
  } else {
    float a = n.x * n.x + n.y * n.y, k = rsqrtf(a);
    p = v3(-n.y * k, n.x * k, 0.f);
    q = v3(-n.z * p.y, n.z * p.x, a * k);
  }
}

// robot primitive K (link7 frame F) vs the object box (centre / rotation / half extents) and vs the table top
template <int K>
__device__ __forceinline__ void prim_contacts(const Frame& F, const WorldParams& W, V3 bx, const Rot& R, const CStore& S, int& nr, int& ns) {
  using Pm = pgm::Prim<K>;
  const float mu_l = Pm::slot == 8 ? W.mu_ee : W.mu_link;  // slot 8 = panda_ee
  V3 c = frame_point(F, Pm::x, Pm::y, Pm::z);
  V3 a = frame_dir(F, Pm::ax, Pm::ay, Pm::az);
  // --- vs object, evaluated in the box frame
  {
    V3 cl = rot_applyT(R, c - bx);
    const float m = fminf(0.04f, 0.1f * fminf(W.half[0], fminf(W.half[1], W.half[2])));  // btBoxShape::setSafeMargin
    V3 core = v3(W.half[0] - m, W.half[1] - m, W.half[2] - m);
    V3 o = v3(0.f, 0.f, 0.f);
    const float bound = Pm::r + (Pm::type == 1 ? Pm::h : 0.f) + 0.001f;
    if (sphere_box<float>(cl, bound, o, core, m).d < kBreaking) {  // bounding-sphere reject
      const float mu = mu_l * W.mu_object;
      float d;
      V3 dl, pal;
      bool done = false;
      if (Pm::type == 0) {
        Closest k = sphere_box<float>(cl, Pm::r, o, core, m);
        d = k.d; dl = k.dir; pal = cl + Pm::r * dl;
      } else {
        V3 al = rot_applyT(R, a);
        GjkIn<float> g; g.c = cl; g.a = al; g.b = o; g.e = core; g.r = Pm::r; g.h = Pm::h;
        V3 qa, qb;
        float dc = gjk_cyl_box(g, &qa, &qb);
        if (dc > 0.f) {
#pragma unroll
          for (int it = 0; it < 4; it++) {  // polish (see cyl_box)
            qb = v3(fminf(fmaxf(qa.x, -core.x), core.x), fminf(fmaxf(qa.y, -core.y), core.y), fminf(fmaxf(qa.z, -core.z), core.z));
            float dd = point_cylinder<float>(qb, cl, al, Pm::r, Pm::h, qa);
            dc = dd > 0.f ? dd : 0.f;
          }
        }
        V3 v = qb - qa;
        float L = norm(v);
        d = dc - kHullMargin - m;
        if (L > 1e-9f) { dl = (1.0f / L) * v; pal = qa + kHullMargin * dl; }
        else { float cn = norm(cl); dl = cn > 1e-12f ? (-1.0f / cn) * cl : v3(0.f, 0.f, -1.f); pal = qa; }  // deep overlap: line of centres
        // cylinder against a box FACE: the closest point of a generator parallel to the face is not unique (rounding
        // noise would pick it), so the contact is spread like a manifold: both ends of the generator facing the box,
        // clipped to the face, each with its own distance
        if (d < kBreaking) {
          const float dls[3] = {dl.x, dl.y, dl.z}, als[3] = {al.x, al.y, al.z}, cls[3] = {cl.x, cl.y, cl.z};
          const int kf = fabsf(dls[0]) > fabsf(dls[1]) ? (fabsf(dls[0]) > fabsf(dls[2]) ? 0 : 2) : (fabsf(dls[1]) > fabsf(dls[2]) ? 1 : 2);
          if (fabsf(dls[kf]) > 0.999f) {
            const float sg = dls[kf] > 0.f ? 1.f : -1.f, rr = Pm::r + kHullMargin;
            float u[3] = {0.f, 0.f, 0.f};
            u[kf] = sg;
            const float ua = u[0] * als[0] + u[1] * als[1] + u[2] * als[2];
            for (int i = 0; i < 3; i++) u[i] -= ua * als[i];
            const float Lu = sqrtf(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
            if (Lu > 1e-6f) {
              float g0[3], t0 = -Pm::h, t1 = Pm::h;
              for (int i = 0; i < 3; i++) g0[i] = cls[i] + rr * u[i] / Lu;
              for (int j = 0; j < 3; j++) {
                if (j == kf) continue;
                if (fabsf(als[j]) < 1e-9f) { if (fabsf(g0[j]) > W.half[j]) t1 = t0 - 1.f; continue; }
                float ta = (-W.half[j] - g0[j]) / als[j], tb = (W.half[j] - g0[j]) / als[j];
                if (ta > tb) { float t = ta; ta = tb; tb = t; }
                t0 = fmaxf(t0, ta); t1 = fminf(t1, tb);
              }
              if (t1 >= t0) {
                done = true;
                for (int end = 0; end < 2; end++) {
                  if (end == 1 && t1 - t0 < 1e-6f) break;
                  const float t = end == 0 ? t0 : t1;
                  float gl[3], nl[3] = {0.f, 0.f, 0.f};
                  for (int i = 0; i < 3; i++) gl[i] = g0[i] + t * als[i];
                  nl[kf] = -sg;
                  const float de = -sg * gl[kf] - W.half[kf];
                  V3 pa = bx + rot_apply(R, v3(gl[0], gl[1], gl[2]));
                  V3 nw = rot_apply(R, v3(nl[0], nl[1], nl[2]));
                  if (de < kBreaking) keep_closest(S, 1, nr, de, mu, nw, pa, pa - de * nw);
                }
              }
            }
          }
        }
      }
      if (!done && d < kBreaking) {
        V3 dw = rot_apply(R, dl);
        V3 pa = bx + rot_apply(R, pal);
        keep_closest(S, 1, nr, d, mu, v3(-dw.x, -dw.y, -dw.z), pa, pa + d * dw);
      }
    }
  }
  // --- vs the table top: lowest point(s) of the primitive; a cylinder contributes both ends of its lowest generator
  {
    const float top = W.table_pos[2] + W.table_half[2];
    const float mu = mu_l * W.mu_static;
#pragma unroll
    for (int end = 0; end < (Pm::type == 0 ? 1 : 2); end++) {
      V3 pa;
      if (Pm::type == 0) pa = v3(c.x, c.y, c.z - Pm::r);
      else {
        const float r = Pm::r + kHullMargin;
        const float t = end == 0 ? -Pm::h : Pm::h;
        V3 u = v3(-a.z * a.x, -a.z * a.y, 1.f - a.z * a.z);
        float L = norm(u);
        pa = c + t * a;
        if (L > 1e-6f) pa = pa - (r / L) * u;
      }
      if (!over_rect(pa, W.table_pos, W.table_half)) continue;
      const float d = pa.z - top;
      if (d < kBreaking) keep_closest(S, 2, ns, d, mu, v3(0.f, 0.f, 1.f), pa, v3(pa.x, pa.y, top));
    }
  }
}

// contact set at the current poses into the store; counts per pair: object-table, robot-object, robot-table
static __device__ __noinline__ void world_contacts(const WorldParams& W, const Frame& F7, const ObjState& B, const CStore S, int& nt, int& nr, int& ns) {
  nt = 0; nr = 0; ns = 0;
  Rot R = quat_rot(B.qx, B.qy, B.qz, B.qw);
  const float top = W.table_pos[2] + W.table_half[2];
#pragma unroll 1
  for (int v = 0; v < 8; v++) {
    V3 l = v3((v & 1) ? W.half[0] : -W.half[0], (v & 2) ? W.half[1] : -W.half[1], (v & 4) ? W.half[2] : -W.half[2]);
    V3 p = B.x + rot_apply(R, l);
    float z = over_rect(p, W.table_pos, W.table_half) ? top : W.plane_top;
    const float d = p.z - z;
    if (!(d < kBreaking)) continue;
    keep_closest(S, 0, nt, d, W.mu_object * W.mu_static, v3(0.f, 0.f, 1.f), p, p);
  }
  // every robot collider is farther from both the object and the table top than the breaking threshold: nothing to generate
  // (the common case -- every collider on link 7 lies within 0.176 m of the point (0, 0, 0.06) of its frame: lowest the
  // r = 0.04 sphere at z = -0.06, highest the finger-bar spheres at z = 0.207, 0.05 off the axis)
  {
    V3 hc = frame_point(F7, 0.f, 0.f, 0.06f);
    const float reach = 0.18f + kBreaking;
    V3 d = B.x - hc;
    const float br = sqrtf(W.half[0] * W.half[0] + W.half[1] * W.half[1] + W.half[2] * W.half[2]);
    if (hc.z - reach > top && dot(d, d) > (reach + br) * (reach + br)) return;
  }
  // model-table order of the oracle: link7 (21-23), link8 (24-26), hand (30-32), ee (27-29)
  prim_contacts<21>(F7, W, B.x, R, S, nr, ns); prim_contacts<22>(F7, W, B.x, R, S, nr, ns); prim_contacts<23>(F7, W, B.x, R, S, nr, ns);
  prim_contacts<24>(F7, W, B.x, R, S, nr, ns); prim_contacts<25>(F7, W, B.x, R, S, nr, ns); prim_contacts<26>(F7, W, B.x, R, S, nr, ns);
  prim_contacts<30>(F7, W, B.x, R, S, nr, ns); prim_contacts<31>(F7, W, B.x, R, S, nr, ns); prim_contacts<32>(F7, W, B.x, R, S, nr, ns);
  prim_contacts<27>(F7, W, B.x, R, S, nr, ns); prim_contacts<28>(F7, W, B.x, R, S, nr, ns); prim_contacts<29>(F7, W, B.x, R, S, nr, ns);
}

// ------------------------------------------------------------------------------------------------ solver rows
struct ArmKin { V3 ax[7], org[7]; Frame F7; };  // world joint axes / origins and the link7 frame at the current q
__device__ __forceinline__ void arm_kin(V3 base, const float* cq, const float* sq, ArmKin& K) {
  Frame F = base_frame(base);
  F = frame_step<1>(F, cq[0], sq[0]); K.ax[0] = F.z; K.org[0] = F.o;
  F = frame_step<2>(F, cq[1], sq[1]); K.ax[1] = F.z; K.org[1] = F.o;
  F = frame_step<3>(F, cq[2], sq[2]); K.ax[2] = F.z; K.org[2] = F.o;
  F = frame_step<4>(F, cq[3], sq[3]); K.ax[3] = F.z; K.org[3] = F.o;
  F = frame_step<5>(F, cq[4], sq[4]); K.ax[4] = F.z; K.org[4] = F.o;
  F = frame_step<6>(F, cq[5], sq[5]); K.ax[5] = F.z; K.org[5] = F.o;
  F = frame_step<7>(F, cq[6], sq[6]); K.ax[6] = F.z; K.org[6] = F.o;
  K.F7 = F;
}

__device__ __forceinline__ V3 inv_inertia_apply(const Rot& R, V3 iinv, V3 x) {  // R diag(iinv) R^T x
  V3 l = rot_applyT(R, x);
  return rot_apply(R, v3(l.x * iinv.x, l.y * iinv.y, l.z * iinv.z));
}

__host__ __device__ constexpr int sym6(int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; }

// ---- rows.  Contacts against the table top / the plane (pairs 0 and 2) have the normal +z and therefore the fixed
// btPlaneSpace1 tangents t1 = -y, t2 = +x: their row directions are signed unit axes and every product with them is a
// component pick (AXIAL = true, folded at compile time).  Robot-object contacts (pair 1) carry a general normal.
template <int AX> __device__ __forceinline__ float ax_dot(V3 v) { return AX == 0 ? v.z : (AX == 1 ? -v.y : v.x); }          // t . v
template <int AX> __device__ __forceinline__ V3 ax_lever(V3 r) {                                                             // r x t
  return AX == 0 ? v3(r.y, -r.x, 0.f) : (AX == 1 ? v3(r.z, 0.f, -r.x) : v3(0.f, r.z, -r.y));
}
template <int AX> __device__ __forceinline__ void ax_add(V3& v, float s) { if (AX == 0) v.z += s; else if (AX == 1) v.y -= s; else v.x += s; }  // v += s t
template <int AX> __device__ __forceinline__ void ax_add6(float* F, float s) { if (AX == 0) F[2] += s; else if (AX == 1) F[1] -= s; else F[0] += s; }
template <int AX> __device__ __forceinline__ float ax_pick6(const float* V) { return AX == 0 ? V[2] : (AX == 1 ? -V[1] : V[0]); }
// column of the packed symmetric G picked by the axis: (G [t; 0])_k
template <int AX> __device__ __forceinline__ float ax_gcol(const float* G, int k) {
  return AX == 0 ? G[sym6(k, 2)] : (AX == 1 ? -G[sym6(k, 1)] : G[sym6(k, 0)]);
}

// the solver's view of one substep: everything a contact row needs besides the contact itself
struct RowCtx {
  V3 o7, bx, bv, bw;       // link-7 origin; object position and predicted velocities
  Rot R; V3 iinv;          // object rotation, inverse principal inertia
  float inv_mass, mass;
  bool iso;                // the three principal inertias are equal (a cube): R diag(i) R^T = i
};
__device__ __forceinline__ V3 ctx_iinv(const RowCtx& C, V3 x) { return C.iso ? C.iinv.x * x : inv_inertia_apply(C.R, C.iinv, x); }

// direction `axis` (0 normal, 1 / 2 the btPlaneSpace1 tangents) of a contact with normal n
__device__ __forceinline__ V3 contact_axis(V3 n, int axis) {
  if (axis == 0) return n;
  V3 t1, t2;
  plane_space(n, t1, t2);
  return axis == 1 ? t1 : t2;
}

// Geometry of row (c, AX) of a pair-KIND contact: t = direction; ma = (pr - o7) x t, the arm-side lever (KIND != 0);
// jw = s (po - bx) x t with s = +1 (object on table) / -1 (robot on object), the object-side lever (KIND != 2); jl = s t.
template <int KIND, int AX>
struct RowGeom {
  V3 t, ma, jw;
  __device__ __forceinline__ RowGeom(const CStore& S, const RowCtx& C, int c) {
    if (KIND == 1) {
      t = contact_axis(S.vec(c, 0), AX);
      ma = cross(S.vec(c, 3) - C.o7, t);
      jw = cross(t, S.vec(c, 6) - C.bx);  // -(rp x t)
    } else {
      t = v3(0.f, 0.f, 0.f);
      if (KIND == 2) ma = ax_lever<AX>(S.vec(c, 3) - C.o7);
      if (KIND == 0) jw = ax_lever<AX>(S.vec(c, 6) - C.bx);
    }
  }
  // w . V for the arm's 6-vector w = [t; ma]
  __device__ __forceinline__ float arm_dot(const float* V) const {
    const float lin = KIND == 1 ? t.x * V[0] + t.y * V[1] + t.z * V[2] : ax_pick6<AX>(V);
    return lin + ma.x * V[3] + ma.y * V[4] + ma.z * V[5];
  }
  // G w (G packed symmetric, in registers): computed once per row at setup and kept in the contact store
  __device__ __forceinline__ void arm_gw(const float* G, float* gw) const {
#pragma unroll
    for (int k = 0; k < 6; k++) {
      const float lin = KIND == 1 ? G[sym6(k, 0)] * t.x + G[sym6(k, 1)] * t.y + G[sym6(k, 2)] * t.z : ax_gcol<AX>(G, k);
      gw[k] = lin + G[sym6(k, 3)] * ma.x + G[sym6(k, 4)] * ma.y + G[sym6(k, 5)] * ma.z;
    }
  }
  __device__ __forceinline__ void arm_acc(float* F, float s) const {  // F += s w
    if (KIND == 1) { F[0] = fmaf(t.x, s, F[0]); F[1] = fmaf(t.y, s, F[1]); F[2] = fmaf(t.z, s, F[2]); } else ax_add6<AX>(F, s);
    F[3] = fmaf(ma.x, s, F[3]); F[4] = fmaf(ma.y, s, F[4]); F[5] = fmaf(ma.z, s, F[5]);
  }
  __device__ __forceinline__ float body_dot(V3 vl, V3 vw) const {  // jl . vl + jw . vw
    const float lin = KIND == 1 ? -dot(t, vl) : ax_dot<AX>(vl);
    return lin + dot(jw, vw);
  }
};

// rhs / dinv of row (c, AX); V7p = predicted spatial velocity of link 7, G = S M^-1 S^T
template <int KIND, int AX>
__device__ __forceinline__ void setup_row(const CStore& S, const RowCtx& C, const float* G, const float* V7p, int c) {
  const RowGeom<KIND, AX> g(S, C, c);
  float denom = 0.f, rel = 0.f;
  if (KIND != 0) {
    float gw[6];
    g.arm_gw(G, gw);
    denom = g.arm_dot(gw);
    rel = g.arm_dot(V7p);
#pragma unroll
    for (int k = 0; k < 6; k++) S.at(c, 23 + 6 * AX + k) = gw[k];
  }
  if (KIND != 2) {
    const V3 rw = ctx_iinv(C, g.jw);
    denom += (KIND == 1 ? dot(g.t, g.t) : 1.0f) / C.mass + dot(g.jw, rw);
    rel += g.body_dot(C.bv, C.bw);
  }
  const float dinv = denom > 1.1920929e-07f ? 1.0f / denom : 0.f;
  float rhs;
  if (AX != 0) rhs = -rel * dinv;
  else {
    const float d = S.at(c, 9);
    float pos_err = 0.f, vel_err = -rel;
    if (d > 0.f) vel_err -= d / kDt; else pos_err = -d * kErp2 / kDt;
    rhs = (pos_err + vel_err) * dinv;
  }
  S.at(c, 11 + AX) = rhs; S.at(c, 14 + AX) = dinv; S.at(c, 17 + AX) = 0.f; S.at(c, 20 + AX) = dinv != 0.f ? denom : 0.f;
}
template <int KIND>
__device__ __forceinline__ void setup_contact(const CStore& S, const RowCtx& C, const float* G, const float* V7p, int c) {
  setup_row<KIND, 0>(S, C, G, V7p, c); setup_row<KIND, 1>(S, C, G, V7p, c); setup_row<KIND, 2>(S, C, G, V7p, c);
}

// One projected Gauss-Seidel update of row (c, AX) with impulse bounds [lo, hi].  V7 = spatial velocity change of link 7
// so far, F7 = contact wrench collected in this sweep (the caller applies it to the joint velocities), dvl / dvw = velocity
// change of the object.  Returns the impulse change; the row's residual (Bullet's early-out test) is delta / dinv.
template <int KIND, int AX>
__device__ __forceinline__ float solve_row(const CStore& S, const RowCtx& C, int c, float lo, float hi, float* V7, float* F7, V3& dvl, V3& dvw,
                                           float& res) {
  const RowGeom<KIND, AX> g(S, C, c);
  float jv = 0.f;
  if (KIND != 0) jv = g.arm_dot(V7);
  if (KIND != 2) jv += g.body_dot(dvl, dvw);
  const float applied = S.at(c, 17 + AX);
  const float now = fminf(fmaxf(applied + (S.at(c, 11 + AX) - jv * S.at(c, 14 + AX)), lo), hi);  // lo <= hi always
  const float delta = now - applied;
  S.at(c, 17 + AX) = now;
  if (KIND != 0) {
    g.arm_acc(F7, delta);
#pragma unroll
    for (int k = 0; k < 6; k++) V7[k] = fmaf(S.at(c, 23 + 6 * AX + k), delta, V7[k]);
  }
  if (KIND != 2) {
    const float s = delta * C.inv_mass;
    if (KIND == 1) dvl = axpy<float>(-s, g.t, dvl); else ax_add<AX>(dvl, s);
    dvw = axpy<float>(delta, ctx_iinv(C, g.jw), dvw);
  }
  res = fmaxf(res, fabsf(delta) * S.at(c, 20 + AX));  // |delta / dinv|
  return delta;
}

// The sweeps stay ROLLED (one copy of each row body per pair kind): the solver loop has to sit in the 32 KB instruction
// cache the SM's four warps share, see profiles/r02_k_step_obj.md.
#ifndef PG_OBJ_UNROLL
#define PG_OBJ_UNROLL 1
#endif
constexpr int kObjUnroll = PG_OBJ_UNROLL;
// normal rows of one pair, in contact order; `act` collects how far the object-table island still moves (pair 0)
template <int KIND>
__device__ __forceinline__ void sweep_normals(const CStore& S, const RowCtx& C, int cnt, float* V7, float* F7, V3& dvl, V3& dvw, float& res,
                                              float& act) {
#pragma unroll kObjUnroll
  for (int i = 0; i < kPairContacts; i++) {
    if (i >= cnt) break;
    const CStore Sc = S.view(KIND * kPairContacts + i);
    const float delta = solve_row<KIND, 0>(Sc, C, 0, 0.f, 1e10f, V7, F7, dvl, dvw, res);
    if (KIND == 0) act = fmaxf(act, fabsf(delta) - 2.4e-7f * fabsf(Sc.at(0, 17)));
  }
}
// friction rows of one pair, bounded by mu x the normal impulse
template <int KIND>
__device__ __forceinline__ void sweep_frictions(const CStore& S, const RowCtx& C, int cnt, float* V7, float* F7, V3& dvl, V3& dvw, float& res,
                                                float& act) {
#pragma unroll kObjUnroll
  for (int i = 0; i < kPairContacts; i++) {
    if (i >= cnt) break;
    const CStore Sc = S.view(KIND * kPairContacts + i);
    const float tot = Sc.at(0, 17);
    if (tot > 0.f) {
      const float lim = Sc.at(0, 10) * tot;
      const float d1 = solve_row<KIND, 1>(Sc, C, 0, -lim, lim, V7, F7, dvl, dvw, res);
      const float d2 = solve_row<KIND, 2>(Sc, C, 0, -lim, lim, V7, F7, dvl, dvw, res);
      if (KIND == 0) act = fmaxf(act, fmaxf(fabsf(d1), fabsf(d2)) - 2.4e-7f * lim);
    }
  }
}

// ------------------------------------------------------------------------------------------------ one stepSimulation()
// q, qd, cq, sq, tgt as in substep() (pg_dynamics.cuh) with stride 1; B = the object; S / XG = this thread's slices of the
// block's shared-memory contact store and X / G matrices.  SYNC: block barriers between the phases (every thread of the
// block calls this, inactive ones with active = false), see substep().
template <bool SYNC>
__device__ __forceinline__ void world_substep(bool active, bool velocity, bool unit_imp, V3 base, const WorldParams& W, float* q, float* qd,
                                              float* cq, float* sq, const float* tgt, ObjState& B, const CStore S, float* XG, int stride) {
  // collision detection sees the poses at the start of the step
  int nt = 0, nr = 0, ns = 0;
  if (active) {
    Frame F7 = link7_frame(base, cq, sq);
    world_contacts(W, F7, B, S, nt, nr, ns);
  }
  if (SYNC) __syncthreads();
  const bool has_arm = (nr + ns) > 0;
  float tau[7], Mi[28];
  if (active) rnea_bias(cq, sq, qd, tau);
  if (SYNC) __syncthreads();
  RowCtx C;
  float V7p[6], G[21];  // G = S M^-1 S^T (packed symmetric): registers, needed by the row setup only
  float rhs[7], dinv[7], max_imp[7], app[7], dv[7], lrhs[7], lsg[7], lapp[7];
  if (active) {
    // 1. unconstrained velocities
    crba(cq, sq, Mi);
    spd_inverse7(Mi);
#pragma unroll
    for (int i = 0; i < 7; i++) {
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 7; j++) acc -= Mi[sym(i, j)] * tau[j];
      qd[i] = fminf(fmaxf(qd[i] + acc * kDt, -kMaxCoordVel), kMaxCoordVel);
    }
    C.R = quat_rot(B.qx, B.qy, B.qz, B.qw);
    // btBoxShape::calculateLocalInertia on the full extents
    const float lx = 2.f * W.half[0], ly = 2.f * W.half[1], lz = 2.f * W.half[2];
    const V3 I = v3(W.mass / 12.f * (ly * ly + lz * lz), W.mass / 12.f * (lx * lx + lz * lz), W.mass / 12.f * (lx * lx + ly * ly));
    C.iinv = v3(1.f / I.x, 1.f / I.y, 1.f / I.z);
    {
      V3 Iw = inv_inertia_apply(C.R, I, B.w);  // R diag(I) R^T w
      V3 gy = cross(B.w, Iw);
      float vn = norm(B.v), wn = norm(B.w);
      float da = kDamp + kDamp * wn, dl = kDamp + kDamp * vn;
      V3 al = inv_inertia_apply(C.R, C.iinv, v3(-gy.x - Iw.x * da, -gy.y - Iw.y * da, -gy.z - Iw.z * da));
      B.v = v3(B.v.x + kDt * (-B.v.x * dl), B.v.y + kDt * (-B.v.y * dl), B.v.z + kDt * (-B.v.z * dl - kGravity));
      B.w = axpy<float>(kDt, al, B.w);
    }
    C.bx = B.x; C.bv = B.v; C.bw = B.w; C.mass = W.mass; C.inv_mass = 1.0f / W.mass;
    C.iso = (I.x == I.y) && (I.y == I.z);
    C.o7 = v3(0.f, 0.f, 0.f);
    // 2. rows: joint limits + motors (arm) ...
    const float lo[7] = {pgm::Link<1>::lo, pgm::Link<2>::lo, pgm::Link<3>::lo, pgm::Link<4>::lo, pgm::Link<5>::lo, pgm::Link<6>::lo, pgm::Link<7>::lo};
    const float hi[7] = {pgm::Link<1>::hi, pgm::Link<2>::hi, pgm::Link<3>::hi, pgm::Link<4>::hi, pgm::Link<5>::hi, pgm::Link<6>::hi, pgm::Link<7>::hi};
#pragma unroll
    for (int d = 0; d < 7; d++) {
      float denom = Mi[sym(d, d)];
      dinv[d] = denom > 1.1920929e-07f ? 1.0f / denom : 0.f;
      float want = velocity ? tgt[d] : 50.0f * (tgt[d] - q[d]);
      rhs[d] = (want - qd[d]) * dinv[d];
      max_imp[d] = unit_imp ? 1.0f : motor_imp(d);
      app[d] = 0.f; dv[d] = 0.f; lsg[d] = 0.f; lrhs[d] = 0.f; lapp[d] = 0.f;
      float pl = q[d] - lo[d], pu = hi[d] - q[d];
      if (pl <= 0.f) { lsg[d] = 1.f; lrhs[d] = (-pl * kLimitErp / kDt - qd[d]) * dinv[d]; }
      else if (pu <= 0.f) { lsg[d] = -1.f; lrhs[d] = (-pu * kLimitErp / kDt + qd[d]) * dinv[d]; }
    }
    // ... and the link-7 task space of the contact rows, only when a robot collider is in contact
    if (has_arm) {
      ArmKin K;
      arm_kin(base, cq, sq, K);
      C.o7 = K.org[6];
      float Sm[6][7];  // S = [ax_d x (o7 - org_d) ; ax_d]
#pragma unroll
      for (int d = 0; d < 7; d++) {
        V3 l = cross(K.ax[d], C.o7 - K.org[d]);
        Sm[0][d] = l.x; Sm[1][d] = l.y; Sm[2][d] = l.z; Sm[3][d] = K.ax[d].x; Sm[4][d] = K.ax[d].y; Sm[5][d] = K.ax[d].z;
      }
#pragma unroll
      for (int k = 0; k < 6; k++) {
        float v = 0.f;
#pragma unroll
        for (int d = 0; d < 7; d++) v = fmaf(Sm[k][d], qd[d], v);
        V7p[k] = v;
      }
      float X[7][6];
#pragma unroll
      for (int i = 0; i < 7; i++)
#pragma unroll
        for (int k = 0; k < 6; k++) {
          float v = 0.f;
#pragma unroll
          for (int j = 0; j < 7; j++) v = fmaf(Mi[sym(i, j)], Sm[k][j], v);
          X[i][k] = v; XG[(i * 6 + k) * stride] = v;
        }
#pragma unroll
      for (int k = 0; k < 6; k++)
#pragma unroll
        for (int l = 0; l <= k; l++) {
          float v = 0.f;
#pragma unroll
          for (int j = 0; j < 7; j++) v = fmaf(Sm[k][j], X[j][l], v);
          G[sym6(k, l)] = v;
        }
    } else {
#pragma unroll
      for (int k = 0; k < 6; k++) V7p[k] = 0.f;
#pragma unroll
      for (int k = 0; k < 21; k++) G[k] = 0.f;
    }
    for (int i = 0; i < nt; i++) setup_contact<0>(S.view(i), C, G, V7p, 0);
    for (int i = 0; i < nr; i++) setup_contact<1>(S.view(kPairContacts + i), C, G, V7p, 0);
    for (int i = 0; i < ns; i++) setup_contact<2>(S.view(2 * kPairContacts + i), C, G, V7p, 0);
  }
  if (SYNC) __syncthreads();
  if (active) {
    // 3. projected Gauss-Seidel.  ONE loop for every lane (arm-only lanes skip the predicated contact parts) so that a warp
    // never runs two solver loops back to back.  dsum = impulses of this sweep's arm rows: they reach the task space once
    // per sweep (V7 += X^T dsum) just as the contact wrench F7 reaches the joints once per sweep (dv += X F7).
    float V7[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, F7[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, dsum[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    bool any_limit = false;
#pragma unroll
    for (int d = 0; d < 7; d++) any_limit |= lsg[d] != 0.f;
#define PG_WM_ROW(d)                                                        \
  {                                                                         \
    float t = fmaf(-dv[d], dinv[d], app[d] + rhs[d]);                       \
    float cl = fminf(fmaxf(t, -max_imp[d]), max_imp[d]);                    \
    float delta = cl - app[d];                                              \
    app[d] = cl;                                                            \
    _Pragma("unroll") for (int i = 0; i < 7; i++) dv[i] = fmaf(Mi[sym(i, d)], delta, dv[i]); \
    dsum[d] += delta;                                                       \
    res = fmaxf(res, fabsf(delta * Mi[sym(d, d)]));                         \
  }
#define PG_WL_ROW(d)                                                        \
  if (lsg[d] != 0.f) {                                                      \
    float delta = lrhs[d] - lsg[d] * dv[d] * dinv[d];                       \
    float sum = lapp[d] + delta;                                            \
    if (sum < 0.f) { delta = -lapp[d]; lapp[d] = 0.f; }                     \
    else if (sum > kLimitMaxImpulse) { delta = kLimitMaxImpulse - lapp[d]; lapp[d] = kLimitMaxImpulse; } \
    else lapp[d] = sum;                                                     \
    float sd = lsg[d] * delta;                                              \
    _Pragma("unroll") for (int i = 0; i < 7; i++) dv[i] = fmaf(Mi[sym(i, d)], sd, dv[i]); \
    dsum[d] += sd;                                                          \
    res = fmaxf(res, fabsf(delta * Mi[sym(d, d)]));                         \
  }
    V3 dvl = v3(0, 0, 0), dvw = v3(0, 0, 0);
    const bool obj_island = nr == 0;  // nothing but the table touches the object: its rows are an independent island
    bool obj_live = true;
    for (int it = 0; it < kSolverIters; it++) {
      float res = 0.f;
      if (it & 1) {  // odd iteration: forward (limits 0..6, then motors 0..6)
        if (any_limit) { PG_WL_ROW(0) PG_WL_ROW(1) PG_WL_ROW(2) PG_WL_ROW(3) PG_WL_ROW(4) PG_WL_ROW(5) PG_WL_ROW(6) }
        PG_WM_ROW(0) PG_WM_ROW(1) PG_WM_ROW(2) PG_WM_ROW(3) PG_WM_ROW(4) PG_WM_ROW(5) PG_WM_ROW(6)
      } else {       // even iteration: backward
        PG_WM_ROW(6) PG_WM_ROW(5) PG_WM_ROW(4) PG_WM_ROW(3) PG_WM_ROW(2) PG_WM_ROW(1) PG_WM_ROW(0)
        if (any_limit) { PG_WL_ROW(6) PG_WL_ROW(5) PG_WL_ROW(4) PG_WL_ROW(3) PG_WL_ROW(2) PG_WL_ROW(1) PG_WL_ROW(0) }
      }
      if (has_arm) {  // the arm rows' impulses reach link 7: V7 += sum_d X[d][:] dsum[d]
#pragma unroll
        for (int d = 0; d < 7; d++) {
#pragma unroll
          for (int k = 0; k < 6; k++) V7[k] = fmaf(XG[(d * 6 + k) * stride], dsum[d], V7[k]);
          dsum[d] = 0.f;
        }
      }
      // Rows of the object-table pair form an independent island while nothing else touches the object.  Once every one
      // of them moves by less than an ulp of its impulse the island sits at its fixed point -- further sweeps would
      // reproduce the same values -- and it is skipped for the rest of the solve (the arm usually needs all 50 sweeps,
      // the resting object 3-5).
      const bool skip_obj = obj_island && !obj_live;
      float act = 0.f;
      if (!skip_obj) sweep_normals<0>(S, C, nt, V7, F7, dvl, dvw, res, act);   // all normals, in contact order
      if (has_arm) {
        sweep_normals<1>(S, C, nr, V7, F7, dvl, dvw, res, act);
        sweep_normals<2>(S, C, ns, V7, F7, dvl, dvw, res, act);
      }
      if (!skip_obj) sweep_frictions<0>(S, C, nt, V7, F7, dvl, dvw, res, act);  // then all frictions
      if (has_arm) {
        sweep_frictions<1>(S, C, nr, V7, F7, dvl, dvw, res, act);
        sweep_frictions<2>(S, C, ns, V7, F7, dvl, dvw, res, act);
        // the sweep's contact wrench reaches the joint velocities: dv += X F7
#pragma unroll
        for (int i = 0; i < 7; i++) {
          float v = dv[i];
#pragma unroll
          for (int k = 0; k < 6; k++) v = fmaf(XG[(i * 6 + k) * stride], F7[k], v);
          dv[i] = v;
        }
#pragma unroll
        for (int k = 0; k < 6; k++) F7[k] = 0.f;
      }
      if (obj_island && obj_live && act <= 0.f) obj_live = false;
      if (res <= 3.1622777e-4f) break;  // sqrt(1e-7)
    }
#undef PG_WM_ROW
#undef PG_WL_ROW
    // 4. apply, clamp, integrate
#pragma unroll
    for (int d = 0; d < 7; d++) {
      qd[d] = fminf(fmaxf(qd[d] + dv[d], -kMaxCoordVel), kMaxCoordVel);
      float dq = qd[d] * kDt;
      q[d] += dq;
      float d2 = dq * dq;
      float cd = fmaf(d2, fmaf(d2, 1.0f / 24.0f, -0.5f), 1.0f);
      float sd = dq * fmaf(d2, fmaf(d2, 1.0f / 120.0f, -1.0f / 6.0f), 1.0f);
      float c0 = cq[d], s0 = sq[d];
      cq[d] = c0 * cd - s0 * sd;
      sq[d] = s0 * cd + c0 * sd;
    }
    B.v = B.v + dvl; B.w = B.w + dvw;
    B.x = axpy<float>(kDt, B.v, B.x);
    {
      float fa = norm(B.w), s;
      if (fa * kDt > 0.78539816f) fa = 0.78539816f / kDt;  // ANGULAR_MOTION_THRESHOLD
      if (fa < 0.001f) s = 0.5f * kDt - kDt * kDt * kDt * 0.020833333333f * fa * fa;
      else s = sinf(0.5f * fa * kDt) / fa;
      V3 a = s * B.w;
      float cw = cosf(fa * kDt * 0.5f);
      float nx = cw * B.qx + a.x * B.qw + a.y * B.qz - a.z * B.qy;
      float ny = cw * B.qy + a.y * B.qw + a.z * B.qx - a.x * B.qz;
      float nz = cw * B.qz + a.z * B.qw + a.x * B.qy - a.y * B.qx;
      float nw = cw * B.qw - a.x * B.qx - a.y * B.qy - a.z * B.qz;
      float L = rsqrtf(nx * nx + ny * ny + nz * nz + nw * nw);
      B.qx = nx * L; B.qy = ny * L; B.qz = nz * L; B.qw = nw * L;
    }
  }
  if (SYNC) __syncthreads();
}

}  // namespace pg
