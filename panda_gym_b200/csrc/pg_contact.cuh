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

// kind 0: object (A) on table / plane (B); 1: robot (A) on object (B); 2: robot (A) on table top (B).
// n = unit normal from B towards A; pa / pb = contact points on A / B
struct Contact { int kind; float d, mu; V3 n, pa, pb; };

__device__ __forceinline__ void keep_closest(Contact* list, int& n, const Contact& c) {  // sorted insert, stable, capacity 4
  int pos = n;
  while (pos > 0 && c.d < list[pos - 1].d) pos--;
  if (pos >= kPairContacts) return;
  int last = n < kPairContacts ? n : kPairContacts - 1;
  for (int i = last; i > pos; i--) list[i] = list[i - 1];
  list[pos] = c;
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
__device__ __forceinline__ void prim_contacts(const Frame& F, const WorldParams& W, V3 bx, const Rot& R, Contact* cr, int& nr, Contact* cs,
                                              int& ns) {
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
      Contact ct; ct.kind = 1; ct.mu = mu_l * W.mu_object;
      V3 dl, pal;
      bool done = false;
      if (Pm::type == 0) {
        Closest k = sphere_box<float>(cl, Pm::r, o, core, m);
        ct.d = k.d; dl = k.dir; pal = cl + Pm::r * dl;
      } else {
        V3 al = rot_applyT(R, a);
        GjkIn<float> g; g.c = cl; g.a = al; g.b = o; g.e = core; g.r = Pm::r; g.h = Pm::h;
        V3 qa, qb;
        float d = gjk_cyl_box(g, &qa, &qb);
        if (d > 0.f) {
#pragma unroll
          for (int it = 0; it < 4; it++) {  // polish (see cyl_box)
            qb = v3(fminf(fmaxf(qa.x, -core.x), core.x), fminf(fmaxf(qa.y, -core.y), core.y), fminf(fmaxf(qa.z, -core.z), core.z));
            float dd = point_cylinder<float>(qb, cl, al, Pm::r, Pm::h, qa);
            d = dd > 0.f ? dd : 0.f;
          }
        }
        V3 v = qb - qa;
        float L = norm(v);
        ct.d = d - kHullMargin - m;
        if (L > 1e-9f) { dl = (1.0f / L) * v; pal = qa + kHullMargin * dl; }
        else { float cn = norm(cl); dl = cn > 1e-12f ? (-1.0f / cn) * cl : v3(0.f, 0.f, -1.f); pal = qa; }  // deep overlap: line of centres
        // cylinder against a box FACE: the closest point of a generator parallel to the face is not unique (rounding
        // noise would pick it), so the contact is spread like a manifold: both ends of the generator facing the box,
        // clipped to the face, each with its own distance
        if (ct.d < kBreaking) {
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
                  ct.d = -sg * gl[kf] - W.half[kf];
                  ct.pa = bx + rot_apply(R, v3(gl[0], gl[1], gl[2]));
                  ct.n = rot_apply(R, v3(nl[0], nl[1], nl[2]));
                  ct.pb = ct.pa - ct.d * ct.n;
                  if (ct.d < kBreaking) keep_closest(cr, nr, ct);
                }
              }
            }
          }
        }
      }
      if (!done && ct.d < kBreaking) {
        V3 dw = rot_apply(R, dl);
        ct.n = v3(-dw.x, -dw.y, -dw.z);
        ct.pa = bx + rot_apply(R, pal);
        ct.pb = ct.pa + ct.d * dw;
        keep_closest(cr, nr, ct);
      }
    }
  }
  // --- vs the table top: lowest point(s) of the primitive; a cylinder contributes both ends of its lowest generator
  {
    const float top = W.table_pos[2] + W.table_half[2];
    Contact ct; ct.kind = 2; ct.mu = mu_l * W.mu_static; ct.n = v3(0.f, 0.f, 1.f);
#pragma unroll
    for (int end = 0; end < (Pm::type == 0 ? 1 : 2); end++) {
      if (Pm::type == 0) ct.pa = v3(c.x, c.y, c.z - Pm::r);
      else {
        const float r = Pm::r + kHullMargin;
        const float t = end == 0 ? -Pm::h : Pm::h;
        V3 u = v3(-a.z * a.x, -a.z * a.y, 1.f - a.z * a.z);
        float L = norm(u);
        ct.pa = c + t * a;
        if (L > 1e-6f) ct.pa = ct.pa - (r / L) * u;
      }
      if (!over_rect(ct.pa, W.table_pos, W.table_half)) continue;
      ct.d = ct.pa.z - top;
      ct.pb = v3(ct.pa.x, ct.pa.y, top);
      if (ct.d < kBreaking) keep_closest(cs, ns, ct);
    }
  }
}

// contact set at the current poses; returns the count (<= 12), order: object-table, robot-object, robot-table
static __device__ __noinline__ int world_contacts(const WorldParams& W, const Frame& F7, const ObjState& B, Contact* out) {
  Contact ct[kPairContacts], cr[kPairContacts], cs[kPairContacts];
  int nt = 0, nr = 0, ns = 0;
  Rot R = quat_rot(B.qx, B.qy, B.qz, B.qw);
  const float top = W.table_pos[2] + W.table_half[2];
#pragma unroll 1
  for (int v = 0; v < 8; v++) {
    V3 l = v3((v & 1) ? W.half[0] : -W.half[0], (v & 2) ? W.half[1] : -W.half[1], (v & 4) ? W.half[2] : -W.half[2]);
    V3 p = B.x + rot_apply(R, l);
    float z = over_rect(p, W.table_pos, W.table_half) ? top : W.plane_top;
    Contact c; c.kind = 0; c.d = p.z - z; c.mu = W.mu_object * W.mu_static;
    if (!(c.d < kBreaking)) continue;
    c.n = v3(0.f, 0.f, 1.f); c.pa = p; c.pb = v3(p.x, p.y, z);
    keep_closest(ct, nt, c);
  }
  // model-table order of the oracle: link7 (21-23), link8 (24-26), hand (30-32), ee (27-29)
  prim_contacts<21>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<22>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<23>(F7, W, B.x, R, cr, nr, cs, ns);
  prim_contacts<24>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<25>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<26>(F7, W, B.x, R, cr, nr, cs, ns);
  prim_contacts<30>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<31>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<32>(F7, W, B.x, R, cr, nr, cs, ns);
  prim_contacts<27>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<28>(F7, W, B.x, R, cr, nr, cs, ns); prim_contacts<29>(F7, W, B.x, R, cr, nr, cs, ns);
  int n = 0;
  for (int i = 0; i < nt; i++) out[n++] = ct[i];
  for (int i = 0; i < nr; i++) out[n++] = cr[i];
  for (int i = 0; i < ns; i++) out[n++] = cs[i];
  return n;
}

// ------------------------------------------------------------------------------------------------ solver rows
struct Row {
  float Ja[7], ra[7];   // arm: joint-space Jacobian and response M^-1 Ja
  V3 jl, jw, rw;        // object: linear / angular Jacobian, angular response I^-1 jw (linear response = jl / mass)
  float rhs, dinv, applied;
  bool arm, body;
};

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

// one row of contact c along direction t (normal or friction axis); velocities are the predicted ones
static __device__ __noinline__ void setup_row(Row& r, const ArmKin& K, const float* Mi, const float* qd, const ObjState& B, const WorldParams& W,
                                       const Rot& R, V3 iinv, const Contact& c, V3 t, bool friction) {
  float denom = 0.f, rel = 0.f;
  r.arm = c.kind != 0; r.body = c.kind != 2; r.applied = 0.f;
  if (r.arm) {
#pragma unroll
    for (int d = 0; d < 7; d++) r.Ja[d] = dot(cross(K.ax[d], c.pa - K.org[d]), t);
#pragma unroll
    for (int i = 0; i < 7; i++) {
      float v = 0.f;
#pragma unroll
      for (int j = 0; j < 7; j++) v = fmaf(Mi[sym(i, j)], r.Ja[j], v);
      r.ra[i] = v;
    }
#pragma unroll
    for (int d = 0; d < 7; d++) { denom = fmaf(r.Ja[d], r.ra[d], denom); rel = fmaf(r.Ja[d], qd[d], rel); }
  } else {
#pragma unroll
    for (int d = 0; d < 7; d++) { r.Ja[d] = 0.f; r.ra[d] = 0.f; }
  }
  if (r.body) {
    const float s = c.kind == 0 ? 1.f : -1.f;
    V3 rp = (c.kind == 0 ? c.pa : c.pb) - B.x;
    r.jl = s * t; r.jw = s * cross(rp, t);
    r.rw = inv_inertia_apply(R, iinv, r.jw);
    denom += dot(r.jl, r.jl) / W.mass + dot(r.jw, r.rw);
    rel += dot(r.jl, B.v) + dot(r.jw, B.w);
  } else { r.jl = v3(0, 0, 0); r.jw = v3(0, 0, 0); r.rw = v3(0, 0, 0); }
  r.dinv = denom > 1.1920929e-07f ? 1.0f / denom : 0.f;
  if (friction) r.rhs = -rel * r.dinv;
  else {
    float pos_err = 0.f, vel_err = -rel;
    if (c.d > 0.f) vel_err -= c.d / kDt; else pos_err = -c.d * kErp2 / kDt;
    r.rhs = (pos_err + vel_err) * r.dinv;
  }
}

// returns the impulse change; the row's residual (what Bullet's early-out tests) is delta / dinv
__device__ __forceinline__ float solve_row(Row& r, float lo, float hi, float inv_mass, float* dva, V3& dvl, V3& dvw) {
  float jv = 0.f;
  if (r.arm) {
#pragma unroll
    for (int d = 0; d < 7; d++) jv = fmaf(r.Ja[d], dva[d], jv);
  }
  if (r.body) jv += dot(r.jl, dvl) + dot(r.jw, dvw);
  float delta = r.rhs - jv * r.dinv;
  float sum = r.applied + delta;
  if (sum < lo) { delta = lo - r.applied; r.applied = lo; }
  else if (sum > hi) { delta = hi - r.applied; r.applied = hi; }
  else r.applied = sum;
  if (r.arm) {
#pragma unroll
    for (int d = 0; d < 7; d++) dva[d] = fmaf(r.ra[d], delta, dva[d]);
  }
  if (r.body) { dvl = axpy<float>(delta * inv_mass, r.jl, dvl); dvw = axpy<float>(delta, r.rw, dvw); }
  return delta;
}
__device__ __forceinline__ float row_residual(const Row& r, float delta) { return r.dinv != 0.f ? fabsf(delta / r.dinv) : 0.f; }

// ------------------------------------------------------------------------------------------------ one stepSimulation()
// q, qd, cq, sq, tgt as in substep() (pg_dynamics.cuh) with stride 1; B = the object.  Returns the PGS iterations used.
static __device__ __noinline__ int world_substep(bool velocity, bool unit_imp, V3 base, const WorldParams& W, float* q, float* qd, float* cq,
                                          float* sq, const float* tgt, ObjState& B) {
  // collision detection sees the poses at the start of the step
  ArmKin K;
  arm_kin(base, cq, sq, K);
  Contact ct[kMaxContacts];
  const int nc = world_contacts(W, K.F7, B, ct);
  const Rot R = quat_rot(B.qx, B.qy, B.qz, B.qw);
  // btBoxShape::calculateLocalInertia on the full extents
  const float lx = 2.f * W.half[0], ly = 2.f * W.half[1], lz = 2.f * W.half[2];
  const V3 I = v3(W.mass / 12.f * (ly * ly + lz * lz), W.mass / 12.f * (lx * lx + lz * lz), W.mass / 12.f * (lx * lx + ly * ly));
  const V3 iinv = v3(1.f / I.x, 1.f / I.y, 1.f / I.z);
  // 1. unconstrained velocities
  float tau[7], Mi[28];
  rnea_bias(cq, sq, qd, tau);
  crba(cq, sq, Mi);
  spd_inverse7(Mi);
#pragma unroll
  for (int i = 0; i < 7; i++) {
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 7; j++) acc -= Mi[sym(i, j)] * tau[j];
    qd[i] = fminf(fmaxf(qd[i] + acc * kDt, -kMaxCoordVel), kMaxCoordVel);
  }
  {
    V3 Iw = inv_inertia_apply(R, I, B.w);  // R diag(I) R^T w
    V3 gy = cross(B.w, Iw);
    float vn = norm(B.v), wn = norm(B.w);
    float da = kDamp + kDamp * wn, dl = kDamp + kDamp * vn;
    V3 al = inv_inertia_apply(R, iinv, v3(-gy.x - Iw.x * da, -gy.y - Iw.y * da, -gy.z - Iw.z * da));
    B.v = v3(B.v.x + kDt * (-B.v.x * dl), B.v.y + kDt * (-B.v.y * dl), B.v.z + kDt * (-B.v.z * dl - kGravity));
    B.w = axpy<float>(kDt, al, B.w);
  }
  // 2. rows: joint limits + motors (arm), contacts
  const float lo[7] = {pgm::Link<1>::lo, pgm::Link<2>::lo, pgm::Link<3>::lo, pgm::Link<4>::lo, pgm::Link<5>::lo, pgm::Link<6>::lo, pgm::Link<7>::lo};
  const float hi[7] = {pgm::Link<1>::hi, pgm::Link<2>::hi, pgm::Link<3>::hi, pgm::Link<4>::hi, pgm::Link<5>::hi, pgm::Link<6>::hi, pgm::Link<7>::hi};
  float rhs[7], dinv[7], max_imp[7], app[7], dv[7], lrhs[7], lsg[7], lapp[7];
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
  Row nrow[kMaxContacts], frow[2 * kMaxContacts];
  for (int i = 0; i < nc; i++) {
    V3 t1, t2;
    setup_row(nrow[i], K, Mi, qd, B, W, R, iinv, ct[i], ct[i].n, false);
    plane_space(ct[i].n, t1, t2);
    setup_row(frow[2 * i], K, Mi, qd, B, W, R, iinv, ct[i], t1, true);
    setup_row(frow[2 * i + 1], K, Mi, qd, B, W, R, iinv, ct[i], t2, true);
  }
  // 3. projected Gauss-Seidel
#define PG_WM_ROW(d)                                                        \
  if (max_imp[d] != 0.f) {                                                  \
    float t = fmaf(-dv[d], dinv[d], app[d] + rhs[d]);                       \
    float cl = fminf(fmaxf(t, -max_imp[d]), max_imp[d]);                    \
    float delta = cl - app[d];                                              \
    app[d] = cl;                                                            \
    _Pragma("unroll") for (int i = 0; i < 7; i++) dv[i] = fmaf(Mi[sym(i, d)], delta, dv[i]); \
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
    res = fmaxf(res, fabsf(delta * Mi[sym(d, d)]));                         \
  }
  V3 dvl = v3(0, 0, 0), dvw = v3(0, 0, 0);
  bool obj_island = true, obj_live = true;
  for (int i = 0; i < nc; i++) obj_island &= ct[i].kind != 1;
  const float inv_mass = 1.0f / W.mass;
  int it = 0;
  for (; it < kSolverIters; it++) {
    float res = 0.f;
    if (it & 1) {  // odd iteration: forward (limits 0..6, then motors 0..6)
      PG_WL_ROW(0) PG_WL_ROW(1) PG_WL_ROW(2) PG_WL_ROW(3) PG_WL_ROW(4) PG_WL_ROW(5) PG_WL_ROW(6)
      PG_WM_ROW(0) PG_WM_ROW(1) PG_WM_ROW(2) PG_WM_ROW(3) PG_WM_ROW(4) PG_WM_ROW(5) PG_WM_ROW(6)
    } else {       // even iteration: backward
      PG_WM_ROW(6) PG_WM_ROW(5) PG_WM_ROW(4) PG_WM_ROW(3) PG_WM_ROW(2) PG_WM_ROW(1) PG_WM_ROW(0)
      PG_WL_ROW(6) PG_WL_ROW(5) PG_WL_ROW(4) PG_WL_ROW(3) PG_WL_ROW(2) PG_WL_ROW(1) PG_WL_ROW(0)
    }
    // Rows of the object-table pair form an independent island while nothing else touches the object.  Once every one
    // of them moves by less than an ulp of its impulse the island sits at its fixed point -- further sweeps would
    // reproduce the same values -- and it is skipped for the rest of the solve (the arm usually needs all 50 sweeps,
    // the resting object 3-5).
    const bool skip_obj = obj_island && !obj_live;
    float act = 0.f;
    for (int i = 0; i < nc; i++) {
      if (skip_obj && ct[i].kind == 0) continue;
      float delta = solve_row(nrow[i], 0.f, 1e10f, inv_mass, dv, dvl, dvw);
      res = fmaxf(res, row_residual(nrow[i], delta));
      if (ct[i].kind == 0) act = fmaxf(act, fabsf(delta) - 2.4e-7f * fabsf(nrow[i].applied));
    }
    for (int i = 0; i < 2 * nc; i++) {
      if (skip_obj && ct[i >> 1].kind == 0) continue;
      float tot = nrow[i >> 1].applied;
      if (tot > 0.f) {
        float lim = ct[i >> 1].mu * tot;
        float delta = solve_row(frow[i], -lim, lim, inv_mass, dv, dvl, dvw);
        res = fmaxf(res, row_residual(frow[i], delta));
        if (ct[i >> 1].kind == 0) act = fmaxf(act, fabsf(delta) - 2.4e-7f * lim);
      }
    }
    if (obj_island && obj_live && act <= 0.f) obj_live = false;
    if (res <= 3.1622777e-4f) { it++; break; }  // sqrt(1e-7)
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
  return it;
}

}  // namespace pg
