// pg_scene.cuh -- per-environment collision / observation pass over the 9 ReachAO collision links.
//
// Replaces (reference) ReachAO.check_collided (reach_ao.py:896-900) in its flag-only form and
// ReachAO.get_obs / get_vector_obs (reach_ao.py:902-959) in its full form; both go through the pyb_utils-fork
// CollisionDetector in the reference, i.e. 9 links x M obstacles getClosestPoints calls per query.
#pragma once
#include "pg_geom.cuh"

namespace pg {

struct SceneParams {
  int nobs;
  Obst obst[kMaxObst];
  float obst_bound[kMaxObst];  // bounding radius of each obstacle shape
  float table_pos[3], table_core[3], table_margin;
  int has_table;
  int has_box;                 // any cuboid obstacle in the scene (selects the barrier pattern of the flag pass)
  int skip_far;                // d_link of every environment is current: the flag pass may cull links by it (scene_skip_mask)
  // static scene shared by all environments (benchmark scenes, pg_set_scene): axis-aligned boxes, two float4 per box
  // = (centre xyz, collision margin), (half extents xyz, bounding radius)
  int nstatic;
  const float4* sbox;
  // per-environment cuboid sizes (reachao_rand_shape), else nullptr: f32 [kMaxObst][5][n] = half extents xyz, margin, bounding radius
  const float* env_box;
};
// A translation unit compiled with PG_ENVBOX 0 ignores env_box (the default step kernel: its flag pass reads the obstacle
// shapes straight from the kernel parameters); the launcher picks the PG_ENVBOX 1 twin when the stage has per-environment sizes.
#ifndef PG_ENVBOX
#define PG_ENVBOX 1
#endif
__device__ __forceinline__ Obst scene_obst(const SceneParams& P, int j, int e, int n, float& bound) {
  PG_CHECK(j >= 0 && j < kMaxObst && e >= 0 && e < n);
  Obst ob = P.obst[j];
  bound = P.obst_bound[j];
#if PG_ENVBOX
  if (P.env_box && ob.type == 1) {
    const float* b = P.env_box + (size_t)(j * 5) * n + e;
    ob.sx = b[0]; ob.sy = b[n]; ob.sz = b[(size_t)2 * n]; ob.margin = b[(size_t)3 * n]; bound = b[(size_t)4 * n];
  }
#endif
  return ob;
}
constexpr int kMaxStaticBoxes = 64;
__device__ __forceinline__ Obst static_obst(const SceneParams& P, int j, V3& pos) {
  const float4 a = __ldg(P.sbox + 2 * j), b = __ldg(P.sbox + 2 * j + 1);
  pos = v3(a.x, a.y, a.z);
  Obst ob; ob.type = 1; ob.sx = b.x; ob.sy = b.y; ob.sz = b.z; ob.margin = a.w;
  return ob;
}

// run-time copy of the compile-time primitive table (render, neutral-pose table of a custom reset pose)
struct PrimRT { int frame, type; float x, y, z, ax, ay, az, r, h; };
template <int K> __host__ __device__ constexpr PrimRT prim_rt() {
  using P = pgm::Prim<K>;
  return PrimRT{P::frame, P::type, P::x, P::y, P::z, P::ax, P::ay, P::az, P::r, P::h};
}
static __device__ __constant__ PrimRT kPrimRT[pgm::kNumPrim] = {
    prim_rt<0>(),  prim_rt<1>(),  prim_rt<2>(),  prim_rt<3>(),  prim_rt<4>(),  prim_rt<5>(),  prim_rt<6>(),  prim_rt<7>(),  prim_rt<8>(),
    prim_rt<9>(),  prim_rt<10>(), prim_rt<11>(), prim_rt<12>(), prim_rt<13>(), prim_rt<14>(), prim_rt<15>(), prim_rt<16>(), prim_rt<17>(),
    prim_rt<18>(), prim_rt<19>(), prim_rt<20>(), prim_rt<21>(), prim_rt<22>(), prim_rt<23>(), prim_rt<24>(), prim_rt<25>(), prim_rt<26>(),
    prim_rt<27>(), prim_rt<28>(), prim_rt<29>(), prim_rt<30>(), prim_rt<31>(), prim_rt<32>()};

// primitives of collision-link slot S: [first, first + count)
__host__ __device__ constexpr int slot_first(int s) { return s <= 4 ? 3 * s : 3 * s + 3; }
__host__ __device__ constexpr int slot_count(int s) { return s == 4 ? 6 : 3; }
__host__ __device__ constexpr float cabs(float x) { return x < 0 ? -x : x; }
__host__ __device__ constexpr float cmax(float a, float b) { return a > b ? a : b; }

// conservative bounding sphere of a slot: centre = its first primitive, radius via L1 offsets
template <int K0, int K> struct BoundAcc {
  using A = pgm::Prim<K0>; using B = pgm::Prim<K>;
  static constexpr float ext = B::r + (B::type == 1 ? B::h : 0.f);
  static constexpr float off = cabs(B::x - A::x) + cabs(B::y - A::y) + cabs(B::z - A::z);
  static constexpr float value = off + ext;
};
template <int K0, int N> struct SlotBound { static constexpr float value = cmax(SlotBound<K0, N - 1>::value, BoundAcc<K0, K0 + N - 1>::value); };
template <int K0> struct SlotBound<K0, 0> { static constexpr float value = 0.f; };

// Deferred cylinder-box queries of the observation pass.  Which (primitive, box) pairs survive the analytic lower
// bound differs from lane to lane; calling GJK at each of the 30 call sites would serialise the warp 30 times with a
// few lanes active.  Instead every lane queues its survivors and the warp drains the queues together, one
// convergent GJK call per round (rounds = longest queue in the warp, typically 2-4).
struct CylBoxCand { V3 c, a; float r, h, lb; int j, slot; };
struct CylBoxQueue { static constexpr int kCap = 12; CylBoxCand q[kCap]; int n; };

// Measured on B200 (profiles/r02_variants.txt), 65 536 envs: with box obstacles the per-substep flag pass is a long,
// divergently-culled code path and the warps of a block drift apart in it (ncu: 56 % stall_no_inst); block barriers between
// its link groups keep them inside one instruction-cache window: reachao_rand 3.28 -> 1.90 ms per step.  Sphere-only scenes
// run 2 % slower with those barriers, so they are taken only when the scene has a box (a kernel parameter: uniform over the
// block).  One shared copy of the cylinder-box sign test instead of ten inlined ones: reachao3 0.748 -> 0.717 ms.
#ifndef PG_SCENE_SYNC
#define PG_SCENE_SYNC 2   // 0 never, 1 always, 2 when the scene has box obstacles
#endif
#ifndef PG_CYLBOX_NOINLINE
#define PG_CYLBOX_NOINLINE 1
#endif
// cylinder (centre rel to the box centre, unit axis a, radius r, half length h) vs box core, margin m: is the distance <= 0 ?
// Decided in closed form except in the sliver between the two bounds (see cyl_box_bound).
#if PG_CYLBOX_NOINLINE
static __device__ __noinline__
#else
__device__ __forceinline__
#endif
bool cyl_box_hits(V3 rel, V3 a, float r, float h, V3 core, float margin) {
  const float thr = kHullMargin + margin;
  {  // bounding sphere of the cylinder against the core
    V3 v = rel - box_clamp<float>(rel, core);
    const float reach = r + h + thr;
    if (dot(v, v) > reach * reach) return false;
  }
  CylBoxBoundT<float> bd = cyl_box_bound<float>(rel, a, r, h, core);
  if (bd.lb > thr) return false;
  if (bd.exact) return true;
  V3 on;
  if (point_cylinder<float>(bd.pb, rel, a, r, h, on) <= thr) return true;  // the box's closest point already reaches the cylinder
#ifndef PG_FLAG_FULL_GJK
  // only the sign is wanted: GJK stops as soon as one of its two bounds crosses the threshold (gjk_cyl_box_decide) instead of
  // converging (20 iterations + 4 polish projections) while the rest of the block waits at the pass's barriers
  GjkIn<float> g; g.c = rel; g.a = a; g.b = v3(0.f, 0.f, 0.f); g.e = core; g.r = r; g.h = h;
  return gjk_cyl_box_decide<float>(g, thr) <= thr;
#else
  return cyl_box<float>(rel, a, r, h, v3(0.f, 0.f, 0.f), core, margin).d <= 0.f;
#endif
}

template <int K>
__device__ __forceinline__ void prim_vs_obstacle(const Frame& F, V3 o, const Obst& ob, float& best, V3& bdir, CylBoxQueue* Q = nullptr,
                                                 int j = 0) {
  using Pm = pgm::Prim<K>;
  V3 c = frame_point(F, Pm::x, Pm::y, Pm::z);
  Closest k;
  if (Pm::type == 0) {
    if (ob.type == 0) k = sphere_sphere<float>(c, Pm::r, o, ob.sx);
    else k = sphere_box<float>(c, Pm::r, o, v3(ob.sx - ob.margin, ob.sy - ob.margin, ob.sz - ob.margin), ob.margin);
  } else {
    V3 a = frame_dir(F, Pm::ax, Pm::ay, Pm::az);
    if (ob.type == 0) k = cyl_sphere<float>(c, a, Pm::r, Pm::h, o, ob.sx);
    else {
      // cylinder-box: closed form whenever the closest feature of the cylinder is its lateral surface (cyl_box_bound);
      // a pair whose lower bound cannot beat the running minimum of this link is skipped; only a slanted end-cap
      // contact that could still win goes to GJK (deferred to the warp-convergent queue when there is one)
      V3 core = v3(ob.sx - ob.margin, ob.sy - ob.margin, ob.sz - ob.margin);
      CylBoxBoundT<float> bd = cyl_box_bound<float>(c - o, a, Pm::r, Pm::h, core);
      const float lb = bd.lb - kHullMargin - ob.margin;
      if (lb >= best) return;
      if (bd.exact) { best = lb; bdir = bd.lb > 0.f ? bd.u : v3(0.f, 0.f, 0.f); return; }
      if (Q && Q->n < CylBoxQueue::kCap) {
        CylBoxCand& cd = Q->q[Q->n++];
        cd.c = c; cd.a = a; cd.r = Pm::r; cd.h = Pm::h; cd.lb = lb; cd.j = j; cd.slot = Pm::slot;
        return;
      }
      k = cyl_box<float>(c, a, Pm::r, Pm::h, o, core, ob.margin);
    }
  }
  if (k.d < best) { best = k.d; bdir = k.dir; }
}
// flag-only form: "is the distance <= 0", squared-distance tests for the sphere pairs (no sqrt / divide)
template <int K>
__device__ __forceinline__ bool prim_hits_obstacle(const Frame& F, V3 o, const Obst& ob) {
  using Pm = pgm::Prim<K>;
  V3 c = frame_point(F, Pm::x, Pm::y, Pm::z);
  if (Pm::type == 0) {
    if (ob.type == 0) { V3 d = o - c; float rr = Pm::r + ob.sx; return dot(d, d) <= rr * rr; }
    V3 rel = c - o;
    float ex = ob.sx - ob.margin, ey = ob.sy - ob.margin, ez = ob.sz - ob.margin;
    V3 v = v3(rel.x - fminf(fmaxf(rel.x, -ex), ex), rel.y - fminf(fmaxf(rel.y, -ey), ey), rel.z - fminf(fmaxf(rel.z, -ez), ez));
    float rr = Pm::r + ob.margin;
    return dot(v, v) <= rr * rr;  // centre inside the core gives v = 0: a hit
  }
  V3 a = frame_dir(F, Pm::ax, Pm::ay, Pm::az);
  if (ob.type == 0) { float best = 1e-6f; V3 dir; prim_vs_obstacle<K>(F, o, ob, best, dir); return best <= 0.f; }
  return cyl_box_hits(c - o, a, Pm::r, Pm::h, v3(ob.sx - ob.margin, ob.sy - ob.margin, ob.sz - ob.margin), ob.margin);
}

template <int K>
__device__ __forceinline__ float prim_vs_table(const Frame& F, const SceneParams& S) {
  using Pm = pgm::Prim<K>;
  V3 c = frame_point(F, Pm::x, Pm::y, Pm::z);
  V3 b = v3(S.table_pos[0], S.table_pos[1], S.table_pos[2]), e = v3(S.table_core[0], S.table_core[1], S.table_core[2]);
  if (Pm::type == 0) return sphere_box<float>(c, Pm::r, b, e, S.table_margin).d;
  return cyl_table_distance(c, frame_dir(F, Pm::ax, Pm::ay, Pm::az), Pm::r, Pm::h, b, e, S.table_margin);
}

template <int K0, int N> struct PrimLoop {
  static __device__ __forceinline__ void obstacle(const Frame& F, V3 o, const Obst& ob, float& best, V3& dir, CylBoxQueue* Q, int j) {
    // spheres first (analytic, they tighten the running minimum), cylinders afterwards
    PrimLoop<K0, N - 1>::obstacle(F, o, ob, best, dir, Q, j);
    prim_vs_obstacle<K0 + N - 1>(F, o, ob, best, dir, Q, j);
  }
  static __device__ __forceinline__ float table(const Frame& F, const SceneParams& S) {
    return fminf(PrimLoop<K0, N - 1>::table(F, S), prim_vs_table<K0 + N - 1>(F, S));
  }
  static __device__ __forceinline__ bool hits(const Frame& F, V3 o, const Obst& ob) {
    return PrimLoop<K0, N - 1>::hits(F, o, ob) | prim_hits_obstacle<K0 + N - 1>(F, o, ob);  // straight-line, no branches
  }
  // box obstacle: the spheres of the link first (cheap); the link colliders are capsules, so a slanted end contact reaches a
  // sphere before the cylinder and the cylinder-box test is only needed while no sphere hits
  static __device__ __forceinline__ bool hits_kind(const Frame& F, V3 o, const Obst& ob, int type) {
    bool h = PrimLoop<K0, N - 1>::hits_kind(F, o, ob, type);
    if (pgm::Prim<K0 + N - 1>::type == type) h |= prim_hits_obstacle<K0 + N - 1>(F, o, ob);
    return h;
  }
};
template <int K0> struct PrimLoop<K0, 0> {
  static __device__ __forceinline__ void obstacle(const Frame&, V3, const Obst&, float&, V3&, CylBoxQueue*, int) {}
  static __device__ __forceinline__ float table(const Frame&, const SceneParams&) { return 3.0e38f; }
  static __device__ __forceinline__ bool hits(const Frame&, V3, const Obst&) { return false; }
  static __device__ __forceinline__ bool hits_kind(const Frame&, V3, const Obst&, int) { return false; }
};

// ---- temporal culling of the per-substep flag pass.
// Obstacles do not move during a step, and the observation pass that ended the previous step (or the reset) left the exact
// minimum distance d0 of every collision link to the obstacles at the pose q0 this step starts from.  Slot S lies inside the
// ball of radius R_S about its centre c_S (first primitive), rigid with joint frame k.  Going from q0 to q, a point
// c_S + v of it is displaced by at most
//     M_S = |c_S(q) - c_S(q0)| + R_S * sum_{j <= k} |q_j - q0_j|
// (translation of the centre plus |v| <= R_S times the angle the frame has turned by, which the joint angles bound).  While
// d0_S - M_S stays positive no collider of the slot can touch an obstacle and its obstacle tests are skipped for this
// substep: the flags are the ones the full test gives (tests/test_gpu_parity.py::test_flag_pass_culling_changes_nothing
// compares whole rollouts bit for bit).  The table test has its own height cull and is not affected.
constexpr float kSkipMargin = 2e-3f;  // fp32 error of d0 (GJK / closed forms: ~1e-5) with two orders of magnitude to spare
struct CullState {                    // per-thread slices of the block's shared memory, element i at [i * stride]
  const float* q; const float* q0;    // joint angles now / at the start of the step
  const float* slack;                 // [9] d0_S - kSkipMargin (negative: do not cull)
  const float* ctr0;                  // [9][3] slot centres at q0
  int stride;
};
template <int S> __device__ __forceinline__ V3 slot_centre(const Frame& F) {
  using P0 = pgm::Prim<slot_first(S)>;
  return frame_point(F, P0::x, P0::y, P0::z);
}
// slot centres at the pose (cq, sq) into ctr0 (start of a step)
__device__ __forceinline__ void scene_centres(V3 base, const float* cq, const float* sq, float* ctr0, int stride) {
  Frame F = base_frame(base);
  V3 c;
#define PG_PUT(S) c = slot_centre<S>(F); ctr0[(3 * S) * stride] = c.x; ctr0[(3 * S + 1) * stride] = c.y; ctr0[(3 * S + 2) * stride] = c.z;
  F = frame_step<1>(F, cq[0], sq[0]); PG_PUT(0)
  F = frame_step<2>(F, cq[1], sq[1]); PG_PUT(1)
  F = frame_step<3>(F, cq[2], sq[2]); PG_PUT(2)
  F = frame_step<4>(F, cq[3], sq[3]); PG_PUT(3)
  F = frame_step<5>(F, cq[4], sq[4]); PG_PUT(4)
  F = frame_step<6>(F, cq[5], sq[5]); PG_PUT(5)
  F = frame_step<7>(F, cq[6], sq[6]); PG_PUT(6) PG_PUT(7) PG_PUT(8)
#undef PG_PUT
}

// One collision link.  FULL: exact per-link minimum distance + closest pair (observation); otherwise only
// "is any distance <= 0", with bounding-sphere culling.  obs_pos: f32 [kMaxObst][3][n] (SoA), e = env.
template <int S, bool FULL>
__device__ __forceinline__ void slot_eval(const SceneParams& P, const float* __restrict__ obs_pos, int n, int e, const Frame& F,
                                          bool& hit, float& d_out, V3& u_out, CylBoxQueue* Q, const CullState* cull = nullptr, float turned = 0.f) {
  constexpr int K0 = slot_first(S), N = slot_count(S);
  using P0 = pgm::Prim<K0>;
  constexpr float R = SlotBound<K0, N>::value + 0.002f;
  float best = FULL ? 999.0f : 1e-6f;  // FULL: pyb_utils saturates at max_distance; flag-only: just the sign matters
  V3 dir = v3(0, 0, 0);
  V3 ctr = frame_point(F, P0::x, P0::y, P0::z);
  int nobs = P.nobs;
  bool culled = false;
  if (!FULL && cull) {  // temporal culling: the slot cannot have reached any obstacle yet
    const int st = cull->stride;
    const V3 d = ctr - v3(cull->ctr0[(3 * S) * st], cull->ctr0[(3 * S + 1) * st], cull->ctr0[(3 * S + 2) * st]);
    if (cull->slack[S * st] > norm(d) + R * turned) { nobs = 0; culled = true; }
  }
  for (int j = 0; j < nobs; j++) {
    V3 o = v3(obs_pos[(size_t)(j * 3 + 0) * n + e], obs_pos[(size_t)(j * 3 + 1) * n + e], obs_pos[(size_t)(j * 3 + 2) * n + e]);
#if PG_ENVBOX
    float bound;
    const Obst ob = scene_obst(P, j, e, n, bound);
#define PG_OB ob
#define PG_BOUND bound
#else  // the default step kernel: shapes read from the kernel parameters where they are used (no copies: measured, 1.7 % of k_step)
#define PG_OB P.obst[j]
#define PG_BOUND P.obst_bound[j]
#endif
    if (!FULL) {
      V3 d = o - ctr;
      float rr = R + PG_BOUND;
      if (dot(d, d) > rr * rr) continue;
      if (PG_OB.type == 0) { if (PrimLoop<K0, N>::hits(F, o, PG_OB)) hit = true; }
      else if (PrimLoop<K0, N>::hits_kind(F, o, PG_OB, 0) || PrimLoop<K0, N>::hits_kind(F, o, PG_OB, 1)) hit = true;
      continue;
    }
    PrimLoop<K0, N>::obstacle(F, o, PG_OB, best, dir, Q, j);
#undef PG_OB
#undef PG_BOUND
  }
  const int nst = culled ? 0 : P.nstatic;  // culled slots skip the static boxes too (d_link covers them)
  for (int j = 0; j < nst; j++) {
    V3 o;
    const Obst ob = static_obst(P, j, o);
    {  // the slot's bounding sphere against the box itself (a long thin shelf has a useless bounding sphere)
      const V3 rel = ctr - o;
      const V3 v = v3(rel.x - fminf(fmaxf(rel.x, -ob.sx), ob.sx), rel.y - fminf(fmaxf(rel.y, -ob.sy), ob.sy), rel.z - fminf(fmaxf(rel.z, -ob.sz), ob.sz));
      const float reach = FULL ? R + best : R;  // FULL: a box farther than the running minimum cannot lower it
      if (dot(v, v) > reach * reach) continue;
    }
    if (!FULL) {
      if (PrimLoop<K0, N>::hits_kind(F, o, ob, 0) || PrimLoop<K0, N>::hits_kind(F, o, ob, 1)) hit = true;
      continue;
    }
    PrimLoop<K0, N>::obstacle(F, o, ob, best, dir, Q, kMaxObst + j);
  }
  if (S >= 1 && P.has_table) {  // table: panda_link0 / panda_link1 ignored (reach_ao.py:898)
    bool near = true;
    if (!FULL) near = (ctr.z - R) <= (P.table_pos[2] + P.table_core[2] + P.table_margin + 0.002f);  // else: above the top plane
    if (near && PrimLoop<K0, N>::table(F, P) <= 0.f) hit = true;
  }
  if (FULL) { d_out = best; u_out = dir; }  // sign / saturation applied after the deferred queries (scene_pass)
}

// all 9 collision links at joint cosines/sines (cq, sq).  FULL fills d_link[9], u_link[9].
// SYNC: called by every thread of the block (see substep) with block barriers between groups of links.
template <bool FULL, bool SYNC>
__device__ __forceinline__ bool scene_pass(bool active, const SceneParams& P, const float* __restrict__ obs_pos, int n, int e, V3 base,
                                           const float* cq, const float* sq, float* d_link, V3* u_link, const CullState* cull = nullptr) {
  bool hit = false;
  float turned = 0.f;  // sum of |q_j - q0_j| over the joints up to the current frame
#define PG_TURN(j) if (!FULL && cull) turned += fabsf(cull->q[(j) * cull->stride] - cull->q0[(j) * cull->stride]);
  float d; V3 u;
  CylBoxQueue queue; queue.n = 0;
  CylBoxQueue* Q = FULL ? &queue : nullptr;
  Frame F = base_frame(base);
  if (active) {
    F = frame_step<1>(F, cq[0], sq[0]); PG_TURN(0) slot_eval<0, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[0] = d; u_link[0] = u; }
    F = frame_step<2>(F, cq[1], sq[1]); PG_TURN(1) slot_eval<1, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[1] = d; u_link[1] = u; }
    F = frame_step<3>(F, cq[2], sq[2]); PG_TURN(2) slot_eval<2, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[2] = d; u_link[2] = u; }
  }
  if (SYNC && (FULL || PG_SCENE_SYNC == 1 || (PG_SCENE_SYNC == 2 && P.has_box))) __syncthreads();
  if (active) {
    F = frame_step<4>(F, cq[3], sq[3]); PG_TURN(3) slot_eval<3, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[3] = d; u_link[3] = u; }
    F = frame_step<5>(F, cq[4], sq[4]); PG_TURN(4) slot_eval<4, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[4] = d; u_link[4] = u; }
  }
  if (SYNC && (FULL || PG_SCENE_SYNC == 1 || (PG_SCENE_SYNC == 2 && P.has_box))) __syncthreads();
  if (active) {
    F = frame_step<6>(F, cq[5], sq[5]); PG_TURN(5) slot_eval<5, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[5] = d; u_link[5] = u; }
    F = frame_step<7>(F, cq[6], sq[6]); PG_TURN(6)
    slot_eval<6, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[6] = d; u_link[6] = u; }
  }
  if (SYNC && (FULL || PG_SCENE_SYNC == 1 || (PG_SCENE_SYNC == 2 && P.has_box))) __syncthreads();
  if (active) {
    slot_eval<7, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[7] = d; u_link[7] = u; }
    slot_eval<8, FULL>(P, obs_pos, n, e, F, hit, d, u, Q, cull, turned); if (FULL) { d_link[8] = d; u_link[8] = u; }
  }
  if (FULL && active) {
    // drain the deferred cylinder-box queries, the whole warp together
    for (int it = 0;; it++) {
      const bool has = it < queue.n;
      if (!__any_sync(__activemask(), has)) break;
      if (has) {
        const CylBoxCand cd = queue.q[it];
        if (cd.lb < d_link[cd.slot]) {
          Obst ob; V3 o;
          if (cd.j >= kMaxObst) ob = static_obst(P, cd.j - kMaxObst, o);
          else {
#if PG_ENVBOX
            float bound;
            ob = scene_obst(P, cd.j, e, n, bound);
#else
            ob = P.obst[cd.j];
#endif
            o = v3(obs_pos[(size_t)(cd.j * 3 + 0) * n + e], obs_pos[(size_t)(cd.j * 3 + 1) * n + e], obs_pos[(size_t)(cd.j * 3 + 2) * n + e]);
          }
          Closest k = cyl_box<float>(cd.c, cd.a, cd.r, cd.h, o, v3(ob.sx - ob.margin, ob.sy - ob.margin, ob.sz - ob.margin), ob.margin);
          if (k.d < d_link[cd.slot]) { d_link[cd.slot] = k.d; u_link[cd.slot] = k.dir; }
        }
      }
    }
    // utils.unit_vector(pA, pB) = nan_to_num((pB - pA) / |pB - pA|) with pB - pA = d * dir
#pragma unroll
    for (int i = 0; i < 9; i++) {
      float b = d_link[i]; V3 dr = u_link[i];
      u_link[i] = b > 0.f ? dr : (b < 0.f ? v3(-dr.x, -dr.y, -dr.z) : v3(0, 0, 0));
      if (b <= 0.f) hit = true;
    }
  }
#undef PG_TURN
  if (SYNC && PG_SYNC_E) __syncthreads();
  return hit;
}

}  // namespace pg
