// pg_reset.cuh -- device code of the domain-randomised reset (ReachAO.reset and helpers, reach_ao.py:965-1211).
#pragma once
#include "pg_kernels.cuh"
#include "pg_neutral.h"

namespace pg {

// ------------------------------------------------------------------------------------------------ reset
struct Body { int type; double px, py, pz, sx, sy, sz; };  // sphere (sx = r) or axis-aligned box (half extents)

__device__ inline double box_margin_d(double sx, double sy, double sz) { double m = fmin(sx, fmin(sy, sz)); return fmin(0.04, 0.1 * m); }

// min distance robot (all 36 primitives at the neutral pose, base offset) vs body: body-level getClosestPoints
// cutoff: pairs whose analytic lower bound exceeds it are skipped (the caller only compares the result to it)
// WARP: the 32 lanes of a warp work on the same environment and split the primitives (then min-reduce)
// (c0, s0) = cos / sin of joint 1 relative to the neutral pose (set_random_robot_base): joint 1 turns about the base's
// vertical axis and the base link's primitives are symmetric about it, so the turned robot against a SPHERE body equals
// the neutral robot against the body turned back by the same angle
template <bool WARP>
__device__ __noinline__ double robot_body_distance(const Params& P, const Body& b, double cutoff, double c0 = 1.0, double s0 = 0.0) {
  double best = 1e300;
  const int k0 = WARP ? (threadIdx.x & 31) : 0, kstep = WARP ? 32 : 1;
  const double ox = b.px - P.based[0], oy = b.py - P.based[1];
  D3 o = v3t<double>(c0 * ox + s0 * oy, -s0 * ox + c0 * oy, b.pz - P.based[2]);
  double m = b.type == 1 ? box_margin_d(b.sx, b.sy, b.sz) : 0.0;
  D3 core = v3t<double>(b.sx - m, b.sy - m, b.sz - m);
  for (int k = k0; k < pgm::kNumNeutralPrim; k += kstep) {
    const pgm::NeutralPrim& np = P.nprims[k];
    D3 c = v3t<double>(np.c[0], np.c[1], np.c[2]);
    double d;
    if (b.type == 0) {
      // sphere body: the caller only compares the result with `cutoff`, so a primitive whose bounding sphere (radius r for a
      // sphere, r + h + hull margin >= the true bound for a cylinder) stays farther than that is skipped on SQUARED distances --
      // no fp64 sqrt / divide on the common path (they were 35 % of this kernel's stall samples)
      const double dx = c.x - o.x, dy = c.y - o.y, dz = c.z - o.z;
      const double reach = cutoff + b.sx + np.r + (np.type == 0 ? 0.0 : np.h + (double)kHullMargin);
      if (dx * dx + dy * dy + dz * dz > reach * reach) continue;
    }
    if (np.type == 0) d = b.type == 0 ? sphere_sphere<double>(c, np.r, o, b.sx).d : sphere_box<double>(c, np.r, o, core, m).d;
    else {
      D3 a = v3t<double>(np.a[0], np.a[1], np.a[2]);
      if (b.type == 0) d = cyl_sphere<double>(c, a, np.r, np.h, o, b.sx).d;
      else {
        // closed form unless the closest feature is a slanted end cap (cyl_box_bound), GJK only there
        CylBoxBoundT<double> bd = cyl_box_bound<double>(c - o, a, np.r, np.h, core);
        const double lb = bd.lb - (double)kHullMargin - m;
        if (lb > cutoff || lb >= best) continue;
        if (bd.exact) { best = lb; continue; }
        d = cyl_box<double>(c, a, np.r, np.h, o, core, m).d;
      }
    }
    best = fmin(best, d);
  }
  if (WARP) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
  }
  return best;
}
// check_collision(bodyA, bodyB, margin) for the sphere / box bodies of the scene (reach_ao.py:1169-1181)
__device__ inline bool body_body_collision(const Body& a, const Body& b, double margin) {
  if (a.type == 0 && b.type == 0) {
    double dx = a.px - b.px, dy = a.py - b.py, dz = a.pz - b.pz;
    return margin >= sqrt(dx * dx + dy * dy + dz * dz) - a.sx - b.sx;
  }
  if (a.type == 1 && b.type == 1)  // Bullet's SAT box-box detector reports nothing for separated boxes
    return fabs(a.px - b.px) <= a.sx + b.sx && fabs(a.py - b.py) <= a.sy + b.sy && fabs(a.pz - b.pz) <= a.sz + b.sz;
  const Body& s = a.type == 0 ? a : b;
  const Body& x = a.type == 0 ? b : a;
  double m = box_margin_d(x.sx, x.sy, x.sz);
  double d = sphere_box<double>(v3t<double>(s.px, s.py, s.pz), s.sx, v3t<double>(x.px, x.py, x.pz), v3t<double>(x.sx - m, x.sy - m, x.sz - m), m).d;
  return margin >= d;
}

__device__ inline void sample_hollow(Pcg& g, double rmin, double rmax, int upper, int front, int q34, double* out) {
  const double PI = 3.141592653589793;
  double phi = pcg_uniform(g, 0.0, 2 * PI);
  double theta = upper ? pcg_uniform(g, 0.0, 0.5 * PI) : pcg_uniform(g, 0.0, PI);
  if (front) phi = pcg_uniform(g, -0.5 * PI, 0.5 * PI);
  if (q34) phi = pcg_uniform(g, -0.75 * PI, 0.75 * PI);
  double r = cbrt(pcg_uniform(g, rmin * rmin * rmin, rmax * rmax * rmax));
  out[0] = r * sin(theta) * cos(phi); out[1] = r * sin(theta) * sin(phi); out[2] = r * cos(theta);
}
__device__ inline void sample_goal(const StageDev& st, Pcg& g, double* out) {
  if (st.goal_sampler == 0) for (int k = 0; k < 3; k++) out[k] = pcg_uniform(g, st.goal_lo[k], st.goal_hi[k]);
  else sample_hollow(g, st.goal_rmin, st.goal_rmax, st.goal_upper, st.goal_front, st.goal_34, out);
}
__device__ inline void sample_obstacle(const Params& P, Pcg& g, const double* goal, const double* ee, double* out) {
  const StageDev& st = P.st;
  double s[3], c[3];
  if (st.obs_sampler == 0) { for (int k = 0; k < 3; k++) out[k] = pcg_uniform(g, st.goal_lo[k], st.goal_hi[k]); return; }
  double r = pcg_double(g);
  if (st.obs_sampler == 1) {
    if (r > 0.3) { sample_hollow(g, 0.1, 0.5, 0, 0, 0, s); for (int k = 0; k < 3; k++) c[k] = goal[k]; }
    else if (r > 0.1) { sample_hollow(g, 0.1, 0.4, 0, 0, 0, s); for (int k = 0; k < 3; k++) c[k] = ee[k]; }
    else { sample_hollow(g, 0.3, 0.6, 1, 0, 0, s); for (int k = 0; k < 3; k++) c[k] = ee[3 + k]; }  // get_link_position(0) = COM of panda_link1
  } else if (st.obs_sampler == 2) {
    sample_hollow(g, 0.1, 0.5, 0, 0, 0, s);
    for (int k = 0; k < 3; k++) c[k] = r > 0.5 ? goal[k] : ee[k];
  } else {
    if (r > 0.3) { sample_hollow(g, 0.2, 0.6, 0, 0, 0, s); for (int k = 0; k < 3; k++) c[k] = goal[k]; }
    else { sample_hollow(g, 0.2, 0.4, 0, 0, 0, s); for (int k = 0; k < 3; k++) c[k] = ee[k]; }
  }
  for (int k = 0; k < 3; k++) out[k] = s[k] + c[k];
}

__device__ inline Body obstacle_body(const Params& P, int j, int e) {
  Body b; b.type = P.st.obs_type[j];
  b.px = P.obs_pos[(size_t)(j * 3 + 0) * P.n + e]; b.py = P.obs_pos[(size_t)(j * 3 + 1) * P.n + e]; b.pz = P.obs_pos[(size_t)(j * 3 + 2) * P.n + e];
  b.sx = P.st.obs_size[j][0]; b.sy = P.st.obs_size[j][1]; b.sz = P.st.obs_size[j][2];
  return b;
}

// RobotTaskEnv.reset (core.py:298-308) = Panda.reset (panda.py:290-298) + Reach.reset (reach.py:63-78) /
// ReachAO.reset (reach_ao.py:965-1033) with its rejection-sampling helpers (:1101-1181)
//
// WARP = false: one thread per environment (full / masked resets).  WARP = true: auto-reset of the few environments
// k_step appended to the compact list, one WARP per environment: every lane runs the same (cheap, serial) sampling
// code on the same PCG64 state, the 36-primitive robot distance query is split across the lanes.  Grid-stride
// over the list.
}  // namespace pg
