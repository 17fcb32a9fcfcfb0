// pg_render.cuh -- headless render("rgb_array") as a CUDA ray-caster.
//
// Replaces PyBullet.render (panda_gym/pybullet.py:152-180) -> getCameraImage with the OpenGL renderer, which cannot exist
// on the target machines (no graphics stack).  Same camera: computeViewMatrixFromYawPitchRoll(target, distance, yaw,
// pitch, roll, upAxisIndex = 2) and computeProjectionMatrixFOV(fov 60, aspect width / height, near 0.1, far 100).  The
// scene is analytic: the plane and table boxes (pybullet.py:759-815), the robot's collision primitives (spheres and
// cylinders of franka_panda_custom_0; its visual meshes are not loaded), the obstacles (red spheres / orange cuboids,
// reach_ao.py:819-860), the Push object (green box) and the translucent goal marker.  One thread per pixel, nearest hit,
// Lambert shading with a fixed light; output RGBA8 [height][width][4] like getCameraImage's pixel array.
// Oracle: oracle/pgo_render.c (fp64 restatement of this file's mathematics).
#pragma once
#include "pg_kernels.cuh"
#include "pg_neutral.h"
#include "pg_launch.h"

namespace pg {


struct RPrim {
  int type;              // 0 sphere, 1 cylinder, 2 box
  int rot;               // box: 0 axis aligned, 1 rotated by the object's quaternion
  float c[3], a[3];      // centre; cylinder axis
  float r, h;            // radius, cylinder half length
  float e[3];            // box half extents
  float rgb[3];
};
constexpr int kMaxRPrims = 5 + pgm::kNumPrim + kMaxObst + kMaxStaticBoxes + 2;

constexpr float kLight[3] = {0.40824829f, -0.40824829f, 0.81649658f};  // normalize(1, -1, 2)
constexpr float kAmbient = 0.35f, kDiffuse = 0.65f;
constexpr float kBackground[3] = {0.7f, 0.7f, 0.8f};

struct Hit { float t; V3 n; };

__device__ __forceinline__ bool hit_sphere(V3 o, V3 d, V3 c, float r, float tmin, Hit& h) {
  V3 oc = o - c;
  float b = dot(oc, d), cc = dot(oc, oc) - r * r, disc = b * b - cc;
  if (disc < 0.f) return false;
  float sq = sqrtf(disc), t = -b - sq;
  if (t < tmin) t = -b + sq;
  if (t < tmin || t >= h.t) return false;
  h.t = t; h.n = (1.0f / r) * (oc + t * d);
  return true;
}
__device__ __forceinline__ bool hit_cylinder(V3 o, V3 d, V3 c, V3 a, float r, float hl, float tmin, Hit& h) {
  V3 oc = o - c;
  float dpar = dot(d, a), opar = dot(oc, a);
  V3 dperp = d - dpar * a, operp = oc - opar * a;
  float A = dot(dperp, dperp), B = dot(operp, dperp), C = dot(operp, operp) - r * r;
  bool any = false;
  if (A > 1e-12f) {
    float disc = B * B - A * C;
    if (disc >= 0.f) {
      float t = (-B - sqrtf(disc)) / A, z = opar + t * dpar;
      if (t >= tmin && t < h.t && fabsf(z) <= hl) { h.t = t; h.n = (1.0f / r) * (operp + t * dperp); any = true; }
    }
  }
  if (fabsf(dpar) > 1e-12f) {
#pragma unroll
    for (int s = 0; s < 2; s++) {
      float sg = s ? 1.f : -1.f, t = (sg * hl - opar) / dpar;
      V3 q = operp + t * dperp;
      if (t >= tmin && t < h.t && dot(q, q) <= r * r) { h.t = t; h.n = sg * a; any = true; }
    }
  }
  return any;
}
// box in its own frame: ol / dl = ray origin / direction in box coordinates; returns the local normal
__device__ __forceinline__ bool hit_box_local(V3 ol, V3 dl, V3 e, float tmin, float tmax, float& t_out, V3& n_out) {
  const float o3[3] = {ol.x, ol.y, ol.z}, d3[3] = {dl.x, dl.y, dl.z}, e3[3] = {e.x, e.y, e.z};
  float tn = -3.0e38f, tf = 3.0e38f;
  int an = 0, af = 0; float sn = 0.f, sf = 0.f;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    if (fabsf(d3[k]) < 1e-12f) { if (fabsf(o3[k]) > e3[k]) return false; continue; }
    float inv = 1.0f / d3[k], t1 = (-e3[k] - o3[k]) * inv, t2 = (e3[k] - o3[k]) * inv, s1 = -1.f, s2 = 1.f;
    if (t1 > t2) { float t = t1; t1 = t2; t2 = t; s1 = 1.f; s2 = -1.f; }
    if (t1 > tn) { tn = t1; an = k; sn = s1; }
    if (t2 < tf) { tf = t2; af = k; sf = s2; }
  }
  if (tn > tf) return false;
  float t = tn; int ax = an; float sg = sn;
  if (t < tmin) { t = tf; ax = af; sg = sf; }
  if (t < tmin || t >= tmax) return false;
  t_out = t;
  n_out = v3(ax == 0 ? sg : 0.f, ax == 1 ? sg : 0.f, ax == 2 ? sg : 0.f);
  return true;
}

__device__ __forceinline__ bool hit_prim(const RPrim& p, const Rot& R, V3 o, V3 d, float tmin, Hit& h) {
  V3 c = v3(p.c[0], p.c[1], p.c[2]);
  if (p.type == 0) return hit_sphere(o, d, c, p.r, tmin, h);
  if (p.type == 1) return hit_cylinder(o, d, c, v3(p.a[0], p.a[1], p.a[2]), p.r, p.h, tmin, h);
  V3 ol = o - c, dl = d;
  if (p.rot) { ol = rot_applyT(R, ol); dl = rot_applyT(R, d); }
  float t; V3 n;
  if (!hit_box_local(ol, dl, v3(p.e[0], p.e[1], p.e[2]), tmin, h.t, t, n)) return false;
  h.t = t; h.n = p.rot ? rot_apply(R, n) : n;
  return true;
}

__device__ __forceinline__ void shade(const float* rgb, V3 n, float* out) {
  float l = fmaxf(0.f, n.x * kLight[0] + n.y * kLight[1] + n.z * kLight[2]);
  float k = kAmbient + kDiffuse * l;
  out[0] = rgb[0] * k; out[1] = rgb[1] * k; out[2] = rgb[2] * k;
}

// RGBA8 image of environment `env`: [height][width][4]
__global__ void __launch_bounds__(256) k_render(const Params P, const Camera cam, int env, uint8_t* __restrict__ out) {
  __shared__ RPrim s_prims[kMaxRPrims];
  __shared__ RPrim s_goal;
  __shared__ float s_goal_alpha;
  __shared__ int s_n;
  __shared__ Frame s_F[8];
  __shared__ float s_q[4];
  const int t = threadIdx.y * blockDim.x + threadIdx.x, n = P.n;
  // ---- scene: the first threads place the primitives of environment `env`
  if (t < 7) {  // world frame of link t + 1
    float cq[7], sq[7];
    for (int d = 0; d < 7; d++) sincosf(P.q[(size_t)d * n + env], &sq[d], &cq[d]);
    Frame F = base_frame(v3(P.base[0], P.base[1], P.base[2]));
    F = frame_step<1>(F, cq[0], sq[0]); if (t >= 1) F = frame_step<2>(F, cq[1], sq[1]); if (t >= 2) F = frame_step<3>(F, cq[2], sq[2]);
    if (t >= 3) F = frame_step<4>(F, cq[3], sq[3]); if (t >= 4) F = frame_step<5>(F, cq[4], sq[4]); if (t >= 5) F = frame_step<6>(F, cq[5], sq[5]);
    if (t >= 6) F = frame_step<7>(F, cq[6], sq[6]);
    s_F[t + 1] = F;
  }
  __syncthreads();
  auto set = [](RPrim& p, int type, V3 c, V3 a, float r, float h, V3 e, float cr, float cg, float cb, int rot) {
    p.type = type; p.rot = rot; p.c[0] = c.x; p.c[1] = c.y; p.c[2] = c.z; p.a[0] = a.x; p.a[1] = a.y; p.a[2] = a.z; p.r = r; p.h = h;
    p.e[0] = e.x; p.e[1] = e.y; p.e[2] = e.z; p.rgb[0] = cr; p.rgb[1] = cg; p.rgb[2] = cb;
  };
  const V3 z3 = v3(0.f, 0.f, 0.f);
  // fixed layout: 0 plane, 1 table, 2..4 base link, 5..37 moving links, then obstacles / object
  if (t == 0) set(s_prims[0], 2, v3(0.f, 0.f, -0.41f), z3, 0.f, 0.f, v3(3.f, 3.f, 0.01f), 0.15f, 0.15f, 0.15f, 0);  // create_plane(z_offset = -0.4)
  if (t == 1) set(s_prims[1], 2, v3((float)P.table_pos[0], (float)P.table_pos[1], (float)P.table_pos[2]), z3, 0.f, 0.f,
                  v3((float)P.table_half[0], (float)P.table_half[1], (float)P.table_half[2]), 0.3f, 0.3f, 0.3f, 0);
  if (t >= 2 && t < 5) {  // panda_link0 (fixed to the base): the last three rows of the neutral-pose table
    const pgm::NeutralPrim& np = pgm::kNeutralPrims[pgm::kNumPrim + t - 2];
    set(s_prims[t], np.type, v3((float)np.c[0] + P.base[0], (float)np.c[1] + P.base[1], (float)np.c[2] + P.base[2]),
        v3((float)np.a[0], (float)np.a[1], (float)np.a[2]), (float)np.r, (float)np.h, z3, 1.f, 1.f, 1.f, 0);
  }
  if (t >= 5 && t < 5 + pgm::kNumPrim) {
    const PrimRT& pr = kPrimRT[t - 5];
    const Frame& F = s_F[pr.frame];
    set(s_prims[t], pr.type, frame_point(F, pr.x, pr.y, pr.z), frame_dir(F, pr.ax, pr.ay, pr.az), pr.r, pr.h, z3, 1.f, 1.f, 1.f, 0);
  }
  if (t == 255) {
    int k = 5 + pgm::kNumPrim;
    if (P.task == 1) {
      for (int j = 0; j < P.scene.nobs; j++) {
        V3 o = v3(P.obs_pos_f[(size_t)(j * 3 + 0) * n + env], P.obs_pos_f[(size_t)(j * 3 + 1) * n + env], P.obs_pos_f[(size_t)(j * 3 + 2) * n + env]);
        if (o.z < -50.f) continue;  // parked by set_random_num_obs (reach_ao.py:1062-1082)
        float bound;
        const Obst ob = scene_obst(P.scene, j, env, n, bound);  // per-environment cuboid sizes where the stage has them
        if (ob.type == 0) set(s_prims[k++], 0, o, z3, ob.sx, 0.f, z3, 1.f, 0.f, 0.f, 0);
        else set(s_prims[k++], 2, o, z3, 0.f, 0.f, v3(ob.sx, ob.sy, ob.sz), 1.f, 0.5f, 0.f, 0);
      }
    }
    if (P.task == 1) {  // benchmark scene (pg_set_scene): URDF default grey
      for (int j = 0; j < P.scene.nstatic; j++) {
        V3 o;
        const Obst ob = static_obst(P.scene, j, o);
        set(s_prims[k++], 2, o, z3, 0.f, 0.f, v3(ob.sx, ob.sy, ob.sz), 0.6f, 0.6f, 0.6f, 0);
      }
    }
    if (P.task >= 2) {
      set(s_prims[k++], 2, v3(P.obj[env], P.obj[(size_t)n + env], P.obj[(size_t)2 * n + env]), z3, 0.f, 0.f,
          v3(P.world.half[0], P.world.half[1], P.world.half[2]), 0.1f, 0.9f, 0.1f, 1);
      for (int c = 0; c < 4; c++) s_q[c] = P.obj[(size_t)(3 + c) * n + env];
    } else { s_q[0] = s_q[1] = s_q[2] = 0.f; s_q[3] = 1.f; }
    s_n = k;
    // the goal marker: a translucent sphere (reach.py:36-41, reach_ao.py:276-282) / box of the object's size (push.py:40-47)
    V3 g = v3((float)P.goal[env], (float)P.goal[(size_t)n + env], (float)P.goal[(size_t)2 * n + env]);
    if (P.task >= 2) set(s_goal, 2, g, z3, 0.f, 0.f, v3(P.world.half[0], P.world.half[1], P.world.half[2]), 0.1f, 0.9f, 0.1f, 0);
    else set(s_goal, 0, g, z3, 0.02f, 0.f, z3, 0.1f, 0.9f, 0.1f, 0);
    s_goal_alpha = P.task == 1 ? 0.8f : 0.3f;
  }
  __syncthreads();
  const int px = blockIdx.x * blockDim.x + threadIdx.x, py = blockIdx.y * blockDim.y + threadIdx.y;
  if (px >= cam.width || py >= cam.height) return;
  // ---- primary ray through the pixel centre
  const float sx = (2.f * (px + 0.5f) / cam.width - 1.f) * cam.aspect * cam.tan_half_fov;
  const float sy = (1.f - 2.f * (py + 0.5f) / cam.height) * cam.tan_half_fov;
  V3 d = v3(cam.fwd[0] + sx * cam.right[0] + sy * cam.up[0], cam.fwd[1] + sx * cam.right[1] + sy * cam.up[1],
            cam.fwd[2] + sx * cam.right[2] + sy * cam.up[2]);
  const float dl = norm(d);
  d = (1.0f / dl) * d;
  const V3 o = v3(cam.eye[0], cam.eye[1], cam.eye[2]);
  const float tmin = cam.near_plane * dl, tmax = cam.far_plane * dl;  // depth along the view axis in [near, far]
  const Rot R = quat_rot(s_q[0], s_q[1], s_q[2], s_q[3]);
  Hit h; h.t = tmax; h.n = v3(0.f, 0.f, 1.f);
  int best = -1;
  for (int k = 0; k < s_n; k++)
    if (hit_prim(s_prims[k], R, o, d, tmin, h)) best = k;
  float col[3] = {kBackground[0], kBackground[1], kBackground[2]};
  if (best >= 0) shade(s_prims[best].rgb, h.n, col);
  Hit hg; hg.t = h.t; hg.n = v3(0.f, 0.f, 1.f);
  const Rot I = quat_rot(0.f, 0.f, 0.f, 1.f);
  if (hit_prim(s_goal, I, o, d, tmin, hg)) {
    float gc[3];
    shade(s_goal.rgb, hg.n, gc);
    const float al = s_goal_alpha;
    for (int c = 0; c < 3; c++) col[c] = al * gc[c] + (1.f - al) * col[c];
  }
  uint8_t* p = out + ((size_t)py * cam.width + px) * 4;
  for (int c = 0; c < 3; c++) p[c] = (uint8_t)(fminf(fmaxf(col[c], 0.f), 1.f) * 255.f + 0.5f);
  p[3] = 255;
}

}  // namespace pg
