// pg_reset.cuh -- device code of the domain-randomised reset (ReachAO.reset and helpers, reach_ao.py:965-1211).
#pragma once
#include "pg_kernels.cuh"
#include "pg_neutral.h"
#include "pg_ziggurat.h"
#include "../../include/panda_b200.h"

namespace pg {

// ------------------------------------------------------------------------------------------------ reset
// sphere (sx = r) or axis-aligned box (half extents); margin < 0: Bullet's safe margin rule (createCollisionShape bodies), else
// the given collision margin (boxes imported from a scene URDF: 0.001)
struct Body { int type; double px, py, pz, sx, sy, sz; double margin = -1.0; };

__device__ inline double box_margin_d(double sx, double sy, double sz) { double m = fmin(sx, fmin(sy, sz)); return fmin(0.04, 0.1 * m); }

// min distance robot (all 36 primitives at the neutral pose, base offset) vs body: body-level getClosestPoints
// cutoff: pairs whose analytic lower bound exceeds it are skipped (the caller only compares the result to it)
// WARP: the 32 lanes of a warp work on the same environment and split the primitives (then min-reduce)
// (c0, s0) = cos / sin of joint 1 relative to the neutral pose (set_random_robot_base): joint 1 turns about the base's
// vertical axis and the base link's primitives are symmetric about it, so the turned robot against a SPHERE body equals
// the neutral robot against the body turned back by the same angle
// fp64 forward kinematics of the seven moving links at arbitrary joint values (robot frame: base at the origin), for the
// start-pose randomisers that put the arm anywhere (reach_ao.py:782-817).  Frame 0 = base, i = child of revolute joint i;
// R row-major link -> robot.
struct PoseD { double R[8][9]; double o[8][3]; };
__device__ inline void pose_fk(const double* q, PoseD& F) {
  for (int k = 0; k < 9; k++) F.R[0][k] = (k % 4 == 0) ? 1.0 : 0.0;
  F.o[0][0] = F.o[0][1] = F.o[0][2] = 0.0;
  for (int i = 0; i < 7; i++) {
    const double* J = pgm::kJointOrigin[i];
    const double c = cos(q[i]), s = sin(q[i]);
    double E[9];  // child -> parent: origin rotation * Rz(q)
    for (int r = 0; r < 3; r++) {
      const double a = J[3 + 3 * r], b = J[4 + 3 * r];
      E[3 * r] = a * c + b * s; E[3 * r + 1] = b * c - a * s; E[3 * r + 2] = J[5 + 3 * r];
    }
    const double* Rp = F.R[i];
    double* Rc = F.R[i + 1];
    for (int r = 0; r < 3; r++)
      for (int k = 0; k < 3; k++) Rc[3 * r + k] = Rp[3 * r] * E[k] + Rp[3 * r + 1] * E[3 + k] + Rp[3 * r + 2] * E[6 + k];
    for (int r = 0; r < 3; r++) F.o[i + 1][r] = F.o[i][r] + Rp[3 * r] * J[0] + Rp[3 * r + 1] * J[1] + Rp[3 * r + 2] * J[2];
  }
}
__device__ inline void pose_point(const PoseD& F, int frame, const double* l, double* out) {
  const double* R = F.R[frame];
  for (int r = 0; r < 3; r++) out[r] = F.o[frame][r] + R[3 * r] * l[0] + R[3 * r + 1] * l[1] + R[3 * r + 2] * l[2];
}

// min distance robot (all 36 primitives, base offset) vs body: body-level getClosestPoints
// cutoff: pairs whose analytic lower bound exceeds it are skipped (the caller only compares the result to it)
// WARP: the 32 lanes of a warp work on the same environment and split the primitives (then min-reduce)
// pose == nullptr: the robot stands at the reset pose of the stage (Params::nprims), turned about joint 1 by (c0, s0) =
// cos / sin of joint 1 relative to that pose (set_random_robot_base): joint 1 turns about the base's vertical axis and the
// base link's primitives are symmetric about it, so the turned robot against a SPHERE body equals the neutral robot against
// the body turned back by the same angle.  pose != nullptr: the primitives are placed by the given link frames, and only those
// carried by frames 0 .. frame_max count (frame_max = 1: the base link and panda_link1, the column no start pose moves).
// GPOSE = false compiles the general-pose path out (the reset kernels of stages without such a randomizer: the hot auto-reset).
template <bool WARP, bool GPOSE = false>
__device__ __noinline__ double robot_body_distance(const Params& P, const Body& b, double cutoff, double c0 = 1.0, double s0 = 0.0,
                                                   const PoseD* pose = nullptr, int frame_max = 7) {
  double best = 1e300;
  const int k0 = WARP ? (threadIdx.x & 31) : 0, kstep = WARP ? 32 : 1;
  const double ox = b.px - P.based[0], oy = b.py - P.based[1];
  const D3 o_turned = v3t<double>(c0 * ox + s0 * oy, -s0 * ox + c0 * oy, b.pz - P.based[2]);
  const D3 o_fixed = v3t<double>(ox, oy, b.pz - P.based[2]);
  double m = b.type == 1 ? (b.margin >= 0.0 ? b.margin : box_margin_d(b.sx, b.sy, b.sz)) : 0.0;
  D3 core = v3t<double>(b.sx - m, b.sy - m, b.sz - m);
  for (int k = k0; k < pgm::kNumNeutralPrim; k += kstep) {
    // table order: primitives of the moving links (panda_link1's first: 0..2), then the base link's (kNumPrim ..): the base does
    // not turn with joint 1, so it meets the body where it is
    const bool base_prim = k >= pgm::kNumPrim;
    if (!(GPOSE && pose) && frame_max < 7 && !(k < 3 || base_prim)) continue;
    const D3 o = base_prim ? o_fixed : o_turned;
    pgm::NeutralPrim np;
    if (GPOSE && pose) {
      const pgm::LocalPrim& lp = pgm::kLocalPrims[k];
      if (lp.frame > frame_max) continue;
      np.type = lp.type; np.r = lp.r; np.h = lp.h;
      pose_point(*pose, lp.frame, lp.c, np.c);
      const double* R = pose->R[lp.frame];
      for (int r = 0; r < 3; r++) np.a[r] = R[3 * r] * lp.a[0] + R[3 * r + 1] * lp.a[1] + R[3 * r + 2] * lp.a[2];
    } else np = P.nprims[k];
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
        d = cyl_box_decide<double>(c, a, np.r, np.h, o, core, m, cutoff);  // every caller only compares the result with `cutoff`
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
  double m = x.margin >= 0.0 ? x.margin : box_margin_d(x.sx, x.sy, x.sz);
  double d = sphere_box<double>(v3t<double>(s.px, s.py, s.pz), s.sx, v3t<double>(x.px, x.py, x.pz), v3t<double>(x.sx - m, x.sy - m, x.sz - m), m).d;
  return margin >= d;
}

// The part of the robot no start pose moves: the base link, panda_link1 (colliders on joint 1's axis) and the ball of radius 0.06
// about the shoulder (origin of panda_link2: panda_link2's cylinder, r 0.06 / half length 0.06 centred there, contains it whatever
// joints 1 and 2 do).  An obstacle within `margin` of this set is within `margin` of the robot in EVERY pose (oracle twin:
// pgo_env.c robot_core_within).
template <bool WARP, bool GPOSE>
__device__ inline bool robot_core_within(const Params& P, const Body& b, double margin, const PoseD* pose) {
  if (margin >= robot_body_distance<WARP, GPOSE>(P, b, margin, 1.0, 0.0, pose, 1)) return true;
  Body sh; sh.type = 0; sh.sx = 0.06; sh.sy = sh.sz = 0.0;
  sh.px = P.based[0]; sh.py = P.based[1]; sh.pz = P.based[2] + pgm::kJointOrigin[0][2];
  return body_body_collision(sh, b, margin);
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
  if (P.st.random_size && b.type == 1) {  // reachao_rand_shape: this environment's own cuboid
    b.sx = P.env_size[(size_t)(j * 3 + 0) * P.n + e]; b.sy = P.env_size[(size_t)(j * 3 + 1) * P.n + e]; b.sz = P.env_size[(size_t)(j * 3 + 2) * P.n + e];
  }
  return b;
}

// ------------------------------------------------------------------------------------------------ numpy pieces
// Generator.standard_exponential(): numpy's 256-layer ziggurat; tables recovered from / verified against numpy itself
// (tools/gen_ziggurat.py).  No FMA contraction in the wedge test (parity with the host generator).
__device__ inline double pcg_std_exponential(Pcg& g) {
  for (;;) {
    uint64_t ri = pcg_next(g) >> 3;
    const int idx = (int)(ri & 0xff);
    ri >>= 8;
    const double x = __dmul_rn((double)ri, pgz::ZIG_we[idx]);
    if (ri < pgz::ZIG_ke[idx]) return x;
    const double u = pcg_double(g);
    if (idx == 0) return pgz::ZIG_R - log1p(-u);
    if (__dadd_rn(__dmul_rn(pgz::ZIG_fe[idx - 1] - pgz::ZIG_fe[idx], u), pgz::ZIG_fe[idx]) < exp(-x)) return x;
  }
}
// Generator.dirichlet(alpha=[1, 1, 1]): three standard_gamma(1.0) = standard_exponential draws times the reciprocal of their sum
__device__ inline void pcg_dirichlet111(Pcg& g, double* out) {
  double acc = 0.0;
  for (int j = 0; j < 3; j++) { out[j] = pcg_std_exponential(g); acc = acc + out[j]; }
  const double inv = 1.0 / acc;
  for (int j = 0; j < 3; j++) out[j] = out[j] * inv;
}
// sample_inside_torus (reach_ao.py:1213-1236): R 0.5, r 0.05, centre height 0.5
__device__ inline void sample_torus(Pcg& g, int front, double* out) {
  const double PI = 3.141592653589793;
  double theta = __dmul_rn(2.0 * PI, pcg_double(g));
  if (front) theta = pcg_uniform(g, -0.5 * PI, 0.5 * PI);
  const double phi = __dmul_rn(2.0 * PI, pcg_double(g));
  const double rad = __dmul_rn(0.05, sqrt(pcg_double(g)));
  const double ring = __dadd_rn(0.5, __dmul_rn(rad, cos(phi)));
  out[0] = __dmul_rn(ring, cos(theta)); out[1] = __dmul_rn(ring, sin(theta)); out[2] = __dadd_rn(__dmul_rn(rad, sin(phi)), 0.5);
}

// ------------------------------------------------------------------------------------------------ rtb_ik
// glibc rand() of the environment's worker process: r[k] = r[k-31] + r[k-3] (TYPE_3), 31-word ring + front / back index
struct CRand { uint32_t r[31]; int f, b; };
__device__ inline int crand_next(CRand& c) {
  PG_CHECK(c.f >= 0 && c.f < 31 && c.b >= 0 && c.b < 31);
  c.r[c.f] += c.r[c.b];
  const int res = (int)(c.r[c.f] >> 1);
  c.f = c.f == 30 ? 0 : c.f + 1; c.b = c.b == 30 ? 0 : c.b + 1;
  return res;
}
// n x n linear solve, Gaussian elimination with partial pivoting, in place (n <= 7)
__device__ inline void solve_gauss(int n, double* A, double* b) {
  for (int c = 0; c < n; c++) {
    int piv = c;
    for (int r = c + 1; r < n; r++) if (fabs(A[r * n + c]) > fabs(A[piv * n + c])) piv = r;
    if (piv != c) { for (int k = 0; k < n; k++) { double t = A[c * n + k]; A[c * n + k] = A[piv * n + k]; A[piv * n + k] = t; } double t = b[c]; b[c] = b[piv]; b[piv] = t; }
    for (int r = c + 1; r < n; r++) {
      const double f = A[r * n + c] / A[c * n + c];
      for (int k = c; k < n; k++) A[r * n + k] -= f * A[c * n + k];
      b[r] -= f * b[c];
    }
  }
  for (int r = n - 1; r >= 0; r--) { double s = b[r]; for (int k = r + 1; k < n; k++) s -= A[r * n + k] * b[k]; b[r] = s / A[r * n + r]; }
}
// Link frames down the chain with everything at compile-time offsets (registers, no indexed local memory): joint axes
// ax[i] (z of frame i + 1), joint origins org[i], the final link-7 rotation R and origin o.
__device__ __forceinline__ void chain7(const double* q, double (&ax)[7][3], double (&org)[7][3], double (&R)[9], double (&o)[3]) {
  R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
  o[0] = o[1] = o[2] = 0.0;
#pragma unroll
  for (int i = 0; i < 7; i++) {
    const double* J = pgm::kJointOrigin[i];
    double s, c;
    sincos(q[i], &s, &c);
    double E[9], Rn[9];
#pragma unroll
    for (int r = 0; r < 3; r++) { E[3 * r] = J[3 + 3 * r] * c + J[4 + 3 * r] * s; E[3 * r + 1] = J[4 + 3 * r] * c - J[3 + 3 * r] * s; E[3 * r + 2] = J[5 + 3 * r]; }
#pragma unroll
    for (int r = 0; r < 3; r++) {
      o[r] += R[3 * r] * J[0] + R[3 * r + 1] * J[1] + R[3 * r + 2] * J[2];
#pragma unroll
      for (int k = 0; k < 3; k++) Rn[3 * r + k] = R[3 * r] * E[k] + R[3 * r + 1] * E[3 + k] + R[3 * r + 2] * E[6 + k];
    }
#pragma unroll
    for (int k = 0; k < 9; k++) R[k] = Rn[k];
    ax[i][0] = R[2]; ax[i][1] = R[5]; ax[i][2] = R[8];
    org[i][0] = o[0]; org[i][1] = o[1]; org[i][2] = o[2];
  }
}
// One search of ik_lm_chan from q (in / out): up to 100 Levenberg-Marquardt steps q += (J^T J + E I)^-1 J^T e.  Returns 1 when it
// stopped with E < 0.01 inside the toolbox model's joint limits (q wrapped into (-pi, pi]), 0 otherwise.  The 7 x 7 system is
// symmetric positive definite: solved by an unrolled Cholesky factorisation (the oracle eliminates with partial pivoting -- the
// same solution to rounding).
static __device__ __noinline__ int lm_search(double* q, const double* target, const double* Rt) {
  const double lo[7] = {-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973};
  const double hi[7] = {2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973};
  const double PI = 3.141592653589793;
  for (int it = 1; it <= 100; it++) {
    double ax[7][3], org[7][3], R7[9], o7[3];
    chain7(q, ax, org, R7, o7);
    double fl[3], e[6], R[9];
#pragma unroll
    for (int r = 0; r < 3; r++) fl[r] = o7[r] + R7[3 * r] * pgm::kLocalFlange[0] + R7[3 * r + 1] * pgm::kLocalFlange[1] + R7[3 * r + 2] * pgm::kLocalFlange[2];
#pragma unroll
    for (int k = 0; k < 3; k++) e[k] = target[k] - fl[k];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int j = 0; j < 3; j++) R[3 * i + j] = Rt[3 * i] * R7[3 * j] + Rt[3 * i + 1] * R7[3 * j + 1] + Rt[3 * i + 2] * R7[3 * j + 2];
    const double li[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
    const double ln = sqrt(li[0] * li[0] + li[1] * li[1] + li[2] * li[2]), tr = R[0] + R[4] + R[8];
    if (ln < 1e-6) { e[3] = tr > 0 ? 0.0 : PI / 2 * (R[0] + 1); e[4] = tr > 0 ? 0.0 : PI / 2 * (R[4] + 1); e[5] = tr > 0 ? 0.0 : PI / 2 * (R[8] + 1); }
    else { const double ang = atan2(ln, tr - 1) / ln; e[3] = ang * li[0]; e[4] = ang * li[1]; e[5] = ang * li[2]; }
    double E = 0;
#pragma unroll
    for (int k = 0; k < 6; k++) E += 0.5 * e[k] * e[k];
    if (E < 0.01) {
      int ok = 1;
#pragma unroll
      for (int d = 0; d < 7; d++) { q[d] = fmod(q[d] + PI, 2 * PI) - PI; if (q[d] < lo[d] || q[d] > hi[d]) ok = 0; }
      return ok;
    }
    double J[6][7];
#pragma unroll
    for (int d = 0; d < 7; d++) {
      const double r0 = fl[0] - org[d][0], r1 = fl[1] - org[d][1], r2 = fl[2] - org[d][2];
      J[0][d] = ax[d][1] * r2 - ax[d][2] * r1; J[1][d] = ax[d][2] * r0 - ax[d][0] * r2; J[2][d] = ax[d][0] * r1 - ax[d][1] * r0;
      J[3][d] = ax[d][0]; J[4][d] = ax[d][1]; J[5][d] = ax[d][2];
    }
    double L[7][7], g[7];  // lower triangle of J^T J + E I, factorised in place
#pragma unroll
    for (int i = 0; i < 7; i++) {
#pragma unroll
      for (int j = 0; j <= i; j++) {
        double s = (i == j) ? E : 0.0;
#pragma unroll
        for (int k = 0; k < 6; k++) s += J[k][i] * J[k][j];
        L[i][j] = s;
      }
      double s = 0;
#pragma unroll
      for (int k = 0; k < 6; k++) s += J[k][i] * e[k];
      g[i] = s;
    }
#pragma unroll
    for (int j = 0; j < 7; j++) {
      double d = L[j][j];
#pragma unroll
      for (int k = 0; k < j; k++) d -= L[j][k] * L[j][k];
      const double inv = rsqrt(d);
      L[j][j] = inv;  // 1 / L_jj
#pragma unroll
      for (int i = j + 1; i < 7; i++) {
        double v = L[i][j];
#pragma unroll
        for (int k = 0; k < j; k++) v -= L[i][k] * L[j][k];
        L[i][j] = v * inv;
      }
    }
#pragma unroll
    for (int i = 0; i < 7; i++) {  // L y = g
      double v = g[i];
#pragma unroll
      for (int k = 0; k < i; k++) v -= L[i][k] * g[k];
      g[i] = v * L[i][i];
    }
#pragma unroll
    for (int i = 6; i >= 0; i--) {  // L^T x = y
      double v = g[i];
#pragma unroll
      for (int k = i + 1; k < 7; k++) v -= L[k][i] * g[k];
      g[i] = v * L[i][i];
    }
#pragma unroll
    for (int d = 0; d < 7; d++) q[d] += g[d];
  }
  return 0;
}
// Panda.rtb_ik (panda.py:220-224) = roboticstoolbox 1.1.0 ERobot.ik_lm_chan(Tep, ilimit=100, slimit=100, tol=0.01, q0=neutral):
// Tep = fkine(q_seed) moved to `target`; up to 100 searches (lm_search), the first from q_seed, each later one from a random
// configuration q_i = lo_i + rand() / RAND_MAX * (hi_i - lo_i) of the worker process's C rand() stream (7 draws per restart, also
// after the last failed search).  Robot frame (the toolbox model has no base offset).  Returns 0 solved, 1 solved after a
// restart, 2 unsolved (q = the 100th random configuration).  [UPSTREAM-RECALL, parity unpinned: oracle/pgo_dyn.c]
// WARP: the 32 lanes of a warp hold the same arguments; restart s draws words 7 (s - 2) .. 7 (s - 2) + 6 of the stream whatever
// happened before it, so the lanes run 32 restarts at a time and the lowest-numbered success wins -- the sequential result at
// 1 / 32 of the latency (an unreachable target costs 5 search rounds instead of 100).
__device__ inline void restart_config(CRand& c, double* q) {
  const double lo[7] = {-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973};
  const double hi[7] = {2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973};
  for (int d = 0; d < 7; d++) q[d] = lo[d] + (double)crand_next(c) / 2147483647.0 * (hi[d] - lo[d]);
}
template <bool WARP>
__device__ __noinline__ int ik_lm_chan(const double* q_seed, const double* target, CRand& cr, double* q) {
  double Rt[9];
  {
    double ax[7][3], org[7][3], o7[3];
    chain7(q_seed, ax, org, Rt, o7);  // panda_link8 shares link 7's orientation
  }
  for (int d = 0; d < 7; d++) q[d] = q_seed[d];
  if (lm_search(q, target, Rt)) return 0;
  if (!WARP) {
    for (int search = 2; search <= 100; search++) {
      restart_config(cr, q);
      if (lm_search(q, target, Rt)) return 1;
    }
    restart_config(cr, q);
    return 2;
  }
  const int lane = threadIdx.x & 31;
  for (int base = 2; base <= 100; base += 32) {
    const int s = base + lane;  // this lane's search
    CRand c2 = cr;
    for (int k = 0; k < 7 * lane; k++) crand_next(c2);
    double ql[7];
    int ok = 0;
    if (s <= 100) { restart_config(c2, ql); ok = lm_search(ql, target, Rt); }
    const unsigned won = __ballot_sync(0xffffffffu, ok);
    if (won) {
      const int w = __ffs(won) - 1;
      for (int d = 0; d < 7; d++) q[d] = __shfl_sync(0xffffffffu, ql[d], w);
      for (int k = 0; k < 7 * (w + 1); k++) crand_next(cr);
      return 1;
    }
    const int used = (100 - base + 1) < 32 ? (100 - base + 1) : 32;
    for (int k = 0; k < 7 * used; k++) crand_next(cr);
  }
  restart_config(cr, q);
  return 2;
}
// PyBullet.inverse_kinematics(link 11, position) (pybullet.py:465-493, position only): Bullet's damped least squares
// dq = (J^T J + 0.5 I)^-1 J^T e on the link-11 FRAME origin, <= 20 iterations, residual 1e-4, steps clamped to 45 degrees,
// seeded by the current joint angles; fp64 twin of oracle/pgo_dyn.c pgo_arm_ik.  target in the robot frame.
static __device__ __noinline__ void ik_bullet_position(const double* q_seed, const double* target, double* q) {
  const double lee[3] = {pgm::kLocalEEOrigin[0], pgm::kLocalEEOrigin[1], pgm::kLocalEEOrigin[2]};
  PoseD F;
  for (int d = 0; d < 7; d++) q[d] = q_seed[d];
  double diff = 1e30;
  for (int it = 0; it < 20 && diff > 1e-4; it++) {
    pose_fk(q, F);
    double p[3], e[3], J[3][7];
    pose_point(F, 7, lee, p);
    for (int k = 0; k < 3; k++) e[k] = target[k] - p[k];
    diff = sqrt(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
    for (int d = 0; d < 7; d++) {
      const double ax[3] = {F.R[d + 1][2], F.R[d + 1][5], F.R[d + 1][8]};
      const double r[3] = {p[0] - F.o[d + 1][0], p[1] - F.o[d + 1][1], p[2] - F.o[d + 1][2]};
      J[0][d] = ax[1] * r[2] - ax[2] * r[1]; J[1][d] = ax[2] * r[0] - ax[0] * r[2]; J[2][d] = ax[0] * r[1] - ax[1] * r[0];
    }
    double A[49], b[7];
    for (int i = 0; i < 7; i++) {
      for (int j = 0; j < 7; j++) A[i * 7 + j] = J[0][i] * J[0][j] + J[1][i] * J[1][j] + J[2][i] * J[2][j] + (i == j ? 0.5 : 0.0);
      b[i] = J[0][i] * e[0] + J[1][i] * e[1] + J[2][i] * e[2];
    }
    solve_gauss(7, A, b);
    double mx = 0;
    for (int i = 0; i < 7; i++) mx = fmax(mx, fabs(b[i]));
    const double cap = 45.0 * 3.141592653589793 / 180.0;
    if (mx > cap) for (int i = 0; i < 7; i++) b[i] *= cap / mx;
    for (int i = 0; i < 7; i++) q[i] += b[i];
  }
}

// The scenario's robot_pose_randomizer other than set_random_robot_base (pg_pose_randomizer, reach_ao.py:782-817 and the
// scene builders): joint values into qpose, link frames into pose.  Every lane of a warp-cooperative reset runs it on the same
// streams (same result in every lane).  Returns reset-status bits: 8 = an unbounded rejection loop of the reference stopped
// after 1000 tries, 16 = rtb_ik found no solution.
template <bool WARP>
__device__ __noinline__ int randomize_pose(const Params& P, int e, Pcg& g, CRand& cr, double* qpose, PoseD& pose, const double* q_cur,
                                           int tries_done = 0) {  // tries_done: set_robot_random_pose tries this call has already made
  const StageDev& st = P.st;
  const double* seed = P.nextra + 6;  // Panda.neutral_joint_values of the stage (q0 and the orientation target of rtb_ik)
  int status = 0;
  double t[3];
  // rtb_ik takes the world-frame target as is: the toolbox model has no base offset (ReachAO's base sits at the origin anyway)
  switch (st.pose_kind) {
    case PG_POSE_IK_GOAL_SPACE:
      sample_goal(st, g, t);
      if (ik_lm_chan<WARP>(seed, t, cr, qpose) == 2) status |= 16;
      break;
    case PG_POSE_IK_BOX:
      for (int k = 0; k < 3; k++) t[k] = pcg_uniform(g, st.pose_lo[k], st.pose_hi[k]);
      if (ik_lm_chan<WARP>(seed, t, cr, qpose) == 2) status |= 16;
      break;
    case PG_POSE_BULLET_IK_BOX: {  // calculateInverseKinematics starts from the joints the robot has right now (q_cur)
      for (int k = 0; k < 3; k++) t[k] = pcg_uniform(g, st.pose_lo[k], st.pose_hi[k]);
      double q_now[7];
      for (int d = 0; d < 7; d++) q_now[d] = q_cur[d];
      for (int k = 0; k < 3; k++) t[k] -= P.based[k];
      ik_bullet_position(q_now, t, qpose);
      break; }
    case PG_POSE_IK_SPHERE: {      // until check_collided(0.0) is false, against the obstacles where the last episode left them
      for (int i = 0;; i++) {
        sample_hollow(g, st.pose_lo[0], st.pose_hi[0], 1, 0, 0, t);
        if (ik_lm_chan<WARP>(seed, t, cr, qpose) == 2) status |= 16;
        float cq[7], sq[7];
        for (int d = 0; d < 7; d++) { cq[d] = (float)cos(qpose[d]); sq[d] = (float)sin(qpose[d]); }
        const bool hit = scene_pass<false, false>(true, P.scene, P.obs_pos_f, P.n, e, v3(P.base[0], P.base[1], P.base[2]), cq, sq, nullptr, nullptr);
        if (!hit) break;
        if (i >= 999) { status |= 8; break; }
      }
      break; }
    default: {                     // set_robot_random_pose: until the EE height is inside [0.4, 0.6]
      for (int i = tries_done;; i++) {
        sample_torus(g, st.pose_kind == PG_POSE_TORUS_FRONT, t);
        if (ik_lm_chan<WARP>(seed, t, cr, qpose) == 2) status |= 16;
        pose_fk(qpose, pose);
        double ee[3];
        pose_point(pose, 7, pgm::kLocalEE, ee);
        const double z = ee[2] + P.based[2];
        if (0.6 >= z && z >= 0.4) break;
        if (i >= 999) { status |= 8; break; }
      }
      break; }
  }
  pose_fk(qpose, pose);
  return status;
}

// ------------------------------------------------------------------------------------------------ reset observation, one warp
// The observation a warp-cooperative reset ends with (ReachAO with obstacles; arm at rest, cached pose = current pose).  observe()
// runs the unrolled per-link scene pass on every lane alike: 52 of the auto-reset kernel's 100 us, all of it instruction fetch
// (~2.5 k instructions walked once).  Here the 9 links x M obstacles are dealt out to the lanes, one (link, obstacle) pair each,
// and ONE loop over the run-time copy of the primitive table does the closest-distance work of every pair at once; the per-link
// minimum over the obstacles is gathered with shuffles.  Same closed forms as prim_vs_obstacle<K> (no deferred queue: a pair's
// cylinder-box GJK runs in place), so the numbers equal the unrolled pass to fp32 rounding.
__device__ __forceinline__ void prim_vs_obstacle_rt(const Frame& F, const PrimRT& pr, V3 o, const Obst& ob, float& best, V3& bdir) {
  const V3 c = frame_point(F, pr.x, pr.y, pr.z);
  Closest k;
  if (pr.type == 0) {
    if (ob.type == 0) k = sphere_sphere<float>(c, pr.r, o, ob.sx);
    else k = sphere_box<float>(c, pr.r, o, v3(ob.sx - ob.margin, ob.sy - ob.margin, ob.sz - ob.margin), ob.margin);
  } else {
    const V3 a = frame_dir(F, pr.ax, pr.ay, pr.az);
    if (ob.type == 0) k = cyl_sphere<float>(c, a, pr.r, pr.h, o, ob.sx);
    else {
      const V3 core = v3(ob.sx - ob.margin, ob.sy - ob.margin, ob.sz - ob.margin);
      CylBoxBoundT<float> bd = cyl_box_bound<float>(c - o, a, pr.r, pr.h, core);
      const float lb = bd.lb - kHullMargin - ob.margin;
      if (lb >= best) return;
      if (bd.exact) { best = lb; bdir = bd.lb > 0.f ? bd.u : v3(0.f, 0.f, 0.f); return; }
      k = cyl_box<float>(c, a, pr.r, pr.h, o, core, ob.margin);
    }
  }
  if (k.d < best) { best = k.d; bdir = k.dir; }
}
// all 32 lanes call it with the same arguments; returns the achieved goal (EE position) in ag on every lane
static __device__ __noinline__ void observe_reset_warp(const Params& P, int e, const float* q, const float* cq, const float* sq, float* ag) {
  const int n = P.n, lane = threadIdx.x & 31;
  const V3 base = v3(P.base[0], P.base[1], P.base[2]);
  Frame Fr[8];
  Fr[0] = base_frame(base);
  Fr[1] = frame_step<1>(Fr[0], cq[0], sq[0]); Fr[2] = frame_step<2>(Fr[1], cq[1], sq[1]); Fr[3] = frame_step<3>(Fr[2], cq[2], sq[2]);
  Fr[4] = frame_step<4>(Fr[3], cq[3], sq[3]); Fr[5] = frame_step<5>(Fr[4], cq[4], sq[4]); Fr[6] = frame_step<6>(Fr[5], cq[5], sq[5]);
  Fr[7] = frame_step<7>(Fr[6], cq[6], sq[6]);
  const V3 ee = frame_point(Fr[7], pgm::kEEx, pgm::kEEy, pgm::kEEz);
  ag[0] = ee.x; ag[1] = ee.y; ag[2] = ee.z;
  const int nb = P.scene.nobs + P.scene.nstatic, np = 9 * nb;
  float mybest = 999.0f;  // lanes 0..8: minimum of link `lane` over the obstacles (pyb_utils saturates at max_distance)
  V3 mydir = v3(0.f, 0.f, 0.f);
  for (int p0 = 0; p0 < np; p0 += 32) {
    const int p = p0 + lane;
    float best = 999.0f;
    V3 dir = v3(0.f, 0.f, 0.f);
    if (p < np) {
      const int s = p / nb, j = p - s * nb;
      Obst ob; V3 o;
      if (j < P.scene.nobs) {
        float bound;
        ob = scene_obst(P.scene, j, e, n, bound);
        o = v3(P.obs_pos_f[(size_t)(j * 3 + 0) * n + e], P.obs_pos_f[(size_t)(j * 3 + 1) * n + e], P.obs_pos_f[(size_t)(j * 3 + 2) * n + e]);
      } else ob = static_obst(P.scene, j - P.scene.nobs, o);
      const int k0 = slot_first(s), k1 = k0 + slot_count(s);
      PG_CHECK(s >= 0 && s < 9 && k0 >= 0 && k1 <= pgm::kNumPrim);
      for (int k = k0; k < k1; k++) {
        const PrimRT& pr = kPrimRT[k];
        PG_CHECK(pr.frame >= 0 && pr.frame < 8);
        prim_vs_obstacle_rt(Fr[pr.frame], pr, o, ob, best, dir);
      }
    }
    const int live = np - p0 < 32 ? np - p0 : 32;
    for (int src = 0; src < live; src++) {  // pair p0 + src belongs to link (p0 + src) / nb: its lane takes it if it is closer
      const float d = __shfl_sync(0xffffffffu, best, src);
      const float dx = __shfl_sync(0xffffffffu, dir.x, src), dy = __shfl_sync(0xffffffffu, dir.y, src), dz = __shfl_sync(0xffffffffu, dir.z, src);
      if ((p0 + src) / nb == lane && d < mybest) { mybest = d; mydir = v3(dx, dy, dz); }
    }
  }
  float* o = P.obs + (size_t)e * P.obs_dim;
  const int off = P.obs_js ? 20 : 6;
  if (lane == 0) {
    o[0] = ee.x; o[1] = ee.y; o[2] = ee.z; o[3] = 0.f; o[4] = 0.f; o[5] = 0.f;
    if (P.obs_js) for (int d = 0; d < 7; d++) { o[6 + d] = q[d]; o[13 + d] = 0.f; }
    for (int d = 0; d < 7; d++) P.q_dl[(size_t)d * n + e] = q[d];
    float* a = P.achieved + (size_t)e * 3;
    float* g = P.desired + (size_t)e * 3;
    a[0] = ee.x; a[1] = ee.y; a[2] = ee.z;
    g[0] = (float)P.goal[e]; g[1] = (float)P.goal[(size_t)n + e]; g[2] = (float)P.goal[(size_t)2 * n + e];
  }
  // utils.unit_vector(pA, pB) = nan_to_num((pB - pA) / |pB - pA|) with pB - pA = d * dir
  const V3 u = mybest > 0.f ? mydir : (mybest < 0.f ? v3(-mydir.x, -mydir.y, -mydir.z) : v3(0.f, 0.f, 0.f));
  float mn = lane < 9 ? mybest : 3.0e38f;
  for (int k = 16; k > 0; k >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, k));
  if (lane < 9) {
    const int i = lane;
    P.d_link[(size_t)i * n + e] = mybest;
    if (P.obstacle_obs == 1) { o[off + i] = mybest; o[off + 9 + 3 * i] = u.x; o[off + 10 + 3 * i] = u.y; o[off + 11 + 3 * i] = u.z; }
    else if (P.obstacle_obs == 2) { o[off + 3 * i] = u.x; o[off + 1 + 3 * i] = u.y; o[off + 2 + 3 * i] = u.z; }
    else if (P.obstacle_obs == 3) o[off + i] = mybest;
    else if (P.obstacle_obs == 5) {  // "vectors+past", first observation of the episode: the history is filled with it
      const float cur[3] = {u.x, u.y, u.z};
      for (int k = 0; k < 3; k++) {
        const int c = 3 * i + k;
        o[off + c] = cur[k]; o[off + 27 + c] = cur[k]; o[off + 54 + c] = cur[k];
        P.past_vec[(size_t)c * n + e] = cur[k]; P.past_vec[(size_t)(27 + c) * n + e] = cur[k];
      }
    } else if (i == 0) o[off] = mn;
  }
  __syncwarp();
}

// RobotTaskEnv.reset (core.py:298-308) = Panda.reset (panda.py:290-298) + Reach.reset (reach.py:63-78) /
// ReachAO.reset (reach_ao.py:965-1033) with its rejection-sampling helpers (:1101-1181)
//
// WARP = false: one thread per environment (full / masked resets).  WARP = true: auto-reset of the few environments
// k_step appended to the compact list, one WARP per environment: every lane runs the same (cheap, serial) sampling
// code on the same PCG64 state, the 36-primitive robot distance query is split across the lanes.  Grid-stride
// over the list.
}  // namespace pg
