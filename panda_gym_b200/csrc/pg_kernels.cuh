// pg_kernels.cuh -- the batched environment kernels: step (set_action -> n_substeps x stepSimulation with
// per-substep collision check -> observation / success / truncation / reward), domain-randomised reset,
// observation refresh, vectorised compute_reward.  One environment per thread, SoA state in HBM.
#pragma once
#include <stdint.h>
#include "pg_scene.cuh"
#include "pg_contact.cuh"
#include "pg_neutral.h"

namespace pg {

constexpr int kBlock = 128;
// k_step launch shape.  BASELINE configs[1] runs 65 536 environments per GPU = 443 threads per SM on 148 SMs: the
// kernel must keep >= 14 warps resident per SM or a second, nearly empty wave doubles its latency-bound runtime.
// Default: 448 threads (14 warps) x 128 registers = one block per SM, 148 x 448 = 66 304 >= 65 536 -> a single
// wave.  (Each SM sub-partition owns 16 K registers: 4 resident warps per sub-partition cap a thread at 128.)
#ifndef PG_STEP_BLOCK
#define PG_STEP_BLOCK 448
#endif
#ifndef PG_STEP_MINB
#define PG_STEP_MINB 1
#endif
constexpr int kStepBlock = PG_STEP_BLOCK;
#ifdef PG_STEP_MAXREG
#define PG_STEP_BOUNDS __maxnreg__(PG_STEP_MAXREG)
#else
#define PG_STEP_BOUNDS __launch_bounds__(PG_STEP_BLOCK, PG_STEP_MINB)
#endif
#ifndef PG_STEP_SYNC
#define PG_STEP_SYNC 1
#endif
constexpr bool kStepSync = PG_STEP_SYNC != 0;  // block barriers between the phases of a substep (I-cache locality)

struct StageDev {  // device copy of pg_stage (scenario parameter block)
  int goal_sampler;
  double goal_lo[3], goal_hi[3], goal_rmin, goal_rmax;
  int goal_upper, goal_front, goal_34;
  int nobs, obs_type[kMaxObst];
  double obs_size[kMaxObst][3];
  int obs_sampler, randomize_obstacles, allow_overlap, random_num_obs, size_lo, size_hi;
  int random_base;   // randomize_robot_pose with set_random_robot_base (reach_ao.py:1238-1241)
  int max_episode_steps;
  float ee_thr, speed_thr;
};

struct Params {
  int n, task, control, reward_type, goal_condition, action_limiter, obs_js, obstacle_obs, obs_dim, act_dim, task_obs_dim;
  int n_substeps, truncate_on_collision, terminate_on_success, auto_reset;
  float collision_reward;
  double safety_distance;
  float base[3];
  double based[3];
  double table_pos[3], table_half[3];
  StageDev st;
  SceneParams scene;
  WorldParams world;             // Push / PickAndPlace (tasks 2, 3)
  double obj_half_z;             // object_size / 2 in fp64 for the reset arithmetic (push.py:22, :73)
  uint64_t base_seed;
  // SoA state
  float *q, *qd;                 // [7][n]
  double* goal;                  // [3][n]
  double* obs_pos;               // [kMaxObst][3][n]
  float* obs_pos_f;              // fp32 mirror read by the per-substep collision pass (refreshed by reset / set_obstacles)
  float *recent, *prev;          // [7][n] action history (panda.py:138-139)
  float *cur_vel, *cur_acc, *jerk; // [7][n] (panda.py:167-172)
  float* d_link;                 // [9][n]
  float* obj;                    // [13][n] object position, quaternion xyzw, linear / angular velocity (tasks 2, 3)
  float* past_vec;               // [2][27][n] previous two vector observations ("vectors+past", reach_ao.py:918-922)
  uint8_t *have_recent, *have_prev, *collided, *goal_reached, *need_reset;
  int *steps, *ep_len, *last_len;
  float *ep_ret, *last_ret;
  uint64_t* rng;                 // [4][n] PCG64 state hi, lo, inc hi, inc lo
  uint32_t* episode;             // resets so far
  // outputs
  float *obs, *achieved, *desired, *reward, *term_obs, *term_achieved;
  uint8_t *done, *flags;
  double* stats;                 // [8]
  int *reset_list, *reset_count;   // compacted indices of environments that finished in the last step
  // evaluation metrics of perform_benchmark (evaluation/evaluate.py:131-300), filled by k_eval_metrics when enabled
  float* ev_acc;                 // [4][n] running sums of the current episode: effort, manipulability, jerk norm, ee speed
  int* ev_len;                   // [n] steps of the current episode
  float* ev_ret;                 // [n] return of the current episode
  double* ev;                    // [12] see pg_get_eval_metrics
};

// panda.py:63 joint_forces
__device__ __forceinline__ float max_force(int d) { return d < 4 ? 87.0f : (d == 4 ? 12.0f : 120.0f); }

// ------------------------------------------------------------------------------------------------ RNG
// numpy Generator(PCG64(SeedSequence(seed))) == gymnasium seeding.np_random(seed) (core.py:302)
typedef unsigned __int128 u128;
struct Pcg { uint64_t s_hi, s_lo, i_hi, i_lo; int has32; uint32_t buf32; };
__device__ __forceinline__ uint32_t ss_hashmix(uint32_t v, uint32_t& hc) { v ^= hc; hc *= 0x931e8875u; v *= hc; v ^= v >> 16; return v; }
__device__ __forceinline__ uint32_t ss_mix(uint32_t x, uint32_t y) { uint32_t r = 0xca01f9ddu * x - 0x4973f715u * y; r ^= r >> 16; return r; }
__device__ __forceinline__ u128 pcg_mult() { return ((u128)0x2360ed051fc65da4ULL << 64) | 0x4385df649fccf645ULL; }
__device__ inline void pcg_seed(Pcg& g, uint64_t seed) {
  uint32_t ent0 = (uint32_t)seed, ent1 = (uint32_t)(seed >> 32);
  uint32_t pool[4], hc = 0x43b0d7e5u;
  pool[0] = ss_hashmix(ent0, hc); pool[1] = ss_hashmix(ent1, hc); pool[2] = ss_hashmix(0u, hc); pool[3] = ss_hashmix(0u, hc);
#pragma unroll
  for (int s = 0; s < 4; s++)
#pragma unroll
    for (int d = 0; d < 4; d++)
      if (s != d) pool[d] = ss_mix(pool[d], ss_hashmix(pool[s], hc));
  uint32_t w[8], hb = 0x8b51f9ddu;
#pragma unroll
  for (int i = 0; i < 8; i++) { uint32_t v = pool[i & 3]; v ^= hb; hb *= 0x58f38dedu; v *= hb; v ^= v >> 16; w[i] = v; }
  uint64_t st0 = (uint64_t)w[0] | ((uint64_t)w[1] << 32), st1 = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
  uint64_t st2 = (uint64_t)w[4] | ((uint64_t)w[5] << 32), st3 = (uint64_t)w[6] | ((uint64_t)w[7] << 32);
  u128 initstate = ((u128)st0 << 64) | st1, initseq = ((u128)st2 << 64) | st3;
  u128 inc = (initseq << 1) | 1, state = 0;
  state = state * pcg_mult() + inc;
  state += initstate;
  state = state * pcg_mult() + inc;
  g.s_hi = (uint64_t)(state >> 64); g.s_lo = (uint64_t)state; g.i_hi = (uint64_t)(inc >> 64); g.i_lo = (uint64_t)inc;
  g.has32 = 0; g.buf32 = 0;
}
__device__ __forceinline__ uint64_t pcg_next(Pcg& g) {
  u128 state = ((u128)g.s_hi << 64) | g.s_lo, inc = ((u128)g.i_hi << 64) | g.i_lo;
  state = state * pcg_mult() + inc;
  g.s_hi = (uint64_t)(state >> 64); g.s_lo = (uint64_t)state;
  uint64_t x = g.s_hi ^ g.s_lo;
  unsigned rot = (unsigned)(g.s_hi >> 58);
  return (x >> rot) | (x << ((64 - rot) & 63));
}
__device__ __forceinline__ double pcg_double(Pcg& g) { return (double)(pcg_next(g) >> 11) * (1.0 / 9007199254740992.0); }
// numpy random_uniform: off + range * next_double, no FMA contraction (bit parity with the host generator)
__device__ __forceinline__ double pcg_uniform(Pcg& g, double lo, double hi) { return __dadd_rn(lo, __dmul_rn(hi - lo, pcg_double(g))); }
__device__ __forceinline__ uint32_t pcg_u32(Pcg& g) {
  if (g.has32) { g.has32 = 0; return g.buf32; }
  uint64_t v = pcg_next(g);
  g.has32 = 1; g.buf32 = (uint32_t)(v >> 32);
  return (uint32_t)v;
}
__device__ inline uint32_t pcg_lemire32(Pcg& g, uint32_t rng) {  // Generator.integers, range rng inclusive
  uint32_t excl = rng + 1;
  uint64_t m = (uint64_t)pcg_u32(g) * excl;
  uint32_t left = (uint32_t)m;
  if (left < excl) {
    uint32_t thr = (0xFFFFFFFFu - rng) % excl;
    while (left < thr) { m = (uint64_t)pcg_u32(g) * excl; left = (uint32_t)m; }
  }
  return (uint32_t)(m >> 32);
}
__device__ inline uint32_t pcg_interval32(Pcg& g, uint32_t mx) {  // random_interval (Generator.shuffle)
  if (mx == 0) return 0;
  uint32_t mask = mx, v;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  while ((v = (pcg_u32(g) & mask)) > mx) {}
  return v;
}
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ULL; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL; x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL; return x ^ (x >> 31);
}
// seed of episode k of global environment g:  k = 0 -> base + g (SB3 VecEnv.seed convention: numpy-compatible
// first reset); later episodes (the reference re-seeds from OS entropy, core.py:302 with seed=None) -> hashed
__host__ __device__ __forceinline__ uint64_t episode_seed(uint64_t base, uint64_t g, uint32_t k) {
  return k == 0 ? base + g : splitmix64(base + g + 0x9E3779B97F4A7C15ULL * (uint64_t)k);
}

// ------------------------------------------------------------------------------------------------ observation
// utils.distance on float32 operands: float32 norm without FMA contraction, np.round(., 6) in float32
__device__ __forceinline__ float f32_distance(const float* a, const float* b) {
  float dx = __fsub_rn(a[0], b[0]), dy = __fsub_rn(a[1], b[1]), dz = __fsub_rn(a[2], b[2]);
  float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
  float d = __fsqrt_rn(s);
  return __fdiv_rn(rintf(__fmul_rn(d, 1e6f)), 1e6f);
}

// joint-space velocity recursion: angular velocity of link7 and linear velocity of its origin, link7 frame
template <int I>
__device__ __forceinline__ void vel_step(V3& w, V3& v, float c, float s, float qd) {
  using L = pgm::Link<I>;
  V3 vp = v + cross(w, v3(L::px, L::py, L::pz));
  w = to_child<L::rot>(w, c, s); w.z += qd;
  v = to_child<L::rot>(vp, c, s);
}

struct StepOut { float ag[3]; bool success; };

// RobotTaskEnv._get_obs (core.py:286-296) + Panda.get_obs (panda.py:264-288) + ReachAO.get_obs (reach_ao.py:902-959).
// q_fk: joint angles behind PyBullet's cached link transforms (one substep old, see oracle/pgo_dyn.c);
// cq/sq: cos/sin of the CURRENT q.
// fresh: first observation of an episode (the reference pre-fills the "vectors+past" history at reset).
__device__ __forceinline__ void observe(const Params& P, int e, const float* q, const float* qd, const float* q_fk, const float* cq,
                                        const float* sq, float* ag_out, bool fresh) {
  const int n = P.n;
  V3 base = v3(P.base[0], P.base[1], P.base[2]);
  // end-effector pose from the cached transforms
  float cf[7], sf[7];
#pragma unroll
  for (int d = 0; d < 7; d++) sincosf(q_fk[d], &sf[d], &cf[d]);
  Frame Fc = link7_frame(base, cf, sf);
  V3 ee = frame_point(Fc, pgm::kEEx, pgm::kEEy, pgm::kEEz);
  // link-local COM velocity of link 11 from the current state, rotated by the cached basis
  V3 w = v3(0, 0, 0), v = v3(0, 0, 0);
  vel_step<1>(w, v, cq[0], sq[0], qd[0]); vel_step<2>(w, v, cq[1], sq[1], qd[1]); vel_step<3>(w, v, cq[2], sq[2], qd[2]);
  vel_step<4>(w, v, cq[3], sq[3], qd[3]); vel_step<5>(w, v, cq[4], sq[4], qd[4]); vel_step<6>(w, v, cq[5], sq[5], qd[5]);
  vel_step<7>(w, v, cq[6], sq[6], qd[6]);
  V3 vl = v + cross(w, v3(pgm::kEEx, pgm::kEEy, pgm::kEEz));
  V3 ev = frame_dir(Fc, vl.x, vl.y, vl.z);
  float* o = P.obs + (size_t)e * P.obs_dim;
  o[0] = ee.x; o[1] = ee.y; o[2] = ee.z; o[3] = ev.x; o[4] = ev.y; o[5] = ev.z;
  int off = 6;
  if (P.obs_js) {
#pragma unroll
    for (int d = 0; d < 7; d++) { o[6 + d] = q[d]; o[13 + d] = qd[d]; }
    off = 20;
  }
  ag_out[0] = ee.x; ag_out[1] = ee.y; ag_out[2] = ee.z;
  if (P.task >= 2) {
    // fingers opening (panda.py:282-284): the loaded URDF has fixed finger joints, getJointState reports 0
    if (P.task == 3) o[off++] = 0.f;
    // Push.get_obs (push.py:49-63): object position, euler rotation (getEulerFromQuaternion), velocity, angular velocity
    float b[13];
#pragma unroll
    for (int k = 0; k < 13; k++) b[k] = P.obj[(size_t)k * n + e];
    float x = b[3], y = b[4], z = b[5], w = b[6], roll, pitch, yaw;
    float sarg = -2.f * (x * z - w * y);
    if (sarg <= -0.99999f) { roll = 0.f; pitch = -1.57079632679f; yaw = 2.f * atan2f(x, -y); }
    else if (sarg >= 0.99999f) { roll = 0.f; pitch = 1.57079632679f; yaw = 2.f * atan2f(-x, y); }
    else {
      roll = atan2f(2.f * (y * z + w * x), w * w - x * x - y * y + z * z);
      pitch = asinf(sarg);
      yaw = atan2f(2.f * (x * y + w * z), w * w + x * x - y * y - z * z);
    }
    o[off] = b[0]; o[off + 1] = b[1]; o[off + 2] = b[2]; o[off + 3] = roll; o[off + 4] = pitch; o[off + 5] = yaw;
#pragma unroll
    for (int k = 0; k < 6; k++) o[off + 6 + k] = b[7 + k];
    ag_out[0] = b[0]; ag_out[1] = b[1]; ag_out[2] = b[2];  // push.py:65-67
  }
  if (P.task == 1 && P.obstacle_obs != 0) {
    if (P.scene.nobs == 0) {  // reach_ao.py:904
      for (int k = 0; k < 27; k++) o[off + k] = 1.0f;
    } else {
      float dl[9]; V3 ul[9];
      scene_pass<true, false>(true, P.scene, P.obs_pos_f, n, e, base, cq, sq, dl, ul);
      float mn = dl[0];
#pragma unroll
      for (int i = 0; i < 9; i++) { P.d_link[(size_t)i * n + e] = dl[i]; mn = fminf(mn, dl[i]); }
      if (P.obstacle_obs == 1) {
#pragma unroll
        for (int i = 0; i < 9; i++) { o[off + i] = dl[i]; o[off + 9 + 3 * i] = ul[i].x; o[off + 10 + 3 * i] = ul[i].y; o[off + 11 + 3 * i] = ul[i].z; }
      } else if (P.obstacle_obs == 2) {
#pragma unroll
        for (int i = 0; i < 9; i++) { o[off + 3 * i] = ul[i].x; o[off + 1 + 3 * i] = ul[i].y; o[off + 2 + 3 * i] = ul[i].z; }
      } else if (P.obstacle_obs == 3) {
#pragma unroll
        for (int i = 0; i < 9; i++) o[off + i] = dl[i];
      } else if (P.obstacle_obs == 5) {  // [v(t-2) | v(t-1) | v(t)], history shifted by one
#pragma unroll
        for (int i = 0; i < 9; i++) {
          const float cur[3] = {ul[i].x, ul[i].y, ul[i].z};
#pragma unroll
          for (int k = 0; k < 3; k++) {
            const int c = 3 * i + k;
            float h2 = fresh ? cur[k] : P.past_vec[(size_t)c * n + e], h1 = fresh ? cur[k] : P.past_vec[(size_t)(27 + c) * n + e];
            o[off + c] = h2; o[off + 27 + c] = h1; o[off + 54 + c] = cur[k];
            P.past_vec[(size_t)c * n + e] = h1; P.past_vec[(size_t)(27 + c) * n + e] = cur[k];
          }
        }
      } else o[off] = mn;
    }
  }
  float* a = P.achieved + (size_t)e * 3;
  float* g = P.desired + (size_t)e * 3;
  a[0] = ag_out[0]; a[1] = ag_out[1]; a[2] = ag_out[2];
  g[0] = (float)P.goal[e]; g[1] = (float)P.goal[(size_t)n + e]; g[2] = (float)P.goal[(size_t)2 * n + e];
}

__device__ __forceinline__ float ee_speed(const Params& P, int e) {
  const float* o = P.obs + (size_t)e * P.obs_dim;
  return sqrtf(o[3] * o[3] + o[4] * o[4] + o[5] * o[5]);
}

// is_success: reach.py:80-82, reach_ao.py:1243-1261
__device__ __forceinline__ bool is_success(const Params& P, int e, const float* ag, const float* dg) {
  float d = f32_distance(ag, dg);
  if (P.task == 1 && P.goal_condition == 1) {
    bool gr = P.goal_reached[e] != 0;
    if (!gr) { gr = (d < P.st.ee_thr) && (ee_speed(P, e) < P.st.speed_thr); P.goal_reached[e] = gr; }
    return gr;
  }
  return d < P.st.ee_thr;
}

// compute_reward: reach.py:84-89, reach_ao.py:1308-1383; state-dependent terms read environment `src`
__device__ __forceinline__ float reward_fn(const Params& P, int src, const float* ag, const float* dg) {
  const int n = P.n;
  float d = f32_distance(ag, dg);
  float thr = P.st.ee_thr;
  if (P.task != 1) return P.reward_type == 0 ? -(float)(d > thr) : -d;  // reach.py:84-89, push.py:93-98, pick_and_place.py:91-96
  float coll = (float)P.collided[src];
  float r;
  if (P.reward_type == 0) {
    if (P.goal_condition == 0) r = -1.0f + (float)((d + coll) < thr);
    else r = 1.0f - (float)((d < thr) && (ee_speed(P, src) < P.st.speed_thr));
  } else {
    float effort = 0.f, jerk = 0.f, obs = 0.f;
    for (int k = 0; k < 7; k++) { float a = P.cur_acc[(size_t)k * n + src], j = P.jerk[(size_t)k * n + src]; effort += a * a; jerk += j * j; }
    effort = sqrtf(effort); jerk = sqrtf(jerk);
    for (int i = 0; i < 9; i++) obs += fmaxf(0.f, 1.f - P.d_link[(size_t)i * n + src] / 0.05f);
    if (P.reward_type == 1) r = -(10e-3f * d * d + logf(d * d + 10e-4f) + 0.1f * obs);
    else if (P.reward_type == 2) r = P.goal_condition == 0 ? -((float)(d > thr) * jerk) : (float)((d < thr) && (ee_speed(P, src) < P.st.speed_thr)) - jerk;
    else if (P.reward_type == 3) r = -(float)(d > thr) - effort;
    else r = expf(-20.f * d * d) - 0.005f * effort - 0.1f * obs;
  }
  if (P.truncate_on_collision && (P.reward_type == 0 || P.reward_type == 2 || P.reward_type == 3)) r += coll * P.collision_reward;
  return r;
}

// ------------------------------------------------------------------------------------------------ IK (ee control)
// calculateInverseKinematics(link 11, pos, quat (1,0,0,0)): damped least squares on the link-11 FRAME origin
// (not its COM), joint damping 0.5, <= 20 iterations, stop when |pos error| <= 1e-4 (oracle: pgo_arm_ik).
__device__ __forceinline__ void ik_dls(V3 base, const float* q0, V3 target, float* out) {
  float q[7];
#pragma unroll
  for (int d = 0; d < 7; d++) q[d] = q0[d];
  bool live = true;
  for (int it = 0; it < 20; it++) {
    float c[7], s[7];
#pragma unroll
    for (int d = 0; d < 7; d++) sincosf(q[d], &s[d], &c[d]);
    // joint axes (world z columns) and origins along the chain
    V3 ax[7], org[7];
    Frame F = base_frame(base);
    F = frame_step<1>(F, c[0], s[0]); ax[0] = F.z; org[0] = F.o;
    F = frame_step<2>(F, c[1], s[1]); ax[1] = F.z; org[1] = F.o;
    F = frame_step<3>(F, c[2], s[2]); ax[2] = F.z; org[2] = F.o;
    F = frame_step<4>(F, c[3], s[3]); ax[3] = F.z; org[3] = F.o;
    F = frame_step<5>(F, c[4], s[4]); ax[4] = F.z; org[4] = F.o;
    F = frame_step<6>(F, c[5], s[5]); ax[5] = F.z; org[5] = F.o;
    F = frame_step<7>(F, c[6], s[6]); ax[6] = F.z; org[6] = F.o;
    V3 p = frame_point(F, pgm::kEEox, pgm::kEEoy, pgm::kEEoz);
    V3 ep = target - p;
    if (!(norm(ep) > 1e-4f)) live = false;
    if (!__any_sync(__activemask(), live)) break;
    // orientation error: deltaQ = target * current^-1, target = (1,0,0,0) xyzw; link 11 axes = link7 * Rz(-45deg)
    V3 ex = pgm::kEErot_c * F.x + pgm::kEErot_s * F.y;   // link-11 x axis in world
    V3 ey = pgm::kEErot_c * F.y - pgm::kEErot_s * F.x;   // link-11 y axis
    V3 ez = F.z;
    // rotation matrix R (columns ex, ey, ez) -> quaternion (x, y, z, w)
    float qx, qy, qz, qw;
    {
      float m00 = ex.x, m11 = ey.y, m22 = ez.z, tr = m00 + m11 + m22;
      if (tr > 0.f) { float S = sqrtf(tr + 1.f) * 2.f; qw = 0.25f * S; qx = (ey.z - ez.y) / S; qy = (ez.x - ex.z) / S; qz = (ex.y - ey.x) / S; }
      else if (m00 > m11 && m00 > m22) { float S = sqrtf(1.f + m00 - m11 - m22) * 2.f; qw = (ey.z - ez.y) / S; qx = 0.25f * S; qy = (ey.x + ex.y) / S; qz = (ez.x + ex.z) / S; }
      else if (m11 > m22) { float S = sqrtf(1.f + m11 - m00 - m22) * 2.f; qw = (ez.x - ex.z) / S; qx = (ey.x + ex.y) / S; qy = 0.25f * S; qz = (ez.y + ey.z) / S; }
      else { float S = sqrtf(1.f + m22 - m00 - m11) * 2.f; qw = (ex.y - ey.x) / S; qx = (ez.x + ex.z) / S; qy = (ez.y + ey.z) / S; qz = 0.25f * S; }
    }
    // (1,0,0,0) * conj(q):  a = (1,0,0,0) xyzw, b = (-qx,-qy,-qz,qw)
    float dx = qw, dy = -qz, dz = qy, dw = qx;  // aw*b + ... with ax = 1: x = bw, y = -bz... see oracle formula
    // oracle: dq = {aw*bx + ax*bw + ay*bz - az*by, aw*by + ay*bw + az*bx - ax*bz, aw*bz + az*bw + ax*by - ay*bx, aw*bw - ax*bx - ay*by - az*bz}
    // with a = (1,0,0,0): dq = {bw, -bz, by, -bx} = {qw, qz, -qy, qx}
    dx = qw; dy = qz; dz = -qy; dw = qx;
    // angle * axis with angle = 2 acos(w) (btQuaternion::getAngle/getAxis); evaluated as 2 atan2(|xyz|, w), which
    // is the same number in exact arithmetic but well conditioned in fp32 for the small angles met here
    float sn = sqrtf(dx * dx + dy * dy + dz * dz);
    float angle = 2.f * atan2f(sn, dw);
    if (angle > 3.14159265358979f) angle -= 6.28318530717959f;
    V3 er = sn > 1e-12f ? (angle / sn) * v3(dx, dy, dz) : v3(0, 0, 0);
    // Jacobian columns: Jt_d = ax_d x (p - org_d), Jr_d = ax_d ; A = J^T J + 0.5 I ; b = J^T e
    V3 jt[7];
#pragma unroll
    for (int d = 0; d < 7; d++) jt[d] = cross(ax[d], p - org[d]);
    float A[28], b[7];
#pragma unroll
    for (int i = 0; i < 7; i++) {
#pragma unroll
      for (int j = 0; j <= i; j++) A[sym(i, j)] = dot(jt[i], jt[j]) + dot(ax[i], ax[j]) + (i == j ? 0.5f : 0.f);
      b[i] = dot(jt[i], ep) + dot(ax[i], er);
    }
    spd_inverse7(A);
    float dth[7], mx = 0.f;
#pragma unroll
    for (int i = 0; i < 7; i++) {
      float v = 0.f;
#pragma unroll
      for (int j = 0; j < 7; j++) v += A[sym(i, j)] * b[j];
      dth[i] = v; mx = fmaxf(mx, fabsf(v));
    }
    const float max_dls = 0.78539816339744831f;
    float sc = mx > max_dls ? max_dls / mx : 1.f;
#pragma unroll
    for (int i = 0; i < 7; i++) if (live) q[i] += dth[i] * sc;
  }
#pragma unroll
  for (int d = 0; d < 7; d++) out[d] = q[d];
}

// ------------------------------------------------------------------------------------------------ step
// RobotTaskEnv.step (core.py:352-368): Panda.set_action (panda.py:120-172) -> sim.step()
// (pybullet.py:68-71 / reach_ao.py:182-188) -> _get_obs -> is_success / is_truncated / compute_reward,
// then gymnasium TimeLimit and Monitor bookkeeping.
// First half of a step for environment e: action limiter, action history, motor targets (written to
// tgt[d * tstride]), jerk / effort bookkeeping.  q / qd return the state the simulation starts from.
__device__ __forceinline__ void step_begin(const Params& P, const float* __restrict__ actions, int e, float* q, float* qd, float* tgt,
                                           int tstride, bool& velocity_mode, bool& unit_imp) {
  const int n = P.n;
  float a[7];
#pragma unroll
  for (int d = 0; d < 7; d++) { q[d] = P.q[(size_t)d * n + e]; qd[d] = P.qd[(size_t)d * n + e]; a[d] = 0.f; }
  // --- set_action: limiter
  const int na = P.act_dim;
  for (int k = 0; k < 7; k++) if (k < na) a[k] = actions[(size_t)e * na + k];
  if (P.action_limiter == 1) {
    bool any = false; float mx = 0.f;
    for (int k = 0; k < 7; k++) if (k < na) { any |= (a[k] < -1.f || a[k] > 1.f); mx = fmaxf(mx, fabsf(a[k])); }
    if (na > 7) { float a7 = actions[(size_t)e * na + 7]; any |= (a7 < -1.f || a7 > 1.f); mx = fmaxf(mx, fabsf(a7)); }  // PickAndPlace joints: finger entry
    if (any) for (int k = 0; k < 7; k++) a[k] = a[k] / mx;
  } else {
#pragma unroll
    for (int k = 0; k < 7; k++) a[k] = fminf(fmaxf(a[k], -1.f), 1.f);
  }
  // --- action history (previous_action <- recent_action <- action)
#pragma unroll
  for (int k = 0; k < 7; k++) { P.prev[(size_t)k * n + e] = P.recent[(size_t)k * n + e]; P.recent[(size_t)k * n + e] = a[k]; }
  P.have_prev[e] = P.have_recent[e]; P.have_recent[e] = 1;
  // --- targets
  if (P.control == 1) {
    // ee_displacement_to_target_arm_angles (panda.py:226-246): uses the CACHED end-effector position (last observation)
    const float* lo = P.obs + (size_t)e * P.obs_dim;  // observation[0:3] = get_ee_position()
    V3 t = v3(lo[0] + __fmul_rn(a[0], 0.05f), lo[1] + __fmul_rn(a[1], 0.05f), fmaxf(0.f, lo[2] + __fmul_rn(a[2], 0.05f)));
    float tq[7];
    ik_dls(v3(P.base[0], P.base[1], P.base[2]), q, t, tq);
#pragma unroll
    for (int d = 0; d < 7; d++) tgt[d * tstride] = tq[d];
  } else {
#pragma unroll
    for (int d = 0; d < 7; d++) tgt[d * tstride] = q[d] + __fmul_rn(a[d], 0.05f);  // panda.py:257-261
  }
  if (P.control == 2) { velocity_mode = true;
#pragma unroll
    for (int d = 0; d < 7; d++) tgt[d * tstride] = a[d]; }
  if (P.control == 3) {  // teleport (resetJointState), motors stay at their load-time default: v* = 0, impulse 1
    velocity_mode = true; unit_imp = true;
#pragma unroll
    for (int d = 0; d < 7; d++) { q[d] = tgt[d * tstride]; qd[d] = 0.f; tgt[d * tstride] = 0.f; }
  }
  // --- joint velocity / acceleration / jerk bookkeeping (panda.py:167-172), sampled before the sim step
#pragma unroll
  for (int d = 0; d < 7; d++) {
    float pv = P.cur_vel[(size_t)d * n + e], pa = P.cur_acc[(size_t)d * n + e];
    float ca = pv - qd[d];
    P.cur_vel[(size_t)d * n + e] = qd[d]; P.cur_acc[(size_t)d * n + e] = ca; P.jerk[(size_t)d * n + e] = fabsf(pa - ca);
  }
}

// Second half of a step: state write-back, observation, success / truncation / reward, TimeLimit, Monitor statistics.
__device__ __forceinline__ void step_end(const Params& P, int e, const float* q, const float* qd, bool collided) {
  const int n = P.n;
  float cq[7], sq[7], q_fk[7];
  // joint angles behind PyBullet's cached link transforms = q at the start of the last simulated substep
#pragma unroll
  for (int d = 0; d < 7; d++) { q_fk[d] = q[d] - qd[d] * kDt; sincosf(q[d], &sq[d], &cq[d]); }
#pragma unroll
  for (int d = 0; d < 7; d++) { P.q[(size_t)d * n + e] = q[d]; P.qd[(size_t)d * n + e] = qd[d]; }
  if (P.task == 1) P.collided[e] = collided;
  // --- observation, success, truncation, reward
  float ag[3], dg[3];
  observe(P, e, q, qd, q_fk, cq, sq, ag, false);
  dg[0] = P.desired[(size_t)e * 3]; dg[1] = P.desired[(size_t)e * 3 + 1]; dg[2] = P.desired[(size_t)e * 3 + 2];
  bool succ = is_success(P, e, ag, dg);
  bool terminated = P.terminate_on_success ? succ : false;
  bool truncated = (P.task == 1 && P.truncate_on_collision) ? collided : false;
  float r = reward_fn(P, e, ag, dg);
  int steps = P.steps[e] + 1;
  P.steps[e] = steps;
  bool timeout = steps >= P.st.max_episode_steps;
  bool done = terminated || truncated || timeout;
  P.reward[e] = r;
  P.done[e] = done;
  P.flags[e] = (uint8_t)((succ ? 1 : 0) | (terminated ? 2 : 0) | (truncated ? 4 : 0) | ((timeout && !terminated) ? 8 : 0));
  float ret = P.ep_ret[e] + r; int len = P.ep_len[e] + 1;
  P.ep_ret[e] = ret; P.ep_len[e] = len;
  if (done) {
    P.last_ret[e] = ret; P.last_len[e] = len;
    const float* o = P.obs + (size_t)e * P.obs_dim;
    float* t = P.term_obs + (size_t)e * P.obs_dim;
    for (int k = 0; k < P.obs_dim; k++) t[k] = o[k];
    P.term_achieved[(size_t)e * 3] = ag[0]; P.term_achieved[(size_t)e * 3 + 1] = ag[1]; P.term_achieved[(size_t)e * 3 + 2] = ag[2];
    P.need_reset[e] = 1;
    if (P.auto_reset) P.reset_list[atomicAdd(P.reset_count, 1)] = e;
    atomicAdd(&P.stats[0], 1.0);
    if (succ) atomicAdd(&P.stats[1], 1.0);
    if (truncated) atomicAdd(&P.stats[2], 1.0);
    if (timeout && !terminated && !truncated) atomicAdd(&P.stats[3], 1.0);
    atomicAdd(&P.stats[4], (double)ret);
    atomicAdd(&P.stats[5], (double)len);
  }
}

// One environment per thread (fallback / A-B reference for k_step2 below).
__global__ void PG_STEP_BOUNDS k_step(const Params P, const float* __restrict__ actions) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = P.n;
  const bool valid = e < n;  // every thread of the block stays for the barriers of the substep loop
  __shared__ float s_tgt[7 * kStepBlock], s_q[7 * kStepBlock];  // touched twice per substep: kept out of the register file
  float* tgt = s_tgt + threadIdx.x;
  float* qs = s_q + threadIdx.x;
  float q[7], qd[7], cq[7], sq[7];
  bool velocity_mode = false, unit_imp = false;
  if (valid) {
    step_begin(P, actions, e, q, qd, tgt, kStepBlock, velocity_mode, unit_imp);
#pragma unroll
    for (int d = 0; d < 7; d++) { sincosf(q[d], &sq[d], &cq[d]); qs[d * kStepBlock] = q[d]; }
  }
  // --- sim.step(): n_substeps x stepSimulation, collision check after each (ReachAO); a collided environment
  //     stops simulating (reach_ao.py:186-188) but its thread keeps meeting the block barriers
  bool collided = false, active = valid;
  const V3 base = v3(P.base[0], P.base[1], P.base[2]);
  for (int k = 0; k < P.n_substeps; k++) {
    substep<kStepSync>(active, velocity_mode, unit_imp, qs, qd, cq, sq, tgt, kStepBlock);
    if (P.task == 1) {
      bool hit = scene_pass<false, kStepSync>(active, P.scene, P.obs_pos_f, n, e, base, cq, sq, nullptr, nullptr);
      if (active && hit) { collided = true; active = false; }
    }
  }
  if (!valid) return;
#pragma unroll
  for (int d = 0; d < 7; d++) q[d] = qs[d * kStepBlock];
  step_end(P, e, q, qd, collided);
}

// Push / PickAndPlace (tasks 2, 3): the same step with the object world (pg_contact.cuh) in place of the arm-only
// substep.  One environment per thread; rows and contacts live in local memory.
__device__ __forceinline__ ObjState load_obj(const Params& P, int e) {
  const size_t n = P.n;
  ObjState B;
  B.x = v3(P.obj[e], P.obj[n + e], P.obj[2 * n + e]);
  B.qx = P.obj[3 * n + e]; B.qy = P.obj[4 * n + e]; B.qz = P.obj[5 * n + e]; B.qw = P.obj[6 * n + e];
  B.v = v3(P.obj[7 * n + e], P.obj[8 * n + e], P.obj[9 * n + e]);
  B.w = v3(P.obj[10 * n + e], P.obj[11 * n + e], P.obj[12 * n + e]);
  return B;
}
__device__ __forceinline__ void store_obj(const Params& P, int e, const ObjState& B) {
  const size_t n = P.n;
  P.obj[e] = B.x.x; P.obj[n + e] = B.x.y; P.obj[2 * n + e] = B.x.z;
  P.obj[3 * n + e] = B.qx; P.obj[4 * n + e] = B.qy; P.obj[5 * n + e] = B.qz; P.obj[6 * n + e] = B.qw;
  P.obj[7 * n + e] = B.v.x; P.obj[8 * n + e] = B.v.y; P.obj[9 * n + e] = B.v.z;
  P.obj[10 * n + e] = B.w.x; P.obj[11 * n + e] = B.w.y; P.obj[12 * n + e] = B.w.z;
}
// Launch shape: BASELINE configs[3] runs 16 384 environments per GPU = 512 warps for 592 warp schedulers, so the kernel is
// latency-bound whatever the block size; 32-thread blocks at least spread the warps over all 148 SMs (128-thread blocks
// left 20 SMs idle) and give each warp a whole L1 for its solver rows.
constexpr int kObjBlock = 32;
__global__ void __launch_bounds__(kObjBlock) k_step_obj(const Params P, const float* __restrict__ actions) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= P.n) return;
  float q[7], qd[7], cq[7], sq[7], tgt[7];
  bool velocity_mode = false, unit_imp = false;
  step_begin(P, actions, e, q, qd, tgt, 1, velocity_mode, unit_imp);
#pragma unroll
  for (int d = 0; d < 7; d++) sincosf(q[d], &sq[d], &cq[d]);
  ObjState B = load_obj(P, e);
  const V3 base = v3(P.base[0], P.base[1], P.base[2]);
  for (int k = 0; k < P.n_substeps; k++) world_substep(velocity_mode, unit_imp, base, P.world, q, qd, cq, sq, tgt, B);
  store_obj(P, e, B);
  step_end(P, e, q, qd, false);
}

#ifdef PG_PAIR_KERNEL
// Two environments per thread: thread t of block b owns environments b*448 + t (half A) and b*448 + 224 + t (half B)
// and runs the dynamics of both on the packed fp32 pipe (pg_packed.cuh, substep2).  The per-step prologue / epilogue
// and the collision pass are the scalar code, run once per half.
constexpr int kPairBlock = 224;                  // threads per block (7 warps, up to 255 registers each)
constexpr int kPairEnvs = 2 * kPairBlock;        // environments per block
constexpr int kPairSmemBytes = (5 * 7 * kPairBlock * 2 + 14 * kPairBlock) * 4;  // q, tgt, qd, cos, sin (P2) + snapshot
__global__ void __launch_bounds__(kPairBlock, 1) k_step2(const Params P, const float* __restrict__ actions) {
  extern __shared__ __align__(16) float s_dyn[];
  const int n = P.n, t = threadIdx.x;
  const int e0 = blockIdx.x * kPairEnvs + t;
  P2* s_q = reinterpret_cast<P2*>(s_dyn) + t;     // element d at [d * kPairBlock]
  P2* s_tgt = s_q + 7 * kPairBlock;
  P2* s_qd = s_tgt + 7 * kPairBlock;
  P2* s_c = s_qd + 7 * kPairBlock;
  P2* s_s = s_c + 7 * kPairBlock;
  float* snap = s_dyn + 5 * 7 * kPairBlock * 2 + t;  // element k at [k * kPairBlock]
  bool velocity_mode = false, unit_imp = false;
#pragma unroll 1
  for (int h = 0; h < 2; h++) {
    const int e = e0 + h * kPairBlock;
    float q[7], qd[7], c[7], s[7];
    float* tg = reinterpret_cast<float*>(s_tgt) + h;  // this half of element d at [d * 2 * kPairBlock]
    if (e < n) {
      step_begin(P, actions, e, q, qd, tg, 2 * kPairBlock, velocity_mode, unit_imp);
#pragma unroll
      for (int d = 0; d < 7; d++) sincosf(q[d], &s[d], &c[d]);
    } else {
#pragma unroll
      for (int d = 0; d < 7; d++) { q[d] = 0.f; qd[d] = 0.f; c[d] = 1.f; s[d] = 0.f; tg[d * 2 * kPairBlock] = 0.f; }
    }
#pragma unroll
    for (int d = 0; d < 7; d++) {
      (reinterpret_cast<float*>(s_q) + h)[d * 2 * kPairBlock] = q[d];
      (reinterpret_cast<float*>(s_qd) + h)[d * 2 * kPairBlock] = qd[d];
      (reinterpret_cast<float*>(s_c) + h)[d * 2 * kPairBlock] = c[d];
      (reinterpret_cast<float*>(s_s) + h)[d * 2 * kPairBlock] = s[d];
    }
  }
  P2 qd[7], cq[7], sq[7];
#pragma unroll
  for (int d = 0; d < 7; d++) { qd[d] = s_qd[d * kPairBlock]; cq[d] = s_c[d * kPairBlock]; sq[d] = s_s[d * kPairBlock]; }
  bool actA = e0 < n, actB = e0 + kPairBlock < n, collA = false, collB = false;
  const V3 base = v3(P.base[0], P.base[1], P.base[2]);
  for (int k = 0; k < P.n_substeps; k++) {
    substep2<kStepSync>(actA, actB, velocity_mode, unit_imp, s_q, qd, cq, sq, s_tgt, kPairBlock, snap, kPairBlock);
    if (P.task == 1) {
#pragma unroll 1
      for (int h = 0; h < 2; h++) {
        const bool act = h ? actB : actA;
        float c[7], s[7];
#pragma unroll
        for (int d = 0; d < 7; d++) { c[d] = half_of(cq[d], h); s[d] = half_of(sq[d], h); }
        bool hit = scene_pass<false, false>(act, P.scene, P.obs_pos_f, n, e0 + h * kPairBlock, base, c, s, nullptr, nullptr);
        if (act && hit) { if (h) { collB = true; actB = false; } else { collA = true; actA = false; } }
      }
    }
  }
#pragma unroll
  for (int d = 0; d < 7; d++) s_qd[d * kPairBlock] = qd[d];
#pragma unroll 1
  for (int h = 0; h < 2; h++) {
    const int e = e0 + h * kPairBlock;
    if (e >= n) continue;
    float q[7], qv[7];
#pragma unroll
    for (int d = 0; d < 7; d++) { q[d] = (reinterpret_cast<float*>(s_q) + h)[d * 2 * kPairBlock]; qv[d] = (reinterpret_cast<float*>(s_qd) + h)[d * 2 * kPairBlock]; }
    step_end(P, e, q, qv, h ? collB : collA);
  }
}
#endif  // PG_PAIR_KERNEL

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
    const pgm::NeutralPrim& np = pgm::kNeutralPrims[k];
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
        // analytic lower bound: cover the cylinder with ns spheres along its axis
        int ns = np.h <= np.r ? 1 : (np.h <= 2 * np.r ? 2 : (np.h <= 3 * np.r ? 3 : 4));
        double seg = np.h / ns, cover = sqrt(np.r * np.r + seg * seg) + (double)kHullMargin, lb = 1e300;
        for (int i = 0; i < ns; i++) lb = fmin(lb, sphere_box<double>(c + (-np.h + (2 * i + 1) * seg) * a, cover, o, core, m).d);
        if (lb > cutoff || lb >= best) continue;
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
template <bool WARP>
__global__ void __launch_bounds__(kBlock) k_reset(const __grid_constant__ Params P, const uint8_t* __restrict__ mask, const uint64_t* __restrict__ seeds,
                                                  uint64_t first_env) {
  const int n = P.n;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int count = WARP ? *P.reset_count : n;
  const int stride = WARP ? (gridDim.x * blockDim.x) >> 5 : gridDim.x * blockDim.x;
  const bool leader = WARP ? (threadIdx.x & 31) == 0 : true;
  for (int item = WARP ? tid >> 5 : tid; item < count; item += stride) {
  int e = item;
  if (WARP) e = P.reset_list[item];
  else if (mask && !mask[e]) continue;
  P.need_reset[e] = 0;
  uint32_t ep = P.episode[e];
  P.episode[e] = ep + 1;
  uint64_t seed = seeds ? seeds[e] : episode_seed(P.base_seed, first_env + (uint64_t)e, ep);
  Pcg g; pcg_seed(g, seed);
  const StageDev& st = P.st;
  // Panda.reset (panda.py:290-298) -- or, with randomize_robot_pose, the scenario's randomizer in its place
  // (reach_ao.py:132-133): set_random_robot_base draws joint 1 from its limits (panda.py:77-79) and keeps the rest neutral
  const bool rbase = P.task == 1 && st.random_base;
  double q0 = pgm::kNeutralQ[0], c0 = 1.0, s0 = 0.0;
  if (rbase) { q0 = pcg_uniform(g, -2.7437, 2.7437); c0 = cos(q0 - pgm::kNeutralQ[0]); s0 = sin(q0 - pgm::kNeutralQ[0]); }
  float q[7], qd[7], cq[7], sq[7];
  P.steps[e] = 0; P.ep_ret[e] = 0.f; P.ep_len[e] = 0;
  double goal[3], ee[6];  // ee[0..2] = get_ee_position(), ee[3..5] = get_link_position(0), both at the reset pose
  int status = 0;
#define PG_EE_AT_POSE()                                                                                   \
  do {                                                                                                    \
    ee[0] = c0 * pgm::kNeutralEE[0] - s0 * pgm::kNeutralEE[1] + P.based[0];                               \
    ee[1] = s0 * pgm::kNeutralEE[0] + c0 * pgm::kNeutralEE[1] + P.based[1];                               \
    ee[2] = pgm::kNeutralEE[2] + P.based[2];                                                              \
    ee[3] = c0 * pgm::kNeutralLink0[0] - s0 * pgm::kNeutralLink0[1] + P.based[0];                         \
    ee[4] = s0 * pgm::kNeutralLink0[0] + c0 * pgm::kNeutralLink0[1] + P.based[1];                         \
    ee[5] = pgm::kNeutralLink0[2] + P.based[2];                                                           \
  } while (0)
  PG_EE_AT_POSE();
  if (P.task == 0) {
    for (int k = 0; k < 3; k++) goal[k] = pcg_uniform(g, st.goal_lo[k], st.goal_hi[k]);
  } else if (P.task >= 2) {
    // Push.reset (push.py:69-87) / PickAndPlace.reset (pick_and_place.py:65-85): goal, then object; xy range 0.3;
    // PickAndPlace lifts the goal by U(0, 0.2) and flattens it again with probability 0.3
    const double half = P.obj_half_z;
    double nx = pcg_uniform(g, -0.15, 0.15), ny = pcg_uniform(g, -0.15, 0.15), nz = pcg_uniform(g, 0.0, P.task == 3 ? 0.2 : 0.0);
    if (P.task == 3 && pcg_double(g) < 0.3) nz = 0.0;
    goal[0] = 0.0 + nx; goal[1] = 0.0 + ny; goal[2] = half + nz;
    nx = pcg_uniform(g, -0.15, 0.15); ny = pcg_uniform(g, -0.15, 0.15); nz = pcg_uniform(g, 0.0, 0.0);
    ObjState B;
    B.x = v3((float)(0.0 + nx), (float)(0.0 + ny), (float)(half + nz));
    B.qx = B.qy = B.qz = 0.f; B.qw = 1.f; B.v = v3(0, 0, 0); B.w = v3(0, 0, 0);  // set_base_pose: identity orientation, velocities zeroed
    store_obj(P, e, B);
    P.have_recent[e] = 0; P.have_prev[e] = 0;
  } else {
    P.goal_reached[e] = 0; P.collided[e] = 0;
    Body table; table.type = 1; table.px = P.table_pos[0]; table.py = P.table_pos[1]; table.pz = P.table_pos[2];
    table.sx = P.table_half[0]; table.sy = P.table_half[1]; table.sz = P.table_half[2];
    Body probe; probe.type = 0; probe.sx = 0.05; probe.sy = probe.sz = 0;
    // set_coll_free_goal: pass 0 = (["table","robot"], 0.1); pass 1 (only without obstacle randomisation) = all bodies, 0.03
    const int passes = st.randomize_obstacles ? 1 : 2;
    for (int pass = 0; pass < passes; pass++) {
      const double margin = pass == 0 ? 0.1 : 0.03;
      int i = 0;
      for (;;) {
        sample_goal(st, g, goal);
        if (i > 9999) { for (int k = 0; k < 3; k++) goal[k] = ee[k]; status = 1; break; }
        i++;
        probe.px = goal[0]; probe.py = goal[1]; probe.pz = goal[2];
        bool coll = false;
        if (pass == 1) for (int j = 0; j < st.nobs; j++) coll |= body_body_collision(probe, obstacle_body(P, j, e), margin);
        coll |= body_body_collision(probe, table, margin);
        if (!coll) coll |= (margin >= robot_body_distance<WARP>(P, probe, margin, c0, s0));
        if (!coll) break;
      }
    }
    if (st.randomize_obstacles) {  // set_coll_free_obs(0.03)
      const double margin = 0.03;
      for (int j = 0; j < st.nobs && status < 2; j++) {
        int i = 0;
        for (;;) {
          if (i > 9999) { status = 2; break; }
          i++;
          double pos[3];
          sample_obstacle(P, g, goal, ee, pos);
          for (int k = 0; k < 3; k++) P.obs_pos[(size_t)(j * 3 + k) * n + e] = pos[k];
          Body ob = obstacle_body(P, j, e);
          bool coll = body_body_collision(table, ob, margin);
          coll |= body_body_collision(probe, ob, margin);
          if (!st.allow_overlap) {
            for (int k = 0; k < st.nobs; k++) if (k != j) coll |= body_body_collision(obstacle_body(P, k, e), ob, 0.0);
            probe.px = probe.py = probe.pz = 0.0;  // the probe sphere is moved to the origin and left there
            coll |= !body_body_collision(probe, ob, 1.0);
          }
          if (!coll) coll |= (margin + P.safety_distance >= robot_body_distance<WARP>(P, ob, margin + P.safety_distance, c0, s0));
          if (!coll) break;
        }
      }
    }
    if (rbase) {  // set_coll_free_robot (reach_ao.py:1035-1060): per obstacle, re-draw the pose while the robot is within 0.05
      for (int j = 0; j < st.nobs; j++) {
        Body ob = obstacle_body(P, j, e);
        int i = 0;
        while (0.05 >= robot_body_distance<WARP>(P, ob, 0.05, c0, s0)) {
          if (i > 9999) { status |= 4; break; }  // set_joint_neutral: neutral_joint_values[0] already holds the last draw
          i++;
          q0 = pcg_uniform(g, -2.7437, 2.7437); c0 = cos(q0 - pgm::kNeutralQ[0]); s0 = sin(q0 - pgm::kNeutralQ[0]);
        }
      }
    }
    if (st.random_num_obs) {  // set_random_num_obs (reach_ao.py:1062-1082)
      int nb = st.nobs;
      uint32_t span = (uint32_t)(st.size_hi - 1 - st.size_lo);
      int k = st.size_lo + (span == 0 ? 0 : (int)pcg_lemire32(g, span));
      int to_move = abs(k - nb), order[kMaxObst];
      for (int i = 0; i < nb; i++) order[i] = i;
      for (int i = nb - 1; i >= 1; i--) { int j = (int)pcg_interval32(g, (uint32_t)i); int t = order[i]; order[i] = order[j]; order[j] = t; }
      for (int i = 0; i < nb && to_move > 0; i++, to_move--) {
        P.obs_pos[(size_t)(order[i] * 3 + 0) * n + e] = 99.9; P.obs_pos[(size_t)(order[i] * 3 + 1) * n + e] = 99.9; P.obs_pos[(size_t)(order[i] * 3 + 2) * n + e] = -99.9;
      }
    }
    P.have_recent[e] = 0; P.have_prev[e] = 0;
    for (int k = 0; k < st.nobs * 3; k++) P.obs_pos_f[(size_t)k * n + e] = (float)P.obs_pos[(size_t)k * n + e];
  }
#undef PG_EE_AT_POSE
#pragma unroll
  for (int d = 0; d < 7; d++) {
    q[d] = (float)(d == 0 ? q0 : pgm::kNeutralQ[d]); qd[d] = 0.f;
    P.q[(size_t)d * n + e] = q[d]; P.qd[(size_t)d * n + e] = 0.f;
    sincosf(q[d], &sq[d], &cq[d]);
  }
  for (int k = 0; k < 3; k++) P.goal[(size_t)k * n + e] = goal[k];
  P.rng[e] = g.s_hi; P.rng[(size_t)n + e] = g.s_lo; P.rng[(size_t)2 * n + e] = g.i_hi; P.rng[(size_t)3 * n + e] = g.i_lo;
  if (status && leader) atomicAdd(&P.stats[6], 1.0);
  float ag[3], dg[3];
  observe(P, e, q, qd, q, cq, sq, ag, true);
  dg[0] = (float)goal[0]; dg[1] = (float)goal[1]; dg[2] = (float)goal[2];
  bool succ = is_success(P, e, ag, dg);
  if (!WARP) { P.done[e] = 0; P.flags[e] = succ ? 1 : 0; P.reward[e] = 0.f; }
  }
}

// refresh observation / goals at the current state (no stepping)
__global__ void __launch_bounds__(kBlock) k_observe(const Params P) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= P.n) return;
  float q[7], qd[7], cq[7], sq[7], ag[3];
#pragma unroll
  for (int d = 0; d < 7; d++) { q[d] = P.q[(size_t)d * P.n + e]; qd[d] = P.qd[(size_t)d * P.n + e]; sincosf(q[d], &sq[d], &cq[d]); }
  observe(P, e, q, qd, q, cq, sq, ag, true);
  if (P.task == 1) {
    bool hit = scene_pass<false, false>(true, P.scene, P.obs_pos_f, P.n, e, v3(P.base[0], P.base[1], P.base[2]), cq, sq, nullptr, nullptr);
    P.flags[e] = (uint8_t)((P.flags[e] & ~4) | (hit ? 4 : 0));
  }
}

__global__ void k_mirror_obstacles(const Params P) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (int64_t)kMaxObst * 3 * P.n) P.obs_pos_f[i] = (float)P.obs_pos[i];
}

__global__ void k_reward(const Params P, const float* __restrict__ ag, const float* __restrict__ dg, float* __restrict__ out, int64_t m, int src) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  float a[3] = {ag[3 * i], ag[3 * i + 1], ag[3 * i + 2]}, g[3] = {dg[3 * i], dg[3 * i + 1], dg[3 * i + 2]};
  out[i] = reward_fn(P, src, a, g);
}

__global__ void k_random_actions(float* __restrict__ out, int64_t m, uint64_t seed, uint64_t counter) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  uint64_t h = splitmix64(seed ^ splitmix64(counter * 0x100000001B3ULL + (uint64_t)i));
  out[i] = (float)((double)(h >> 11) * (2.0 / 9007199254740992.0) - 1.0);
}

// ------------------------------------------------------------------------------------------------ debug / parity kernels
__global__ void PG_STEP_BOUNDS k_debug_substeps(int n, float* q_io, float* qd_io, const float* tgt_in, float f0, float f1, float f2,
                                                           float f3, float f4, float f5, float f6, int nsub, int velocity) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = e < n;
  __shared__ float s_tgt[7 * kStepBlock], s_q[7 * kStepBlock];
  float* t = s_tgt + threadIdx.x;
  float* q = s_q + threadIdx.x;  // element d at q[d * kStepBlock]
  float qd[7], c[7], s[7];
  if (valid) {
#pragma unroll
    for (int d = 0; d < 7; d++) { q[d * kStepBlock] = q_io[(size_t)d * n + e]; qd[d] = qd_io[(size_t)d * n + e]; t[d * kStepBlock] = tgt_in[(size_t)d * n + e]; sincosf(q[d * kStepBlock], &s[d], &c[d]); }
  }
  for (int k = 0; k < nsub; k++) {
    // parity kernel: re-evaluate cos/sin exactly every 20 substeps like the env step does
    if (valid && k > 0 && (k % 20) == 0) {
#pragma unroll
      for (int d = 0; d < 7; d++) sincosf(q[d * kStepBlock], &s[d], &c[d]);
    }
    substep<kStepSync>(valid, velocity != 0, false, q, qd, c, s, t, kStepBlock);
  }
  if (!valid) return;
#pragma unroll
  for (int d = 0; d < 7; d++) { q_io[(size_t)d * n + e] = q[d * kStepBlock]; qd_io[(size_t)d * n + e] = qd[d]; }
}
#ifdef PG_PAIR_KERNEL
__global__ void __launch_bounds__(kPairBlock, 1) k_debug_substeps2(int n, float* q_io, float* qd_io, const float* tgt_in, int nsub, int velocity) {
  extern __shared__ __align__(16) float s_dyn[];
  const int t = threadIdx.x, e0 = blockIdx.x * kPairEnvs + t, e1 = e0 + kPairBlock;
  P2* q = reinterpret_cast<P2*>(s_dyn) + t;
  P2* tg = q + 7 * kPairBlock;
  float* snap = s_dyn + 5 * 7 * kPairBlock * 2 + t;
  const bool vA = e0 < n, vB = e1 < n;
  P2 qd[7], c[7], s[7];
#pragma unroll
  for (int d = 0; d < 7; d++) {
    q[d * kPairBlock] = P2(vA ? q_io[(size_t)d * n + e0] : 0.f, vB ? q_io[(size_t)d * n + e1] : 0.f);
    qd[d] = P2(vA ? qd_io[(size_t)d * n + e0] : 0.f, vB ? qd_io[(size_t)d * n + e1] : 0.f);
    tg[d * kPairBlock] = P2(vA ? tgt_in[(size_t)d * n + e0] : 0.f, vB ? tgt_in[(size_t)d * n + e1] : 0.f);
  }
  for (int k = 0; k < nsub; k++) {
    if ((k % 20) == 0) {
#pragma unroll
      for (int d = 0; d < 7; d++) { P2 v = q[d * kPairBlock]; sincosf(v.v.x, &s[d].v.x, &c[d].v.x); sincosf(v.v.y, &s[d].v.y, &c[d].v.y); }
    }
    substep2<kStepSync>(vA, vB, velocity != 0, false, q, qd, c, s, tg, kPairBlock, snap, kPairBlock);
  }
#pragma unroll
  for (int d = 0; d < 7; d++) {
    P2 v = q[d * kPairBlock];
    if (vA) { q_io[(size_t)d * n + e0] = v.v.x; qd_io[(size_t)d * n + e0] = qd[d].v.x; }
    if (vB) { q_io[(size_t)d * n + e1] = v.v.y; qd_io[(size_t)d * n + e1] = qd[d].v.y; }
  }
}
#endif
__global__ void k_debug_minv(int n, const float* q_in, float* out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float c[7], s[7], M[28];
#pragma unroll
  for (int d = 0; d < 7; d++) sincosf(q_in[(size_t)d * n + e], &s[d], &c[d]);
  crba(c, s, M);
  spd_inverse7(M);
#pragma unroll
  for (int k = 0; k < 28; k++) out[(size_t)k * n + e] = M[k];
}
// Yoshikawa index at the flange for joint cosines / sines (see k_manipulability)
__device__ __forceinline__ float manipulability_cs(const float* c, const float* s) {
  ArmKin K;
  arm_kin(v3(0.f, 0.f, 0.f), c, s, K);
  const V3 p = frame_point(K.F7, 0.f, 0.f, 0.107f);
  V3 J[7];
#pragma unroll
  for (int d = 0; d < 7; d++) J[d] = cross(K.ax[d], p - K.org[d]);
  float det = 0.f;
#pragma unroll
  for (int i = 0; i < 5; i++)
#pragma unroll
    for (int j = i + 1; j < 6; j++) {
      V3 cij = cross(J[i], J[j]);
#pragma unroll
      for (int k = j + 1; k < 7; k++) { float m = dot(cij, J[k]); det = fmaf(m, m, det); }
    }
  return det > 0.f ? sqrtf(det) : 0.f;
}

// The per-step bookkeeping of evaluation/evaluate.py:perform_benchmark (:236-262) and its episode-end accounting
// (:264-283), for every environment, on the state k_step left behind and BEFORE the auto-reset overwrites it:
//   per step     total_effort += ||current_joint_acceleration||   (reach_ao.py:1299-1302)
//                jerk.append(||current_joint_jerk||)              (:1304-1305)
//                total_manipulability += task.manipulability      (:1311, :1381 -> panda.py:314-317 at the new q)
//                ee_speed.append(||get_ee_velocity()||)           (observation[3:6])
//   episode end  is_success -> success bucket (length, effort / length, manipulability / length, all jerk and speed
//                samples); else is_truncated -> collision; else timeout; reward and length always
__global__ void __launch_bounds__(128) k_eval_metrics(const Params P) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x, n = P.n;
  if (e >= n) return;
  float c[7], s[7], a2 = 0.f, j2 = 0.f;
#pragma unroll
  for (int d = 0; d < 7; d++) {
    sincosf(P.q[(size_t)d * n + e], &s[d], &c[d]);
    float a = P.cur_acc[(size_t)d * n + e], j = P.jerk[(size_t)d * n + e];
    a2 = fmaf(a, a, a2); j2 = fmaf(j, j, j2);
  }
  const float* o = P.obs + (size_t)e * P.obs_dim;
  const float speed = sqrtf(o[3] * o[3] + o[4] * o[4] + o[5] * o[5]);
  float effort = P.ev_acc[e] + sqrtf(a2), manip = P.ev_acc[(size_t)n + e] + manipulability_cs(c, s);
  float jerk = P.ev_acc[(size_t)2 * n + e] + sqrtf(j2), sp = P.ev_acc[(size_t)3 * n + e] + speed;
  int len = P.ev_len[e] + 1;
  float ret = P.ev_ret[e] + P.reward[e];
  if (P.done[e]) {
    const uint8_t f = P.flags[e];
    atomicAdd(&P.ev[6], 1.0); atomicAdd(&P.ev[7], (double)len); atomicAdd(&P.ev[9], (double)ret);
    if (f & 1) {
      atomicAdd(&P.ev[0], 1.0); atomicAdd(&P.ev[1], (double)len);
      atomicAdd(&P.ev[2], (double)effort / len); atomicAdd(&P.ev[3], (double)manip / len);
      atomicAdd(&P.ev[4], (double)jerk); atomicAdd(&P.ev[5], (double)sp);
    } else if (f & 4) atomicAdd(&P.ev[8], 1.0);
    else atomicAdd(&P.ev[10], 1.0);
    effort = manip = jerk = sp = ret = 0.f; len = 0;
  }
  P.ev_acc[e] = effort; P.ev_acc[(size_t)n + e] = manip; P.ev_acc[(size_t)2 * n + e] = jerk; P.ev_acc[(size_t)3 * n + e] = sp;
  P.ev_len[e] = len; P.ev_ret[e] = ret;
}

// Panda.get_manipulability (panda.py:314-317): roboticstoolbox Panda().manipulability(q, axes="trans") = Yoshikawa index
// sqrt(det(Jv Jv^T)) of the 3 x 7 translational Jacobian at the toolbox model's end-effector link panda_link8 (flange =
// link7 frame + (0, 0, 0.107)).  compute_reward refreshes it every step (reach_ao.py:1311, :1381); it feeds no reward.
// q f32 [7][n] -> out f32 [n].  Same formula as oracle pgo_arm_manipulability.
__global__ void k_manipulability(int n, const float* __restrict__ q_in, float* __restrict__ out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float c[7], s[7];
#pragma unroll
  for (int d = 0; d < 7; d++) sincosf(q_in[(size_t)d * n + e], &s[d], &c[d]);
  ArmKin K;
  arm_kin(v3(0.f, 0.f, 0.f), c, s, K);
  const V3 p = frame_point(K.F7, 0.f, 0.f, 0.107f);
  // det(Jv Jv^T) by Cauchy-Binet = sum over the 35 column triples of det(J_S)^2: no cancellation between the terms, so the
  // index keeps its relative accuracy in fp32 near singular poses (the 3 x 3 Gram determinant does not)
  V3 J[7];
#pragma unroll
  for (int d = 0; d < 7; d++) J[d] = cross(K.ax[d], p - K.org[d]);
  float det = 0.f;
#pragma unroll
  for (int i = 0; i < 5; i++)
#pragma unroll
    for (int j = i + 1; j < 6; j++) {
      V3 cij = cross(J[i], J[j]);
#pragma unroll
      for (int k = j + 1; k < 7; k++) { float m = dot(cij, J[k]); det = fmaf(m, m, det); }
    }
  out[e] = det > 0.f ? sqrtf(det) : 0.f;
}
__global__ void k_debug_ik(int n, const float* q_in, const float* tgt, float bx, float by, float bz, float* out) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float q[7], o[7];
#pragma unroll
  for (int d = 0; d < 7; d++) q[d] = q_in[(size_t)d * n + e];
  ik_dls(v3(bx, by, bz), q, v3(tgt[e], tgt[(size_t)n + e], tgt[(size_t)2 * n + e]), o);
#pragma unroll
  for (int d = 0; d < 7; d++) out[(size_t)d * n + e] = o[d];
}

__global__ void __launch_bounds__(256) k_fma_peak(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
    x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
  }
  float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
  if (s == 123.456f) out[0] = s;
}

}  // namespace pg
