// pg_kernels.cuh -- the batched environment kernels: step (set_action -> n_substeps x stepSimulation with
// per-substep collision check -> observation / success / truncation / reward), domain-randomised reset,
// observation refresh, vectorised compute_reward.  One environment per thread, SoA state in HBM.
#pragma once
#include <stdint.h>
#include "pg_scene.cuh"
#include "pg_neutral.h"
#include "pg_contact.cuh"

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
  int pose_active, pose_kind;      // the scenario's robot_pose_randomizer replaces Panda.reset (pg_pose_randomizer)
  double pose_lo[3], pose_hi[3];
  int random_size;   // reachao_rand_shape: cuboid half extents re-drawn at every reset (reach_ao.py:1084-1089)
  int max_episode_steps;
  float ee_thr, speed_thr;
};

struct Params {
  int n, task, control, reward_type, goal_condition, action_limiter, obs_js, obstacle_obs, obs_dim, act_dim, task_obs_dim;
  int n_substeps, truncate_on_collision, terminate_on_success, auto_reset;
  float collision_reward;
  double safety_distance;
  int has_fixed_target; double fixed_target[3];  // train_config.py:29
  int pose_retry_limit;          // tries of set_coll_free_robot per obstacle, 0 = the reference's 10 000
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
  double* env_size;              // [kMaxObst][3][n] per-environment cuboid half extents (random_size_cuboids only)
  float* env_box;                // [kMaxObst][5][n] fp32 mirror: half extents, collision margin, bounding radius (scene.env_box)
  uint32_t* crand;               // [n][33] glibc rand() state of each environment's "worker process": ik_lm_chan's restart stream
  float *recent, *prev;          // [7][n] action history (panda.py:138-139)
  float *cur_vel, *cur_acc, *jerk; // [7][n] (panda.py:167-172)
  float* d_link;                 // [9][n]
  // reset pose: collision primitives of the robot at Panda.neutral_joint_values (base at the origin), EE position, COM of
  // panda_link1 and the joint values themselves -- the generated constants of pg_neutral.h, or the same quantities for the
  // neutral pose a benchmark-scene builder installs (pg_stage.custom_neutral), filled by k_neutral_table
  pgm::NeutralPrim* nprims;      // [36]
  double* nextra;                // ee[3] | link0[3] | q[7]
  float* q_dl;                   // [7][n] joint angles d_link was evaluated at (scene_skip_mask trusts d_link only at exactly this pose)
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
  int *hard_list, *hard_count;     // environments a one-thread reset deferred to the warp-cooperative pass (pg_reset_body.cuh kDeferTries)
  // compact rows of the episodes that ended in this step, in the order of reset_list, term_w floats each:
  //   [0] env index (int bits) [1] episode return [2] episode length (int bits) [3] flags (int bits) [4..6] achieved goal [7] -
  //   [8 ..] terminal observation.  One slot of a 3-deep ring (pg_step rotates it): the host fetches the rows of a step it is
  //   still interested in with ONE copy of count x term_w floats (LazyInfos), nothing is snapshotted per step.
  float* term_rows;
  int term_w;
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
// BitGenerator.advance(delta): skip delta draws by LCG jump-ahead, O(log delta) 128-bit multiplications
__device__ inline void pcg_advance(Pcg& g, uint64_t delta) {
  u128 state = ((u128)g.s_hi << 64) | g.s_lo, inc = ((u128)g.i_hi << 64) | g.i_lo;
  u128 acc_mult = 1, acc_plus = 0, cur_mult = pcg_mult(), cur_plus = inc;
  while (delta > 0) {
    if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
    cur_plus = (cur_mult + 1) * cur_plus;
    cur_mult *= cur_mult;
    delta >>= 1;
  }
  state = acc_mult * state + acc_plus;
  g.s_hi = (uint64_t)(state >> 64); g.s_lo = (uint64_t)state;
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
// AT_REST (the observation a reset ends with: joints just teleported, velocities zero, cached pose = current pose): the frame
// chain reuses (cq, sq) and the velocity chain is skipped -- same numbers, ~1 k fewer instructions on the auto-reset's
// instruction-fetch-bound path.
template <bool AT_REST = false>
__device__ __forceinline__ void observe(const Params& P, int e, const float* q, const float* qd, const float* q_fk, const float* cq,
                                        const float* sq, float* ag_out, bool fresh) {
  const int n = P.n;
  V3 base = v3(P.base[0], P.base[1], P.base[2]);
  Frame Fc;
  V3 ev = v3(0.f, 0.f, 0.f);
  if (AT_REST) Fc = link7_frame(base, cq, sq);
  else {
    // end-effector pose from the cached transforms
    float cf[7], sf[7];
#pragma unroll
    for (int d = 0; d < 7; d++) sincosf(q_fk[d], &sf[d], &cf[d]);
    Fc = link7_frame(base, cf, sf);
    // link-local COM velocity of link 11 from the current state, rotated by the cached basis
    V3 w = v3(0, 0, 0), v = v3(0, 0, 0);
    vel_step<1>(w, v, cq[0], sq[0], qd[0]); vel_step<2>(w, v, cq[1], sq[1], qd[1]); vel_step<3>(w, v, cq[2], sq[2], qd[2]);
    vel_step<4>(w, v, cq[3], sq[3], qd[3]); vel_step<5>(w, v, cq[4], sq[4], qd[4]); vel_step<6>(w, v, cq[5], sq[5], qd[5]);
    vel_step<7>(w, v, cq[6], sq[6], qd[6]);
    V3 vl = v + cross(w, v3(pgm::kEEx, pgm::kEEy, pgm::kEEz));
    ev = frame_dir(Fc, vl.x, vl.y, vl.z);
  }
  V3 ee = frame_point(Fc, pgm::kEEx, pgm::kEEy, pgm::kEEz);
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
    if (P.scene.nobs + P.scene.nstatic == 0) {  // reach_ao.py:904
      for (int k = 0; k < 27; k++) o[off + k] = 1.0f;
    } else {
      float dl[9]; V3 ul[9];
      scene_pass<true, false>(true, P.scene, P.obs_pos_f, n, e, base, cq, sq, dl, ul);
      float mn = dl[0];
#pragma unroll
      for (int i = 0; i < 9; i++) { P.d_link[(size_t)i * n + e] = dl[i]; mn = fminf(mn, dl[i]); }
#pragma unroll
      for (int d = 0; d < 7; d++) P.q_dl[(size_t)d * n + e] = q[d];
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
    // deltaQ = a * conj(q) with a = (1,0,0,0) xyzw, b = conj(q) = (-qx,-qy,-qz,qw):
    //   {aw*bx + ax*bw + ay*bz - az*by, aw*by + ay*bw + az*bx - ax*bz, aw*bz + az*bw + ax*by - ay*bx, aw*bw - ax*bx - ay*by - az*bz}
    //   = {bw, -bz, by, -bx} = {qw, qz, -qy, qx}   (oracle/pgo_dyn.c pgo_arm_ik)
    const float dx = qw, dy = qz, dz = -qy, dw = qx;
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
    {
      const int slot = atomicAdd(P.reset_count, 1);
      PG_CHECK(slot >= 0 && slot < P.n);
      P.reset_list[slot] = e;
      float* row = P.term_rows + (size_t)slot * P.term_w;
      row[0] = __int_as_float(e); row[1] = ret; row[2] = __int_as_float(len); row[3] = __int_as_float((int)P.flags[e]);
      row[4] = ag[0]; row[5] = ag[1]; row[6] = ag[2]; row[7] = 0.f;
      for (int k = 0; k < P.obs_dim; k++) row[8 + k] = o[k];
    }
    atomicAdd(&P.stats[0], 1.0);
    if (succ) atomicAdd(&P.stats[1], 1.0);
    if (truncated) atomicAdd(&P.stats[2], 1.0);
    if (timeout && !terminated && !truncated) atomicAdd(&P.stats[3], 1.0);
    atomicAdd(&P.stats[4], (double)ret);
    atomicAdd(&P.stats[5], (double)len);
  }
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

}  // namespace pg
