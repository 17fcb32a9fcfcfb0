// pg_tu_reset.cu -- domain-randomised reset (RobotTaskEnv.reset, core.py:298-308) and the observation refresh
#include "pg_reset.cuh"
#include "pg_launch.h"

namespace pg {

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
  const double* __restrict__ NX = P.nextra;  // ee[3] | link0[3] | q[7] of the reset pose
  double q0 = NX[6], c0 = 1.0, s0 = 0.0;
  if (rbase) { q0 = pcg_uniform(g, -2.7437, 2.7437); c0 = cos(q0 - NX[6]); s0 = sin(q0 - NX[6]); }
  float q[7], qd[7], cq[7], sq[7];
  P.steps[e] = 0; P.ep_ret[e] = 0.f; P.ep_len[e] = 0;
  double goal[3], ee[6];  // ee[0..2] = get_ee_position(), ee[3..5] = get_link_position(0), both at the reset pose
  int status = 0;
#define PG_EE_AT_POSE()                                                                                   \
  do {                                                                                                    \
    ee[0] = c0 * NX[0] - s0 * NX[1] + P.based[0];                               \
    ee[1] = s0 * NX[0] + c0 * NX[1] + P.based[1];                               \
    ee[2] = NX[2] + P.based[2];                                                              \
    ee[3] = c0 * NX[3] - s0 * NX[4] + P.based[0];                         \
    ee[4] = s0 * NX[3] + c0 * NX[4] + P.based[1];                         \
    ee[5] = NX[5] + P.based[2];                                                           \
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
    // config.fixed_target (reach_ao.py:981): the first pass is skipped (no draws), the goal keeps the configured value unless
    // the second pass -- taken whenever the obstacles are not re-sampled -- replaces it
    const int passes = st.randomize_obstacles ? 1 : 2;
    if (P.has_fixed_target) for (int k = 0; k < 3; k++) goal[k] = P.fixed_target[k];
    for (int pass = P.has_fixed_target ? 1 : 0; pass < passes; pass++) {
      const double margin = pass == 0 ? 0.1 : 0.03;
      int i = 0;
      for (;;) {
        sample_goal(st, g, goal);
        if (i > 9999) { for (int k = 0; k < 3; k++) goal[k] = ee[k]; status = 1; break; }
        i++;
        probe.px = goal[0]; probe.py = goal[1]; probe.pz = goal[2];
        bool coll = false;
        if (pass == 1) {
          for (int j = 0; j < st.nobs; j++) coll |= body_body_collision(probe, obstacle_body(P, j, e), margin);
          for (int j = 0; j < P.scene.nstatic && !coll; j++) {  // the scene's URDF bodies are obstacles like any other (reach_ao.py:986-988)
            V3 o;
            const Obst ob = static_obst(P.scene, j, o);
            const double m = ob.margin;
            coll |= margin >= sphere_box<double>(v3t<double>(probe.px, probe.py, probe.pz), probe.sx, v3t<double>(o.x, o.y, o.z),
                                                 v3t<double>(ob.sx - m, ob.sy - m, ob.sz - m), m).d;
          }
        }
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
          q0 = pcg_uniform(g, -2.7437, 2.7437); c0 = cos(q0 - NX[6]); s0 = sin(q0 - NX[6]);
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
    q[d] = (float)(d == 0 ? q0 : NX[6 + d]); qd[d] = 0.f;
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

void launch_reset_all(const Params& P, const uint8_t* mask, const uint64_t* seeds, uint64_t first_env, cudaStream_t s) {
  k_reset<false><<<(P.n + kBlock - 1) / kBlock, kBlock, 0, s>>>(P, mask, seeds, first_env);
}
void launch_reset_list(const Params& P, uint64_t first_env, int sm_count, cudaStream_t s) {
  k_reset<true><<<sm_count * 8, kBlock, 0, s>>>(P, nullptr, nullptr, first_env);  // 32 warps per SM, grid-stride over the list
}
// The reset-pose table (Params::nprims / nextra).  q == nullptr: the generated constants of pg_neutral.h (panda.py:67).
// Otherwise the same quantities at the joint values q[7] a benchmark-scene builder installs (reach_ao.py:322, 343, 356,
// 424), by the fp32 forward kinematics of the step kernel -- 1e-7 m from the fp64 oracle, which only matters for a goal
// sample that lands within that distance of its rejection threshold.
struct NeutralQ { double q[7]; int custom; };
__global__ void k_neutral_table(const Params P, const NeutralQ nq) {
  const int k = threadIdx.x;
  if (!nq.custom) {
    if (k < pgm::kNumNeutralPrim) P.nprims[k] = pgm::kNeutralPrims[k];
    if (k < 3) { P.nextra[k] = pgm::kNeutralEE[k]; P.nextra[3 + k] = pgm::kNeutralLink0[k]; }
    if (k < 7) P.nextra[6 + k] = pgm::kNeutralQ[k];
    return;
  }
  if (k >= pgm::kNumNeutralPrim) return;
  float cq[7], sq[7];
  for (int d = 0; d < 7; d++) { cq[d] = (float)cos(nq.q[d]); sq[d] = (float)sin(nq.q[d]); }
  Frame F[8];
  F[0] = base_frame(v3(0.f, 0.f, 0.f));
  F[1] = frame_step<1>(F[0], cq[0], sq[0]); F[2] = frame_step<2>(F[1], cq[1], sq[1]); F[3] = frame_step<3>(F[2], cq[2], sq[2]);
  F[4] = frame_step<4>(F[3], cq[3], sq[3]); F[5] = frame_step<5>(F[4], cq[4], sq[4]); F[6] = frame_step<6>(F[5], cq[5], sq[5]);
  F[7] = frame_step<7>(F[6], cq[6], sq[6]);
  pgm::NeutralPrim np;
  if (k < pgm::kNumPrim) {
    const PrimRT& pr = kPrimRT[k];
    const V3 c = frame_point(F[pr.frame], pr.x, pr.y, pr.z), a = frame_dir(F[pr.frame], pr.ax, pr.ay, pr.az);
    np.type = pr.type; np.c[0] = c.x; np.c[1] = c.y; np.c[2] = c.z; np.a[0] = a.x; np.a[1] = a.y; np.a[2] = a.z; np.r = pr.r; np.h = pr.h;
  } else np = pgm::kNeutralPrims[k];  // panda_link0: fixed to the base
  P.nprims[k] = np;
  if (k == 0) {
    const V3 ee = frame_point(F[7], pgm::kEEx, pgm::kEEy, pgm::kEEz);
    const V3 l0 = frame_point(F[1], pgm::Link<1>::cx, pgm::Link<1>::cy, pgm::Link<1>::cz);
    P.nextra[0] = ee.x; P.nextra[1] = ee.y; P.nextra[2] = ee.z; P.nextra[3] = l0.x; P.nextra[4] = l0.y; P.nextra[5] = l0.z;
    for (int d = 0; d < 7; d++) P.nextra[6 + d] = nq.q[d];
  }
}
void launch_neutral_table(const Params& P, const double* q7, cudaStream_t s) {
  NeutralQ nq;
  nq.custom = q7 != nullptr;
  for (int d = 0; d < 7; d++) nq.q[d] = q7 ? q7[d] : 0.0;
  k_neutral_table<<<1, 64, 0, s>>>(P, nq);
}
void launch_observe(const Params& P, cudaStream_t s) { k_observe<<<(P.n + kBlock - 1) / kBlock, kBlock, 0, s>>>(P); }
}  // namespace pg
