// pg_reset_body.cuh -- bodies of the reset and observation-refresh kernels, included (through pg_reset_variant.inc) by the four
// translation units pg_tu_reset_e{0,1}g{0,1}.cu: e = PG_ENVBOX (0: obstacle shapes from the kernel parameters, the default; 1:
// per-environment cuboid sizes, scenario reachao_rand_shape), g = start-pose randomizer that puts the arm anywhere.  One unit per
// variant so that they compile in parallel (each is ~1 minute of nvcc).

// One environment's reset.  WARP: the 32 lanes of a warp run it together on the same environment (same streams, same control
// flow; the 36-primitive robot distance query is split across the lanes).  AUTO: auto-reset of an environment that finished in
// the step just taken (done / flags / reward of that step stay); otherwise a reset the caller asked for.
// GPOSE: the stage has a start-pose randomizer that puts the arm anywhere (rtb_ik / Bullet IK); false compiles all of that out
// A one-thread reset whose set_coll_free_robot loop runs long (an obstacle most start poses touch) gives up after kDeferTries tries,
// puts the environment on Params::hard_list and leaves it to the warp-cooperative pass that follows (PG_RESET_HARD_KERNEL), which
// redoes the reset from its seed with the 32-tries-per-round forms: at a time-out wave (50 000 one-thread resets) the handful of
// such environments no longer holds the step for 10 000 sequential inverse-kinematics solves.
constexpr int kDeferTries = 64;
// -DPG_RESET_TRACE (tools/trace_reset.py, variant build only): per reset, clock stamps of its phases, SM id and environment
#ifdef PG_RESET_TRACE
static __device__ unsigned long long g_trace[8 * 65536];
static __device__ int g_trace_n;
__device__ __forceinline__ unsigned long long trace_gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define PG_TRACE(k, v) do { if (tslot >= 0) g_trace[tslot * 8 + (k)] = (unsigned long long)(v); } while (0)
#else
#define PG_TRACE(k, v) do {} while (0)
#endif
template <bool WARP, bool GPOSE>
__device__ __forceinline__ void reset_env(const Params& P, const int e, const uint64_t* __restrict__ seeds, const uint64_t first_env, const bool AUTO) {
  const int n = P.n;
  const bool leader = WARP ? (threadIdx.x & 31) == 0 : true;
  PG_CHECK(e >= 0 && e < n);
#ifdef PG_RESET_TRACE
  int tslot = -1;
  if (leader) {
    tslot = atomicAdd(&g_trace_n, 1);
    if (tslot >= 65536) tslot = -1;
    unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    PG_TRACE(0, trace_gt()); PG_TRACE(1, clock64()); PG_TRACE(7, ((unsigned long long)e << 32) | (smid << 8) | (WARP ? 1 : 0));
  }
#endif
  P.need_reset[e] = 0;
  uint32_t ep = P.episode[e];
  P.episode[e] = ep + 1;
  uint64_t seed = seeds ? seeds[e] : episode_seed(P.base_seed, first_env + (uint64_t)e, ep);
  Pcg g; pcg_seed(g, seed);
  const StageDev& st = P.st;
  // Panda.reset (panda.py:290-298) -- or, with randomize_robot_pose, the scenario's randomizer in its place
  // (reach_ao.py:132-133): set_random_robot_base draws joint 1 from its limits (panda.py:77-79) and keeps the rest neutral
  const bool rbase = P.task == 1 && st.random_base;
  const bool gpose = GPOSE;  // (the launcher checks: ReachAO, pose_active, not set_random_robot_base)
  const double* __restrict__ NX = P.nextra;  // ee[3] | link0[3] | q[7] of the reset pose
  double q0 = NX[6], c0 = 1.0, s0 = 0.0;
  if (rbase) { q0 = pcg_uniform(g, -2.7437, 2.7437); c0 = cos(q0 - NX[6]); s0 = sin(q0 - NX[6]); }
  float q[7], qd[7], cq[7], sq[7];
  P.steps[e] = 0; P.ep_ret[e] = 0.f; P.ep_len[e] = 0;
  double goal[3], ee[6];  // ee[0..2] = get_ee_position(), ee[3..5] = get_link_position(0), both at the reset pose
  double old_pos[kMaxObst * 3];  // obstacle positions before this reset, for a one-thread reset that defers itself (kDeferTries)
  if (!WARP && P.task == 1 && (GPOSE || st.random_base))
    for (int k = 0; k < st.nobs * 3; k++) old_pos[k] = P.obs_pos[(size_t)k * n + e];
  int status = 0;
  PoseD pose; double qpose[7]; CRand cr;
  const PoseD* pp = gpose ? &pose : nullptr;
  if (gpose) {
    const uint32_t* cs = P.crand + (size_t)33 * e;
    for (int k = 0; k < 31; k++) cr.r[k] = cs[k];
    cr.f = (int)cs[31]; cr.b = (int)cs[32];
    // the joints the robot has when the randomizer runs: what the last episode left, or, in a fresh environment, the neutral
    // pose its scene builder installed (reach_ao.py:405 set_joint_neutral) -- Bullet's IK (industrial) starts from them
    for (int d = 0; d < 7; d++) qpose[d] = ep == 0 ? NX[6 + d] : (double)P.q[(size_t)d * n + e];
    status |= randomize_pose<WARP>(P, e, g, cr, qpose, pose, qpose);
  }
#define PG_EE_AT_POSE()                                                                                   \
  do {                                                                                                    \
    ee[0] = c0 * NX[0] - s0 * NX[1] + P.based[0];                               \
    ee[1] = s0 * NX[0] + c0 * NX[1] + P.based[1];                               \
    ee[2] = NX[2] + P.based[2];                                                              \
    ee[3] = c0 * NX[3] - s0 * NX[4] + P.based[0];                         \
    ee[4] = s0 * NX[3] + c0 * NX[4] + P.based[1];                         \
    ee[5] = NX[5] + P.based[2];                                                           \
  } while (0)
#define PG_EE_AT_GPOSE()                                                                                  \
  do {                                                                                                    \
    pose_point(pose, 7, pgm::kLocalEE, ee); pose_point(pose, 1, pgm::kLocalLink0, ee + 3);                \
    for (int k = 0; k < 3; k++) { ee[k] += P.based[k]; ee[3 + k] += P.based[k]; }                         \
  } while (0)
  if (gpose) PG_EE_AT_GPOSE(); else PG_EE_AT_POSE();
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
    if (st.random_size) {  // add_random_size_cuboids (reach_ao.py:969-979, 1084-1089): half extents 0.2 * Dirichlet(1, 1, 1), in
      for (int j = 0; j < st.nobs; j++)  // dict order; the re-created body sits at create_obstacle_cuboid's default position
        if (st.obs_type[j] == 1) {
          double f[3];
          pcg_dirichlet111(g, f);
          for (int k = 0; k < 3; k++) P.env_size[(size_t)(j * 3 + k) * n + e] = 0.2 * f[k];
          P.obs_pos[(size_t)(j * 3 + 0) * n + e] = 0.1; P.obs_pos[(size_t)(j * 3 + 1) * n + e] = 0.0; P.obs_pos[(size_t)(j * 3 + 2) * n + e] = 0.1;
        }
    }
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
        if (!coll) coll |= (margin >= robot_body_distance<WARP, GPOSE>(P, probe, margin, c0, s0, pp));
        if (!coll) break;
      }
    }
    PG_TRACE(2, clock64());
    if (st.randomize_obstacles) {  // set_coll_free_obs(0.03)
      const double margin = 0.03;
      bool stop = false;  // StopIteration escapes set_coll_free_obs in the reference (reach_ao.py:1144-1145)
      for (int j = 0; j < st.nobs && !stop; j++) {
        int i = 0;
        for (;;) {
          if (i > 9999) { status |= 2; stop = true; break; }
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
          if (!coll) coll |= (margin + P.safety_distance >= robot_body_distance<WARP, GPOSE>(P, ob, margin + P.safety_distance, c0, s0, pp));
          if (!coll) break;
        }
      }
    }
    PG_TRACE(3, clock64());
    if (gpose && st.pose_kind == PG_POSE_IK_SPHERE)  // its check_collided (fp32 scene pass) must see the obstacles just sampled
      for (int k = 0; k < st.nobs * 3; k++) P.obs_pos_f[(size_t)k * n + e] = (float)P.obs_pos[(size_t)k * n + e];
    if (rbase || gpose) {  // set_coll_free_robot (reach_ao.py:1035-1060): per obstacle, re-draw the pose while the robot is within 0.05
      // self.obstacles = the sampled obstacles, then (benchmark scenes) the scene's bodies (reach_ao.py:300-306)
      const int nbody = st.nobs + (gpose ? P.scene.nstatic : 0);
      const int limit = P.pose_retry_limit > 0 ? P.pose_retry_limit : 10000;  // reach_ao.py:1047: i > 9999
      for (int j = 0; j < nbody; j++) {
        Body ob;
        if (j < st.nobs) ob = obstacle_body(P, j, e);
        else {
          V3 o;
          const Obst so = static_obst(P.scene, j - st.nobs, o);
          ob.type = 1; ob.px = o.x; ob.py = o.y; ob.pz = o.z; ob.sx = so.sx; ob.sy = so.sy; ob.sz = so.sz; ob.margin = so.margin;
        }
        int i = 0;
        // futile case (see oracle/pgo_env.c set_coll_free_robot): the obstacle stands within 0.05 of the robot's fixed core, which
        // no pose moves -- all `limit` tries of the reference fail, then it sets the neutral pose.  Short-circuited.  Random base:
        // one draw per try and the loop ends on its last draw, so the numpy stream jumps limit - 1 draws and draws the last one
        // (identical to running the loop).  rtb_ik randomizers: the numpy stream jumps over the tries' draws where their number is
        // fixed; the C rand() stream is left alone (stated deviation).
        if (0.05 >= robot_body_distance<WARP, GPOSE>(P, ob, 0.05, c0, s0, pp) && robot_core_within<WARP, GPOSE>(P, ob, 0.05, pp)) {
          if (gpose) {
            if (st.pose_kind == PG_POSE_IK_GOAL_SPACE)
              pcg_advance(g, (uint64_t)limit * (uint64_t)(st.goal_sampler == 0 ? 3 : 3 + (st.goal_front ? 1 : 0) + (st.goal_34 ? 1 : 0)));
            if (st.pose_kind == PG_POSE_IK_BOX || st.pose_kind == PG_POSE_BULLET_IK_BOX) pcg_advance(g, 3ull * (uint64_t)limit);
            for (int d = 0; d < 7; d++) qpose[d] = NX[6 + d];
            pose_fk(qpose, pose);
          } else {
            pcg_advance(g, (uint64_t)limit - 1);
            q0 = pcg_uniform(g, -2.7437, 2.7437); c0 = cos(q0 - NX[6]); s0 = sin(q0 - NX[6]);
          }
          status |= 4;
          continue;
        }
        // Warp-cooperative reset, randomizers that draw a fixed number of values per try (ik_goal_space, ik_box): the 32 lanes run
        // the next 32 tries at once -- lane L on the numpy stream advanced by L tries, FIRST search of ik_lm_chan only (it needs no
        // C rand() draw), then the robot-obstacle distance of its own pose.  The lowest lane that either found a free pose or needs
        // a restart decides: all tries before it failed without touching the C rand() stream, exactly as in sequence; a free pose
        // ends the loop, a restart is replayed the sequential way (randomize_pose, which owns the restart stream) and speculation
        // resumes after it.  Same poses, same streams as one try at a time; the 10 000-try loop of an obstacle every pose happens to
        // touch (0.1 % of resets, not seen by the fixed-core test) takes ~1 / 25 of the time.
        const bool spec = WARP && GPOSE && (st.pose_kind == PG_POSE_IK_GOAL_SPACE || st.pose_kind == PG_POSE_IK_BOX);
        const int draws = st.pose_kind == PG_POSE_IK_BOX || st.goal_sampler == 0 ? 3 : 3 + (st.goal_front ? 1 : 0) + (st.goal_34 ? 1 : 0);
        double Rt[9];
        if (spec) { double ax_[7][3], org_[7][3], o7_[3]; chain7(NX + 6, ax_, org_, Rt, o7_); }
        // set_robot_random_pose (torus kinds) draws a fixed number of values per TRY, several tries per call (until the EE height
        // fits): the same speculation one level down -- the lowest lane that needs a restart or whose pose passes the height test
        // decides; lanes before it are tries of the running call that failed the height test.
        const bool spec_torus = WARP && GPOSE && (st.pose_kind == PG_POSE_TORUS || st.pose_kind == PG_POSE_TORUS_FRONT);
        const int draws_torus = st.pose_kind == PG_POSE_TORUS_FRONT ? 4 : 3;
        if (spec_torus) { double ax_[7][3], org_[7][3], o7_[3]; chain7(NX + 6, ax_, org_, Rt, o7_); }
        bool in_call = false;  // a robot.reset() call of the loop below is under way (torus speculation only)
        int inner = 0;         // height-test tries it has made
        while (0.05 >= robot_body_distance<WARP, GPOSE>(P, ob, 0.05, c0, s0, pp)) {
          if (!WARP && i >= kDeferTries && limit > kDeferTries && st.pose_kind != PG_POSE_IK_SPHERE && st.pose_kind != PG_POSE_BULLET_IK_BOX) {
            // as if this reset had not started: the episode counter and the obstacle positions it has overwritten go back (the
            // overlap test of set_coll_free_obs reads the obstacles not yet re-sampled where the last episode left them);
            // nothing else it wrote survives the redo
            P.episode[e] = ep;
            for (int k = 0; k < st.nobs * 3; k++) P.obs_pos[(size_t)k * n + e] = old_pos[k];
            P.hard_list[atomicAdd(P.hard_count, 1)] = e;
            return;
          }
          if (spec_torus && in_call) {}  // the try budget is checked when a call starts
          else
          if (i > limit - 1) {  // set_joint_neutral (with set_random_robot_base, neutral_joint_values[0] already holds the last draw)
            status |= 4;
            if (gpose) { for (int d = 0; d < 7; d++) qpose[d] = NX[6 + d]; pose_fk(qpose, pose); }
            break;
          }
          if (spec_torus) {
            if (!in_call) { i++; in_call = true; inner = 0; }
            const int lane = threadIdx.x & 31, nvalid = 32;
            Pcg gl = g;
            pcg_advance(gl, (uint64_t)lane * (uint64_t)draws_torus);
            double t[3], ql[7];
            sample_torus(gl, st.pose_kind == PG_POSE_TORUS_FRONT, t);
            for (int d = 0; d < 7; d++) ql[d] = NX[6 + d];
            const int ok = lm_search(ql, t, Rt);
            bool zpass = false;
            if (ok) {
              PoseD pl;
              pose_fk(ql, pl);
              double eel[3];
              pose_point(pl, 7, pgm::kLocalEE, eel);
              const double z = eel[2] + P.based[2];
              zpass = 0.6 >= z && z >= 0.4;
            }
            const unsigned m_restart = __ballot_sync(0xffffffffu, !ok);
            const unsigned m_z = __ballot_sync(0xffffffffu, zpass);
            const unsigned first = m_restart | m_z;
            const int w = first ? __ffs(first) - 1 : nvalid;
            if (inner + w >= 990 || (first && ((m_restart >> w) & 1u))) {
              // a try that needs the restart stream (or the call nears its 1 000-try bound): the tries before it failed the height
              // test; the rest of the call runs the sequential way
              const int skip = (first && w < nvalid) ? w : 0;
              pcg_advance(g, (uint64_t)skip * (uint64_t)draws_torus);
              status |= randomize_pose<WARP>(P, e, g, cr, qpose, pose, qpose, inner + skip);
              in_call = false;
              continue;
            }
            if (!first) { pcg_advance(g, (uint64_t)nvalid * (uint64_t)draws_torus); inner += nvalid; continue; }  // 32 more failed tries
            for (int d = 0; d < 7; d++) qpose[d] = __shfl_sync(0xffffffffu, ql[d], w);  // try w passes the height test: the call returns
            pose_fk(qpose, pose);
            pcg_advance(g, (uint64_t)(w + 1) * (uint64_t)draws_torus);
            in_call = false;
            continue;
          }
          if (WARP && !GPOSE && rbase) {  // set_random_robot_base: one draw per try, no other stream -- 32 tries per round, a lane each
            const int lane = threadIdx.x & 31, left = limit - i, nvalid = left < 32 ? left : 32;
            Pcg gl = g;
            pcg_advance(gl, (uint64_t)lane);
            const double ql0 = pcg_uniform(gl, -2.7437, 2.7437);
            const bool freep = lane < nvalid &&
                               !(0.05 >= robot_body_distance<false, GPOSE>(P, ob, 0.05, cos(ql0 - NX[6]), sin(ql0 - NX[6]), nullptr));
            const unsigned m_free = __ballot_sync(0xffffffffu, freep);
            const int w = m_free ? __ffs(m_free) - 1 : nvalid - 1;  // the try whose pose the robot has afterwards
            q0 = __shfl_sync(0xffffffffu, ql0, w); c0 = cos(q0 - NX[6]); s0 = sin(q0 - NX[6]);
            pcg_advance(g, (uint64_t)(w + 1));
            i += w + 1;
            continue;
          }
          if (spec) {
            const int lane = threadIdx.x & 31, left = limit - i, nvalid = left < 32 ? left : 32;
            Pcg gl = g;
            pcg_advance(gl, (uint64_t)lane * (uint64_t)draws);
            double t[3], ql[7];
            if (st.pose_kind == PG_POSE_IK_BOX) { for (int k = 0; k < 3; k++) t[k] = pcg_uniform(gl, st.pose_lo[k], st.pose_hi[k]); }
            else sample_goal(st, gl, t);
            int ok = 0, freep = 0;
            if (lane < nvalid) {
              for (int d = 0; d < 7; d++) ql[d] = NX[6 + d];
              ok = lm_search(ql, t, Rt);
              if (ok) {
                PoseD pl;
                pose_fk(ql, pl);
                freep = !(0.05 >= robot_body_distance<false, GPOSE>(P, ob, 0.05, 1.0, 0.0, &pl));
              }
            }
            const unsigned m_restart = __ballot_sync(0xffffffffu, lane < nvalid && !ok);
            const unsigned m_free = __ballot_sync(0xffffffffu, freep != 0);
            const unsigned first = m_restart | m_free;
            const int w = first ? __ffs(first) - 1 : nvalid - 1;  // the try whose pose the robot has afterwards
            PG_CHECK(w >= 0 && w < nvalid && nvalid >= 1 && nvalid <= 32 && i + w + 1 <= limit);
            if (!first || ((m_free >> w) & 1u)) {                 // tries i+1 .. i+w+1 done: all failed, or the last one is free
              for (int d = 0; d < 7; d++) qpose[d] = __shfl_sync(0xffffffffu, ql[d], w);
              pose_fk(qpose, pose);
              pcg_advance(g, (uint64_t)(w + 1) * (uint64_t)draws);
              i += w + 1;
              continue;
            }
            pcg_advance(g, (uint64_t)w * (uint64_t)draws);        // tries before lane w failed; try w needs the restart stream
            i += w;
          }
          i++;
          if (gpose) status |= randomize_pose<WARP>(P, e, g, cr, qpose, pose, qpose);
          else { q0 = pcg_uniform(g, -2.7437, 2.7437); c0 = cos(q0 - NX[6]); s0 = sin(q0 - NX[6]); }
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
    if (st.random_size)
      for (int j = 0; j < st.nobs; j++)
        if (st.obs_type[j] == 1) {  // fp32 mirror for the collision / observation passes (same arithmetic as k_mirror_obstacles)
          const Body b = obstacle_body(P, j, e);
          float* eb = P.env_box + (size_t)(j * 5) * n + e;
          eb[0] = (float)b.sx; eb[n] = (float)b.sy; eb[(size_t)2 * n] = (float)b.sz;
          eb[(size_t)3 * n] = (float)box_margin_d(b.sx, b.sy, b.sz);
          eb[(size_t)4 * n] = sqrtf((float)b.sx * (float)b.sx + (float)b.sy * (float)b.sy + (float)b.sz * (float)b.sz);
        }
  }
  if (WARP && gpose) __syncwarp();  // every lane has read the stream state before the leader stores the advanced one
  if (gpose && leader) {
    uint32_t* cs = P.crand + (size_t)33 * e;
    for (int k = 0; k < 31; k++) cs[k] = cr.r[k];
    cs[31] = (uint32_t)cr.f; cs[32] = (uint32_t)cr.b;
  }
#undef PG_EE_AT_POSE
#undef PG_EE_AT_GPOSE
#pragma unroll
  for (int d = 0; d < 7; d++) {
    q[d] = (float)(gpose ? qpose[d] : (d == 0 ? q0 : NX[6 + d])); qd[d] = 0.f;
    P.q[(size_t)d * n + e] = q[d]; P.qd[(size_t)d * n + e] = 0.f;
    sincosf(q[d], &sq[d], &cq[d]);
  }
  for (int k = 0; k < 3; k++) P.goal[(size_t)k * n + e] = goal[k];
  P.rng[e] = g.s_hi; P.rng[(size_t)n + e] = g.s_lo; P.rng[(size_t)2 * n + e] = g.i_hi; P.rng[(size_t)3 * n + e] = g.i_lo;
  if (status && leader) atomicAdd(&P.stats[6], 1.0);
  float ag[3], dg[3];
  PG_TRACE(4, clock64());
  if (WARP && P.task == 1 && P.obstacle_obs != 0 && P.scene.nobs + P.scene.nstatic > 0) observe_reset_warp(P, e, q, cq, sq, ag);
  else observe<true>(P, e, q, qd, q, cq, sq, ag, true);
  dg[0] = (float)goal[0]; dg[1] = (float)goal[1]; dg[2] = (float)goal[2];
  bool succ = is_success(P, e, ag, dg);
  if (!AUTO) { P.done[e] = 0; P.flags[e] = succ ? 1 : 0; P.reward[e] = 0.f; }
  PG_TRACE(5, clock64()); PG_TRACE(6, trace_gt());
}

// LIST = false: every environment (or those of `mask`), one thread each.  LIST = true: the environments k_step appended to the
// compact list, grid-stride -- one WARP per environment while the list is short (the usual 1 % of the batch: latency matters,
// the lanes split the distance queries), one THREAD per environment when it is long (all environments were reset together, so
// ~80 % of them time out in the same step every max_episode_steps: 32 lanes repeating the same fp64 sampling code made those
// steps 15 ms; a thread each finishes them in under a millisecond).
constexpr int kWarpResetMax = 2048;
template <bool LIST, bool GPOSE>
__global__ void __launch_bounds__(kBlock) PG_RESET_KERNEL(const __grid_constant__ Params P, const uint8_t* __restrict__ mask, const uint64_t* __restrict__ seeds,
                                                  uint64_t first_env) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int total = gridDim.x * blockDim.x;
  if (LIST) {
    const int count = *P.reset_count;
    if (count <= kWarpResetMax) {
      for (int item = tid >> 5; item < count; item += total >> 5) reset_env<true, GPOSE>(P, P.reset_list[item], nullptr, first_env, true);
    } else {
      for (int item = tid; item < count; item += total) reset_env<false, GPOSE>(P, P.reset_list[item], nullptr, first_env, true);
    }
  } else {
    for (int e = tid; e < P.n; e += total)
      if (!mask || mask[e]) reset_env<false, GPOSE>(P, e, seeds, first_env, false);
  }
}
// the environments the one-thread resets deferred (see kDeferTries), a warp each
template <bool GPOSE>
__global__ void __launch_bounds__(kBlock) PG_RESET_HARD_KERNEL(const __grid_constant__ Params P, const uint64_t* __restrict__ seeds, uint64_t first_env,
                                                               int auto_reset) {
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, total = gridDim.x * blockDim.x;
  const int count = *P.hard_count;
  for (int item = tid >> 5; item < count; item += total >> 5) reset_env<true, GPOSE>(P, P.hard_list[item], seeds, first_env, auto_reset != 0);
}

#ifdef PG_OBSERVE_LAUNCH
// refresh observation / goals at the current state (no stepping)
__global__ void __launch_bounds__(kBlock) PG_OBSERVE_KERNEL(const Params P) {
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
#endif
