/* panda_b200.h -- C ABI of the B200-native batched Panda simulator (libpanda_b200.so).
 *
 * Drop-in boundary for the env-step hot path of RaikoPipe/panda-gym.  Every entry point names the reference
 * interface it replaces (file:line into the reference repo).  Conventions: plain pointers and sizes, no C++ /
 * torch types; every call returns 0 on success or a negative pg_status, with a thread-local message available
 * from pg_last_error(); no exception crosses the boundary.  The library owns all device memory; callers get
 * non-owning device pointers through pg_get_view().  A handle is bound to one CUDA device and must be used from
 * one host thread at a time.  `stream` arguments are cudaStream_t passed as void* (NULL = default stream); all
 * work is stream-ordered and asynchronous unless stated otherwise.
 *
 * There is no CPU fallback: pg_create() fails when no CUDA device is usable.
 */
#ifndef PANDA_B200_H
#define PANDA_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct pg_env_s* pg_handle;

enum pg_status { PG_OK = 0, PG_ERR_ARG = -1, PG_ERR_CUDA = -2, PG_ERR_STATE = -3, PG_ERR_UNSUPPORTED = -4 };

/* envs/tasks/reach.py, reach_ao.py, push.py, pick_and_place.py.  Push / PickAndPlace add one free rigid box with
 * frictional contacts (object-table, robot-object, robot-table); PickAndPlace is simulated as the reference loads it:
 * franka_panda_custom_0 has FIXED finger joints, so the 4th action entry moves nothing and the finger width reads 0. */
enum pg_task { PG_TASK_REACH = 0, PG_TASK_REACH_AO = 1, PG_TASK_PUSH = 2, PG_TASK_PICK_AND_PLACE = 3 };
enum pg_control { PG_CTRL_JOINTS = 0, PG_CTRL_EE = 1, PG_CTRL_JSD = 2, PG_CTRL_PCC = 3 }; /* panda.py:141-165 */
enum pg_reward { PG_REW_SPARSE = 0, PG_REW_DENSE_OR_WANG = 1, PG_REW_KUMAR_HER = 2, PG_REW_KUMAR_OPTIM = 3, PG_REW_KUMAR = 4 };
enum pg_obstacle_obs { PG_OOBS_NONE = 0, PG_OOBS_VECTORS_CLOSEST_PER_LINK = 1, PG_OOBS_VECTORS = 2,
                       PG_OOBS_CLOSEST_PER_LINK = 3, PG_OOBS_CLOSEST = 4, PG_OOBS_VECTORS_PAST = 5 };      /* reach_ao.py:908-932 */
enum pg_goal_sampler { PG_GOAL_BOX = 0, PG_GOAL_SHELL = 1 };                     /* reach_ao.py:1183-1211 */
/* ReachAO.robot_pose_randomizer of the scenario (reach_ao.py:62-63, 310-320, 363-375, 395-403, 521-604, 669, 711).  It
 * replaces Panda.reset when TrainConfig.randomize_robot_pose is set or the scenario forces it (reach_ao.py:132-133), and
 * set_coll_free_robot (:1035-1060) re-runs it while the robot stands within 0.05 of an obstacle.
 *   TORUS / TORUS_FRONT  set_robot_random_pose(sample_inside_torus) (:806-817, 1213-1236): rtb_ik targets until the EE height is in [0.4, 0.6]
 *   RANDOM_BASE          set_random_robot_base (:1238-1241): joint 1 uniform in its limits
 *   IK_GOAL_SPACE        set_robot_random_joint_position_ik_goal_space (:801-804): rtb_ik towards a sample of the goal sampler
 *   IK_SPHERE            set_robot_random_joint_position_ik_sphere(rmin, rmax) (:793-799): rtb_ik targets until check_collided is false
 *   IK_BOX               rtb_ik towards U(pose_lo, pose_hi) (narrow_tunnel :310-320, workshop2 :363-375)
 *   BULLET_IK_BOX        PyBullet calculateInverseKinematics (position only, seeded by the current joints) towards U(pose_lo, pose_hi) (industrial :395-403)
 * rtb_ik = roboticstoolbox ik_lm_chan (panda.py:220-224), restated in fp64 on the device; the two rejection loops that are
 * unbounded in the reference stop after 1000 tries here (counted in episode_stats reset_failures). */
enum pg_pose_randomizer { PG_POSE_TORUS = 0, PG_POSE_RANDOM_BASE = 1, PG_POSE_TORUS_FRONT = 2, PG_POSE_IK_GOAL_SPACE = 3,
                          PG_POSE_IK_SPHERE = 4, PG_POSE_IK_BOX = 5, PG_POSE_BULLET_IK_BOX = 6 };
enum pg_obstacle_sampler { PG_OSMP_GOAL_BOX = 0, PG_OSMP_WANG = 1, PG_OSMP_EXPERIMENTAL = 2, PG_OSMP_WANG_N = 3 }; /* :610-644, 650-660 */

#define PG_MAX_OBSTACLES 8

/* Scenario ("curriculum stage") parameter block.  Replaces ReachAO._create_scenario + the create_scenario_*
 * builders (reach_ao.py:225-266, 518-722) and Reach's goal range (reach.py:24-25).  Swapping it with
 * pg_set_stage() is what setup_training.py:233-252 does by rebuilding the SubprocVecEnv per stage. */
typedef struct {
  int goal_sampler;                 /* pg_goal_sampler */
  double goal_lo[3], goal_hi[3];    /* PG_GOAL_BOX */
  double goal_rmin, goal_rmax;      /* PG_GOAL_SHELL (sample_within_hollow_sphere) */
  int goal_upper_half, goal_front_half, goal_three_quarter_front;
  int n_obstacles;                  /* <= PG_MAX_OBSTACLES */
  int obstacle_type[PG_MAX_OBSTACLES];      /* 0 sphere, 1 cuboid */
  double obstacle_size[PG_MAX_OBSTACLES][3];/* sphere: [0] = radius; cuboid: half extents */
  int obstacle_sampler;             /* pg_obstacle_sampler */
  int randomize_obstacle_position;  /* reach_ao.py:987-992 */
  int allow_overlapping_obstacles;  /* reach_ao.py:1150 */
  int random_num_obs, sample_size_lo, sample_size_hi; /* reach_ao.py:1062-1082 */
  int max_episode_steps;            /* gymnasium TimeLimit (panda_gym/__init__.py:15-20) */
  double ee_error_threshold;        /* reach_ao.py:71 / reach.py:15 */
  double speed_threshold;           /* reach_ao.py:72 */
  int pose_randomizer;              /* pg_pose_randomizer; active with pg_config.randomize_robot_pose or force_randomize_pose */
  int force_randomize_pose;         /* the scenario sets randomize_robot_pose itself (reach_ao.py:319, 374, 402, 603) */
  double pose_lo[3], pose_hi[3];    /* IK_BOX / BULLET_IK_BOX: target range; IK_SPHERE: [0] = minor / major radius */
  int random_size_cuboids;          /* reachao_rand_shape (reach_ao.py:606-608, 969-979, 1084-1089): every reset re-creates the cuboids with
                                       half extents 0.2 * Dirichlet(1, 1, 1) from the episode's numpy stream (per-environment sizes) */
  int custom_neutral;               /* 1: neutral_q replaces Panda.neutral_joint_values (panda.py:67) as the reset pose, the way */
  double neutral_q[7];              /*    some benchmark-scene builders do (reach_ao.py:322, 343, 356-359, 424-425) */
  int needs_scene;                  /* set by pg_stage_from_scenario for a benchmark scene: its boxes come from pg_set_scene */
} pg_stage;

/* Environment configuration.  Replaces the constructor arguments of PandaReachAOEnv / PandaReachEnv
 * (panda_tasks.py:72-88, 142-159) and the TrainConfig fields they read (classes/train_config.py:26-59). */
typedef struct {
  int n_envs;                /* environments on this device */
  int task;                  /* pg_task */
  int control;               /* pg_control  (TrainConfig.control_type) */
  int reward_type;           /* pg_reward   (TrainConfig.reward_type) */
  int goal_condition;        /* 0 "reach", 1 "halt" */
  int action_limiter;        /* 0 "clip", 1 "scale" (panda.py:129-135) */
  int obs_js;                /* "js" in obs_type (panda.py:273-279) */
  int obstacle_obs;          /* pg_obstacle_obs (TrainConfig.task_observations["obstacles"]) */
  int n_substeps;            /* 20 (train_config.py:44) */
  int truncate_on_collision; /* train_config.py:27 */
  int terminate_on_success;  /* train_config.py:28 */
  double collision_reward;   /* -100 (train_config.py:33) */
  double safety_distance;    /* train_config.py:39 */
  double base_position[3];   /* panda_tasks.py:82 / :148 */
  int auto_reset;            /* 1: done environments are reset inside pg_step (SB3 VecEnv semantics) */
  int randomize_robot_pose;  /* train_config.py:26: Panda.reset <- the scenario's robot_pose_randomizer, then
                                set_coll_free_robot (reach_ao.py:991-992, 1035-1060) */
  pg_stage stage;
  int has_fixed_target;      /* train_config.py:29 fixed_target: ReachAO.reset skips its first set_coll_free_goal (reach_ao.py:981) */
  double fixed_target[3];    /*   and the goal keeps this value unless the scenario re-samples it (:986-988, obstacle positions fixed) */
  int pose_retry_limit;      /* tries of set_coll_free_robot per obstacle (reach_ao.py:1047: 10 000); 0 = as the reference.  A knob for
                                large batches: one environment in a few thousand whose every start pose fails (for a reason the
                                fixed-core test of pg_reset_body.cuh does not see) otherwise holds the whole step for 10 000 IK solves */
} pg_config;

/* fill cfg with the reference defaults: task PG_TASK_REACH -> PandaReachJoints-v3 (sparse);
 * PG_TASK_REACH_AO -> PandaReachAO-v3 with TrainConfig() defaults.  scenario: name accepted by
 * ReachAO._create_scenario ("reachao1", "reachao3", "wangexp-3", "reach1", ... ) or NULL. */
int pg_default_config(pg_config* cfg, int task, const char* scenario);
/* scenario name -> stage block (reach_ao.py:229-260); returns PG_ERR_ARG for unknown names */
int pg_stage_from_scenario(pg_stage* stage, const char* scenario);

int pg_create(const pg_config* cfg, int device, pg_handle* out);
int pg_destroy(pg_handle h);                                   /* RobotTaskEnv.close (core.py:370-371) */
/* Curriculum stage switch (setup_training.py:233-252 rebuilds the SubprocVecEnv for the next stage): swaps the scenario
 * block AND resets every environment on `stream` -- all episodes start fresh in the new scene, like the rebuilt VecEnv. */
int pg_set_stage(pg_handle h, const pg_stage* stage, void* stream);
/* VecEnv.seed(seed): environment i draws its first reset from numpy-compatible PCG64(seed + first_env_index + i) */
int pg_seed(pg_handle h, uint64_t base_seed, uint64_t first_env_index);

/* RobotTaskEnv.reset (core.py:298-308).  mask_dev: NULL = all environments, else uint8[n] on the device.
 * seeds_dev: NULL = next seed of each environment's own stream, else uint64[n] explicit seeds (reset(seed=s)). */
int pg_reset(pg_handle h, const uint8_t* mask_dev, const uint64_t* seeds_dev, void* stream);
/* RobotTaskEnv.step (core.py:352-368) for all environments + TimeLimit + (optional) SB3 auto-reset.
 * actions_dev: float32 [n][act_dim] row-major on the device. */
int pg_step(pg_handle h, const float* actions_dev, void* stream);

/* Host-buffer form of the step (what a SubprocVecEnv.step call is to its caller): copies actions_host
 * [n][act_dim] to the device, steps, copies the results back and synchronises the stream.  Any output pointer
 * may be NULL.  Buffers should be page-locked for full PCIe speed. */
typedef struct {
  float* observation;    /* [n][obs_dim] */
  float* achieved_goal;  /* [n][3] */
  float* desired_goal;   /* [n][3] */
  float* reward;         /* [n] */
  uint8_t* done;         /* [n] terminated | truncated | time limit */
  uint8_t* flags;        /* [n] bit0 is_success, bit1 terminated, bit2 truncated (collision), bit3 TimeLimit.truncated */
} pg_host_out;
int pg_step_host(pg_handle h, const float* actions_host, const pg_host_out* out, void* stream);
int pg_reset_host(pg_handle h, const pg_host_out* out, void* stream);
/* Episodes that ended in a step (what SB3 reads out of `infos`: terminal_observation, Monitor's episode r / l,
 * TimeLimit.truncated).  The step kernel writes one compact row per finished episode into a 3-deep device ring:
 *   [0] env index (int bits) [1] episode return [2] episode length (int bits) [3] flags (int bits, pg_host_out.flags)
 *   [4..6] terminal achieved goal [7] - [8 ..] terminal observation;  *row_width floats per row, in completion order.
 * pg_step_id: number of pg_step calls so far (the id of the latest step).  pg_finished_count: episodes that ended in step
 * `step_id` (known on the host after pg_step_host of that step; negative pg_status when older than three steps).
 * pg_fetch_finished: ONE device-to-host copy of those rows (at most max_rows) + a stream synchronise; returns the row count. */
int64_t pg_step_id(pg_handle h);
int pg_finished_count(pg_handle h, int64_t step_id);
int pg_fetch_finished(pg_handle h, int64_t step_id, float* rows_host, int max_rows, int* row_width, void* stream);
/* current results -> host buffers without stepping (synchronises the stream) */
int pg_copy_out(pg_handle h, const pg_host_out* out, void* stream);
/* The six results live in one device slab.  A caller that carves its six host buffers out of ONE (page-locked) block at
 * these byte offsets -- offsets6[] in pg_host_out field order: observation, achieved_goal, desired_goal, reward, done,
 * flags -- gets them with a single device-to-host copy of total_bytes per step instead of six.  The observation offset
 * and total_bytes depend on the current observation width (re-query after pg_set_stage changes it). */
int pg_host_layout(pg_handle h, int64_t* offsets6, int64_t* total_bytes);

/* Task.compute_reward(achieved_goal, desired_goal, info) vectorised over n rows (core.py:282; the HER relabel
 * entry).  Like the reference (reach_ao.py:1319,1336,1377) collision / jerk / effort terms read the LIVE state
 * of environment `env_index` (SB3 passes indices=[0]). */
int pg_compute_reward(pg_handle h, const float* ag_dev, const float* dg_dev, float* out_dev, int64_t n, int env_index, void* stream);

/* synthetic rollouts: fill actions_dev [n][act_dim] with U(-1, 1) from a counter-based generator */
int pg_random_actions(pg_handle h, uint64_t seed, uint64_t counter, float* actions_dev, void* stream);

/* RobotTaskEnv.save_state / restore_state / remove_state (core.py:310-336; pybullet.py:79-102) */
int pg_save_state(pg_handle h, int* state_id, void* stream);
int pg_restore_state(pg_handle h, int state_id, void* stream);
int pg_remove_state(pg_handle h, int state_id);

/* non-owning views of device buffers */
enum pg_view_id {
  PG_V_OBSERVATION = 0,   /* f32 [n][obs_dim]  RobotTaskEnv._get_obs()["observation"] (core.py:286-296) */
  PG_V_ACHIEVED_GOAL,     /* f32 [n][3] */
  PG_V_DESIRED_GOAL,      /* f32 [n][3] */
  PG_V_REWARD,            /* f32 [n] */
  PG_V_DONE,              /* u8  [n] */
  PG_V_FLAGS,             /* u8  [n] bit0 success, bit1 terminated, bit2 truncated, bit3 time limit */
  PG_V_TERMINAL_OBSERVATION, /* f32 [n][obs_dim] infos[i]["terminal_observation"] (valid where done) */
  PG_V_TERMINAL_ACHIEVED, /* f32 [n][3] */
  PG_V_EP_RETURN,         /* f32 [n] Monitor "r" of the episode that just ended */
  PG_V_EP_LENGTH,         /* i32 [n] Monitor "l" */
  PG_V_Q,                 /* f32 [7][n] joint angles (SoA) */
  PG_V_QD,                /* f32 [7][n] joint velocities */
  PG_V_GOAL,              /* f64 [3][n] */
  PG_V_OBSTACLE_POS,      /* f64 [PG_MAX_OBSTACLES][3][n] */
  PG_V_STATS,             /* f64 [8] episodes, successes, collisions, timeouts, sum_return, sum_length, reset_failures, - */
  PG_V_STEP_COUNT,        /* i32 [n] */
  PG_V_D_LINK,            /* f32 [9][n] distances_closest_obstacles (reach_ao.py:897) */
  PG_V_OBJECT,            /* f32 [13][n] object position, quaternion xyzw, linear / angular velocity (get_base_* of pybullet.py:182-247) */
  PG_V_OBSTACLE_SIZE,     /* f64 [PG_MAX_OBSTACLES][3][n] per-environment cuboid half extents (pg_stage.random_size_cuboids; rows of spheres unused) */
  PG_V_JOINT_ACC,         /* f32 [7][n] Panda.current_joint_acceleration (panda.py:167-172: previous velocity - velocity, as written) */
  PG_V_JOINT_JERK,        /* f32 [7][n] Panda.current_joint_jerk */
  PG_V_COUNT
};
typedef struct { void* ptr; int dtype; /* 0 f32, 1 u8, 2 i32, 3 f64, 4 u64 */ int ndim; int64_t shape[3]; } pg_view;
int pg_get_view(pg_handle h, int view_id, pg_view* out);
int pg_obs_dim(pg_handle h);
int pg_act_dim(pg_handle h);
int pg_num_envs(pg_handle h);
/* number of kernels this library launched on behalf of the handle so far */
int64_t pg_launch_count(pg_handle h);

/* Evaluation metrics of evaluation/evaluate.py:perform_benchmark (:131-300) accumulated on the device for every
 * environment (the caller immediately above the step path).  pg_set_eval_metrics(h, 1) zeroes the
 * accumulators and makes every pg_step launch one extra small kernel after the step kernel (before the auto-reset);
 * enable = 0 stops it.  pg_get_eval_metrics copies 12 doubles to the host:
 *   [0] successful episodes            [1] sum of their lengths
 *   [2] sum over them of (sum of ||joint acceleration|| per step / length)      -> mean_effort          (:269, :291)
 *   [3] sum over them of (sum of manipulability per step / length)              -> mean_manipulability  (:270, :292)
 *   [4] sum of ||joint jerk|| over all their steps                              -> mean_norm_jerk       (:268, :293)
 *   [5] sum of ||ee velocity|| over all their steps                             -> mean_ee_speed        (:271, :294)
 *   [6] finished episodes   [7] sum of their lengths   [8] collisions (is_truncated and not is_success)
 *   [9] sum of episode rewards   [10] timeouts   [11] reserved
 * Episode classification follows :264-279: is_success first, then is_truncated, else timeout. */
int pg_set_eval_metrics(pg_handle h, int enable, void* stream);
int pg_get_eval_metrics(pg_handle h, double* out12_host, void* stream);

/* Panda.get_manipulability (panda.py:314-317; called by ReachAO.compute_reward, reach_ao.py:1311): Yoshikawa index
 * sqrt(det(Jv Jv^T)) of the translational Jacobian at roboticstoolbox's Panda end-effector link (panda_link8).
 * q_dev NULL: the handle's current joint angles, out_dev f32 [num_envs]; else q_dev f32 [7][n] -> out_dev f32 [n]. */
int pg_manipulability(pg_handle h, const float* q_dev, int64_t n, float* out_dev, void* stream);

/* RobotTaskEnv.render(mode="rgb_array", width, height, target_position, distance, yaw, pitch, roll) (core.py:373-414 ->
 * pybullet.py:152-180: computeViewMatrixFromYawPitchRoll with upAxisIndex 2, computeProjectionMatrixFOV(60, w / h, 0.1, 100),
 * getCameraImage).  The reference rasterises with OpenGL, which does not exist on the target machines; this is a CUDA
 * ray-caster over the same analytic scene (plane, table, the robot's collision primitives, obstacles, object, translucent
 * goal marker) of environment env_index.  rgba_dev: uint8 [height][width][4] on the device, like getCameraImage's pixels.
 * target_position NULL = (0, 0, 0); angles in degrees. */
/* Static scene shared by all environments: the benchmark scenes of ReachAO.setup_benchmark_scenario (reach_ao.py:292-306,
 * builders :308-516) as a set of axis-aligned boxes -- every collision link of the scene's URDF bodies (box geometry, or the
 * cuboid hull of an OBJ mesh) already placed in the world frame by the host loader (panda_gym_b200/scenes.py).
 * boxes_host: n_boxes x 7 floats = centre xyz, half extents xyz, collision margin (0.001 for URDF-imported shapes);
 * n_boxes <= 64, 0 removes the scene.  The boxes take part in the per-substep collision flag, the per-link distance
 * observation and the collision-free goal sampling of reset exactly like the obstacles of a stage; call pg_set_stage (or
 * pg_reset) afterwards to start episodes in the new scene. */
int pg_set_scene(pg_handle h, int n_boxes, const float* boxes_host, void* stream);

int pg_render(pg_handle h, int env_index, int width, int height, const float* target_position, float distance, float yaw, float pitch,
              float roll, uint8_t* rgba_dev, void* stream);

/* ---- matched-state entry points used by the parity tests (pybullet getters/setters of pybullet.py) ---- */
/* resetJointState for every env (pybullet.py:427-435): q_dev f32 [7][n]; velocities zeroed unless qd_dev given */
int pg_set_joint_state(pg_handle h, const float* q_dev, const float* qd_dev, void* stream);
int pg_set_goal(pg_handle h, const double* goal_dev /* [3][n] */, void* stream);
int pg_set_obstacles(pg_handle h, const double* pos_dev /* [n_obstacles][3][n] */, void* stream);
/* set_base_pose + resetBaseVelocity of the "object" body (pybullet.py:350-361): f32 [13][n] position, quaternion xyzw, linear, angular velocity */
int pg_set_object_state(pg_handle h, const float* obj_dev, void* stream);
/* recompute observation / flags at the current state without stepping (RobotTaskEnv._get_obs) */
int pg_observe(pg_handle h, void* stream);
/* n_sub raw stepSimulation() calls with POSITION_CONTROL targets / max forces, no task logic:
 * q, qd f32 [7][n] in/out, target f32 [7][n], max_force f32 [7] (host) */
int pg_debug_substeps(int device, int n, float* q_dev, float* qd_dev, const float* target_dev, const float* max_force_host,
                      int n_sub, int velocity_mode, void* stream);
/* joint-space inverse mass matrix: q f32 [7][n] -> minv f32 [28][n] packed lower (row-major i*(i+1)/2+j) */
int pg_debug_minv(int device, int n, const float* q_dev, float* minv_dev, void* stream);
/* calculateInverseKinematics (pybullet.py:465-493) for link 11 with quaternion (1,0,0,0): q f32 [7][n],
 * target f32 [3][n] -> out f32 [7][n] */
int pg_debug_ik(int device, int n, const float* q_dev, const float* target_dev, const float* base, float* out_dev, void* stream);

/* roofline denominator: FP32 FMA issue-rate micro-benchmark (8 independent FFMA chains per thread, all SMs).
 * Runs `iters` x 8 x 2 flops per thread on `blocks` x 256 threads; *tflops_out = achieved TFLOP/s (CUDA events). */
int pg_debug_fma_peak(int device, int blocks, int iters, double* tflops_out);

/* launch shape of the step kernel: out[6] = registers/thread, maxThreadsPerBlock, static smem bytes, local bytes,
 * block size, resident blocks per SM */
int pg_debug_step_kernel_attrs(int device, int* out);
/* ---------------------------------------------------------------------------------------------------------
 * On-device Hindsight-Experience-Replay buffer.  Replaces stable_baselines3 HerReplayBuffer (SB3 >= 2.0) as the
 * reference uses it (classes/train_config.py:2,15; default kwargs n_sampled_goal = 4, strategy "future"):
 *   add()    -> DictReplayBuffer.add + episode bookkeeping (ep_start / ep_length, overwrite invalidation)
 *   sample() -> valid = flatnonzero(ep_length > 0); np.random.choice(valid, B); first int(her_ratio * B) samples get a
 *               future achieved goal t = np.random.randint(cur, ep_len) and a recomputed reward; output order
 *               cat(real, virtual); dones * (1 - timeouts).
 * exact_rng = 1 consumes the MT19937 state loaded with pg_her_set_mt_state exactly like numpy's legacy
 * RandomState does for those two calls (indices bit-identical to SB3 for the same np.random state);
 * exact_rng = 0 draws from a parallel counter-based generator (same distribution, not the numpy stream). */
typedef struct pg_her_s* pg_her_handle;
typedef struct {
  float *observation, *achieved_goal, *desired_goal;                 /* [batch][obs_dim], [batch][3], [batch][3] */
  float *next_observation, *next_achieved_goal, *next_desired_goal;
  float *action, *reward, *done;                                     /* [batch][act_dim], [batch], [batch] */
  int64_t batch, n_virtual;
  int64_t row_stride;  /* floats between consecutive rows of EVERY field: the batch is one [batch][row_stride] block (64 B
                          multiples) and the fields are column ranges of it -- what sharding.allgather_samples exchanges */
} pg_her_batch;  /* device pointers owned by the buffer, valid until the next pg_her_sample */
int pg_her_create(int device, int64_t buffer_size, int n_envs, int obs_dim, int act_dim, int n_sampled_goal, int64_t max_batch, pg_her_handle* out);
int pg_her_destroy(pg_her_handle h);
/* all inputs are device pointers laid out [n_envs][...]; done / timeout are uint8 (timeout may be NULL) */
int pg_her_add(pg_her_handle h, const float* obs, const float* ag, const float* dg, const float* next_obs, const float* next_ag,
               const float* next_dg, const float* action, const float* reward, const uint8_t* done, const uint8_t* timeout, void* stream);
/* np.random.get_state()[1:3] -> device / back (host arrays) */
int pg_her_set_mt_state(pg_her_handle h, const uint32_t* key624_host, int pos);
int pg_her_get_mt_state(pg_her_handle h, uint32_t* key624_host, int* pos);
/* env: the simulator handle whose compute_reward relabels (reward terms read environment reward_env, SB3 uses 0) */
int pg_her_sample(pg_her_handle h, pg_handle env, int64_t batch, int exact_rng, uint64_t fast_seed, int reward_env, pg_her_batch* out, void* stream);
/* pg_her_sample for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.
 * Same sampling, relabelling and reward recompute as pg_her_sample, same use of the buffer's MT19937 state / fast counter and of
 * its cached valid list, but (1) the batch is written to the CALLER's device block packed_out_dev, [batch][row_stride] floats in
 * the record's field order (row_stride and the column ranges are those pg_her_sample reports: observation | next observation |
 * achieved | desired | next achieved | next desired | action | reward | done, row_stride = (2*obs_dim + act_dim + 16 + 15) / 16 * 16),
 * and (2) every per-call scratch (index lists, compaction block counts, status word) lives in ONE file-scope static __device__
 * array of the library (1.25 MiB, held by every GPU that loads the library until exit) instead of the buffer's own allocations.
 * The call makes no cudaMalloc / cudaFree and calls no allocating routine; the first call on a device resolves the array's
 * address (cudaGetSymbolAddress, no allocation) -- make that call outside a capture.
 * Limits: batch <= PG_HER_STATIC_MAX_BATCH rows and buffer capacity (rows * n_envs) <= PG_HER_STATIC_MAX_TRANSITIONS; beyond
 * either the call fails with PG_ERR_ARG and pg_last_error() names the limit.
 * All calls on one GPU share the array and it keeps the last call's contents: calls that overlap on the GPU (different streams,
 * different buffers) race; order them on one stream or with events.  The status word is cleared by every call.
 * status_out_dev (device int, may be NULL) receives 0, or 1 when no episode has finished yet (the batch is then left untouched).
 * pg_her_last_indices keeps reporting the last pg_her_sample call, not this one. */
#define PG_HER_STATIC_MAX_BATCH 65536
#define PG_HER_STATIC_MAX_TRANSITIONS 16777216
int pg_her_sample_static(pg_her_handle h, pg_handle env, int64_t batch, int exact_rng, uint64_t fast_seed, int reward_env,
                         float* packed_out_dev, int* status_out_dev, void* stream);
/* (row, env, t) of the last sample() in SB3's draw order (virtual first), for parity tests; synchronises */
int pg_her_last_indices(pg_her_handle h, int64_t batch, int* row_host, int* env_host, int* t_host, int* n_valid_host, int* status_host);
int pg_her_info(pg_her_handle h, int* rows, int* pos, int* full, int64_t* launches);

const char* pg_last_error(void);
const char* pg_version(void);

#ifdef __cplusplus
}
#endif
#endif
