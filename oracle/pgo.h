/* oracle/pgo.h -- CPU restatement ("oracle") of the reference's env-step hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build, load or call it.
 *
 * PARITY STATUS: the arithmetic of this path lives in un-vendored third-party code
 * (pybullet==3.2.6, pyb_utils@34550e0, stable_baselines3@58b1502 -- reference requirements.txt:107,108,142)
 * which is absent from /root/reference and not installable here.  This file restates their published
 * algorithms and is pinned against every known-answer vector the reference's own
 * tests hold for the path (test/pybullet_test.py: dt, free fall, link COM position, the joint-5 motor
 * KAT, the IK KAT) -- see tests/test_oracle_kat.py.  Pieces with no reference vector (ReachAO distance
 * aggregation, reset samplers) are "parity unpinned" and say so where they are defined.
 */
#ifndef PGO_H
#define PGO_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define PGO_MAXL 16 /* links  */
#define PGO_MAXD 9  /* dofs   */
#define PGO_MAXOBS 64 /* obstacle slots per env: <= 8 sampled obstacles of a training stage, or the boxes of a benchmark scene */
#define PGO_NCL 9   /* ReachAO collision links: link1..link8, panda_ee (reach_ao.py:98-99) */

/* ---------------------------------------------------------------- robot + Bullet multibody step */
typedef struct pgo_arm pgo_arm; /* opaque */

/* model: 0 = franka_panda_custom_0 (7 dof, what Panda loads), 1 = stock franka_panda (9 dof) */
pgo_arm* pgo_arm_create(int model, const double base_pos[3]);
void pgo_arm_destroy(pgo_arm*);
int pgo_arm_ndof(const pgo_arm*);
int pgo_arm_nlinks(const pgo_arm*);
/* physics knobs (hypothesis switches of SURVEY.md App. A): */
void pgo_arm_set_damping(pgo_arm*, double lin, double ang);
void pgo_arm_set_solver(pgo_arm*, int iters, double residual_thr);
/* resetJointState (pybullet.py:427-435): q <- value, qd <- 0 */
void pgo_arm_reset_joint(pgo_arm*, int dof, double q);
void pgo_arm_get_q(const pgo_arm*, double* q, double* qd);
void pgo_arm_set_q(pgo_arm*, const double* q, const double* qd);
/* setJointMotorControlArray (pybullet.py:437-463): mode 0 = VELOCITY_CONTROL (target = velocity),
 * 1 = POSITION_CONTROL (target = position, kp 0.1, kd 1.0, target velocity 0). max_force in N(m). */
void pgo_arm_set_motor(pgo_arm*, int dof, int mode, double target, double max_force);
/* one stepSimulation() of dt = 1/500 (pybullet.py:50,71). returns PGS iterations used */
int pgo_arm_substep(pgo_arm*);
/* getLinkState(link): COM world position [0], orientation xyzw [1], COM lin vel [6], ang vel [7] */
void pgo_arm_link_state(const pgo_arm*, int link, double pos[3], double quat_xyzw[4], double lin[3], double ang[3]);
/* world transform of the URDF link frame (row-major R link->world, origin) */
void pgo_arm_link_frame(const pgo_arm*, int link, double R[9], double p[3]);
/* joint-space inverse mass matrix (ndof x ndof, row-major) at the current q, for tests */
void pgo_arm_minv(pgo_arm*, double* minv);
/* calculateInverseKinematics (pybullet.py:465-493): DLS, <=20 iters, residual 1e-4, seed = current q.
 * quat_xyzw may be NULL (position only).  out[ndof]. */
void pgo_arm_ik(const pgo_arm*, int link, const double pos[3], const double* quat_xyzw, double* out);
void pgo_arm_set_ik(pgo_arm*, double joint_damping, int max_iters, double residual_thr);
/* Panda.rtb_ik (panda.py:220-224): roboticstoolbox ik_lm_chan towards fkine(q_seed) moved to `target` (see pgo_dyn.c).
 * returns 0 solved, 1 solved after a random restart, 2 unsolved; q_out[7]. */
typedef struct { uint32_t r[31]; int f, b; } pgo_crand; /* glibc rand() state (the toolbox's restart stream) */
void pgo_crand_seed(pgo_crand*, uint32_t seed);
int pgo_crand_next(pgo_crand*);
int pgo_arm_ik_lm_chan(const pgo_arm*, const double target[3], const double* q_seed, int ilimit, int slimit, double tol,
                       double lambda, pgo_crand* restart, double* q_out, int* iters_out);
/* translational Jacobian (3 x ndof) of a link COM, world frame */
void pgo_arm_jacobian(const pgo_arm*, int link, double* Jt, double* Jr);
/* Panda.get_manipulability (panda.py:314-317): Yoshikawa index of the translational Jacobian at the flange (panda_link8) */
double pgo_arm_manipulability(const pgo_arm*);

/* ---------------------------------------------------------------- closest distances (getClosestPoints) */
typedef struct {
  int type;        /* 0 sphere, 1 box (axis aligned, world; createCollisionShape: safe margin), 2 box imported from a URDF
                      (margin gUrdfDefaultCollisionMargin = 0.001), -1 empty slot */
  double pos[3];
  double size[3];  /* sphere: size[0] = radius; box: half extents */
} pgo_obstacle;
/* Bullet box collision margin: min(0.04, 0.1 * min half extent) (btBoxShape::setSafeMargin) */
double pgo_box_margin(const double half[3]);
double pgo_obstacle_margin(const pgo_obstacle* ob); /* by type: safe margin (1) or 0.001 (2) */
/* signed distance between one robot collision primitive (index into the model's primitive table,
 * placed by the arm's current FK) and an obstacle; pa/pb = closest points on robot / obstacle */
double pgo_prim_obstacle_distance(const pgo_arm*, int prim, const pgo_obstacle*, double pa[3], double pb[3]);
/* per-link minimum over the link's primitives */
double pgo_link_obstacle_distance(const pgo_arm*, int link, const pgo_obstacle*, double pa[3], double pb[3]);
/* generic convex distance used for cylinder-box (GJK on the margin-shrunk cores), exposed for tests:
 * cylinder (centre c, unit axis a, radius r, half length h) vs box (centre b, half extents e) */
double pgo_gjk_cyl_box(const double c[3], const double a[3], double r, double h,
                       const double b[3], const double e[3], double pa[3], double pb[3]);

/* ---------------------------------------------------------------- Push / PickAndPlace world (pgo_world.c; a18) */
typedef struct { double pos[3], quat[4] /* xyzw */, lin[3], ang[3]; } pgo_body; /* free rigid box, world-frame velocities */
typedef struct {
  double half[3], mass;                       /* the "object" box (push.py:33-39) */
  double table_pos[3], table_half[3];         /* create_table (pybullet.py:780-815) */
  double plane_pos[3], plane_half[3];         /* create_plane (pybullet.py:759-778) */
  double mu_object, mu_static, mu_link, mu_ee; /* lateral friction: Bullet default 0.5; panda.py:69-70 sets link 9 to 1.0 */
} pgo_world;
/* kind 0: object (A) on table / plane (B); 1: robot link (A) on object (B); 2: robot link (A) on table top (B).
 * n = unit normal from B towards A; pa / pb = contact points on A / B; d = signed distance; mu = combined friction */
typedef struct { int kind, link; double d, n[3], pa[3], pb[3], mu; } pgo_contact;
void pgo_world_default(pgo_world*);
/* contact set at the current poses (<= 12 entries: 4 per body pair) */
int pgo_world_contacts(const pgo_arm*, const pgo_body*, const pgo_world*, pgo_contact* out);
/* one stepSimulation() of arm + object with contacts; returns the PGS iterations used */
int pgo_world_substep(pgo_arm*, pgo_body*, const pgo_world*);
/* getEulerFromQuaternion (pybullet.py:218) */
void pgo_quat_to_euler(const double quat_xyzw[4], double rpy[3]);

/* ---------------------------------------------------------------- RNG (numpy Generator(PCG64) restated) */
typedef struct { uint64_t s_hi, s_lo, inc_hi, inc_lo; } pgo_pcg64;
void pgo_pcg64_seed(pgo_pcg64*, uint64_t seed); /* == np.random.default_rng(seed) / gymnasium seeding.np_random(seed) */
uint64_t pgo_pcg64_next(pgo_pcg64*);
void pgo_pcg64_advance(pgo_pcg64*, uint64_t delta); /* BitGenerator.advance: skip delta draws */
double pgo_pcg64_double(pgo_pcg64*);            /* Generator.random() */
double pgo_pcg64_uniform(pgo_pcg64*, double lo, double hi); /* Generator.uniform(lo, hi) */
double pgo_pcg64_standard_exponential(pgo_pcg64*);          /* Generator.standard_exponential() (256-layer ziggurat) */
void pgo_pcg64_dirichlet111(pgo_pcg64*, double out[3]);     /* Generator.dirichlet([1, 1, 1]) */

/* ---------------------------------------------------------------- environments */
typedef struct pgo_env pgo_env; /* opaque */

typedef struct {
  int task;              /* 0 = Reach (reach.py), 1 = ReachAO (reach_ao.py), 2 = Push (push.py), 3 = PickAndPlace (pick_and_place.py) */
  int control;           /* 0 = joints ("js"/"joints": q_tgt = q + 0.05 a), 1 = "ee" (IK), 2 = "jsd" velocity, 3 = "pcc" teleport */
  int reward_type;       /* 0 sparse, 1 dense (Reach) / wang, 2 kumar_her, 3 kumar_optim, 4 kumar (ReachAO) */
  int goal_condition;    /* 0 reach, 1 halt */
  int action_limiter;    /* 0 clip, 1 scale */
  int obs_js;            /* 1 if "js" in obs_type (panda.py:273-279) */
  int obstacle_obs;      /* 0 none, 1 vectors+closest_per_link, 2 vectors, 3 closest_per_link, 4 closest */
  int n_substeps;        /* 20 */
  int truncate_on_collision, terminate_on_success;
  double collision_reward;   /* -100 */
  double ee_error_threshold; /* 0.05 */
  double speed_threshold;
  double safety_distance;    /* 0 */
  int max_episode_steps;     /* TimeLimit */
  int scenario;          /* ReachAO: see PGO_SCN_* */
  int n_obstacles;       /* wang-N / wangexp-N */
  double base_pos[3];
  int randomize_robot_pose; /* train_config.py:26: Panda.reset <- the scenario's robot_pose_randomizer (reach_ao.py:132-133) */
  /* PGO_SCN_STATIC: a benchmark scene (reach_ao.py:292-516): goal box and neutral joint values of its builder; the scene's
   * boxes are handed over with pgo_env_set_obstacles (type 2) after pgo_env_create and stay put (no obstacle sampling) */
  double scene_goal_lo[3], scene_goal_hi[3];
  int custom_neutral; double neutral_q[7];
  int has_fixed_target; double fixed_target[3]; /* train_config.py:29, consumed at reach_ao.py:981 */
  /* PGO_SCN_STATIC only: the scene builder's own robot_pose_randomizer (narrow_tunnel :310-320, workshop2 :363-375: an
   * rtb_ik towards U(lo, hi); industrial :395-403: PyBullet's own IK towards U(lo, hi); all with randomize_robot_pose forced on);
   * 0 keeps the task default (set_robot_random_pose, :62-63) */
  int scene_pose_randomizer, scene_force_randomize; double scene_pose_lo[3], scene_pose_hi[3];
  int pose_retry_limit;  /* tries of set_coll_free_robot per obstacle; 0 = the reference's 10 000 (reach_ao.py:1047) */
  int no_shortcuts;      /* 1: run the futile rejection loops as written instead of short-circuiting them (pgo_env.c set_coll_free_robot) */
} pgo_env_config;
/* ReachAO.robot_pose_randomizer kinds (reach_ao.py:62-63, 521-604, 669, 711, 782-817, 1238-1241) */
enum { PGO_POSE_TORUS = 0, PGO_POSE_RANDOM_BASE = 1, PGO_POSE_TORUS_FRONT = 2, PGO_POSE_IK_GOAL_SPACE = 3, PGO_POSE_IK_SPHERE = 4,
       PGO_POSE_IK_BOX = 5, PGO_POSE_BULLET_IK_BOX = 6 };

enum { PGO_SCN_NONE = 0, PGO_SCN_REACH1, PGO_SCN_REACH2, PGO_SCN_REACH3, PGO_SCN_REACHAO1, PGO_SCN_REACHAO2,
       PGO_SCN_REACHAO3, PGO_SCN_REACHAO_RAND, PGO_SCN_WANG, PGO_SCN_WANGEXP, PGO_SCN_STATIC,
       PGO_SCN_REACHAO_RAND_START, PGO_SCN_REACHAO_RAND_SHAPE };

void pgo_env_default_config(pgo_env_config*, int task);
pgo_env* pgo_env_create(const pgo_env_config*);
void pgo_env_destroy(pgo_env*);
int pgo_env_obs_dim(const pgo_env*);
int pgo_env_act_dim(const pgo_env*);
/* RobotTaskEnv.reset(seed) (core.py:298-308).  returns 0 ok, else a bit mask: 1 / 2 / 4 / 8 = the goal / obstacle / set_coll_free_robot /
 * start-pose rejection loop hit its cap, 16 = an rtb_ik call found no solution in its 100 searches. */
int pgo_env_reset(pgo_env*, uint64_t seed, float* obs, float* achieved, float* desired, int* is_success);
/* RobotTaskEnv.step (core.py:352-368) + TimeLimit.  flags out: [is_success, terminated, truncated(collision), time_limit] */
void pgo_env_step(pgo_env*, const float* action, float* obs, float* achieved, float* desired, float* reward, int* flags);
/* direct state access for matched-state parity tests */
pgo_arm* pgo_env_arm(pgo_env*);
pgo_body* pgo_env_object(pgo_env*);   /* tasks 2, 3 */
pgo_world* pgo_env_world(pgo_env*);
void pgo_env_set_goal(pgo_env*, const double goal[3]);
void pgo_env_get_goal(const pgo_env*, double goal[3]);
void pgo_env_set_obstacles(pgo_env*, int n, const pgo_obstacle*);
int pgo_env_get_obstacles(const pgo_env*, pgo_obstacle*);
/* ReachAO.check_collided (reach_ao.py:896-900): fills d_link[9]; returns collided flag */
int pgo_env_check_collided(pgo_env*, double* d_link, double* d_table_min);
/* per-step readings of evaluation/evaluate.py:perform_benchmark (:252-259): effort, jerk norm, manipulability, ee speed */
void pgo_env_step_metrics(pgo_env*, double out[4]);
/* observation pieces at the current state */
void pgo_env_get_obs(pgo_env*, float* obs, float* achieved, float* desired);
/* compute_reward(achieved, desired) vectorised over n rows, the HER relabel entry (core.py:282) */
void pgo_env_compute_reward(pgo_env*, const float* ag, const float* dg, int n, float* out);
int pgo_env_is_collided(const pgo_env*);
void pgo_env_get_config(const pgo_env*, pgo_env_config* out);
void pgo_env_get_table(const pgo_env*, double pos[3], double half[3]);
/* render("rgb_array") restated (pgo_render.c): RGBA8 [height][width][4]; camera arguments as PyBullet.render (pybullet.py:117-180) */
void pgo_render(pgo_env*, int width, int height, const double target[3], double distance, double yaw_deg, double pitch_deg,
                double roll_deg, unsigned char* rgba);

#ifdef __cplusplus
}
#endif
#endif
