/* oracle/pgo_env.c -- CPU restatement of the reference's Python env layer for the hot path:
 *   RobotTaskEnv.reset/step/_get_obs        panda_gym/envs/core.py:286-308, 352-368
 *   Panda.set_action / get_obs / reset      panda_gym/envs/robots/panda.py:120-172, 226-312
 *   Reach                                   panda_gym/envs/tasks/reach.py
 *   ReachAO                                 panda_gym/envs/tasks/reach_ao.py:182-198, 518-722, 896-1264, 1308-1383
 *   distance()                              panda_gym/utils.py:4-16
 * TEST INFRASTRUCTURE ONLY (see oracle/pgo.h).  The pyb_utils-fork distance aggregation behind
 * check_collided / get_obs is reconstructed from its call sites: PARITY UNPINNED.
 */
#define _GNU_SOURCE
#include "pgo.h"
#include "pgo_internal.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define N_ARM 7
static const double NEUTRAL[7] = {0.0, -0.3, 0.0, -2.2, 0.0, 2.0, 0.78539816339744830962}; /* panda.py:67 */
static const double FORCES[7] = {87.0, 87.0, 87.0, 87.0, 12.0, 120.0, 120.0};              /* panda.py:63 */
static const int COLL_LINKS[PGO_NCL] = {0, 1, 2, 3, 4, 5, 6, 7, 9};                        /* reach_ao.py:98-99 */
#define EE_LINK 11                                                                          /* panda.py:68 */
#define NPRIM_ALL 36

struct pgo_env {
  pgo_env_config cfg;
  pgo_arm* arm;
  double goal[3];
  pgo_obstacle obs[PGO_MAXOBS];
  int nobs;
  int is_collided, goal_reached, steps;
  double d_link[PGO_NCL];
  int have_recent, have_prev;
  float recent[8], prev[8];
  double cur_vel[7], prev_vel[7], cur_acc[7], prev_acc[7], jerk[7];
  pgo_pcg64 rng;
  double dummy[3]; /* the collidable r=0.05 probe sphere (reach_ao.py:284-290) */
  float past[3][27]; int npast; /* past_obstacle_observations (reach_ao.py:49-53, 918-922): last three vector observations */
  /* scenario parameter block */
  int goal_sampler;        /* 0 box, 1 hollow sphere */
  double goal_lo[3], goal_hi[3];
  double goal_rmin, goal_rmax; int goal_upper, goal_front, goal_34;
  int obs_sampler;         /* 0 goal box, 1 wang (70/20/10), 2 experimental (50/50), 3 wang-N (70/30) */
  int randomize_obstacles, allow_overlap, random_num_obs, size_lo, size_hi;
  double neutral0; /* Panda.neutral_joint_values[0], mutated by set_random_robot_base */
  double neutral[7]; /* Panda.neutral_joint_values (panda.py:67), replaced by some benchmark-scene builders */
  int pose_kind, pose_active;  /* robot_pose_randomizer of the scenario (PGO_POSE_*) and whether it replaces Panda.reset (:132-133) */
  double pose_lo[3], pose_hi[3]; /* IK_BOX: target range; IK_SPHERE: [0] = minor / major radius */
  int random_size_cuboids;     /* reach_ao.py:105, 608 */
  pgo_crand crand;             /* the worker process's C rand() stream: restart configurations of rtb_ik (pgo_arm_ik_lm_chan) */
  double table_pos[3], table_half[3];
  /* Push / PickAndPlace (tasks 2, 3): the free object and its world */
  pgo_body obj;
  pgo_world world;
};

/* ------------------------------------------------------------------ helpers */
static float f32_distance(const float* a, const float* b) {
  /* utils.distance on float32 inputs: norm in float32 (no FMA), np.round(., 6) in float32 */
  volatile float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
  volatile float xx = dx * dx, yy = dy * dy, zz = dz * dz;
  volatile float s = xx + yy; s = s + zz;
  volatile float d = sqrtf(s);
  volatile float m = d * 1e6f;
  m = rintf(m);
  return m / 1e6f;
}

void pgo_env_default_config(pgo_env_config* c, int task) {
  memset(c, 0, sizeof *c);
  c->task = task;
  c->control = 0; c->reward_type = 0; c->goal_condition = 0; c->action_limiter = 0;
  c->n_substeps = 20; c->collision_reward = -100.0; c->ee_error_threshold = 0.05; c->speed_threshold = 0.1;
  c->safety_distance = 0.0; c->max_episode_steps = 50;
  if (task == 0) { /* PandaReach*: Panda(obs_type=("ee",)), base (-0.6,0,0) (panda_tasks.py:81-88) */
    c->obs_js = 0; c->obstacle_obs = 0; c->truncate_on_collision = 0; c->terminate_on_success = 0;
    c->base_pos[0] = -0.6; c->scenario = PGO_SCN_NONE;
  } else if (task >= 2) { /* PandaPush-v3 / PandaPickAndPlace-v3: ee control, base (-0.6,0,0) (panda_tasks.py:37-72, __init__.py:58-77) */
    c->obs_js = 0; c->obstacle_obs = 0; c->truncate_on_collision = 0; c->terminate_on_success = 0;
    c->base_pos[0] = -0.6; c->scenario = PGO_SCN_NONE; c->control = 1;
  } else {         /* PandaReachAO-v3 with TrainConfig() defaults (train_config.py) */
    c->obs_js = 1; c->obstacle_obs = 1; c->truncate_on_collision = 1; c->terminate_on_success = 1;
    c->scenario = PGO_SCN_REACHAO1;
  }
}

static void add_sphere(pgo_env* e, double r) {
  pgo_obstacle* o = &e->obs[e->nobs++];
  o->type = 0; o->pos[0] = 0.1; o->pos[1] = 0; o->pos[2] = 0.1; o->size[0] = r; o->size[1] = o->size[2] = 0; /* reach_ao.py:819 */
}
static void add_cuboid(pgo_env* e, double h) {
  pgo_obstacle* o = &e->obs[e->nobs++];
  o->type = 1; o->pos[0] = 0.1; o->pos[1] = 0; o->pos[2] = 0.1; o->size[0] = o->size[1] = o->size[2] = h;    /* reach_ao.py:841 */
}
static void goal_box(pgo_env* e, double xl, double xh, double yl, double yh, double zl, double zh) {
  e->goal_sampler = 0;
  e->goal_lo[0] = xl; e->goal_lo[1] = yl; e->goal_lo[2] = zl; e->goal_hi[0] = xh; e->goal_hi[1] = yh; e->goal_hi[2] = zh;
}
static void goal_shell(pgo_env* e, double rmin, double rmax, int upper, int front, int q34) {
  e->goal_sampler = 1; e->goal_rmin = rmin; e->goal_rmax = rmax; e->goal_upper = upper; e->goal_front = front; e->goal_34 = q34;
}

static void setup_scenario(pgo_env* e) {
  const pgo_env_config* c = &e->cfg;
  e->nobs = 0;
  if (c->task == 0) { /* Reach: goal_range 0.3 (reach.py:24-25); table 1.1x0.7x0.4 at x=-0.3 (reach.py:33) */
    goal_box(e, -0.15, 0.15, -0.15, 0.15, 0.0, 0.3);
    e->table_pos[0] = -0.3; e->table_pos[1] = 0; e->table_pos[2] = -0.2;
    e->table_half[0] = 0.55; e->table_half[1] = 0.35; e->table_half[2] = 0.2;
    return;
  }
  /* ReachAO: table 2.0x1.3x0.4 (reach_ao.py:272); default goal range (reach_ao.py:78-82) */
  e->table_pos[0] = 0; e->table_pos[1] = 0; e->table_pos[2] = -0.2;
  e->table_half[0] = 1.0; e->table_half[1] = 0.65; e->table_half[2] = 0.2;
  goal_box(e, -0.3 / 2.5 + 0.6, 0.3 / 2.5 + 0.6, -0.3 / 1.5, 0.3 / 1.5, 0.0, 0.3);
  e->obs_sampler = 0; e->randomize_obstacles = 0; e->allow_overlap = 0; e->random_num_obs = 0;
  e->pose_kind = PGO_POSE_TORUS; e->pose_active = c->randomize_robot_pose; e->random_size_cuboids = 0;       /* :61-63 */
  switch (c->scenario) {
    case PGO_SCN_REACH1: goal_box(e, -0.4 / 2 + 0.6, 0.4 / 2 + 0.6, -0.4 / 2, 0.4 / 2, 0, 0.4);               /* :518-522 */
      e->pose_kind = PGO_POSE_TORUS_FRONT; break;
    case PGO_SCN_REACH2: goal_shell(e, 0.5, 0.85, 1, 0, 1); e->pose_kind = PGO_POSE_TORUS_FRONT; break;      /* :524-531 */
    case PGO_SCN_REACH3: goal_shell(e, 0.5, 0.85, 1, 0, 1); e->pose_kind = PGO_POSE_TORUS; break;            /* :533-539 */
    case PGO_SCN_REACHAO1:                                                                                   /* :541-545 */
      goal_box(e, -0.4 / 2 + 0.6, 0.4 / 2 + 0.6, -0.4 / 2, 0.4 / 2, 0, 0.4);
      e->pose_kind = PGO_POSE_TORUS_FRONT;
      e->randomize_obstacles = 1; add_sphere(e, 0.04); break;
    case PGO_SCN_REACHAO2:                                                                                   /* :547-564 */
      goal_shell(e, 0.5, 0.8, 1, 1, 0); e->randomize_obstacles = 1; e->obs_sampler = 1;
      e->pose_kind = PGO_POSE_IK_GOAL_SPACE;
      add_sphere(e, 0.05); add_sphere(e, 0.05); break;
    case PGO_SCN_REACHAO3:                                                                                   /* :573-585 */
      goal_shell(e, 0.5, 0.8, 1, 0, 0); e->randomize_obstacles = 1; e->obs_sampler = 1;
      e->pose_kind = PGO_POSE_IK_GOAL_SPACE;
      for (int i = 0; i < 3; i++) add_sphere(e, 0.05);
      break;
    case PGO_SCN_REACHAO_RAND: case PGO_SCN_REACHAO_RAND_START: case PGO_SCN_REACHAO_RAND_SHAPE:             /* :587-608 */
      goal_shell(e, 0.5, 0.8, 1, 0, 0); e->randomize_obstacles = 1; e->obs_sampler = 2;
      e->pose_kind = PGO_POSE_IK_GOAL_SPACE;
      for (int i = 0; i < 3; i++) add_sphere(e, 0.05);
      for (int i = 0; i < 3; i++) add_cuboid(e, 0.05);
      e->random_num_obs = 1; e->allow_overlap = 1; e->size_lo = 4; e->size_hi = 6;
      if (c->scenario == PGO_SCN_REACHAO_RAND_START) {                                                       /* :601-604 */
        e->pose_active = 1; e->pose_kind = PGO_POSE_IK_SPHERE; e->pose_lo[0] = 0.45; e->pose_hi[0] = 0.7;
      }
      if (c->scenario == PGO_SCN_REACHAO_RAND_SHAPE) e->random_size_cuboids = 1;                              /* :606-608 */
      break;
    case PGO_SCN_WANG:                                                                                       /* :646-699 */
      goal_shell(e, 0.4, 0.95, 1, 0, 0); e->randomize_obstacles = 1; e->obs_sampler = 3;
      e->pose_kind = PGO_POSE_TORUS;
      for (int i = 0; i < c->n_obstacles; i++) add_sphere(e, 0.05);
      break;
    case PGO_SCN_WANGEXP:                                                                                    /* :701-722 */
      goal_shell(e, 0.5, 0.8, 1, 0, 0); e->randomize_obstacles = 1; e->obs_sampler = 1;
      e->pose_kind = PGO_POSE_RANDOM_BASE;
      for (int i = 0; i < c->n_obstacles; i++) add_sphere(e, 0.05);
      break;
    case PGO_SCN_STATIC:                                                                                     /* :292-516 */
      goal_box(e, c->scene_goal_lo[0], c->scene_goal_hi[0], c->scene_goal_lo[1], c->scene_goal_hi[1], c->scene_goal_lo[2], c->scene_goal_hi[2]);
      if (c->scene_pose_randomizer) { e->pose_kind = c->scene_pose_randomizer; memcpy(e->pose_lo, c->scene_pose_lo, 24); memcpy(e->pose_hi, c->scene_pose_hi, 24); }
      if (c->scene_force_randomize) e->pose_active = 1;
      break;
    default: break;
  }
}

pgo_env* pgo_env_create(const pgo_env_config* c) {
  pgo_env* e = (pgo_env*)calloc(1, sizeof *e);
  e->cfg = *c;
  e->arm = pgo_arm_create(0, c->base_pos);
  for (int d = 0; d < N_ARM; d++) e->neutral[d] = c->custom_neutral ? c->neutral_q[d] : NEUTRAL[d];
  setup_scenario(e);
  pgo_crand_seed(&e->crand, 1); /* a fresh worker process: rand() without srand() */
  pgo_world_default(&e->world);
  e->obj.quat[3] = 1.0; e->obj.pos[2] = e->world.half[2];
  for (int k = 0; k < PGO_NCL; k++) e->d_link[k] = 0.0; /* reach_ao.py:110 np.zeros */
  for (int d = 0; d < N_ARM; d++) pgo_arm_reset_joint(e->arm, d, e->neutral[d]); /* Panda.__init__ / the scene builders end with set_joint_neutral */
  return e;
}
void pgo_env_destroy(pgo_env* e) { if (e) { pgo_arm_destroy(e->arm); free(e); } }
pgo_arm* pgo_env_arm(pgo_env* e) { return e->arm; }
int pgo_env_is_collided(const pgo_env* e) { return e->is_collided; }
void pgo_env_set_goal(pgo_env* e, const double g[3]) { memcpy(e->goal, g, 24); }
void pgo_env_get_goal(const pgo_env* e, double g[3]) { memcpy(g, e->goal, 24); }
void pgo_env_set_obstacles(pgo_env* e, int n, const pgo_obstacle* o) { e->nobs = n; memcpy(e->obs, o, n * sizeof *o); }
int pgo_env_get_obstacles(const pgo_env* e, pgo_obstacle* o) { memcpy(o, e->obs, e->nobs * sizeof *o); return e->nobs; }

static int task_obs_dim(const pgo_env* e) {
  if (e->cfg.task >= 2) return 12; /* push.py:49-63 */
  if (e->cfg.task == 0 || e->cfg.obstacle_obs == 0) return 0;
  if (e->nobs == 0) return 27; /* reach_ao.py:904: np.ones(27) whatever the mode */
  switch (e->cfg.obstacle_obs) { case 1: return 36; case 2: return 27; case 3: return 9; case 5: return 81; default: return 1; }
}
static int robot_obs_dim(const pgo_env* e) { return 6 + (e->cfg.obs_js ? 14 : 0) + (e->cfg.task == 3 ? 1 : 0); } /* panda.py:264-288 */
int pgo_env_obs_dim(const pgo_env* e) { return robot_obs_dim(e) + task_obs_dim(e); }
/* block_gripper=True except PickAndPlace (panda_tasks.py:49), whose extra finger action moves nothing: the loaded
 * custom_0 URDF has FIXED finger joints (panda.urdf:467-479) */
int pgo_env_act_dim(const pgo_env* e) { return (e->cfg.control == 1 ? 3 : 7) + (e->cfg.task == 3 ? 1 : 0); }
pgo_body* pgo_env_object(pgo_env* e) { return &e->obj; }
pgo_world* pgo_env_world(pgo_env* e) { return &e->world; }

/* ------------------------------------------------------------------ collision queries */
int pgo_env_check_collided(pgo_env* e, double* d_link, double* d_table_min) {
  /* reach_ao.py:896-900 */
  pgo_obstacle table; table.type = 1; memcpy(table.pos, e->table_pos, 24); memcpy(table.size, e->table_half, 24);
  double mn = 1e300, tmn = 1e300;
  for (int i = 0; i < PGO_NCL; i++) {
    double best = 999.0; /* pyb_utils saturates at max_distance */
    for (int j = 0; j < e->nobs; j++) {
      double d = pgo_link_obstacle_distance(e->arm, COLL_LINKS[i], &e->obs[j], 0, 0);
      if (d < best) best = d;
    }
    e->d_link[i] = best;
    if (d_link) d_link[i] = best;
    if (best < mn) mn = best;
    if (i >= 1) { /* table: ignore panda_link0, panda_link1 */
      double d = pgo_link_obstacle_distance(e->arm, COLL_LINKS[i], &table, 0, 0);
      if (d < tmn) tmn = d;
    }
  }
  if (d_table_min) *d_table_min = tmn;
  return (mn <= 0.0 || tmn <= 0.0) ? 1 : 0;
}

/* body-level getClosestPoints(robot, body) min distance (reach_ao.py:1169-1181): every link incl. base */
static double robot_body_distance(pgo_env* e, const pgo_obstacle* o) {
  double best = 1e300;
  for (int p = 0; p < NPRIM_ALL; p++) { double d = pgo_prim_obstacle_distance(e->arm, p, o, 0, 0); if (d < best) best = d; }
  return best;
}
/* sphere body vs sphere / box body; box-box: Bullet's SAT box detector only reports overlapping boxes */
static int body_body_collision(const pgo_obstacle* a, const pgo_obstacle* b, double margin) {
  if (a->type == 0 && b->type == 0) {
    double d[3] = {a->pos[0] - b->pos[0], a->pos[1] - b->pos[1], a->pos[2] - b->pos[2]};
    return margin >= sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]) - a->size[0] - b->size[0];
  }
  if (a->type >= 1 && b->type >= 1) {
    for (int k = 0; k < 3; k++) if (fabs(a->pos[k] - b->pos[k]) > a->size[k] + b->size[k]) return 0;
    return 1;
  }
  const pgo_obstacle* s = a->type == 0 ? a : b;
  const pgo_obstacle* bx = a->type == 0 ? b : a;
  double m = pgo_obstacle_margin(bx), v[3], d2 = 0, rel[3];
  for (int k = 0; k < 3; k++) {
    double ek = bx->size[k] - m;
    rel[k] = s->pos[k] - bx->pos[k];
    double cl = rel[k] > ek ? ek : (rel[k] < -ek ? -ek : rel[k]);
    v[k] = rel[k] - cl; d2 += v[k] * v[k];
  }
  double d;
  if (d2 <= 1.1920929e-07) {
    double best = 1e300;
    for (int k = 0; k < 3; k++) { double ek = bx->size[k] - m; if (ek - rel[k] < best) best = ek - rel[k]; if (ek + rel[k] < best) best = ek + rel[k]; }
    d = -best;
  } else d = sqrt(d2);
  return margin >= d - m - s->size[0];
}

/* The part of the robot no start pose moves: the base link, panda_link1 (its colliders lie on joint 1's axis) and the ball of
 * radius 0.06 about the shoulder (the origin of panda_link2, on both joint axes): panda_link2's cylinder (r 0.06, half length 0.06,
 * centred there) contains it whatever joints 1 and 2 do.  An obstacle within `margin` of this set is within `margin` of the robot
 * in EVERY pose. */
static int robot_core_within(pgo_env* e, const pgo_obstacle* o, double margin) {
  for (int p = 0; p < NPRIM_ALL; p++)
    if (pgo__prim_link(e->arm, p) <= 0 && margin >= pgo_prim_obstacle_distance(e->arm, p, o, 0, 0)) return 1;
  pgo_obstacle sh; sh.type = 0; sh.size[0] = 0.06; sh.size[1] = sh.size[2] = 0;
  sh.pos[0] = e->cfg.base_pos[0]; sh.pos[1] = e->cfg.base_pos[1]; sh.pos[2] = e->cfg.base_pos[2] + 0.333;
  return body_body_collision(&sh, o, margin);
}

/* ------------------------------------------------------------------ samplers (reach_ao.py:1183-1211, 610-644) */
static void sample_hollow(pgo_env* e, double rmin, double rmax, int upper, int front, int q34, double out[3]) {
  double phi = pgo_pcg64_uniform(&e->rng, 0, 2 * M_PI);
  double theta = upper ? pgo_pcg64_uniform(&e->rng, 0, 0.5 * M_PI) : pgo_pcg64_uniform(&e->rng, 0, M_PI);
  if (front) phi = pgo_pcg64_uniform(&e->rng, -0.5 * M_PI, 0.5 * M_PI);
  if (q34) phi = pgo_pcg64_uniform(&e->rng, -0.75 * M_PI, 0.75 * M_PI);
  double r = cbrt(pgo_pcg64_uniform(&e->rng, rmin * rmin * rmin, rmax * rmax * rmax));
  out[0] = r * sin(theta) * cos(phi); out[1] = r * sin(theta) * sin(phi); out[2] = r * cos(theta);
}
static void sample_goal(pgo_env* e, double out[3]) {
  if (e->goal_sampler == 0) for (int k = 0; k < 3; k++) out[k] = pgo_pcg64_uniform(&e->rng, e->goal_lo[k], e->goal_hi[k]);
  else sample_hollow(e, e->goal_rmin, e->goal_rmax, e->goal_upper, e->goal_front, e->goal_34, out);
}
static void ee_position(pgo_env* e, double p[3]) { pgo_arm_link_state(e->arm, EE_LINK, p, 0, 0, 0); }
static void sample_obstacle(pgo_env* e, double out[3]) {
  double s[3], c[3];
  switch (e->obs_sampler) {
    case 0: for (int k = 0; k < 3; k++) out[k] = pgo_pcg64_uniform(&e->rng, e->goal_lo[k], e->goal_hi[k]); return;
    case 1: { /* sample_obstacle_wang */
      double r = pgo_pcg64_double(&e->rng);
      if (r > 0.3) { sample_hollow(e, 0.1, 0.5, 0, 0, 0, s); memcpy(c, e->goal, 24); }
      else if (r > 0.1) { sample_hollow(e, 0.1, 0.4, 0, 0, 0, s); ee_position(e, c); }
      else { sample_hollow(e, 0.3, 0.6, 1, 0, 0, s); pgo_arm_link_state(e->arm, 0, c, 0, 0, 0); }
      break; }
    case 2: { /* sample_obstacle_experimental */
      double r = pgo_pcg64_double(&e->rng);
      sample_hollow(e, 0.1, 0.5, 0, 0, 0, s);
      if (r > 0.5) memcpy(c, e->goal, 24); else ee_position(e, c);
      break; }
    default: { /* create_scenario_wang.sample_wang_obstacle */
      double r = pgo_pcg64_double(&e->rng);
      if (r > 0.3) { sample_hollow(e, 0.2, 0.6, 0, 0, 0, s); memcpy(c, e->goal, 24); }
      else { sample_hollow(e, 0.2, 0.4, 0, 0, 0, s); ee_position(e, c); }
      break; }
  }
  for (int k = 0; k < 3; k++) out[k] = s[k] + c[k];
}

/* Generator.integers / shuffle pieces (reach_ao.py:1062-1082), 32-bit draws with PCG64's buffered half */
typedef struct { int has; uint32_t val; } u32buf;
static uint32_t next_u32(pgo_env* e, u32buf* b) {
  if (b->has) { b->has = 0; return b->val; }
  uint64_t n = pgo_pcg64_next(&e->rng);
  b->has = 1; b->val = (uint32_t)(n >> 32);
  return (uint32_t)n;
}
static uint32_t lemire_u32(pgo_env* e, u32buf* b, uint32_t rng) {
  uint32_t excl = rng + 1;
  uint64_t m = (uint64_t)next_u32(e, b) * excl;
  uint32_t left = (uint32_t)m;
  if (left < excl) {
    uint32_t thr = (0xFFFFFFFFu - rng) % excl;
    while (left < thr) { m = (uint64_t)next_u32(e, b) * excl; left = (uint32_t)m; }
  }
  return (uint32_t)(m >> 32);
}
static uint32_t interval_u32(pgo_env* e, u32buf* b, uint32_t max) {
  if (max == 0) return 0;
  uint32_t mask = max, v;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  while ((v = (next_u32(e, b) & mask)) > max) {}
  return v;
}

/* ------------------------------------------------------------------ reset (reach_ao.py:965-1033) */
static int set_coll_free_goal(pgo_env* e, int with_obstacles, double margin) {
  pgo_obstacle table; table.type = 1; memcpy(table.pos, e->table_pos, 24); memcpy(table.size, e->table_half, 24);
  int i = 0, status = 0;
  for (;;) {
    sample_goal(e, e->goal);
    if (i > 9999) { ee_position(e, e->goal); status = 1; break; }
    i++;
    pgo_obstacle probe; probe.type = 0; memcpy(probe.pos, e->goal, 24); probe.size[0] = 0.05;
    memcpy(e->dummy, e->goal, 24);
    int coll = 0;
    if (with_obstacles) for (int j = 0; j < e->nobs; j++) coll |= body_body_collision(&probe, &e->obs[j], margin);
    coll |= body_body_collision(&probe, &table, margin);
    coll |= (margin >= robot_body_distance(e, &probe));
    if (!coll) break;
  }
  return status;
}
static int set_coll_free_obs(pgo_env* e, double margin) {
  pgo_obstacle table; table.type = 1; memcpy(table.pos, e->table_pos, 24); memcpy(table.size, e->table_half, 24);
  for (int j = 0; j < e->nobs; j++) {
    int i = 0;
    for (;;) {
      if (i > 9999) return 2; /* StopIteration escapes in the reference (reach_ao.py:1144-1145) */
      i++;
      double pos[3];
      sample_obstacle(e, pos);
      memcpy(e->obs[j].pos, pos, 24);
      pgo_obstacle probe; probe.type = 0; memcpy(probe.pos, e->dummy, 24); probe.size[0] = 0.05;
      int coll = (margin + e->cfg.safety_distance >= robot_body_distance(e, &e->obs[j]));
      coll |= body_body_collision(&table, &e->obs[j], margin);
      coll |= body_body_collision(&probe, &e->obs[j], margin);
      if (!e->allow_overlap) {
        for (int k = 0; k < e->nobs; k++) if (k != j) coll |= body_body_collision(&e->obs[k], &e->obs[j], 0.0);
        e->dummy[0] = e->dummy[1] = e->dummy[2] = 0.0; /* probe moved to the origin and LEFT there */
        memcpy(probe.pos, e->dummy, 24);
        coll |= !body_body_collision(&probe, &e->obs[j], 1.0);
      }
      if (!coll) break;
    }
  }
  return 0;
}

static void robot_obs(pgo_env* e, float* o, float* ag) {
  double p[3], v[3], q[PGO_MAXD], qd[PGO_MAXD];
  pgo_arm_link_state(e->arm, EE_LINK, p, 0, v, 0);
  for (int k = 0; k < 3; k++) { o[k] = (float)p[k]; o[3 + k] = (float)v[k]; if (ag) ag[k] = (float)p[k]; }
  if (e->cfg.obs_js) {
    pgo_arm_get_q(e->arm, q, qd);
    for (int d = 0; d < 7; d++) { o[6 + d] = (float)q[d]; o[13 + d] = (float)qd[d]; }
  }
  /* fingers opening (panda.py:282-284): getJointState of the fixed joints 9, 10 reports 0 */
  if (e->cfg.task == 3) o[robot_obs_dim(e) - 1] = 0.0f;
  if (e->cfg.task >= 2 && ag) for (int k = 0; k < 3; k++) ag[k] = (float)e->obj.pos[k]; /* push.py:65-67 */
}

static void task_obs(pgo_env* e, float* o) {
  /* reach_ao.py:902-959 */
  int n = task_obs_dim(e);
  if (n == 0) return;
  if (e->cfg.task >= 2) { /* object position, euler rotation, velocity, angular velocity (push.py:49-63) */
    double rpy[3];
    pgo_quat_to_euler(e->obj.quat, rpy);
    for (int k = 0; k < 3; k++) { o[k] = (float)e->obj.pos[k]; o[3 + k] = (float)rpy[k]; o[6 + k] = (float)e->obj.lin[k]; o[9 + k] = (float)e->obj.ang[k]; }
    return;
  }
  if (e->nobs == 0) { for (int k = 0; k < 27; k++) o[k] = 1.0f; return; }
  double dmin[PGO_NCL], u[PGO_NCL][3], all = 1e300;
  for (int i = 0; i < PGO_NCL; i++) {
    double best = 1e300, pa[3] = {0, 0, 0}, pb[3] = {0, 0, 0};
    for (int j = 0; j < e->nobs; j++) {
      double ta[3], tb[3];
      double d = pgo_link_obstacle_distance(e->arm, COLL_LINKS[i], &e->obs[j], ta, tb);
      if (d < best) { best = d; memcpy(pa, ta, 24); memcpy(pb, tb, 24); }
    }
    dmin[i] = best < 999.0 ? best : 999.0;
    if (dmin[i] < all) all = dmin[i];
    double v[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
    double L = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    for (int k = 0; k < 3; k++) { double x = v[k] / L; u[i][k] = isnan(x) ? 0.0 : x; } /* utils.unit_vector */
  }
  switch (e->cfg.obstacle_obs) {
    case 1: for (int i = 0; i < 9; i++) { o[i] = (float)dmin[i]; for (int k = 0; k < 3; k++) o[9 + 3 * i + k] = (float)u[i][k]; } break;
    case 2: for (int i = 0; i < 9; i++) for (int k = 0; k < 3; k++) o[3 * i + k] = (float)u[i][k]; break;
    case 3: for (int i = 0; i < 9; i++) o[i] = (float)dmin[i]; break;
    case 5: { /* "vectors+past": append the current vectors, return the last three (oldest first) */
      float cur[27];
      for (int i = 0; i < 9; i++) for (int k = 0; k < 3; k++) cur[3 * i + k] = (float)u[i][k];
      if (e->npast == 3) { memcpy(e->past[0], e->past[1], sizeof cur); memcpy(e->past[1], e->past[2], sizeof cur); e->npast = 2; }
      memcpy(e->past[e->npast++], cur, sizeof cur);
      for (int h = 0; h < e->npast; h++) memcpy(o + 27 * h, e->past[h], sizeof cur);
      break; }
    default: o[0] = (float)all; break;
  }
}

void pgo_env_get_obs(pgo_env* e, float* obs, float* ag, float* dg) {
  robot_obs(e, obs, ag);
  task_obs(e, obs + robot_obs_dim(e));
  if (dg) for (int k = 0; k < 3; k++) dg[k] = (float)e->goal[k];
}

static int is_success(pgo_env* e, const float* ag, const float* dg) {
  float d = f32_distance(ag, dg);
  float thr = (float)e->cfg.ee_error_threshold;
  if (e->cfg.task == 1 && e->cfg.goal_condition == 1) { /* "halt": latched (reach_ao.py:1253-1261) */
    if (!e->goal_reached) {
      double v[3]; pgo_arm_link_state(e->arm, EE_LINK, 0, 0, v, 0);
      double sp = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
      e->goal_reached = (d < thr) && (sp < (double)(float)e->cfg.speed_threshold);
    }
    return e->goal_reached;
  }
  return d < thr;
}

/* set_random_robot_base (reach_ao.py:1238-1241): neutral_joint_values[0] <- U(joint_lim_min[0], joint_lim_max[0])
 * (panda.py:77-79), then set_joint_neutral.  It stands in for Panda.reset when randomize_robot_pose is set (:132-133). */
static void random_robot_base(pgo_env* e) {
  e->neutral0 = pgo_pcg64_uniform(&e->rng, -2.7437, 2.7437);
  for (int d = 0; d < N_ARM; d++) pgo_arm_reset_joint(e->arm, d, d == 0 ? e->neutral0 : NEUTRAL[d]);
}
/* sample_inside_torus (reach_ao.py:1213-1236): R 0.5, r 0.05, centre height 0.5 */
static void sample_torus(pgo_env* e, int front, double out[3]) {
  double theta = 2.0 * M_PI * pgo_pcg64_double(&e->rng);
  if (front) theta = pgo_pcg64_uniform(&e->rng, -0.5 * M_PI, 0.5 * M_PI);
  double phi = 2.0 * M_PI * pgo_pcg64_double(&e->rng);
  double rad = 0.05 * sqrt(pgo_pcg64_double(&e->rng));
  out[0] = (0.5 + rad * cos(phi)) * cos(theta); out[1] = (0.5 + rad * cos(phi)) * sin(theta); out[2] = rad * sin(phi) + 0.5;
}
/* Panda.rtb_ik + set_joint_angles (panda.py:220-224, core.py:187-193) */
static int ik_to(pgo_env* e, const double target[3]) {
  double q[7];
  int rc = pgo_arm_ik_lm_chan(e->arm, target, e->neutral, 100, 100, 0.01, 1.0, &e->crand, q, 0);
  for (int d = 0; d < N_ARM; d++) pgo_arm_reset_joint(e->arm, d, q[d]);
  return rc == 2 ? 16 : 0;
}
/* the scenario's robot_pose_randomizer, which replaces Panda.reset (reach_ao.py:132-133) */
static int robot_reset(pgo_env* e) {
  int status = 0;
  double t[3];
  switch (e->pose_kind) {
    case PGO_POSE_RANDOM_BASE: random_robot_base(e); break;
    case PGO_POSE_IK_GOAL_SPACE: sample_goal(e, t); status |= ik_to(e, t); break;              /* :801-804 */
    case PGO_POSE_IK_BOX:                                                                      /* :310-320, 363-375 */
      for (int k = 0; k < 3; k++) t[k] = pgo_pcg64_uniform(&e->rng, e->pose_lo[k], e->pose_hi[k]);
      status |= ik_to(e, t); break;
    case PGO_POSE_BULLET_IK_BOX: {                                                             /* :395-403: robot.inverse_kinematics(ee_link, target)[:7] */
      double ik[PGO_MAXD];
      for (int k = 0; k < 3; k++) t[k] = pgo_pcg64_uniform(&e->rng, e->pose_lo[k], e->pose_hi[k]);
      pgo_arm_ik(e->arm, EE_LINK, t, 0, ik);                                                   /* seeded by the joints the last episode left */
      for (int d = 0; d < N_ARM; d++) pgo_arm_reset_joint(e->arm, d, ik[d]);
      break; }
    case PGO_POSE_IK_SPHERE: {                                                                 /* :793-799 (unbounded there) */
      int i = 0;
      for (;;) {
        sample_hollow(e, e->pose_lo[0], e->pose_hi[0], 1, 0, 0, t);
        status |= ik_to(e, t);
        if (!pgo_env_check_collided(e, 0, 0)) break;
        if (++i > 999) { status |= 8; break; } /* unbounded in the reference: 1000 tries here and on the device */
      }
      break; }
    default: {                                                                                 /* set_robot_random_pose :806-817 (unbounded there) */
      int i = 0;
      for (;;) {
        sample_torus(e, e->pose_kind == PGO_POSE_TORUS_FRONT, t);
        status |= ik_to(e, t);
        double p[3]; ee_position(e, p);
        if (0.6 >= p[2] && p[2] >= 0.4) break;
        if (++i > 999) { status |= 8; break; }
      }
      break; }
  }
  return status;
}
/* set_coll_free_robot (reach_ao.py:1035-1060): per obstacle, re-draw the robot pose while the robot is within 0.05.
 * FUTILE CASE: an obstacle within 0.05 of the robot's fixed core (robot_core_within: base link, panda_link1, the shoulder ball) fails
 * all 10 000 tries whatever they draw, after which the reference prints "Couldn't find collision free robot pose!" and sets the
 * neutral pose.  Short-circuited here and on the device (cfg.no_shortcuts = 1 runs the loop as written, for the tests):
 *   set_random_robot_base: one uniform draw per try and the loop ends on its last draw -- the numpy stream jumps 9 999 draws
 *     (BitGenerator.advance) and draws the 10 000th: IDENTICAL to running the loop;
 *   rtb_ik randomizers: 10 000 inverse-kinematics solves to no effect -- neutral pose, flag 4, the numpy stream jumps over the
 *     draws the tries would have made when their number is fixed (ik_goal_space / ik_box: one target sample per try).  Deviations,
 *     stated: the C rand() stream is NOT advanced by the restarts those solves would have made, and for the variable-draw
 *     randomizers (set_robot_random_pose, ..._ik_sphere) the numpy stream is not advanced either -- it only feeds
 *     set_random_num_obs afterwards (reachao_rand_start).
 * cfg.pose_retry_limit > 0 replaces the 10 000 (a knob for large batches, where the one environment in a thousand whose every
 * pose fails for a reason the core test does not see holds a whole step for 10 000 solves; 0 = as the reference). */
static int draws_per_goal_sample(const pgo_env* e) { return e->goal_sampler == 0 ? 3 : 3 + (e->goal_front ? 1 : 0) + (e->goal_34 ? 1 : 0); }
static int set_coll_free_robot(pgo_env* e) {
  int status = 0;
  const int limit = e->cfg.pose_retry_limit > 0 ? e->cfg.pose_retry_limit : 10000; /* reach_ao.py:1047: i > 9999 */
  for (int j = 0; j < e->nobs; j++) {
    int i = 0;
    if (!e->cfg.no_shortcuts && 0.05 >= robot_body_distance(e, &e->obs[j]) && robot_core_within(e, &e->obs[j], 0.05)) {
      if (e->pose_kind == PGO_POSE_RANDOM_BASE) {      /* one draw per try: the pose the loop ends with is its last draw -- EXACT */
        pgo_pcg64_advance(&e->rng, (uint64_t)limit - 1);
        random_robot_base(e);
      } else {
        if (e->pose_kind == PGO_POSE_IK_GOAL_SPACE) pgo_pcg64_advance(&e->rng, (uint64_t)limit * (uint64_t)draws_per_goal_sample(e));
        if (e->pose_kind == PGO_POSE_IK_BOX || e->pose_kind == PGO_POSE_BULLET_IK_BOX) pgo_pcg64_advance(&e->rng, 3ull * (uint64_t)limit);
        for (int d = 0; d < N_ARM; d++) pgo_arm_reset_joint(e->arm, d, e->neutral[d]);
      }
      status |= 4;
      continue;
    }
    while (0.05 >= robot_body_distance(e, &e->obs[j])) {
      if (i > limit - 1) { /* set_joint_neutral */
        for (int d = 0; d < N_ARM; d++) pgo_arm_reset_joint(e->arm, d, d == 0 && e->pose_kind == PGO_POSE_RANDOM_BASE ? e->neutral0 : e->neutral[d]);
        status |= 4; break;
      }
      i++;
      status |= robot_reset(e);
    }
  }
  return status;
}

int pgo_env_reset(pgo_env* e, uint64_t seed, float* obs, float* ag, float* dg, int* success) {
  int status = 0;
  const int rpose = e->cfg.task == 1 && e->pose_active;
  pgo_pcg64_seed(&e->rng, seed);                       /* core.py:302 */
  if (rpose) status |= robot_reset(e);                 /* robot.reset = robot_pose_randomizer (reach_ao.py:132-133) */
  else for (int d = 0; d < N_ARM; d++) pgo_arm_reset_joint(e->arm, d, e->neutral[d]); /* panda.py:290-298 */
  e->steps = 0;
  if (e->cfg.task == 0) {
    for (int k = 0; k < 3; k++) e->goal[k] = pgo_pcg64_uniform(&e->rng, e->goal_lo[k], e->goal_hi[k]); /* reach.py:75-78 */
  } else if (e->cfg.task >= 2) {
    /* Push.reset (push.py:69-87) / PickAndPlace.reset (pick_and_place.py:65-85): goal then object, xy range 0.3;
     * PickAndPlace lifts the goal by U(0, 0.2) and flattens it again with probability 0.3 */
    const double half = e->world.half[2], hi_z = e->cfg.task == 3 ? 0.2 : 0.0;
    double noise[3];
    noise[0] = pgo_pcg64_uniform(&e->rng, -0.15, 0.15); noise[1] = pgo_pcg64_uniform(&e->rng, -0.15, 0.15);
    noise[2] = pgo_pcg64_uniform(&e->rng, 0.0, hi_z);
    if (e->cfg.task == 3 && pgo_pcg64_double(&e->rng) < 0.3) noise[2] = 0.0;
    e->goal[0] = 0.0 + noise[0]; e->goal[1] = 0.0 + noise[1]; e->goal[2] = half + noise[2];
    noise[0] = pgo_pcg64_uniform(&e->rng, -0.15, 0.15); noise[1] = pgo_pcg64_uniform(&e->rng, -0.15, 0.15);
    noise[2] = pgo_pcg64_uniform(&e->rng, 0.0, 0.0);
    memset(&e->obj, 0, sizeof e->obj);                   /* set_base_pose: identity orientation, velocities zeroed */
    e->obj.pos[0] = 0.0 + noise[0]; e->obj.pos[1] = 0.0 + noise[1]; e->obj.pos[2] = half + noise[2];
    e->obj.quat[3] = 1.0;
  } else {
    e->goal_reached = 0; e->is_collided = 0;
    if (e->random_size_cuboids)                          /* reach_ao.py:969-979, 1084-1089: every cuboid is removed and re-created with */
      for (int j = 0; j < e->nobs; j++)                  /* half extents 0.2 * Dirichlet(1, 1, 1) at create_obstacle_cuboid's default position */
        if (e->obs[j].type == 1) {
          double f[3]; pgo_pcg64_dirichlet111(&e->rng, f);
          for (int k = 0; k < 3; k++) e->obs[j].size[k] = 0.2 * f[k];
          e->obs[j].pos[0] = 0.1; e->obs[j].pos[1] = 0.0; e->obs[j].pos[2] = 0.1;
        }
    if (e->cfg.has_fixed_target) memcpy(e->goal, e->cfg.fixed_target, 24); /* self.goal = config.fixed_target (:381, 410, 418) */
    else status |= set_coll_free_goal(e, 0, 0.1);      /* reach_ao.py:981-982 */
    if (e->randomize_obstacles) status |= set_coll_free_obs(e, 0.03);
    else status |= set_coll_free_goal(e, 1, 0.03);
    if (rpose) status |= set_coll_free_robot(e);       /* reach_ao.py:991-992 */
    e->dummy[0] = 0; e->dummy[1] = 0; e->dummy[2] = -5.0;
    if (e->random_num_obs) {                           /* reach_ao.py:1062-1082 */
      u32buf b = {0, 0};
      int n = e->nobs;
      int k = e->size_lo + (e->size_hi - 1 - e->size_lo == 0 ? 0 : (int)lemire_u32(e, &b, (uint32_t)(e->size_hi - 1 - e->size_lo)));
      int to_move = abs(k - n), order[PGO_MAXOBS];
      for (int i = 0; i < n; i++) order[i] = i;
      for (int i = n - 1; i >= 1; i--) { int j = (int)interval_u32(e, &b, (uint32_t)i); int t = order[i]; order[i] = order[j]; order[j] = t; }
      for (int i = 0; i < n && to_move > 0; i++, to_move--) { e->obs[order[i]].pos[0] = 99.9; e->obs[order[i]].pos[1] = 99.9; e->obs[order[i]].pos[2] = -99.9; }
    }
    e->have_recent = e->have_prev = 0;                 /* reach_ao.py:1025-1026 */
    e->npast = 0;                                      /* reach_ao.py:1028-1033 */
    if (e->cfg.obstacle_obs == 5 && e->nobs > 0) { float tmp[81]; task_obs(e, tmp); task_obs(e, tmp); }
  }
  float a[3], g[3];
  pgo_env_get_obs(e, obs, a, g);
  if (ag) memcpy(ag, a, 12);
  if (dg) memcpy(dg, g, 12);
  if (success) *success = is_success(e, a, g);
  return status;
}

/* ------------------------------------------------------------------ step */
static void set_action(pgo_env* e, const float* action) {
  int na = pgo_env_act_dim(e);
  float a[8];
  for (int k = 0; k < na; k++) a[k] = action[k];
  if (e->cfg.action_limiter == 1) { /* "scale" (panda.py:129-133) */
    int any = 0; float mx = 0;
    for (int k = 0; k < na; k++) { if (a[k] < -1 || a[k] > 1) any = 1; if (fabsf(a[k]) > mx) mx = fabsf(a[k]); }
    if (any) for (int k = 0; k < na; k++) a[k] = a[k] / mx;
  } else for (int k = 0; k < na; k++) a[k] = a[k] < -1.f ? -1.f : (a[k] > 1.f ? 1.f : a[k]);
  memcpy(e->prev, e->recent, sizeof e->prev); e->have_prev = e->have_recent;
  memcpy(e->recent, a, sizeof a); e->have_recent = 1;
  double q[PGO_MAXD], qd[PGO_MAXD], tgt[7];
  pgo_arm_get_q(e->arm, q, qd);
  if (e->cfg.control == 1) { /* ee (panda.py:226-246) */
    double p[3], ik[PGO_MAXD];
    ee_position(e, p);
    for (int k = 0; k < 3; k++) { volatile float s = a[k] * 0.05f; p[k] += (double)s; }
    if (p[2] < 0) p[2] = 0;
    const double quat[4] = {1.0, 0.0, 0.0, 0.0};
    pgo_arm_ik(e->arm, EE_LINK, p, quat, ik);
    for (int d = 0; d < 7; d++) tgt[d] = ik[d];
  } else {
    for (int d = 0; d < 7; d++) { volatile float s = a[d] * 0.05f; tgt[d] = q[d] + (double)s; } /* panda.py:257-261 */
  }
  if (e->cfg.control == 2) { for (int d = 0; d < 7; d++) pgo_arm_set_motor(e->arm, d, 0, (double)a[d], FORCES[d]); }
  else if (e->cfg.control == 3) { for (int d = 0; d < 7; d++) pgo_arm_reset_joint(e->arm, d, tgt[d]); }
  else { for (int d = 0; d < 7; d++) pgo_arm_set_motor(e->arm, d, 1, tgt[d], FORCES[d]); }
  /* panda.py:167-172 (velocities sampled AFTER a possible teleport, before the sim step) */
  pgo_arm_get_q(e->arm, q, qd);
  for (int d = 0; d < 7; d++) {
    e->prev_vel[d] = e->cur_vel[d]; e->prev_acc[d] = e->cur_acc[d];
    e->cur_vel[d] = qd[d];
    e->cur_acc[d] = e->prev_vel[d] - e->cur_vel[d];
    e->jerk[d] = fabs(e->prev_acc[d] - e->cur_acc[d]);
  }
}

static double norm7(const double* v) { double s = 0; for (int d = 0; d < 7; d++) s += v[d] * v[d]; return sqrt(s); }

/* what evaluation/evaluate.py:perform_benchmark reads after every env.step (:252-259): task.get_norm_effort()
 * (reach_ao.py:1299-1302), task.get_norm_jerk() (:1304-1305), task.manipulability (:1311, :1381 -> panda.py:314-317 at the
 * current joint angles) and ||robot.get_ee_velocity()|| */
void pgo_env_step_metrics(pgo_env* e, double out[4]) {
  double v[3]; pgo_arm_link_state(e->arm, EE_LINK, 0, 0, v, 0);
  out[0] = norm7(e->cur_acc); out[1] = norm7(e->jerk);
  out[2] = pgo_arm_manipulability(e->arm);
  out[3] = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
}

static float reward_one(pgo_env* e, const float* ag, const float* dg) {
  const pgo_env_config* c = &e->cfg;
  float d = f32_distance(ag, dg);
  float thr = (float)c->ee_error_threshold;
  if (c->task != 1) return c->reward_type == 0 ? -(float)(d > thr) : -d; /* reach.py:84-89, push.py:93-98, pick_and_place.py:91-96 */
  double v[3]; pgo_arm_link_state(e->arm, EE_LINK, 0, 0, v, 0);
  double speed = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
  double effort = norm7(e->cur_acc), jerk = norm7(e->jerk);
  float vthr = (float)c->speed_threshold;
  double r;
  switch (c->reward_type) {
    case 0:
      if (c->goal_condition == 0) { float ee = d + (float)e->is_collided; r = -1.0 + (double)(ee < thr); }
      else r = 1.0 - (double)((d < thr) && (speed < (double)vthr));
      break;
    case 1: { /* wang */
      double dd = (double)d, obs = 0;
      for (int i = 0; i < PGO_NCL; i++) { double t = 1 - e->d_link[i] / 0.05; obs += t > 0 ? t : 0; }
      r = -(double)(float)(10e-3 * dd * dd + log(dd * dd + 10e-4) + 0.1 * obs);
      break; }
    case 2: /* kumar_her */
      if (c->goal_condition == 0) r = -(double)(float)((double)(d > thr) * jerk);
      else r = (double)(float)((d < thr) && (speed < (double)vthr)) - jerk;
      break;
    case 3: r = -(double)(float)(d > thr) - effort; break; /* kumar_optim */
    default: { /* kumar */
      double dd = (double)d, obs = 0;
      for (int i = 0; i < PGO_NCL; i++) { double t = 1 - e->d_link[i] / 0.05; obs += t > 0 ? t : 0; }
      r = (double)(float)(exp(-20.0 * dd * dd) - 0.005 * effort - 0.1 * obs);
      break; }
  }
  if (c->truncate_on_collision && (c->reward_type == 0 || c->reward_type == 2 || c->reward_type == 3))
    r += (double)e->is_collided * c->collision_reward;
  return (float)r;
}

void pgo_env_compute_reward(pgo_env* e, const float* ag, const float* dg, int n, float* out) {
  for (int i = 0; i < n; i++) out[i] = reward_one(e, ag + 3 * i, dg + 3 * i);
}

void pgo_env_step(pgo_env* e, const float* action, float* obs, float* ag, float* dg, float* reward, int* flags) {
  set_action(e, action);
  for (int k = 0; k < e->cfg.n_substeps; k++) {
    if (e->cfg.task >= 2) pgo_world_substep(e->arm, &e->obj, &e->world); else
    pgo_arm_substep(e->arm);
    if (e->cfg.task == 1) { /* step_check_collision (reach_ao.py:182-188) */
      e->is_collided = pgo_env_check_collided(e, 0, 0);
      if (e->is_collided) break;
    }
  }
  float a[3], g[3];
  pgo_env_get_obs(e, obs, a, g);
  if (ag) memcpy(ag, a, 12);
  if (dg) memcpy(dg, g, 12);
  int succ = is_success(e, a, g);
  int terminated = e->cfg.terminate_on_success ? succ : 0;
  int truncated = (e->cfg.task == 1 && e->cfg.truncate_on_collision) ? e->is_collided : 0;
  *reward = reward_one(e, a, g);
  e->steps++;
  flags[0] = succ; flags[1] = terminated; flags[2] = truncated;
  flags[3] = e->steps >= e->cfg.max_episode_steps; /* gymnasium TimeLimit */
}

/* test hook: after `skip` Generator.random() draws, k = Generator.integers(lo, hi) then Generator.shuffle(list(range(n))) */
int pgo_test_integers_shuffle(uint64_t seed, int skip, int lo, int hi, int n, int* perm) {
  pgo_env e; memset(&e, 0, sizeof e);
  pgo_pcg64_seed(&e.rng, seed);
  for (int i = 0; i < skip; i++) pgo_pcg64_double(&e.rng);
  u32buf b = {0, 0};
  int k = lo + (hi - 1 - lo == 0 ? 0 : (int)lemire_u32(&e, &b, (uint32_t)(hi - 1 - lo)));
  for (int i = 0; i < n; i++) perm[i] = i;
  for (int i = n - 1; i >= 1; i--) { int j = (int)interval_u32(&e, &b, (uint32_t)i); int t = perm[i]; perm[i] = perm[j]; perm[j] = t; }
  return k;
}

/* CPU baseline driver: `steps` env-steps of one environment with U(-1,1) actions from PCG64(seed) and
 * SB3-style auto-reset (fresh seed per episode).  Returns the number of episodes finished. */
int pgo_bench_rollout(const pgo_env_config* cfg, uint64_t seed, int steps, double* checksum) {
  pgo_env* e = pgo_env_create(cfg);
  pgo_pcg64 ar; pgo_pcg64_seed(&ar, seed ^ 0xabcdef);
  float obs[128], ag[3], dg[3], r, a[8];
  int flags[4], succ, episodes = 0;
  double cs = 0;
  pgo_env_reset(e, seed, obs, ag, dg, &succ);
  for (int s = 0; s < steps; s++) {
    for (int k = 0; k < 7; k++) a[k] = (float)(2.0 * pgo_pcg64_double(&ar) - 1.0);
    pgo_env_step(e, a, obs, ag, dg, &r, flags);
    cs += r;
    if (flags[1] || flags[2] || flags[3]) { episodes++; pgo_env_reset(e, seed + 7919u * (uint64_t)episodes, obs, ag, dg, &succ); }
  }
  if (checksum) *checksum = cs;
  pgo_env_destroy(e);
  return episodes;
}

/* test hook: utils.distance on float32 rows (panda_gym/utils.py:4-16) */
void pgo_test_distance(const float* a, const float* b, int n, float* out) {
  for (int i = 0; i < n; i++) out[i] = f32_distance(a + 3 * i, b + 3 * i);
}

/* accessors for the render oracle (pgo_render.c) */
void pgo_env_get_config(const pgo_env* e, pgo_env_config* out) { *out = e->cfg; }
void pgo_env_get_table(const pgo_env* e, double pos[3], double half[3]) {
  const double* p = e->cfg.task >= 2 ? e->world.table_pos : e->table_pos;
  const double* h = e->cfg.task >= 2 ? e->world.table_half : e->table_half;
  for (int k = 0; k < 3; k++) { pos[k] = p[k]; half[k] = h[k]; }
}
