// pg_api.cu -- C ABI (include/panda_b200.h) over the kernels of pg_kernels.cuh.  Host side only: handle
// bookkeeping, device allocations (the library owns every buffer), parameter blocks, launches.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/panda_b200.h"
#include "pg_kernels.cuh"
#include "pg_her.cuh"
#include "pg_launch.h"

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define PG_CUDA(call)                                                                                         \
  do {                                                                                                        \
    cudaError_t _e = (call);                                                                                  \
    if (_e != cudaSuccess) return fail(PG_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e));      \
  } while (0)

struct Buffer { void** slot; size_t bytes; };

}  // namespace

struct pg_env_s {
  pg_config cfg;
  int device;
  pg::Params P;
  std::vector<Buffer> state_bufs;   // everything save_state snapshots
  std::vector<void*> all_bufs;
  std::map<int, std::vector<void*>> saved;
  uint64_t first_env;
  int64_t launches;
  float* d_actions;                 // staging for the host-buffer step
  int eval_metrics;                 // pg_set_eval_metrics: launch k_eval_metrics after every step
  int sm_count;                     // cudaDevAttrMultiProcessorCount (grid of the auto-reset kernel)
  uint8_t* out_slab;                // per-step results, one allocation (see pg_create)
  size_t slab_obs_off;
  float* term_ring;                 // [3][n][term_w_max] compact rows of finished episodes (Params::term_rows points into it)
  int term_w_max;
  int64_t step_id;                  // pg_step calls so far; ring slot = step_id % 3
  int* finished_host;               // page-locked [4]: finished-episode count of the last three steps (by ring slot)
  float* sbox;                      // static scene boxes (pg_set_scene), [64][8] floats
  bool dlink_valid;                 // d_link / q_dl of EVERY environment come from an observation pass over the current obstacles
  bool cull_enabled;                // PANDA_B200_SCENE_CULL != 0 (default on): temporal culling of the flag pass, see scene_skip_mask
};

namespace {

template <typename T>
int alloc(pg_env_s* h, T** p, size_t count, bool is_state) {
  size_t bytes = count * sizeof(T);
  PG_CUDA(cudaMalloc((void**)p, bytes));
  PG_CUDA(cudaMemset(*p, 0, bytes));
  h->all_bufs.push_back(*p);
  if (is_state) h->state_bufs.push_back(Buffer{(void**)p, bytes});
  return PG_OK;
}

double box_margin(const double* s) { double m = fmin(s[0], fmin(s[1], s[2])); return fmin(0.04, 0.1 * m); }

void fill_stage(pg_env_s* h, const pg_stage* s) {
  pg::Params& P = h->P;
  pg::StageDev& d = P.st;
  d.goal_sampler = s->goal_sampler;
  for (int k = 0; k < 3; k++) { d.goal_lo[k] = s->goal_lo[k]; d.goal_hi[k] = s->goal_hi[k]; }
  d.goal_rmin = s->goal_rmin; d.goal_rmax = s->goal_rmax;
  d.goal_upper = s->goal_upper_half; d.goal_front = s->goal_front_half; d.goal_34 = s->goal_three_quarter_front;
  d.nobs = s->n_obstacles;
  P.scene.nobs = s->n_obstacles;
  P.scene.has_box = P.scene.nstatic > 0;
  for (int j = 0; j < s->n_obstacles; j++) P.scene.has_box |= s->obstacle_type[j] == 1;
  for (int j = 0; j < pg::kMaxObst; j++) {
    d.obs_type[j] = j < s->n_obstacles ? s->obstacle_type[j] : 0;
    for (int k = 0; k < 3; k++) d.obs_size[j][k] = j < s->n_obstacles ? s->obstacle_size[j][k] : 0.0;
    pg::Obst& o = P.scene.obst[j];
    o.type = d.obs_type[j];
    o.sx = (float)d.obs_size[j][0]; o.sy = (float)d.obs_size[j][1]; o.sz = (float)d.obs_size[j][2];
    o.margin = o.type == 1 ? (float)box_margin(d.obs_size[j]) : 0.f;
    P.scene.obst_bound[j] = o.type == 1 ? sqrtf(o.sx * o.sx + o.sy * o.sy + o.sz * o.sz) : o.sx;
  }
  d.obs_sampler = s->obstacle_sampler; d.randomize_obstacles = s->randomize_obstacle_position;
  d.allow_overlap = s->allow_overlapping_obstacles; d.random_num_obs = s->random_num_obs;
  d.size_lo = s->sample_size_lo; d.size_hi = s->sample_size_hi;
  d.pose_active = (h->cfg.task == PG_TASK_REACH_AO && (h->cfg.randomize_robot_pose || s->force_randomize_pose)) ? 1 : 0;
  d.pose_kind = s->pose_randomizer;
  d.random_base = (d.pose_active && s->pose_randomizer == PG_POSE_RANDOM_BASE) ? 1 : 0;
  for (int k = 0; k < 3; k++) { d.pose_lo[k] = s->pose_lo[k]; d.pose_hi[k] = s->pose_hi[k]; }
  d.random_size = (h->cfg.task == PG_TASK_REACH_AO && s->random_size_cuboids) ? 1 : 0;
  P.scene.env_box = d.random_size ? P.env_box : nullptr;
  d.max_episode_steps = s->max_episode_steps;
  d.ee_thr = (float)s->ee_error_threshold; d.speed_thr = (float)s->speed_threshold;
  // observation width (reach_ao.py:902-932: np.ones(27) when the scenario has no obstacles)
  int t = 0;
  if (P.task == 1 && P.obstacle_obs != 0) {
    if (s->n_obstacles == 0 && !s->needs_scene) t = 27;  // a benchmark scene's bodies are obstacles (reach_ao.py:300-306)
    else t = P.obstacle_obs == 1 ? 36 : (P.obstacle_obs == 2 ? 27 : (P.obstacle_obs == 3 ? 9 : (P.obstacle_obs == 5 ? 81 : 1)));
  }
  if (P.task >= 2) t = 12;  // push.py:49-63
  P.task_obs_dim = t;
  P.obs_dim = 6 + (P.obs_js ? 14 : 0) + (P.task == PG_TASK_PICK_AND_PLACE ? 1 : 0) + t;  // + fingers width (panda.py:282-284)
}

// randomize_robot_pose (train_config.py:26): every scenario randomizer is built; the combinations that are not are rejected here
int check_pose_randomizer(const pg_config* cfg, const pg_stage* s) {
  if (cfg->task != PG_TASK_REACH_AO || !(cfg->randomize_robot_pose || s->force_randomize_pose)) return PG_OK;
  if (s->pose_randomizer < PG_POSE_TORUS || s->pose_randomizer > PG_POSE_BULLET_IK_BOX) return fail(PG_ERR_ARG, "unknown pose_randomizer");
  if (s->pose_randomizer == PG_POSE_RANDOM_BASE)  // the neutral-pose table is turned about joint 1: exact for sphere bodies only
    for (int j = 0; j < s->n_obstacles; j++)
      if (s->obstacle_type[j] != 0) return fail(PG_ERR_UNSUPPORTED, "set_random_robot_base with cuboid obstacles is not built");
  return PG_OK;
}

void set_goal_box(pg_stage* s, double xl, double xh, double yl, double yh, double zl, double zh) {
  s->goal_sampler = PG_GOAL_BOX;
  s->goal_lo[0] = xl; s->goal_lo[1] = yl; s->goal_lo[2] = zl; s->goal_hi[0] = xh; s->goal_hi[1] = yh; s->goal_hi[2] = zh;
}
void set_goal_shell(pg_stage* s, double rmin, double rmax, int upper, int front, int q34) {
  s->goal_sampler = PG_GOAL_SHELL; s->goal_rmin = rmin; s->goal_rmax = rmax;
  s->goal_upper_half = upper; s->goal_front_half = front; s->goal_three_quarter_front = q34;
}
void add_sphere(pg_stage* s, double r) { int j = s->n_obstacles++; s->obstacle_type[j] = 0; s->obstacle_size[j][0] = r; s->obstacle_size[j][1] = s->obstacle_size[j][2] = 0; }
void add_cuboid(pg_stage* s, double hx) { int j = s->n_obstacles++; s->obstacle_type[j] = 1; s->obstacle_size[j][0] = s->obstacle_size[j][1] = s->obstacle_size[j][2] = hx; }

}  // namespace

extern "C" {

const char* pg_last_error(void) { return g_err.c_str(); }
const char* pg_version(void) { return "panda_b200 0.1 (sm_100a)"; }

int pg_stage_from_scenario(pg_stage* s, const char* scenario) {
  if (!s) return fail(PG_ERR_ARG, "stage is NULL");
  memset(s, 0, sizeof *s);
  s->max_episode_steps = 50; s->ee_error_threshold = 0.05; s->speed_threshold = 0.1;
  // ReachAO defaults (reach_ao.py:73-82): goal box x 0.6 +- 0.12, y +- 0.2, z [0, 0.3]
  set_goal_box(s, -0.3 / 2.5 + 0.6, 0.3 / 2.5 + 0.6, -0.3 / 1.5, 0.3 / 1.5, 0.0, 0.3);
  if (!scenario || !*scenario) return PG_OK;
  std::string full(scenario), name = full;
  int count = 0;
  size_t dash = full.find('-');  // reach_ao.py:226-227
  if (dash != std::string::npos) { name = full.substr(0, dash); count = atoi(full.c_str() + dash + 1); }
  if (name == "reach_classic") {          // Reach (reach.py:24-25), goal_range 0.3
    set_goal_box(s, -0.15, 0.15, -0.15, 0.15, 0.0, 0.3);
  } else if (name == "reach1") {          // reach_ao.py:518-522
    set_goal_box(s, -0.4 / 2 + 0.6, 0.4 / 2 + 0.6, -0.4 / 2, 0.4 / 2, 0, 0.4);
    s->pose_randomizer = PG_POSE_TORUS_FRONT;
  } else if (name == "reach2" || name == "reach3") {  // :524-539
    set_goal_shell(s, 0.5, 0.85, 1, 0, 1);
    s->pose_randomizer = name == "reach2" ? PG_POSE_TORUS_FRONT : PG_POSE_TORUS;
  } else if (name == "reachao1") {        // :541-545
    set_goal_box(s, -0.4 / 2 + 0.6, 0.4 / 2 + 0.6, -0.4 / 2, 0.4 / 2, 0, 0.4);
    s->pose_randomizer = PG_POSE_TORUS_FRONT;  // inherited from create_scenario_reach1
    s->randomize_obstacle_position = 1; add_sphere(s, 0.04);
  } else if (name == "reachao2") {        // :547-564
    set_goal_shell(s, 0.5, 0.8, 1, 1, 0); s->randomize_obstacle_position = 1; s->obstacle_sampler = PG_OSMP_WANG;
    s->pose_randomizer = PG_POSE_IK_GOAL_SPACE;
    add_sphere(s, 0.05); add_sphere(s, 0.05);
  } else if (name == "reachao3") {        // :573-585
    set_goal_shell(s, 0.5, 0.8, 1, 0, 0); s->randomize_obstacle_position = 1; s->obstacle_sampler = PG_OSMP_WANG;
    s->pose_randomizer = PG_POSE_IK_GOAL_SPACE;
    for (int i = 0; i < 3; i++) add_sphere(s, 0.05);
  } else if (name == "reachao_rand" || name == "reachao_rand_start" || name == "reachao_rand_shape") {    // :587-608
    set_goal_shell(s, 0.5, 0.8, 1, 0, 0); s->randomize_obstacle_position = 1; s->obstacle_sampler = PG_OSMP_EXPERIMENTAL;
    s->pose_randomizer = PG_POSE_IK_GOAL_SPACE;  // inherited from create_scenario_reachao3
    for (int i = 0; i < 3; i++) add_sphere(s, 0.05);
    for (int i = 0; i < 3; i++) add_cuboid(s, 0.05);
    s->random_num_obs = 1; s->allow_overlapping_obstacles = 1; s->sample_size_lo = 4; s->sample_size_hi = 6;
    if (name == "reachao_rand_start") {   // :601-604
      s->force_randomize_pose = 1; s->pose_randomizer = PG_POSE_IK_SPHERE; s->pose_lo[0] = 0.45; s->pose_hi[0] = 0.7;
    }
    if (name == "reachao_rand_shape") s->random_size_cuboids = 1;  // :606-608
  } else if (name == "wang") {            // :646-699
    if (count < 0 || count > PG_MAX_OBSTACLES) return fail(PG_ERR_ARG, "wang-N: N out of range");
    set_goal_shell(s, 0.4, 0.95, 1, 0, 0); s->randomize_obstacle_position = 1; s->obstacle_sampler = PG_OSMP_WANG_N;
    for (int i = 0; i < count; i++) add_sphere(s, 0.05);
  } else if (name == "wangexp") {         // :701-722
    if (count < 0 || count > PG_MAX_OBSTACLES) return fail(PG_ERR_ARG, "wangexp-N: N out of range");
    set_goal_shell(s, 0.5, 0.8, 1, 0, 0); s->randomize_obstacle_position = 1; s->obstacle_sampler = PG_OSMP_WANG;
    for (int i = 0; i < count; i++) add_sphere(s, 0.05);
    s->pose_randomizer = PG_POSE_RANDOM_BASE;  // :711
  } else {
    // benchmark scenes (reach_ao.py:308-516): goal box and, where the builder replaces it, the neutral pose; the scene's
    // boxes are loaded on the host (panda_gym_b200/scenes.py) and handed over with pg_set_scene
    struct SceneDef { const char* name; double lo[3], hi[3]; int neutral; };
    static const double kNeutralSets[5][7] = {
        {0, 0, 0, 0, 0, 0, 0},
        {-1.0, -0.3, 0.0, -2.2, 0.0, 2.0, 0.7853981633974483},                                        // tunnel / narrow_tunnel
        {0.00887326, -0.05377409, -0.03621967, -1.9094068, 0.08791409, 2.00265486, 0.76681184},        // workshop*
        {0.0, 0.12001979, 0.0, -1.64029458, 0.02081271, 3.1, 0.77979846},                              // library*
        {0.94551719, 0.65262327, 0.12742699, -1.74347465, -0.16996126, 1.97424632, 0.88058222}};       // wall (create_stage_wall :457-468)
    static const SceneDef kScenes[] = {
        {"bookshelves", {0.6, -0.35, 0.2}, {0.7, 0.35, 0.8}, 0},      {"warehouse", {0.5, -0.3, 0.0}, {0.85, 0.3, 0.3}, 0},
        {"countertop", {0.5, -0.3, 0.0}, {0.85, 0.3, 0.3}, 0},        {"kitchen", {0.5, -0.3, 0.0}, {0.85, 0.3, 0.3}, 0},
        {"raised_shelves", {0.5, -0.3, 0.0}, {0.85, 0.3, 0.3}, 0},    {"tabletop", {0.5, -0.3, 0.0}, {0.85, 0.3, 0.3}, 0},
        {"tabletop2", {0.5, -0.3, 0.0}, {0.85, 0.3, 0.3}, 0},         {"industrial", {0.5, -0.1, 0.55}, {0.6, 0.1, 0.75}, 0},
        {"tunnel", {0.55, 0.2, 0.2}, {0.75, 0.4, 0.75}, 1},           {"narrow_tunnel", {0.55, 0.2, 0.2}, {0.75, 0.4, 0.75}, 1},
        {"workshop", {-0.7, -0.7, 0.4}, {0.1, -0.4, 0.7}, 2},         {"workshop2", {0.5, -0.15, 0.4}, {0.6, 0.15, 0.5}, 2},
        {"workshop3", {-0.7, -0.7, 0.4}, {0.1, -0.4, 0.7}, 2},        {"library", {0.2, -0.3, 0.0}, {0.7, 0.3, 0.6}, 3},
        {"library1", {0.5, -0.3, 0.0}, {0.85, 0.3, 0.3}, 3},          {"library2", {-0.7, -0.4, 0.4}, {-0.55, 0.4, 0.85}, 3},
        {"wall", {0.45, -0.6, 0.1}, {0.7, -0.1, 0.3}, 4}};  // its one cuboid stays where create_obstacle_cuboid puts it: a static box
    const SceneDef* def = nullptr;
    for (const SceneDef& d : kScenes) if (full == d.name) def = &d;
    if (full == "kasys") return fail(PG_ERR_UNSUPPORTED, "Scenario kasys: its meshes (trolley.obj, wall.obj) are not part of the reference repository");
    if (!def) return fail(PG_ERR_ARG, std::string("Scenario ") + scenario + " not found!");  // reach_ao.py:262-264
    set_goal_box(s, def->lo[0], def->hi[0], def->lo[1], def->hi[1], def->lo[2], def->hi[2]);
    s->needs_scene = 1;
    // scene builders that install their own start-pose randomizer and switch it on (reach_ao.py:310-320, 363-375, 395-403)
    struct PoseDef { const char* name; int kind; double lo[3], hi[3]; };
    static const PoseDef kPoses[] = {{"narrow_tunnel", PG_POSE_IK_BOX, {0.0, -0.6, 0.2}, {0.5, -0.5, 0.7}},
                                     {"workshop2", PG_POSE_IK_BOX, {-0.5, -0.6, 0.6}, {0.2, -0.3, 0.7}},
                                     {"industrial", PG_POSE_BULLET_IK_BOX, {-0.5, -0.8, 0.4}, {0.2, -0.4, 0.7}}};
    for (const PoseDef& pd : kPoses)
      if (full == pd.name) {
        s->force_randomize_pose = 1; s->pose_randomizer = pd.kind;
        for (int k = 0; k < 3; k++) { s->pose_lo[k] = pd.lo[k]; s->pose_hi[k] = pd.hi[k]; }
      }
    if (def->neutral) { s->custom_neutral = 1; for (int d = 0; d < 7; d++) s->neutral_q[d] = kNeutralSets[def->neutral][d]; }
  }
  return PG_OK;
}

int pg_default_config(pg_config* c, int task, const char* scenario) {
  if (!c) return fail(PG_ERR_ARG, "cfg is NULL");
  memset(c, 0, sizeof *c);
  c->n_envs = 1; c->task = task; c->control = PG_CTRL_JOINTS; c->reward_type = PG_REW_SPARSE;
  c->n_substeps = 20; c->collision_reward = -100.0; c->auto_reset = 1;
  if (task == PG_TASK_REACH) {  // PandaReachEnv (panda_tasks.py:72-88): obs_type ("ee",), base (-0.6, 0, 0), no termination
    c->obs_js = 0; c->obstacle_obs = PG_OOBS_NONE; c->base_position[0] = -0.6;
    return pg_stage_from_scenario(&c->stage, "reach_classic");
  }
  if (task == PG_TASK_PUSH || task == PG_TASK_PICK_AND_PLACE) {
    // PandaPush-v3 / PandaPickAndPlace-v3 (panda_tasks.py:37-72, __init__.py:58-77): ee control, obs_type ("ee",), base (-0.6, 0, 0)
    c->control = PG_CTRL_EE; c->obs_js = 0; c->obstacle_obs = PG_OOBS_NONE; c->base_position[0] = -0.6;
    return pg_stage_from_scenario(&c->stage, "reach_classic");
  }
  if (task != PG_TASK_REACH_AO) return fail(PG_ERR_ARG, "unknown task");
  c->obs_js = 1; c->obstacle_obs = PG_OOBS_VECTORS_CLOSEST_PER_LINK; c->truncate_on_collision = 1; c->terminate_on_success = 1;
  return pg_stage_from_scenario(&c->stage, scenario ? scenario : "reachao1");
}

int pg_create(const pg_config* cfg, int device, pg_handle* out) {
  if (!cfg || !out) return fail(PG_ERR_ARG, "NULL argument");
  if (cfg->n_envs <= 0) return fail(PG_ERR_ARG, "n_envs must be positive");
  if (cfg->stage.n_obstacles < 0 || cfg->stage.n_obstacles > PG_MAX_OBSTACLES) return fail(PG_ERR_ARG, "n_obstacles out of range");
  if (int rc = check_pose_randomizer(cfg, &cfg->stage)) return rc;
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev == 0) return fail(PG_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(PG_ERR_ARG, "bad device index");
  PG_CUDA(cudaSetDevice(device));
  int sms = 148;
  PG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  pg_env_s* h = new pg_env_s();
  h->sm_count = sms;
  h->dlink_valid = false; h->sbox = nullptr;
  { const char* v = getenv("PANDA_B200_SCENE_CULL"); h->cull_enabled = !(v && v[0] == '0'); }
  h->cfg = *cfg; h->device = device; h->launches = 0; h->first_env = 0; h->d_actions = nullptr; h->eval_metrics = 0;
  h->finished_host = nullptr; h->term_ring = nullptr; h->out_slab = nullptr;
  pg::Params& P = h->P;
  memset(&P, 0, sizeof P);
  const int n = cfg->n_envs;
  P.n = n; P.task = cfg->task; P.control = cfg->control; P.reward_type = cfg->reward_type; P.goal_condition = cfg->goal_condition;
  P.action_limiter = cfg->action_limiter; P.obs_js = cfg->obs_js; P.obstacle_obs = cfg->obstacle_obs;
  if (cfg->task < 0 || cfg->task > PG_TASK_PICK_AND_PLACE) { delete h; return fail(PG_ERR_ARG, "unknown task"); }
  // block_gripper = True except PickAndPlace (panda_tasks.py:49): its extra finger action drives nothing, the loaded
  // URDF has fixed finger joints
  P.act_dim = (cfg->control == PG_CTRL_EE ? 3 : 7) + (cfg->task == PG_TASK_PICK_AND_PLACE ? 1 : 0);
  P.n_substeps = cfg->n_substeps; P.truncate_on_collision = cfg->truncate_on_collision; P.terminate_on_success = cfg->terminate_on_success;
  P.auto_reset = cfg->auto_reset; P.collision_reward = (float)cfg->collision_reward; P.safety_distance = cfg->safety_distance;
  P.pose_retry_limit = cfg->pose_retry_limit;
  P.has_fixed_target = cfg->has_fixed_target; for (int k = 0; k < 3; k++) P.fixed_target[k] = cfg->fixed_target[k];
  for (int k = 0; k < 3; k++) { P.base[k] = (float)cfg->base_position[k]; P.based[k] = cfg->base_position[k]; }
  if (cfg->task != PG_TASK_REACH_AO) {  // table 1.1 x 0.7 x 0.4 at x = -0.3 (reach.py:33, push.py:31); Reach never queries it
    P.table_pos[0] = -0.3; P.table_half[0] = 0.55; P.table_half[1] = 0.35;
  } else {                           // table 2.0 x 1.3 x 0.4 (reach_ao.py:272)
    P.table_pos[0] = 0.0; P.table_half[0] = 1.0; P.table_half[1] = 0.65;
  }
  P.table_pos[1] = 0.0; P.table_pos[2] = -0.2; P.table_half[2] = 0.2;
  double tm = box_margin(P.table_half);
  P.scene.has_table = cfg->task == PG_TASK_REACH_AO;
  P.scene.table_margin = (float)tm;
  for (int k = 0; k < 3; k++) { P.scene.table_pos[k] = (float)P.table_pos[k]; P.scene.table_core[k] = (float)(P.table_half[k] - tm); }
  {  // Push / PickAndPlace world (push.py:29-47, pybullet.py:759-815)
    pg::WorldParams& W = P.world;
    W.half[0] = W.half[1] = W.half[2] = 0.02f; W.mass = 1.0f; P.obj_half_z = 0.04 / 2;
    for (int k = 0; k < 3; k++) { W.table_pos[k] = (float)P.table_pos[k]; W.table_half[k] = (float)P.table_half[k]; }
    W.plane_top = -0.4f;
    W.mu_object = 0.5f; W.mu_static = 0.5f; W.mu_link = 0.5f; W.mu_ee = 1.0f;  // Bullet default lateral friction; panda.py:69-70
  }
  fill_stage(h, &cfg->stage);
  // observation width is fixed at creation (gym spaces); allocate for the widest layout
  const int max_obs = 6 + 14 + 81;
  int rc = PG_OK;
#define A(ptr, count, st) if ((rc = alloc(h, &ptr, (size_t)(count), st)) != PG_OK) { pg_destroy(h); return rc; }
  A(P.q, 7 * n, true) A(P.qd, 7 * n, true) A(P.goal, 3 * n, true) A(P.obs_pos, (size_t)pg::kMaxObst * 3 * n, true) A(P.obs_pos_f, (size_t)pg::kMaxObst * 3 * n, true)
  A(P.recent, 7 * n, true) A(P.prev, 7 * n, true) A(P.cur_vel, 7 * n, true) A(P.cur_acc, 7 * n, true) A(P.jerk, 7 * n, true)
  A(P.nprims, pgm::kNumNeutralPrim, false) A(P.nextra, 13, false)
  A(P.d_link, 9 * n, true) A(P.q_dl, 7 * n, true) A(P.past_vec, 54 * n, true) A(P.have_recent, n, true) A(P.have_prev, n, true) A(P.collided, n, true) A(P.goal_reached, n, true)
  A(P.need_reset, n, true) A(P.steps, n, true) A(P.ep_len, n, true) A(P.last_len, n, true) A(P.ep_ret, n, true) A(P.last_ret, n, true)
  A(P.rng, 4 * n, true) A(P.episode, n, true) A(P.obj, 13 * n, true)
  A(P.env_size, (size_t)pg::kMaxObst * 3 * n, true) A(P.env_box, (size_t)pg::kMaxObst * 5 * n, true) A(P.crand, (size_t)33 * n, false)
  {  // the six per-step results live in ONE slab so that the host-buffer step moves them with a single device-to-host copy:
     // [achieved 3n f32 | desired 3n f32 | reward n f32 | done n u8 | flags n u8 | pad to 256 B | observation n x obs_dim f32]
     // (the observation comes last: its width depends on the stage, the copy stops after n x obs_dim)
    h->slab_obs_off = ((size_t)30 * n + 255) / 256 * 256;
    const size_t bytes = h->slab_obs_off + (size_t)max_obs * n * sizeof(float);
    A(h->out_slab, bytes, true)
    P.achieved = (float*)h->out_slab; P.desired = P.achieved + 3 * (size_t)n; P.reward = P.desired + 3 * (size_t)n;
    P.done = h->out_slab + (size_t)28 * n; P.flags = h->out_slab + (size_t)29 * n;
    P.obs = (float*)(h->out_slab + h->slab_obs_off);
  }
  A(P.term_obs, (size_t)max_obs * n, true) A(P.term_achieved, 3 * n, true)
  A(P.stats, 8, false) A(P.reset_list, n, false) A(P.reset_count, 1, false) A(P.hard_list, n, false) A(P.hard_count, 1, false)
  h->term_w_max = (8 + max_obs + 3) / 4 * 4;
  A(h->term_ring, (size_t)3 * n * h->term_w_max, false)
  P.term_rows = h->term_ring; P.term_w = (8 + P.obs_dim + 3) / 4 * 4;
  h->step_id = 0;
  if (cudaHostAlloc((void**)&h->finished_host, 4 * sizeof(int), cudaHostAllocDefault) != cudaSuccess) { pg_destroy(h); return fail(PG_ERR_CUDA, "cudaHostAlloc"); }
  memset(h->finished_host, 0, 4 * sizeof(int));
  A(P.ev_acc, 4 * n, true) A(P.ev_len, n, true) A(P.ev_ret, n, true) A(P.ev, 12, false)
  A(h->d_actions, 8 * n, false)
#undef A
  // obstacles start where create_obstacle_* puts them: (0.1, 0, 0.1) (reach_ao.py:819, 841)
  {
    std::vector<double> init((size_t)pg::kMaxObst * 3 * n);
    for (int j = 0; j < pg::kMaxObst; j++)
      for (int e = 0; e < n; e++) { init[(size_t)(j * 3 + 0) * n + e] = 0.1; init[(size_t)(j * 3 + 1) * n + e] = 0.0; init[(size_t)(j * 3 + 2) * n + e] = 0.1; }
    cudaError_t e2 = cudaMemcpy(P.obs_pos, init.data(), init.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (e2 != cudaSuccess) { pg_destroy(h); return fail(PG_ERR_CUDA, cudaGetErrorString(e2)); }
    pg::launch_mirror_obstacles(P, 0);
  }
  {  // every environment stands for one SubprocVecEnv worker process: its C rand() stream starts at glibc's srand(1) state
     // (the restart stream of roboticstoolbox's ik_lm_chan, pg_reset.cuh); r[31] | front index | back index
    uint32_t st[33];
    int32_t word = 1;
    st[0] = 1;
    for (int i = 1; i < 31; i++) {
      const int32_t hi = word / 127773, lo = word % 127773;
      word = 16807 * lo - 2836 * hi;
      if (word < 0) word += 2147483647;
      st[i] = (uint32_t)word;
    }
    int f = 3, b = 0;
    for (int i = 0; i < 310; i++) { st[f] += st[b]; f = (f + 1) % 31; b = (b + 1) % 31; }
    st[31] = (uint32_t)f; st[32] = (uint32_t)b;
    std::vector<uint32_t> init((size_t)33 * n);
    for (int e = 0; e < n; e++) memcpy(&init[(size_t)33 * e], st, sizeof st);
    cudaError_t e2 = cudaMemcpy(P.crand, init.data(), init.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e2 != cudaSuccess) { pg_destroy(h); return fail(PG_ERR_CUDA, cudaGetErrorString(e2)); }
  }
  fill_stage(h, &cfg->stage);  // again, now that the per-environment size buffer exists (scene.env_box)
  pg::launch_neutral_table(P, cfg->stage.custom_neutral ? cfg->stage.neutral_q : nullptr, 0);
  P.base_seed = 0;
  *out = h;
  return PG_OK;
}

int pg_destroy(pg_handle h) {
  if (!h) return PG_OK;
  cudaSetDevice(h->device);
  for (void* p : h->all_bufs) cudaFree(p);
  for (auto& kv : h->saved) for (void* p : kv.second) cudaFree(p);
  if (h->finished_host) cudaFreeHost(h->finished_host);
  delete h;
  return PG_OK;
}

int pg_set_stage(pg_handle h, const pg_stage* s, void* stream) {
  if (!h || !s) return fail(PG_ERR_ARG, "NULL argument");
  if (s->n_obstacles < 0 || s->n_obstacles > PG_MAX_OBSTACLES) return fail(PG_ERR_ARG, "n_obstacles out of range");
  if (int rc = check_pose_randomizer(&h->cfg, s)) return rc;
  h->cfg.stage = *s;
  fill_stage(h, s);
  pg::launch_neutral_table(h->P, s->custom_neutral ? s->neutral_q : nullptr, (cudaStream_t)stream);
  // The reference rebuilds the whole VecEnv for a new stage (setup_training.py:233-252), so every environment starts a
  // fresh episode in the new scene.  Same here: running episodes would otherwise meet newly enabled obstacle slots at
  // stale positions, goals of the old sampler and (when the observation width changes) a cached EE at the old stride.
  return pg_reset(h, nullptr, nullptr, stream);
}

int pg_seed(pg_handle h, uint64_t base_seed, uint64_t first_env_index) {
  if (!h) return fail(PG_ERR_ARG, "NULL handle");
  PG_CUDA(cudaSetDevice(h->device));
  h->P.base_seed = base_seed; h->first_env = first_env_index;
  PG_CUDA(cudaMemset(h->P.episode, 0, sizeof(uint32_t) * h->P.n));
  return PG_OK;
}

int pg_reset(pg_handle h, const uint8_t* mask_dev, const uint64_t* seeds_dev, void* stream) {
  if (!h) return fail(PG_ERR_ARG, "NULL handle");
  PG_CUDA(cudaSetDevice(h->device));
  pg::launch_reset_all(h->P, mask_dev, seeds_dev, h->first_env, (cudaStream_t)stream);
  if (!mask_dev) h->dlink_valid = true;
  h->launches++;
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_step(pg_handle h, const float* actions_dev, void* stream) {
  if (!h || !actions_dev) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  PG_CUDA(cudaMemsetAsync(h->P.reset_count, 0, sizeof(int), (cudaStream_t)stream));
  h->step_id++;
  h->P.term_w = (8 + h->P.obs_dim + 3) / 4 * 4;
  h->P.term_rows = h->term_ring + (size_t)(h->step_id % 3) * h->P.n * h->term_w_max;
  h->P.scene.skip_far = h->cull_enabled && h->dlink_valid && h->P.task == 1 && h->P.obstacle_obs != 0 && h->P.scene.nobs + h->P.scene.nstatic > 0;
  if (h->P.task >= 2) pg::launch_step_obj(h->P, actions_dev, (cudaStream_t)stream);
  else pg::launch_step(h->P, actions_dev, (cudaStream_t)stream);
  h->dlink_valid = true;  // the step ends with the observation pass of every environment
  h->launches++;
  if (h->eval_metrics) {  // perform_benchmark's bookkeeping, on the post-step state and before the auto-reset
    pg::launch_eval_metrics(h->P, (cudaStream_t)stream);
    h->launches++;
  }
  if (h->P.auto_reset) {
    pg::launch_reset_list(h->P, h->first_env, h->sm_count, (cudaStream_t)stream);
    h->launches++;
  }
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

static int copy_out(pg_handle h, const pg_host_out* o, cudaStream_t s) {
  const pg::Params& P = h->P;
  const size_t n = P.n;
  // host buffers laid out like the device slab (pg_host_layout): ONE copy instead of six
  uint8_t* base = (uint8_t*)o->achieved_goal;
  if (base && (uint8_t*)o->desired_goal == base + 12 * n && (uint8_t*)o->reward == base + 24 * n && o->done == base + 28 * n &&
      o->flags == base + 29 * n && (uint8_t*)o->observation == base + h->slab_obs_off) {
    PG_CUDA(cudaMemcpyAsync(base, h->out_slab, h->slab_obs_off + n * P.obs_dim * sizeof(float), cudaMemcpyDeviceToHost, s));
    PG_CUDA(cudaStreamSynchronize(s));
    return PG_OK;
  }
  if (o->observation) PG_CUDA(cudaMemcpyAsync(o->observation, P.obs, n * P.obs_dim * sizeof(float), cudaMemcpyDeviceToHost, s));
  if (o->achieved_goal) PG_CUDA(cudaMemcpyAsync(o->achieved_goal, P.achieved, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
  if (o->desired_goal) PG_CUDA(cudaMemcpyAsync(o->desired_goal, P.desired, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
  if (o->reward) PG_CUDA(cudaMemcpyAsync(o->reward, P.reward, n * sizeof(float), cudaMemcpyDeviceToHost, s));
  if (o->done) PG_CUDA(cudaMemcpyAsync(o->done, P.done, n, cudaMemcpyDeviceToHost, s));
  if (o->flags) PG_CUDA(cudaMemcpyAsync(o->flags, P.flags, n, cudaMemcpyDeviceToHost, s));
  PG_CUDA(cudaStreamSynchronize(s));
  return PG_OK;
}

int pg_host_layout(pg_handle h, int64_t* offsets6, int64_t* total_bytes) {
  if (!h || !offsets6 || !total_bytes) return fail(PG_ERR_ARG, "NULL argument");
  const int64_t n = h->P.n;
  offsets6[0] = (int64_t)h->slab_obs_off; offsets6[1] = 0; offsets6[2] = 12 * n; offsets6[3] = 24 * n; offsets6[4] = 28 * n; offsets6[5] = 29 * n;
  *total_bytes = (int64_t)h->slab_obs_off + n * h->P.obs_dim * (int64_t)sizeof(float);
  return PG_OK;
}

int64_t pg_step_id(pg_handle h) { return h ? h->step_id : 0; }

int pg_finished_count(pg_handle h, int64_t step_id) {
  if (!h || step_id > h->step_id || h->step_id - step_id >= 3 || step_id <= 0) return PG_ERR_STATE;
  return h->finished_host[step_id % 3];
}

int pg_fetch_finished(pg_handle h, int64_t step_id, float* rows_host, int max_rows, int* row_width, void* stream) {
  if (!h || !rows_host || !row_width) return fail(PG_ERR_ARG, "NULL argument");
  if (step_id > h->step_id || step_id <= 0 || h->step_id - step_id >= 3)
    return fail(PG_ERR_STATE, "the finished-episode rows of that step have been overwritten (the ring keeps the last three steps)");
  PG_CUDA(cudaSetDevice(h->device));
  int count = h->finished_host[step_id % 3];
  if (count > max_rows) count = max_rows;
  *row_width = h->P.term_w;
  if (count > 0) {
    const float* src = h->term_ring + (size_t)(step_id % 3) * h->P.n * h->term_w_max;
    PG_CUDA(cudaMemcpyAsync(rows_host, src, (size_t)count * h->P.term_w * sizeof(float), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    PG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  }
  return count;
}

int pg_copy_out(pg_handle h, const pg_host_out* out, void* stream) {
  if (!h || !out) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  return copy_out(h, out, (cudaStream_t)stream);
}

int pg_step_host(pg_handle h, const float* actions_host, const pg_host_out* out, void* stream) {
  if (!h || !actions_host) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  PG_CUDA(cudaMemcpyAsync(h->d_actions, actions_host, (size_t)h->P.n * h->P.act_dim * sizeof(float), cudaMemcpyHostToDevice, s));
  int rc = pg_step(h, h->d_actions, stream);
  if (rc != PG_OK) return rc;
  // how many episodes ended in this step: rides along with the results (page-locked word, valid after the synchronise)
  PG_CUDA(cudaMemcpyAsync(&h->finished_host[h->step_id % 3], h->P.reset_count, sizeof(int), cudaMemcpyDeviceToHost, s));
  if (out) return copy_out(h, out, s);
  PG_CUDA(cudaStreamSynchronize(s));
  return PG_OK;
}

int pg_reset_host(pg_handle h, const pg_host_out* out, void* stream) {
  int rc = pg_reset(h, nullptr, nullptr, stream);
  if (rc != PG_OK) return rc;
  if (out) return copy_out(h, out, (cudaStream_t)stream);
  PG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return PG_OK;
}

int pg_compute_reward(pg_handle h, const float* ag, const float* dg, float* out, int64_t n, int env_index, void* stream) {
  if (!h || !ag || !dg || !out) return fail(PG_ERR_ARG, "NULL argument");
  if (env_index < 0 || env_index >= h->P.n) return fail(PG_ERR_ARG, "env_index out of range");
  if (n <= 0) return PG_OK;
  PG_CUDA(cudaSetDevice(h->device));
  pg::launch_reward(h->P, ag, dg, out, n, env_index, (cudaStream_t)stream);
  h->launches++;
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_random_actions(pg_handle h, uint64_t seed, uint64_t counter, float* actions_dev, void* stream) {
  if (!h || !actions_dev) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  int64_t m = (int64_t)h->P.n * h->P.act_dim;
  pg::launch_random_actions(actions_dev, m, seed, counter, (cudaStream_t)stream);
  h->launches++;
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_save_state(pg_handle h, int* state_id, void* stream) {
  if (!h || !state_id) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  int id = 0;
  while (h->saved.count(id)) id++;  // first free non-negative integer, like Bullet's saveState
  std::vector<void*> copies;
  for (const Buffer& b : h->state_bufs) {
    void* p = nullptr;
    PG_CUDA(cudaMalloc(&p, b.bytes));
    copies.push_back(p);
    PG_CUDA(cudaMemcpyAsync(p, *b.slot, b.bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  }
  h->saved[id] = copies;
  *state_id = id;
  return PG_OK;
}
int pg_restore_state(pg_handle h, int state_id, void* stream) {
  if (!h) return fail(PG_ERR_ARG, "NULL handle");
  auto it = h->saved.find(state_id);
  if (it == h->saved.end()) return fail(PG_ERR_STATE, "Couldn't restore state.");  // pybullet.error text
  PG_CUDA(cudaSetDevice(h->device));
  h->dlink_valid = false;
  for (size_t i = 0; i < h->state_bufs.size(); i++)
    PG_CUDA(cudaMemcpyAsync(*h->state_bufs[i].slot, it->second[i], h->state_bufs[i].bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return PG_OK;
}
int pg_remove_state(pg_handle h, int state_id) {
  if (!h) return fail(PG_ERR_ARG, "NULL handle");
  auto it = h->saved.find(state_id);
  if (it == h->saved.end()) return fail(PG_ERR_STATE, "unknown state id");
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  for (void* p : it->second) cudaFree(p);
  h->saved.erase(it);
  return PG_OK;
}

int pg_get_view(pg_handle h, int id, pg_view* v) {
  if (!h || !v) return fail(PG_ERR_ARG, "NULL argument");
  const pg::Params& P = h->P;
  const int64_t n = P.n;
  memset(v, 0, sizeof *v);
  auto set = [&](void* p, int dt, int nd, int64_t a, int64_t b, int64_t c) { v->ptr = p; v->dtype = dt; v->ndim = nd; v->shape[0] = a; v->shape[1] = b; v->shape[2] = c; };
  switch (id) {
    case PG_V_OBSERVATION: set(P.obs, 0, 2, n, P.obs_dim, 0); break;
    case PG_V_ACHIEVED_GOAL: set(P.achieved, 0, 2, n, 3, 0); break;
    case PG_V_DESIRED_GOAL: set(P.desired, 0, 2, n, 3, 0); break;
    case PG_V_REWARD: set(P.reward, 0, 1, n, 0, 0); break;
    case PG_V_DONE: set(P.done, 1, 1, n, 0, 0); break;
    case PG_V_FLAGS: set(P.flags, 1, 1, n, 0, 0); break;
    case PG_V_TERMINAL_OBSERVATION: set(P.term_obs, 0, 2, n, P.obs_dim, 0); break;
    case PG_V_TERMINAL_ACHIEVED: set(P.term_achieved, 0, 2, n, 3, 0); break;
    case PG_V_EP_RETURN: set(P.last_ret, 0, 1, n, 0, 0); break;
    case PG_V_EP_LENGTH: set(P.last_len, 2, 1, n, 0, 0); break;
    case PG_V_Q: set(P.q, 0, 2, 7, n, 0); break;
    case PG_V_QD: set(P.qd, 0, 2, 7, n, 0); break;
    case PG_V_GOAL: set(P.goal, 3, 2, 3, n, 0); break;
    case PG_V_OBSTACLE_POS: set(P.obs_pos, 3, 3, pg::kMaxObst, 3, n); break;
    case PG_V_STATS: set(P.stats, 3, 1, 8, 0, 0); break;
    case PG_V_STEP_COUNT: set(P.steps, 2, 1, n, 0, 0); break;
    case PG_V_D_LINK: set(P.d_link, 0, 2, 9, n, 0); break;
    case PG_V_OBJECT: set(P.obj, 0, 2, 13, n, 0); break;
    case PG_V_OBSTACLE_SIZE: set(P.env_size, 3, 3, pg::kMaxObst, 3, n); break;
    case PG_V_JOINT_ACC: set(P.cur_acc, 0, 2, 7, n, 0); break;
    case PG_V_JOINT_JERK: set(P.jerk, 0, 2, 7, n, 0); break;
    default: return fail(PG_ERR_ARG, "unknown view id");
  }
  return PG_OK;
}
int pg_obs_dim(pg_handle h) { return h ? h->P.obs_dim : PG_ERR_ARG; }
int pg_act_dim(pg_handle h) { return h ? h->P.act_dim : PG_ERR_ARG; }
int pg_num_envs(pg_handle h) { return h ? h->P.n : PG_ERR_ARG; }
int64_t pg_launch_count(pg_handle h) { return h ? h->launches : 0; }

int pg_set_joint_state(pg_handle h, const float* q_dev, const float* qd_dev, void* stream) {
  if (!h || !q_dev) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  size_t bytes = (size_t)7 * h->P.n * sizeof(float);
  PG_CUDA(cudaMemcpyAsync(h->P.q, q_dev, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  if (qd_dev) PG_CUDA(cudaMemcpyAsync(h->P.qd, qd_dev, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  else PG_CUDA(cudaMemsetAsync(h->P.qd, 0, bytes, (cudaStream_t)stream));
  return PG_OK;
}
int pg_set_goal(pg_handle h, const double* goal_dev, void* stream) {
  if (!h || !goal_dev) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  PG_CUDA(cudaMemcpyAsync(h->P.goal, goal_dev, (size_t)3 * h->P.n * sizeof(double), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return PG_OK;
}
int pg_set_object_state(pg_handle h, const float* obj_dev, void* stream) {
  if (!h || !obj_dev) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  PG_CUDA(cudaMemcpyAsync(h->P.obj, obj_dev, (size_t)13 * h->P.n * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return PG_OK;
}
int pg_set_obstacles(pg_handle h, const double* pos_dev, void* stream) {
  if (!h || !pos_dev) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  PG_CUDA(cudaMemcpyAsync(h->P.obs_pos, pos_dev, (size_t)h->P.st.nobs * 3 * h->P.n * sizeof(double), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  pg::launch_mirror_obstacles(h->P, (cudaStream_t)stream);
  h->dlink_valid = false;  // until the next observation pass (pg_observe / pg_step / full pg_reset)
  return PG_OK;
}
int pg_observe(pg_handle h, void* stream) {
  if (!h) return fail(PG_ERR_ARG, "NULL handle");
  PG_CUDA(cudaSetDevice(h->device));
  pg::launch_observe(h->P, (cudaStream_t)stream);
  h->dlink_valid = true;
  h->launches++;
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_set_scene(pg_handle h, int n_boxes, const float* boxes_host, void* stream) {
  if (!h || n_boxes < 0 || n_boxes > pg::kMaxStaticBoxes || (n_boxes > 0 && !boxes_host)) return fail(PG_ERR_ARG, "pg_set_scene: 0..64 boxes");
  if (h->P.task != PG_TASK_REACH_AO) return fail(PG_ERR_UNSUPPORTED, "pg_set_scene: benchmark scenes belong to ReachAO");
  PG_CUDA(cudaSetDevice(h->device));
  std::vector<float> tab((size_t)pg::kMaxStaticBoxes * 8, 0.f);
  for (int j = 0; j < n_boxes; j++) {
    const float* b = boxes_host + 7 * j;
    if (!(b[3] > 0.f && b[4] > 0.f && b[5] > 0.f) || !(b[6] >= 0.f) || b[6] >= fminf(b[3], fminf(b[4], b[5])))
      return fail(PG_ERR_ARG, "pg_set_scene: half extents must be positive and larger than the margin");
    float* t = &tab[(size_t)8 * j];
    t[0] = b[0]; t[1] = b[1]; t[2] = b[2]; t[3] = b[6];
    t[4] = b[3]; t[5] = b[4]; t[6] = b[5]; t[7] = sqrtf(b[3] * b[3] + b[4] * b[4] + b[5] * b[5]);
  }
  if (!h->sbox) { PG_CUDA(cudaMalloc((void**)&h->sbox, tab.size() * sizeof(float))); h->all_bufs.push_back(h->sbox); }
  PG_CUDA(cudaMemcpyAsync(h->sbox, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice, (cudaStream_t)stream));
  PG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));  // tab is a local
  h->P.scene.nstatic = n_boxes;
  h->P.scene.sbox = (const float4*)h->sbox;
  if (n_boxes > 0) h->P.scene.has_box = 1;
  h->dlink_valid = false;
  return PG_OK;
}

// computeViewMatrixFromYawPitchRoll(upAxisIndex = 2) + computeProjectionMatrixFOV(60, w / h, 0.1, 100) as Bullet builds them:
// eye = target + Rz(yaw) Ry(roll) Rx(pitch) (0, -distance, 0), up = the same rotation of (0, 0, 1)
static void build_camera(pg::Camera& cam, int width, int height, const float* target, float distance, float yaw_deg, float pitch_deg, float roll_deg) {
  const double D2R = 0.017453292519943295;
  const double y = yaw_deg * D2R, p = pitch_deg * D2R, r = roll_deg * D2R;
  const double cy = cos(y), sy = sin(y), cp = cos(p), sp = sin(p), cr = cos(r), sr = sin(r);
  // R = Rz(y) Ry(r) Rx(p)
  const double R[3][3] = {{cy * cr, cy * sr * sp - sy * cp, cy * sr * cp + sy * sp},
                          {sy * cr, sy * sr * sp + cy * cp, sy * sr * cp - cy * sp},
                          {-sr, cr * sp, cr * cp}};
  double eye[3], up[3], f[3], s[3], u[3];
  for (int k = 0; k < 3; k++) { eye[k] = target[k] + R[k][1] * (-(double)distance); up[k] = R[k][2]; }
  double fl = 0, sl = 0;
  for (int k = 0; k < 3; k++) { f[k] = target[k] - eye[k]; fl += f[k] * f[k]; }
  fl = sqrt(fl);
  for (int k = 0; k < 3; k++) f[k] /= fl;
  s[0] = f[1] * up[2] - f[2] * up[1]; s[1] = f[2] * up[0] - f[0] * up[2]; s[2] = f[0] * up[1] - f[1] * up[0];
  for (int k = 0; k < 3; k++) sl += s[k] * s[k];
  sl = sqrt(sl);
  for (int k = 0; k < 3; k++) s[k] /= sl;
  u[0] = s[1] * f[2] - s[2] * f[1]; u[1] = s[2] * f[0] - s[0] * f[2]; u[2] = s[0] * f[1] - s[1] * f[0];
  for (int k = 0; k < 3; k++) { cam.eye[k] = (float)eye[k]; cam.fwd[k] = (float)f[k]; cam.right[k] = (float)s[k]; cam.up[k] = (float)u[k]; }
  cam.tan_half_fov = (float)tan(0.5 * 60.0 * D2R); cam.aspect = (float)width / (float)height;
  cam.near_plane = 0.1f; cam.far_plane = 100.0f; cam.width = width; cam.height = height;
}

int pg_render(pg_handle h, int env_index, int width, int height, const float* target_position, float distance, float yaw, float pitch,
              float roll, uint8_t* rgba_dev, void* stream) {
  if (!h || !rgba_dev) return fail(PG_ERR_ARG, "NULL argument");
  if (env_index < 0 || env_index >= h->P.n) return fail(PG_ERR_ARG, "env_index out of range");
  if (width <= 0 || height <= 0 || width > 8192 || height > 8192) return fail(PG_ERR_ARG, "bad image size");
  if (!(distance > 0.f)) return fail(PG_ERR_ARG, "distance must be positive");
  PG_CUDA(cudaSetDevice(h->device));
  const float zero[3] = {0.f, 0.f, 0.f};
  pg::Camera cam;
  build_camera(cam, width, height, target_position ? target_position : zero, distance, yaw, pitch, roll);
  pg::launch_render(h->P, cam, env_index, rgba_dev, (cudaStream_t)stream);
  h->launches++;
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_set_eval_metrics(pg_handle h, int enable, void* stream) {
  if (!h) return fail(PG_ERR_ARG, "NULL handle");
  PG_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  const size_t n = h->P.n;
  PG_CUDA(cudaMemsetAsync(h->P.ev, 0, 12 * sizeof(double), s));
  PG_CUDA(cudaMemsetAsync(h->P.ev_acc, 0, 4 * n * sizeof(float), s));
  PG_CUDA(cudaMemsetAsync(h->P.ev_len, 0, n * sizeof(int), s));
  PG_CUDA(cudaMemsetAsync(h->P.ev_ret, 0, n * sizeof(float), s));
  h->eval_metrics = enable ? 1 : 0;
  return PG_OK;
}
int pg_get_eval_metrics(pg_handle h, double* out12_host, void* stream) {
  if (!h || !out12_host) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  PG_CUDA(cudaMemcpyAsync(out12_host, h->P.ev, 12 * sizeof(double), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  PG_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return PG_OK;
}

int pg_manipulability(pg_handle h, const float* q_dev, int64_t n, float* out_dev, void* stream) {
  if (!h || !out_dev) return fail(PG_ERR_ARG, "NULL handle or output");
  if (q_dev && n <= 0) return fail(PG_ERR_ARG, "n must be positive when q_dev is given");
  PG_CUDA(cudaSetDevice(h->device));
  const int m = q_dev ? (int)n : h->P.n;
  pg::launch_manipulability(m, q_dev ? q_dev : h->P.q, out_dev, (cudaStream_t)stream);
  h->launches++;
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_debug_substeps(int device, int n, float* q, float* qd, const float* tgt, const float* f, int n_sub, int velocity_mode, void* stream) {
  if (!q || !qd || !tgt || !f || n <= 0) return fail(PG_ERR_ARG, "bad argument");
  static const float kForces[7] = {87.f, 87.f, 87.f, 87.f, 12.f, 120.f, 120.f};  // panda.py:63, compiled into the kernels
  for (int d = 0; d < 7; d++) if (f[d] != kForces[d]) return fail(PG_ERR_UNSUPPORTED, "max_force must equal Panda.joint_forces (panda.py:63)");
  PG_CUDA(cudaSetDevice(device));
  pg::launch_debug_substeps(n, q, qd, tgt, f, n_sub, velocity_mode, (cudaStream_t)stream);
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}
int pg_debug_minv(int device, int n, const float* q, float* out, void* stream) {
  if (!q || !out || n <= 0) return fail(PG_ERR_ARG, "bad argument");
  PG_CUDA(cudaSetDevice(device));
  pg::launch_debug_minv(n, q, out, (cudaStream_t)stream);
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}
int pg_debug_ik(int device, int n, const float* q, const float* tgt, const float* base, float* out, void* stream) {
  if (!q || !tgt || !out || !base || n <= 0) return fail(PG_ERR_ARG, "bad argument");
  PG_CUDA(cudaSetDevice(device));
  pg::launch_debug_ik(n, q, tgt, base[0], base[1], base[2], out, (cudaStream_t)stream);
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_debug_step_kernel_attrs(int device, int* out /* [6]: regs, maxThreadsPerBlock, static smem, local bytes, block, blocks/SM */) {
  if (!out) return fail(PG_ERR_ARG, "bad argument");
  PG_CUDA(cudaSetDevice(device));
  PG_CUDA(pg::step_kernel_attrs(out));
  return PG_OK;
}

int pg_debug_fma_peak(int device, int blocks, int iters, double* tflops_out) {
  if (!tflops_out || blocks <= 0 || iters <= 0) return fail(PG_ERR_ARG, "bad argument");
  PG_CUDA(cudaSetDevice(device));
  float* d = nullptr;
  PG_CUDA(cudaMalloc(&d, 4));
  cudaEvent_t e0, e1;
  PG_CUDA(cudaEventCreate(&e0)); PG_CUDA(cudaEventCreate(&e1));
  pg::launch_fma_peak(blocks, d, iters, 0);  // warm-up
  float best = 1e30f;
  for (int r = 0; r < 5; r++) {
    PG_CUDA(cudaEventRecord(e0));
    pg::launch_fma_peak(blocks, d, iters, 0);
    PG_CUDA(cudaEventRecord(e1));
    PG_CUDA(cudaEventSynchronize(e1));
    float ms = 0; PG_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  *tflops_out = (double)blocks * 256.0 * (double)iters * 16.0 / (best * 1e-3) / 1e12;
  return PG_OK;
}

}  // extern "C"

// =============================================================================================== HER buffer
struct pg_her_s {
  int device;
  pg::HerParams H;
  std::vector<void*> bufs;
  int64_t capacity;      // rows * n_envs
  int64_t max_batch;
  pg::HerBatch out;      // library-owned output batch (max_batch rows)
  int64_t launches;
  uint64_t fast_counter;
  int sm_count;
  double her_ratio;      // SB3 her_ratio in double: nb_virtual = int(her_ratio * batch_size) must round like Python does
  bool valid_stale;      // the compacted valid-index list must be rebuilt (something was added since the last sample)
};

extern "C" {

int pg_her_create(int device, int64_t buffer_size, int n_envs, int obs_dim, int act_dim, int n_sampled_goal, int64_t max_batch, pg_her_handle* out) {
  if (!out || buffer_size <= 0 || n_envs <= 0 || obs_dim <= 0 || act_dim <= 0 || n_sampled_goal < 0 || max_batch <= 0) return fail(PG_ERR_ARG, "bad argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(PG_ERR_CUDA, "no CUDA device: this library has no CPU fallback");
  PG_CUDA(cudaSetDevice(device));
  int her_sms = 148;
  PG_CUDA(cudaDeviceGetAttribute(&her_sms, cudaDevAttrMultiProcessorCount, device));
  if ((pg::her_used(obs_dim, act_dim) + 15) / 16 * 16 > 256) return fail(PG_ERR_ARG, "record wider than 256 floats (obs_dim too large)");
  pg_her_s* h = new pg_her_s();
  h->sm_count = her_sms;
  h->device = device; h->launches = 0; h->fast_counter = 0; h->max_batch = max_batch; h->valid_stale = true;
  pg::HerParams& H = h->H;
  memset(&H, 0, sizeof H);
  H.rows = (int)(buffer_size / n_envs > 0 ? buffer_size / n_envs : 1);  // SB3: buffer_size // n_envs
  H.n_envs = n_envs; H.obs_dim = obs_dim; H.act_dim = act_dim;
  h->her_ratio = 1.0 - (1.0 / (double)(n_sampled_goal + 1));  // SB3: 1 - (1.0 / (n_sampled_goal + 1)), a Python float
  h->capacity = (int64_t)H.rows * n_envs;
  const size_t cap = (size_t)h->capacity;
  auto A = [&](void** p, size_t bytes) -> int {
    PG_CUDA(cudaMalloc(p, bytes)); PG_CUDA(cudaMemset(*p, 0, bytes)); h->bufs.push_back(*p); return PG_OK; };
#define HA(ptr, count, type) { int rc = A((void**)&ptr, (size_t)(count) * sizeof(type)); if (rc != PG_OK) { pg_her_destroy(h); return rc; } }
  H.rec_w = (pg::her_used(obs_dim, act_dim) + 15) / 16 * 16;  // whole 64 B lines per record
  HA(H.rec, cap * H.rec_w, float) HA(H.ep_start, cap, int) HA(H.ep_length, cap, int)
  HA(H.cur_ep_start, n_envs, int) HA(H.cursor, 2, int) HA(H.valid, cap, int) HA(H.n_valid, 1, int)
  HA(H.block_counts, (cap + pg::kHerBlock - 1) / pg::kHerBlock + 1, int) HA(H.mt, 625, uint32_t)
  HA(H.s_row, max_batch, int) HA(H.s_env, max_batch, int) HA(H.s_t, max_batch, int) HA(H.status, 1, int)
  HA(H.s_trow, max_batch, int)
  HA(h->out.packed, (size_t)max_batch * H.rec_w, float)
#undef HA
  // default stream state = np.random.seed(0)-like is NOT assumed: callers load a state with pg_her_set_mt_state
  *out = h;
  return PG_OK;
}

int pg_her_destroy(pg_her_handle h) {
  if (!h) return PG_OK;
  cudaSetDevice(h->device);
  for (void* p : h->bufs) cudaFree(p);
  delete h;
  return PG_OK;
}

int pg_her_set_mt_state(pg_her_handle h, const uint32_t* key624_host, int pos) {
  if (!h || !key624_host || pos < 0 || pos > 624) return fail(PG_ERR_ARG, "bad argument");
  PG_CUDA(cudaSetDevice(h->device));
  uint32_t st[625];
  memcpy(st, key624_host, 624 * sizeof(uint32_t)); st[624] = (uint32_t)pos;
  PG_CUDA(cudaMemcpy(h->H.mt, st, sizeof st, cudaMemcpyHostToDevice));
  return PG_OK;
}
int pg_her_get_mt_state(pg_her_handle h, uint32_t* key624_host, int* pos) {
  if (!h || !key624_host || !pos) return fail(PG_ERR_ARG, "bad argument");
  PG_CUDA(cudaSetDevice(h->device));
  uint32_t st[625];
  PG_CUDA(cudaMemcpy(st, h->H.mt, sizeof st, cudaMemcpyDeviceToHost));
  memcpy(key624_host, st, 624 * sizeof(uint32_t)); *pos = (int)st[624];
  return PG_OK;
}

int pg_her_add(pg_her_handle h, const float* obs, const float* ag, const float* dg, const float* next_obs, const float* next_ag,
               const float* next_dg, const float* action, const float* reward, const uint8_t* done, const uint8_t* timeout, void* stream) {
  if (!h || !obs || !ag || !dg || !next_obs || !next_ag || !next_dg || !action || !reward || !done) return fail(PG_ERR_ARG, "NULL argument");
  PG_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  const pg::HerParams& H = h->H;
  pg::launch_her_add(H, obs, ag, dg, next_obs, next_ag, next_dg, action, reward, done, timeout, s);
  h->valid_stale = true;
  h->launches += 3;
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_her_sample(pg_her_handle h, pg_handle env, int64_t batch, int exact_rng, uint64_t fast_seed, int reward_env, pg_her_batch* out, void* stream) {
  if (!h || !env || !out || batch <= 0 || batch > h->max_batch) return fail(PG_ERR_ARG, "bad argument (batch must be in [1, max_batch])");
  if (reward_env < 0 || reward_env >= env->P.n) return fail(PG_ERR_ARG, "reward_env out of range");
  PG_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  const pg::HerParams& H = h->H;
  const int total = (int)h->capacity;
  const int B = (int)batch, nbv = (int)(h->her_ratio * (double)B);  // int(her_ratio * batch_size), double like SB3
  if (h->valid_stale) {  // valid = flatnonzero(ep_length > 0): unchanged until the next add(), so several gradient steps per
                         // environment step (SB3 gradient_steps > 1) pay the 1 M-entry compaction once
    pg::launch_her_compact(H, total, s);
    h->launches += 3;
    h->valid_stale = false;
  }
  pg::launch_her_indices(H, B, nbv, exact_rng != 0, fast_seed, h->fast_counter++, s);
  pg::launch_her_gather(H, env->P, h->out, B, nbv, reward_env, h->sm_count, s);
  h->launches += 2;
  PG_CUDA(cudaGetLastError());
  // the batch is ONE [batch][row_stride] block in the record's field order; the nine fields are column ranges of it
  float* pk = h->out.packed;
  const int od = H.obs_dim;
  out->observation = pk; out->next_observation = pk + od;
  out->achieved_goal = pk + 2 * od; out->desired_goal = pk + 2 * od + 3;
  out->next_achieved_goal = pk + 2 * od + 6; out->next_desired_goal = pk + 2 * od + 9;
  out->action = pk + 2 * od + 12; out->reward = pk + 2 * od + 12 + H.act_dim; out->done = out->reward + 1;
  out->batch = batch; out->n_virtual = nbv; out->row_stride = H.rec_w;
  return PG_OK;
}

// ------------------------------------------------------------------------------------------ sample() without allocation
// Scratch of pg_her_sample_static: the index lists the index kernel hands to the gather kernel, the per-block counts of the
// valid-index compaction and the status word.  File-scope device memory: every GPU that loads the library holds these 1.25 MiB
// until exit, ALL calls (of all buffers on that GPU) share it, and it keeps the last call's contents.  In host code the name
// `g_her_static` is only the host-side stand-in of the symbol: its device address comes from cudaGetSymbolAddress, once per
// device with that device current (static_scratch below).
constexpr int kStaticMaxBatch = PG_HER_STATIC_MAX_BATCH;            // rows per call
constexpr int kStaticMaxBlocks = PG_HER_STATIC_MAX_TRANSITIONS / pg::kHerBlock + 1;  // compaction blocks (buffer capacity / 256)
constexpr int kStaticStatusAt = 4 * kStaticMaxBatch;               // [0] status, rest of the 256 B line unused
constexpr int kStaticCountsAt = kStaticStatusAt + 64;
constexpr int kStaticInts = kStaticCountsAt + (kStaticMaxBlocks + 63) / 64 * 64;
static __device__ __align__(256) int g_her_static[kStaticInts];

static std::mutex g_static_mu;
static int* g_static_ptr[64] = {nullptr};
// device address of g_her_static on `device` (which the caller has made current); resolved once per device
static int static_scratch(int device, int** out) {
  if (device < 0 || device >= 64) return fail(PG_ERR_ARG, "device index out of range");
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    int* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_her_static);
    if (e != cudaSuccess || !p) { cudaGetLastError(); return fail(PG_ERR_CUDA, std::string("cudaGetSymbolAddress(g_her_static): ") + cudaGetErrorString(e)); }
    g_static_ptr[device] = p;
  }
  *out = g_static_ptr[device];
  return PG_OK;
}

int pg_her_sample_static(pg_her_handle h, pg_handle env, int64_t batch, int exact_rng, uint64_t fast_seed, int reward_env,
                         float* packed_out_dev, int* status_out_dev, void* stream) {
  if (!h || !env || !packed_out_dev || batch <= 0) return fail(PG_ERR_ARG, "bad argument (NULL handle / output, or batch <= 0)");
  if (batch > kStaticMaxBatch)
    return fail(PG_ERR_ARG, "pg_her_sample_static: batch exceeds PG_HER_STATIC_MAX_BATCH (" + std::to_string(kStaticMaxBatch) + " rows): use pg_her_sample");
  if (h->capacity > (int64_t)PG_HER_STATIC_MAX_TRANSITIONS)
    return fail(PG_ERR_ARG, "pg_her_sample_static: buffer capacity exceeds PG_HER_STATIC_MAX_TRANSITIONS (" +
                std::to_string(PG_HER_STATIC_MAX_TRANSITIONS) + "): use pg_her_sample");
  if (reward_env < 0 || reward_env >= env->P.n) return fail(PG_ERR_ARG, "reward_env out of range");
  PG_CUDA(cudaSetDevice(h->device));
  int* scratch = nullptr;
  int rc = static_scratch(h->device, &scratch);
  if (rc != PG_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  pg::HerParams H = h->H;  // the buffer's storage and valid list, with every per-call scratch pointer moved into g_her_static
  H.s_row = scratch; H.s_env = scratch + kStaticMaxBatch; H.s_t = scratch + 2 * kStaticMaxBatch; H.s_trow = scratch + 3 * kStaticMaxBatch;
  H.status = scratch + kStaticStatusAt;
  H.block_counts = scratch + kStaticCountsAt;
  const int total = (int)h->capacity;
  const int B = (int)batch, nbv = (int)(h->her_ratio * (double)B);
  // nothing here has to start at zero for the kernels (counts, index lists and status are fully written before they are read),
  // but the status word is cleared anyway so that a call never reports the previous call's outcome
  PG_CUDA(cudaMemsetAsync(H.status, 0, sizeof(int), s));
  if (h->valid_stale) {  // same cache as pg_her_sample: the valid list belongs to the buffer, only the block counts are scratch
    pg::launch_her_compact(H, total, s);
    h->launches += 3;
    h->valid_stale = false;
  }
  pg::launch_her_indices(H, B, nbv, exact_rng != 0, fast_seed, h->fast_counter++, s);
  pg::HerBatch O; O.packed = packed_out_dev;
  pg::launch_her_gather(H, env->P, O, B, nbv, reward_env, h->sm_count, s);
  h->launches += 2;
  if (status_out_dev) PG_CUDA(cudaMemcpyAsync(status_out_dev, H.status, sizeof(int), cudaMemcpyDeviceToDevice, s));
  PG_CUDA(cudaGetLastError());
  return PG_OK;
}

int pg_her_last_indices(pg_her_handle h, int64_t batch, int* row_host, int* env_host, int* t_host, int* n_valid_host, int* status_host) {
  if (!h || batch <= 0 || batch > h->max_batch) return fail(PG_ERR_ARG, "bad argument");
  PG_CUDA(cudaSetDevice(h->device));
  PG_CUDA(cudaDeviceSynchronize());
  if (row_host) PG_CUDA(cudaMemcpy(row_host, h->H.s_row, batch * sizeof(int), cudaMemcpyDeviceToHost));
  if (env_host) PG_CUDA(cudaMemcpy(env_host, h->H.s_env, batch * sizeof(int), cudaMemcpyDeviceToHost));
  if (t_host) PG_CUDA(cudaMemcpy(t_host, h->H.s_t, batch * sizeof(int), cudaMemcpyDeviceToHost));
  if (n_valid_host) PG_CUDA(cudaMemcpy(n_valid_host, h->H.n_valid, sizeof(int), cudaMemcpyDeviceToHost));
  if (status_host) PG_CUDA(cudaMemcpy(status_host, h->H.status, sizeof(int), cudaMemcpyDeviceToHost));
  return PG_OK;
}

int pg_her_info(pg_her_handle h, int* rows, int* pos, int* full, int64_t* launches) {
  if (!h) return fail(PG_ERR_ARG, "NULL handle");
  PG_CUDA(cudaSetDevice(h->device));
  int cur[2];
  PG_CUDA(cudaMemcpy(cur, h->H.cursor, sizeof cur, cudaMemcpyDeviceToHost));
  if (rows) *rows = h->H.rows;
  if (pos) *pos = cur[0];
  if (full) *full = cur[1];
  if (launches) *launches = h->launches;
  return PG_OK;
}

}  // extern "C"
