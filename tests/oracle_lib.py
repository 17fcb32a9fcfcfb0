"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "_build", "liboracle.so")

c_int, c_double, c_void_p, c_u64, c_float = ctypes.c_int, ctypes.c_double, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_float


class EnvConfig(ctypes.Structure):
    _fields_ = [
        ("task", c_int), ("control", c_int), ("reward_type", c_int), ("goal_condition", c_int),
        ("action_limiter", c_int), ("obs_js", c_int), ("obstacle_obs", c_int), ("n_substeps", c_int),
        ("truncate_on_collision", c_int), ("terminate_on_success", c_int),
        ("collision_reward", c_double), ("ee_error_threshold", c_double), ("speed_threshold", c_double),
        ("safety_distance", c_double), ("max_episode_steps", c_int), ("scenario", c_int), ("n_obstacles", c_int),
        ("base_pos", c_double * 3), ("randomize_robot_pose", c_int),
        ("scene_goal_lo", c_double * 3), ("scene_goal_hi", c_double * 3), ("custom_neutral", c_int), ("neutral_q", c_double * 7),
        ("has_fixed_target", c_int), ("fixed_target", c_double * 3),
        ("scene_pose_randomizer", c_int), ("scene_force_randomize", c_int), ("scene_pose_lo", c_double * 3), ("scene_pose_hi", c_double * 3),
        ("pose_retry_limit", c_int), ("no_shortcuts", c_int),
    ]


class Obstacle(ctypes.Structure):
    _fields_ = [("type", c_int), ("pos", c_double * 3), ("size", c_double * 3)]


class Body(ctypes.Structure):
    _fields_ = [("pos", c_double * 3), ("quat", c_double * 4), ("lin", c_double * 3), ("ang", c_double * 3)]


class World(ctypes.Structure):
    _fields_ = [("half", c_double * 3), ("mass", c_double), ("table_pos", c_double * 3), ("table_half", c_double * 3),
                ("plane_pos", c_double * 3), ("plane_half", c_double * 3),
                ("mu_object", c_double), ("mu_static", c_double), ("mu_link", c_double), ("mu_ee", c_double)]


class Contact(ctypes.Structure):
    _fields_ = [("kind", c_int), ("link", c_int), ("d", c_double), ("n", c_double * 3), ("pa", c_double * 3),
                ("pb", c_double * 3), ("mu", c_double)]


class Pcg64(ctypes.Structure):
    _fields_ = [("s_hi", c_u64), ("s_lo", c_u64), ("inc_hi", c_u64), ("inc_lo", c_u64)]


SCENARIOS = {"none": 0, "reach1": 1, "reach2": 2, "reach3": 3, "reachao1": 4, "reachao2": 5, "reachao3": 6,
             "reachao_rand": 7, "wang": 8, "wangexp": 9, "static": 10, "reachao_rand_start": 11, "reachao_rand_shape": 12}


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    L = ctypes.CDLL(LIB_PATH)
    P = c_void_p

    def sig(name, res, *args):
        f = getattr(L, name)
        f.restype = res
        f.argtypes = list(args)

    sig("pgo_arm_create", P, c_int, P)
    sig("pgo_arm_destroy", None, P)
    sig("pgo_arm_ndof", c_int, P)
    sig("pgo_arm_nlinks", c_int, P)
    sig("pgo_arm_set_damping", None, P, c_double, c_double)
    sig("pgo_arm_set_solver", None, P, c_int, c_double)
    sig("pgo_arm_reset_joint", None, P, c_int, c_double)
    sig("pgo_arm_get_q", None, P, P, P)
    sig("pgo_arm_set_q", None, P, P, P)
    sig("pgo_arm_set_motor", None, P, c_int, c_int, c_double, c_double)
    sig("pgo_arm_substep", c_int, P)
    sig("pgo_arm_link_state", None, P, c_int, P, P, P, P)
    sig("pgo_arm_link_frame", None, P, c_int, P, P)
    sig("pgo_arm_minv", None, P, P)
    sig("pgo_arm_ik", None, P, c_int, P, P, P)
    sig("pgo_arm_set_ik", None, P, c_double, c_int, c_double)
    sig("pgo_arm_jacobian", None, P, c_int, P, P)
    sig("pgo_arm_manipulability", c_double, P)
    sig("pgo_box_margin", c_double, P)
    sig("pgo_prim_obstacle_distance", c_double, P, c_int, P, P, P)
    sig("pgo_link_obstacle_distance", c_double, P, c_int, P, P, P)
    sig("pgo_gjk_cyl_box", c_double, P, P, c_double, c_double, P, P, P, P)
    sig("pgo_pcg64_seed", None, P, c_u64)
    sig("pgo_pcg64_next", c_u64, P)
    sig("pgo_pcg64_double", c_double, P)
    sig("pgo_pcg64_uniform", c_double, P, c_double, c_double)
    sig("pgo_pcg64_advance", None, P, c_u64)
    sig("pgo_pcg64_standard_exponential", c_double, P)
    sig("pgo_pcg64_dirichlet111", None, P, P)
    sig("pgo_arm_ik_lm_chan", c_int, P, P, P, c_int, c_int, c_double, c_double, P, P, P)
    sig("pgo_env_default_config", None, P, c_int)
    sig("pgo_env_create", P, P)
    sig("pgo_env_destroy", None, P)
    sig("pgo_env_obs_dim", c_int, P)
    sig("pgo_env_act_dim", c_int, P)
    sig("pgo_env_reset", c_int, P, c_u64, P, P, P, P)
    sig("pgo_env_step", None, P, P, P, P, P, P, P)
    sig("pgo_env_arm", P, P)
    sig("pgo_env_set_goal", None, P, P)
    sig("pgo_env_get_goal", None, P, P)
    sig("pgo_env_set_obstacles", None, P, c_int, P)
    sig("pgo_env_get_obstacles", c_int, P, P)
    sig("pgo_env_check_collided", c_int, P, P, P)
    sig("pgo_env_step_metrics", None, P, P)
    sig("pgo_env_get_obs", None, P, P, P, P)
    sig("pgo_env_compute_reward", None, P, P, P, c_int, P)
    sig("pgo_env_is_collided", c_int, P)
    sig("pgo_render", None, P, c_int, c_int, P, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_double, P)
    sig("pgo_bench_rollout", c_int, P, c_u64, c_int, P)
    sig("pgo_env_object", ctypes.POINTER(Body), P)
    sig("pgo_env_world", ctypes.POINTER(World), P)
    sig("pgo_world_default", None, P)
    sig("pgo_world_contacts", c_int, P, P, P, P)
    sig("pgo_world_substep", c_int, P, P, P)
    sig("pgo_quat_to_euler", None, P, P)
    sig("pgo_test_distance", None, P, P, c_int, P)
    sig("pgo_test_integers_shuffle", c_int, c_u64, c_int, c_int, c_int, c_int, P)
    _lib = L
    return L


def dp(a):
    return a.ctypes.data_as(c_void_p) if a is not None else None


class Arm:
    """pybullet-body-like handle on the oracle's multibody (model 0 = custom_0, 1 = stock)."""

    def __init__(self, model=0, base=(0.0, 0.0, 0.0), handle=None):
        self.L = lib()
        self._own = handle is None
        self.h = handle if handle is not None else self.L.pgo_arm_create(model, dp(np.asarray(base, dtype=np.float64)))
        self.nd = self.L.pgo_arm_ndof(self.h)

    def __del__(self):
        if getattr(self, "_own", False) and self.h:
            self.L.pgo_arm_destroy(self.h)
            self.h = None

    def q(self):
        q = np.zeros(self.nd); qd = np.zeros(self.nd)
        self.L.pgo_arm_get_q(self.h, dp(q), dp(qd))
        return q, qd

    def set_q(self, q=None, qd=None):
        q = None if q is None else np.ascontiguousarray(q, dtype=np.float64)
        qd = None if qd is None else np.ascontiguousarray(qd, dtype=np.float64)
        self.L.pgo_arm_set_q(self.h, dp(q), dp(qd))

    def reset_joint(self, d, v):
        self.L.pgo_arm_reset_joint(self.h, d, float(v))

    def set_motor(self, d, mode, target, force):
        self.L.pgo_arm_set_motor(self.h, d, mode, float(target), float(force))

    def substep(self):
        return self.L.pgo_arm_substep(self.h)

    def link_state(self, link):
        p = np.zeros(3); q = np.zeros(4); v = np.zeros(3); w = np.zeros(3)
        self.L.pgo_arm_link_state(self.h, link, dp(p), dp(q), dp(v), dp(w))
        return p, q, v, w

    def link_frame(self, link):
        R = np.zeros(9); p = np.zeros(3)
        self.L.pgo_arm_link_frame(self.h, link, dp(R), dp(p))
        return R.reshape(3, 3), p

    def minv(self):
        M = np.zeros((self.nd, self.nd))
        self.L.pgo_arm_minv(self.h, dp(M))
        return M

    def ik(self, link, pos, quat=None):
        out = np.zeros(self.nd)
        pos = np.asarray(pos, dtype=np.float64)
        quat = None if quat is None else np.asarray(quat, dtype=np.float64)
        self.L.pgo_arm_ik(self.h, link, dp(pos), dp(quat), dp(out))
        return out

    def jacobian(self, link):
        Jt = np.zeros((3, self.nd)); Jr = np.zeros((3, self.nd))
        self.L.pgo_arm_jacobian(self.h, link, dp(Jt), dp(Jr))
        return Jt, Jr

    def manipulability(self):
        return float(self.L.pgo_arm_manipulability(self.h))

    def link_obstacle_distance(self, link, obstacle):
        pa = np.zeros(3); pb = np.zeros(3)
        d = self.L.pgo_link_obstacle_distance(self.h, link, ctypes.byref(obstacle), dp(pa), dp(pb))
        return d, pa, pb


def make_obstacle(kind, pos, size):
    o = Obstacle()
    o.type = {"sphere": 0, "urdf_box": 2}.get(kind, 1)
    for k in range(3):
        o.pos[k] = float(pos[k])
    if kind == "sphere":
        o.size[0] = float(size); o.size[1] = 0.0; o.size[2] = 0.0
    else:
        for k in range(3):
            o.size[k] = float(size[k])
    return o


class Env:
    """Single reference-shaped env (RobotTaskEnv semantics) on the oracle."""

    def __init__(self, task=1, **kw):
        self.L = lib()
        self.cfg = EnvConfig()
        self.L.pgo_env_default_config(ctypes.byref(self.cfg), task)
        for k, v in kw.items():
            if k == "scenario" and isinstance(v, str):
                name, _, n = v.partition("-")
                self.cfg.scenario = SCENARIOS[name]
                if n:
                    self.cfg.n_obstacles = int(n)
            elif k in ("base_pos", "scene_goal_lo", "scene_goal_hi", "neutral_q", "fixed_target", "scene_pose_lo", "scene_pose_hi"):
                for i in range(len(v)):
                    getattr(self.cfg, k)[i] = float(v[i])
            else:
                setattr(self.cfg, k, v)
        self.h = self.L.pgo_env_create(ctypes.byref(self.cfg))
        self.obs_dim = self.L.pgo_env_obs_dim(self.h)
        self.act_dim = self.L.pgo_env_act_dim(self.h)
        self.arm = Arm(handle=self.L.pgo_env_arm(self.h))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.pgo_env_destroy(self.h)
            self.h = None

    def reset(self, seed):
        obs = np.zeros(self.obs_dim, np.float32); ag = np.zeros(3, np.float32); dg = np.zeros(3, np.float32)
        succ = c_int(0)
        status = self.L.pgo_env_reset(self.h, int(seed), dp(obs), dp(ag), dp(dg), ctypes.byref(succ))
        return dict(observation=obs, achieved_goal=ag, desired_goal=dg), dict(is_success=bool(succ.value), status=status)

    def step(self, action):
        a = np.ascontiguousarray(action, dtype=np.float32)
        obs = np.zeros(self.obs_dim, np.float32); ag = np.zeros(3, np.float32); dg = np.zeros(3, np.float32)
        r = c_float(0); flags = (c_int * 4)()
        self.L.pgo_env_step(self.h, dp(a), dp(obs), dp(ag), dp(dg), ctypes.byref(r), flags)
        return (dict(observation=obs, achieved_goal=ag, desired_goal=dg), float(r.value),
                dict(is_success=bool(flags[0]), terminated=bool(flags[1]), truncated=bool(flags[2]), time_limit=bool(flags[3])))

    def get_obs(self):
        obs = np.zeros(self.obs_dim, np.float32); ag = np.zeros(3, np.float32); dg = np.zeros(3, np.float32)
        self.L.pgo_env_get_obs(self.h, dp(obs), dp(ag), dp(dg))
        return dict(observation=obs, achieved_goal=ag, desired_goal=dg)

    def goal(self):
        g = np.zeros(3)
        self.L.pgo_env_get_goal(self.h, dp(g))
        return g

    def set_goal(self, g):
        self.L.pgo_env_set_goal(self.h, dp(np.asarray(g, dtype=np.float64)))

    def obstacles(self):
        arr = (Obstacle * 64)()
        n = self.L.pgo_env_get_obstacles(self.h, arr)
        return [(arr[i].type, np.array(arr[i].pos[:]), np.array(arr[i].size[:])) for i in range(n)]

    def set_obstacles(self, obs_list):
        arr = (Obstacle * max(1, len(obs_list)))(*obs_list)
        self.L.pgo_env_set_obstacles(self.h, len(obs_list), arr)
        self.obs_dim = self.L.pgo_env_obs_dim(self.h)  # reach_ao.py:904: the obstacle part is np.ones(27) only while there is no obstacle

    def step_metrics(self):
        out = np.zeros(4)
        self.L.pgo_env_step_metrics(self.h, dp(out))
        return out

    def check_collided(self):
        d = np.zeros(9); t = c_double(0)
        c = self.L.pgo_env_check_collided(self.h, dp(d), ctypes.byref(t))
        return bool(c), d, t.value

    # ---- Push / PickAndPlace (tasks 2, 3): the free object
    def object_state(self):
        b = self.L.pgo_env_object(self.h).contents
        return np.array(b.pos[:]), np.array(b.quat[:]), np.array(b.lin[:]), np.array(b.ang[:])

    def set_object_state(self, pos, quat=(0, 0, 0, 1), lin=(0, 0, 0), ang=(0, 0, 0)):
        b = self.L.pgo_env_object(self.h).contents
        for k in range(3):
            b.pos[k] = float(pos[k]); b.lin[k] = float(lin[k]); b.ang[k] = float(ang[k])
        for k in range(4):
            b.quat[k] = float(quat[k])

    def world_substep(self):
        return self.L.pgo_world_substep(self.L.pgo_env_arm(self.h), self.L.pgo_env_object(self.h), self.L.pgo_env_world(self.h))

    def contacts(self):
        arr = (Contact * 12)()
        n = self.L.pgo_world_contacts(self.L.pgo_env_arm(self.h), self.L.pgo_env_object(self.h), self.L.pgo_env_world(self.h), arr)
        return [dict(kind=c.kind, link=c.link, d=c.d, n=np.array(c.n[:]), pa=np.array(c.pa[:]), pb=np.array(c.pb[:]), mu=c.mu)
                for c in arr[:n]]

    def render(self, width=720, height=480, target_position=(0.0, 0.0, 0.0), distance=1.4, yaw=45.0, pitch=-30.0, roll=0.0):
        out = np.zeros((height, width, 4), np.uint8)
        tp = np.asarray(target_position, np.float64)
        self.L.pgo_render(self.h, width, height, dp(tp), distance, yaw, pitch, roll, dp(out))
        return out

    def compute_reward(self, ag, dg):
        ag = np.ascontiguousarray(ag, dtype=np.float32).reshape(-1, 3); dg = np.ascontiguousarray(dg, dtype=np.float32).reshape(-1, 3)
        out = np.zeros(len(ag), np.float32)
        self.L.pgo_env_compute_reward(self.h, dp(ag), dp(dg), len(ag), dp(out))
        return out


def cpu_rollout(steps_per_env, n_threads, scenario="reachao3", max_episode_steps=100, seed=0):
    """Time the oracle as a CPU baseline: one environment per host thread (ctypes releases the GIL)."""
    import threading
    import time
    L = lib()
    cfg = EnvConfig()
    L.pgo_env_default_config(ctypes.byref(cfg), 1)
    name, _, n = scenario.partition("-")
    cfg.scenario = SCENARIOS[name]
    if n:
        cfg.n_obstacles = int(n)
    cfg.max_episode_steps = max_episode_steps
    res = [0] * n_threads

    def work(i):
        res[i] = L.pgo_bench_rollout(ctypes.byref(cfg), seed + 1000003 * i, steps_per_env, None)

    th = [threading.Thread(target=work, args=(i,)) for i in range(n_threads)]
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    return steps_per_env * n_threads / dt, dt, sum(res)
