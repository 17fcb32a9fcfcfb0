"""``PandaBatch``: thin host object over one ``pg_handle`` (one GPU, N environments).

This is the B200 replacement of the reference's L0-L2 stack (``panda_gym/pybullet.py::PyBullet``,
``envs/robots/panda.py::Panda``, ``envs/tasks/reach*.py``) for the env-step hot path.  PyTorch is used only to
view library-owned device buffers and for streams; all arithmetic happens in the CUDA kernels.
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib
from .config import CONTROL, OBSTACLE_OBS, REWARD_AO, REWARD_CLASSIC, TrainConfig

_TORCH_DTYPES = {0: torch.float32, 1: torch.uint8, 2: torch.int32, 3: torch.float64, 4: torch.int64}
_TYPESTR = {0: "<f4", 1: "|u1", 2: "<i4", 3: "<f8", 4: "<i8"}


class _CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch can view a library-owned buffer without copying."""

    def __init__(self, ptr, shape, dtype_code, owner, strides=None):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": _TYPESTR[dtype_code], "data": (ptr, False),
                                         "version": 2, "strides": None if strides is None else tuple(strides)}
        self._owner = owner


def make_config(task, n_envs, scenario=None, config=None, control_type=None, reward_type=None,
                max_episode_steps=50, ee_error_threshold=0.05, speed_threshold=0.1, auto_reset=True, base_position=None,
                pose_retry_limit=0):
    """Build a pg_config from reference-style arguments (TrainConfig + gym kwargs).  pose_retry_limit: tries of
    set_coll_free_robot per obstacle, 0 = the reference's 10 000 (pg_config.pose_retry_limit)."""
    lib = _lib.load()
    cfg = _lib.Config()
    _lib.check(lib.pg_default_config(ctypes.byref(cfg), task, scenario.encode() if scenario else None))
    cfg.n_envs = int(n_envs)
    cfg.auto_reset = 1 if auto_reset else 0
    cfg._scenario = scenario  # remembered for PandaBatch: a benchmark scene needs its boxes (pg_set_scene)
    if task == 1:
        c = config if config is not None else TrainConfig()
        cfg.control = CONTROL[control_type or c.control_type]
        cfg.reward_type = REWARD_AO[reward_type or c.reward_type]
        cfg.goal_condition = 1 if c.goal_condition == "halt" else 0
        cfg.action_limiter = 1 if c.action_limiter == "scale" else 0
        cfg.obs_js = 1 if "js" in c.obs_type else 0
        if "ee" not in c.obs_type:
            raise _lib.PandaB200Error("obs_type without 'ee' is not supported")
        cfg.obstacle_obs = OBSTACLE_OBS[c.task_observations["obstacles"]]
        if c.task_observations.get("prior") is not None:
            raise _lib.PandaB200Error("task_observations['prior'] (NEO prior, needs Gurobi) is out of scope")
        cfg.randomize_robot_pose = int(bool(c.randomize_robot_pose))
        if c.fixed_target is not None:
            cfg.has_fixed_target = 1
            for k in range(3):
                cfg.fixed_target[k] = float(c.fixed_target[k])
        cfg.n_substeps = int(c.n_substeps)
        cfg.truncate_on_collision = int(bool(c.truncate_on_collision))
        cfg.terminate_on_success = int(bool(c.terminate_on_success))
        cfg.collision_reward = float(c.collision_reward)
        cfg.safety_distance = float(c.safety_distance)
    else:
        # PandaReach* / PandaPush* / PandaPickAndPlace* ids pass control_type "ee" or "joints" (panda_gym/__init__.py:23-77)
        cfg.control = CONTROL[control_type or ("js" if task == 0 else "ee")]
        cfg.reward_type = REWARD_CLASSIC[reward_type or "sparse"]
    if base_position is not None:
        for k in range(3):
            cfg.base_position[k] = float(base_position[k])
    cfg.stage.max_episode_steps = int(max_episode_steps)
    cfg.stage.ee_error_threshold = float(ee_error_threshold)
    cfg.stage.speed_threshold = float(speed_threshold)
    cfg.pose_retry_limit = int(pose_retry_limit)
    return cfg


def stage_from_scenario(scenario, max_episode_steps=50, ee_error_threshold=0.05, speed_threshold=0.1):
    lib = _lib.load()
    st = _lib.Stage()
    _lib.check(lib.pg_stage_from_scenario(ctypes.byref(st), scenario.encode()))
    st.max_episode_steps = int(max_episode_steps)
    st.ee_error_threshold = float(ee_error_threshold)
    st.speed_threshold = float(speed_threshold)
    st._scenario = scenario
    return st


class PandaBatch:
    def __init__(self, cfg, device=0):
        if not torch.cuda.is_available():
            raise _lib.PandaB200Error("no CUDA device visible: panda_gym_b200 runs on the GPU only (no CPU fallback)")
        self.lib = _lib.load()
        self.cfg = cfg
        self.device_index = int(device)
        self.device = torch.device("cuda", self.device_index)
        h = ctypes.c_void_p()
        _lib.check(self.lib.pg_create(ctypes.byref(cfg), self.device_index, ctypes.byref(h)))
        self.h = h
        self._scene_boxes = None
        if cfg.stage.needs_scene:
            from . import scenes
            self.set_scene(scenes.scene_boxes(getattr(cfg, "_scenario", None), os.environ.get("PANDA_GYM_SCENARIOS")))
        self.n = self.lib.pg_num_envs(h)
        self.act_dim = self.lib.pg_act_dim(h)
        self._views = {}

    # ------------------------------------------------------------------ lifetime
    def close(self):
        if getattr(self, "h", None):
            self.lib.pg_destroy(self.h)
            self.h = None
            self._views = {}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def obs_dim(self):
        return self.lib.pg_obs_dim(self.h)

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def view(self, name):
        """Zero-copy torch view of a library-owned device buffer (include/panda_b200.h pg_view_id)."""
        key = (name, self.obs_dim)
        if key not in self._views:
            v = _lib.View()
            _lib.check(self.lib.pg_get_view(self.h, _lib.VIEW_IDS[name], ctypes.byref(v)))
            shape = [int(v.shape[i]) for i in range(v.ndim)]
            with torch.cuda.device(self.device):
                t = torch.as_tensor(_CudaArray(v.ptr, shape, v.dtype, self), device=self.device)
            self._views[key] = t
        return self._views[key]

    # ------------------------------------------------------------------ env API (device buffers)
    def seed(self, seed, first_env_index=0):
        _lib.check(self.lib.pg_seed(self.h, int(seed), int(first_env_index)))

    def set_scene(self, boxes):
        """Static boxes shared by all environments (pg_set_scene): float [n, 7] = centre, half extents, margin, e.g.
        ``scenes.scene_boxes("kitchen")``; None or an empty array removes the scene.  Takes effect with the next reset."""
        b = np.ascontiguousarray(boxes if boxes is not None else np.zeros((0, 7)), dtype=np.float32).reshape(-1, 7)
        _lib.check(self.lib.pg_set_scene(self.h, int(b.shape[0]), b.ctypes.data_as(ctypes.c_void_p) if len(b) else None, self._stream()))
        self._scene_boxes = b if len(b) else None

    def set_stage(self, stage):
        """Swap the scenario block; every environment is reset into the new scene (pg_set_stage).  A benchmark-scene stage
        (``stage_from_scenario("kitchen")``) brings its boxes along; any other stage removes them."""
        if stage.needs_scene:
            from . import scenes
            self.set_scene(scenes.scene_boxes(getattr(stage, "_scenario", None), os.environ.get("PANDA_GYM_SCENARIOS")))
        elif self._scene_boxes is not None:
            self.set_scene(None)
        _lib.check(self.lib.pg_set_stage(self.h, ctypes.byref(stage), self._stream()))
        self._views = {}

    def reset(self, mask=None, seeds=None):
        m = ctypes.c_void_p(mask.data_ptr()) if mask is not None else None
        s = ctypes.c_void_p(seeds.data_ptr()) if seeds is not None else None
        if mask is not None:
            assert mask.dtype == torch.uint8 and mask.is_cuda and mask.numel() == self.n
        if seeds is not None:
            assert seeds.dtype in (torch.int64, torch.uint64) and seeds.is_cuda and seeds.numel() == self.n
        _lib.check(self.lib.pg_reset(self.h, m, s, self._stream()))

    def step(self, actions):
        assert actions.is_cuda and actions.dtype == torch.float32 and actions.is_contiguous()
        assert actions.numel() == self.n * self.act_dim
        _lib.check(self.lib.pg_step(self.h, ctypes.c_void_p(actions.data_ptr()), self._stream()))

    def step_host(self, actions_np, out):
        _lib.check(self.lib.pg_step_host(self.h, actions_np.ctypes.data_as(ctypes.c_void_p), ctypes.byref(out), self._stream()))

    def reset_host(self, out):
        _lib.check(self.lib.pg_reset_host(self.h, ctypes.byref(out), self._stream()))

    def copy_out(self, out):
        _lib.check(self.lib.pg_copy_out(self.h, ctypes.byref(out), self._stream()))

    def step_id(self):
        return int(self.lib.pg_step_id(self.h))

    def finished_count(self, step_id):
        return int(self.lib.pg_finished_count(self.h, int(step_id)))

    def fetch_finished(self, step_id):
        """Compact rows of the episodes that ended in step ``step_id`` (one device-to-host copy): returns
        ``(env_index[int32 k], episode_return[k], episode_length[int32 k], flags[int32 k], achieved[k, 3], observation[k, obs_dim])``
        sorted by environment index."""
        if getattr(self, "_fin_stage", None) is None or self._fin_stage.shape[1] < 8 + self.obs_dim:
            w = (8 + 6 + 14 + 81 + 3) // 4 * 4
            self._fin_stage = torch.empty((self.n, w), dtype=torch.float32).pin_memory()
            self._fin_np = self._fin_stage.numpy().reshape(-1)
        width = ctypes.c_int()
        k = self.lib.pg_fetch_finished(self.h, int(step_id), ctypes.c_void_p(self._fin_stage.data_ptr()), self.n, ctypes.byref(width),
                                       self._stream())
        if k < 0:
            _lib.check(k)
        w = width.value
        rows = self._fin_np[:k * w].reshape(k, w)
        order = np.argsort(rows[:, 0].view(np.int32), kind="stable")
        rows = rows[order]  # fancy indexing copies out of the staging buffer
        iv = rows[:, :4].view(np.int32)
        return iv[:, 0].copy(), rows[:, 1].copy(), iv[:, 2].copy(), iv[:, 3].copy(), rows[:, 4:7].copy(), rows[:, 8:8 + self.obs_dim].copy()

    def host_layout(self):
        """Byte offsets (observation, achieved_goal, desired_goal, reward, done, flags) and total size of the one-block
        host layout that pg_step_host fills with a single device-to-host copy."""
        off = (ctypes.c_int64 * 6)()
        tot = ctypes.c_int64()
        _lib.check(self.lib.pg_host_layout(self.h, off, ctypes.byref(tot)))
        return [int(x) for x in off], int(tot.value)

    def random_actions(self, seed, counter, out=None):
        if out is None:
            out = torch.empty((self.n, self.act_dim), dtype=torch.float32, device=self.device)
        _lib.check(self.lib.pg_random_actions(self.h, int(seed), int(counter), ctypes.c_void_p(out.data_ptr()), self._stream()))
        return out

    def compute_reward(self, achieved_goal, desired_goal, env_index=0):
        ag = achieved_goal.reshape(-1, 3).contiguous()
        dg = desired_goal.reshape(-1, 3).contiguous()
        out = torch.empty(ag.shape[0], dtype=torch.float32, device=self.device)
        _lib.check(self.lib.pg_compute_reward(self.h, ctypes.c_void_p(ag.data_ptr()), ctypes.c_void_p(dg.data_ptr()),
                                              ctypes.c_void_p(out.data_ptr()), ag.shape[0], int(env_index), self._stream()))
        return out

    def observe(self):
        _lib.check(self.lib.pg_observe(self.h, self._stream()))

    def set_eval_metrics(self, enable=True):
        """Start (and zero) / stop the on-device accumulation of perform_benchmark's metrics (evaluate.py:131-300)."""
        _lib.check(self.lib.pg_set_eval_metrics(self.h, 1 if enable else 0, self._stream()))

    def eval_metrics_raw(self):
        out = np.zeros(12, np.float64)
        _lib.check(self.lib.pg_get_eval_metrics(self.h, out.ctypes.data_as(ctypes.c_void_p), self._stream()))
        return out

    def manipulability(self, q=None):
        """Panda.get_manipulability (panda.py:314-317) for every environment (q None) or for q f32 [7][m] on the device."""
        if q is None:
            out = torch.empty(self.n, dtype=torch.float32, device=self.device)
            _lib.check(self.lib.pg_manipulability(self.h, None, 0, ctypes.c_void_p(out.data_ptr()), self._stream()))
        else:
            assert q.is_cuda and q.dtype == torch.float32 and q.dim() == 2 and q.shape[0] == 7 and q.is_contiguous()
            out = torch.empty(q.shape[1], dtype=torch.float32, device=q.device)
            _lib.check(self.lib.pg_manipulability(self.h, ctypes.c_void_p(q.data_ptr()), q.shape[1], ctypes.c_void_p(out.data_ptr()), self._stream()))
        return out

    def render(self, env_index=0, width=720, height=480, target_position=None, distance=1.4, yaw=45.0, pitch=-30.0, roll=0.0):
        """PyBullet.render(mode="rgb_array") (pybullet.py:117-180) for one environment: uint8 [height][width][4] on the
        device, ray-cast by k_render (no graphics stack involved)."""
        out = torch.empty((int(height), int(width), 4), dtype=torch.uint8, device=self.device)
        tp = (ctypes.c_float * 3)(*(target_position if target_position is not None else (0.0, 0.0, 0.0)))
        _lib.check(self.lib.pg_render(self.h, int(env_index), int(width), int(height), tp, float(distance), float(yaw), float(pitch),
                                      float(roll), ctypes.c_void_p(out.data_ptr()), self._stream()))
        return out

    def save_state(self):
        sid = ctypes.c_int(-1)
        _lib.check(self.lib.pg_save_state(self.h, ctypes.byref(sid), self._stream()))
        return sid.value

    def restore_state(self, sid):
        _lib.check(self.lib.pg_restore_state(self.h, int(sid), self._stream()))

    def remove_state(self, sid):
        _lib.check(self.lib.pg_remove_state(self.h, int(sid)))

    def launch_count(self):
        return int(self.lib.pg_launch_count(self.h))

    # ------------------------------------------------------------------ matched-state setters (parity tests)
    def set_joint_state(self, q, qd=None):
        q = q.to(self.device, torch.float32).contiguous()
        qd_p = None
        if qd is not None:
            qd = qd.to(self.device, torch.float32).contiguous()
            qd_p = ctypes.c_void_p(qd.data_ptr())
        _lib.check(self.lib.pg_set_joint_state(self.h, ctypes.c_void_p(q.data_ptr()), qd_p, self._stream()))

    def set_goal(self, goal):
        g = goal.to(self.device, torch.float64).contiguous()
        _lib.check(self.lib.pg_set_goal(self.h, ctypes.c_void_p(g.data_ptr()), self._stream()))

    def set_object_state(self, obj):
        """[13][n]: position, quaternion xyzw, linear velocity, angular velocity of the Push / PickAndPlace object."""
        o = obj.to(self.device, torch.float32).contiguous()
        assert o.numel() == 13 * self.n
        _lib.check(self.lib.pg_set_object_state(self.h, ctypes.c_void_p(o.data_ptr()), self._stream()))

    def set_obstacles(self, pos):
        p = pos.to(self.device, torch.float64).contiguous()
        _lib.check(self.lib.pg_set_obstacles(self.h, ctypes.c_void_p(p.data_ptr()), self._stream()))


def debug_substeps(q, qd, target, max_force, n_sub, velocity_mode=False):
    """n_sub raw stepSimulation() calls on [7][n] SoA tensors (in place)."""
    lib = _lib.load()
    f = np.ascontiguousarray(max_force, dtype=np.float32)
    _lib.check(lib.pg_debug_substeps(q.device.index or 0, q.shape[1], ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(qd.data_ptr()),
                                     ctypes.c_void_p(target.data_ptr()), f.ctypes.data_as(ctypes.c_void_p), int(n_sub),
                                     int(velocity_mode), ctypes.c_void_p(torch.cuda.current_stream(q.device).cuda_stream)))


def debug_minv(q):
    lib = _lib.load()
    out = torch.empty((28, q.shape[1]), dtype=torch.float32, device=q.device)
    _lib.check(lib.pg_debug_minv(q.device.index or 0, q.shape[1], ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(out.data_ptr()),
                                 ctypes.c_void_p(torch.cuda.current_stream(q.device).cuda_stream)))
    return out


def debug_ik(q, target, base=(0.0, 0.0, 0.0)):
    lib = _lib.load()
    out = torch.empty_like(q)
    b = np.asarray(base, dtype=np.float32)
    _lib.check(lib.pg_debug_ik(q.device.index or 0, q.shape[1], ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(target.data_ptr()),
                               b.ctypes.data_as(ctypes.c_void_p), ctypes.c_void_p(out.data_ptr()),
                               ctypes.c_void_p(torch.cuda.current_stream(q.device).cuda_stream)))
    return out
