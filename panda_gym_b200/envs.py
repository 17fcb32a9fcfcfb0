"""Single-environment, gymnasium-shaped front (N = 1 on the GPU) mirroring the reference's env classes.

Reference: ``panda_gym/envs/core.py::RobotTaskEnv`` (:255-414), ``panda_gym/envs/panda_tasks.py`` (PandaReachEnv
:72-88, PandaReachAOEnv :133-159) and the registration helpers of ``panda_gym/__init__.py`` (:15-91).
``make(id, **kwargs)`` plays the role of ``gymnasium.make`` including its ``TimeLimit(max_episode_steps)``
wrapper.  Semantics follow gymnasium: ``reset(seed) -> (obs, info)``, ``step(a) -> (obs, reward, terminated,
truncated, info)``; no auto-reset.
"""
import numpy as np
import torch

from . import _lib, sim, spaces
from .config import TrainConfig

_MAX_EPISODE_STEPS = {"PandaReachAO-v3": 50}
_REGISTERED = {}


def register_reach_ao(max_ep_steps):
    """panda_gym.register_reach_ao (reference __init__.py:15-20)."""
    _REGISTERED["PandaReachAO-v3"] = dict(cls="ao", max_episode_steps=max_ep_steps, kwargs={})


def register_envs(max_ep_steps):
    """panda_gym.register_envs (reference __init__.py:23-91) for the env ids on the supported hot path."""
    for reward_type in ("sparse", "dense"):
        for control_type in ("ee", "joints"):
            rs = "Dense" if reward_type == "dense" else ""
            cs = "Joints" if control_type == "joints" else ""
            kw = dict(reward_type=reward_type, control_type=control_type)
            _REGISTERED["PandaReach%s%s-v3" % (cs, rs)] = dict(cls="reach", max_episode_steps=max_ep_steps, kwargs=dict(kw))
            _REGISTERED["PandaPush%s%s-v3" % (cs, rs)] = dict(cls="push", max_episode_steps=max_ep_steps, kwargs=dict(kw))
            _REGISTERED["PandaPickAndPlace%s%s-v3" % (cs, rs)] = dict(cls="pnp", max_episode_steps=max_ep_steps, kwargs=dict(kw))
    register_reach_ao(max_ep_steps)


def make(env_id, **kwargs):
    if env_id not in _REGISTERED:
        raise _lib.PandaB200Error("env id %r is not registered (call register_envs / register_reach_ao first, as with the reference)" % env_id)
    spec = _REGISTERED[env_id]
    kw = dict(spec["kwargs"])
    kw.update(kwargs)
    cls = {"ao": PandaReachAOEnv, "reach": PandaReachEnv, "push": PandaPushEnv, "pnp": PandaPickAndPlaceEnv}[spec["cls"]]
    return cls(max_episode_steps=spec["max_episode_steps"], **kw)


class _Namespace:
    pass


class _Robot:
    """The part of ``Panda`` (envs/robots/panda.py) that callers above the step path read: evaluation/evaluate.py:246-259,
    the success / collision callbacks.  Getters read the device state of environment 0 of the (one-environment) batch."""

    ee_link = 11                                                  # panda.py:68
    joint_forces = np.array([87.0, 87.0, 87.0, 87.0, 12.0, 120.0, 120.0, 170.0, 170.0])  # panda.py:63
    joint_lim_min = np.array([-2.7437, -1.7837, -2.9007, -3.0421, -2.8065, 0.5445, -3.0159])  # panda.py:77
    joint_lim_max = np.array([2.7437, 1.7837, 2.9007, -0.1518, 2.8065, 4.5169, 3.0159])       # panda.py:78

    def __init__(self, env):
        self._env = env
        nq = env.batch.cfg.stage.neutral_q
        self.neutral_joint_values = (np.array(list(nq) + [0.0, 0.0]) if env.batch.cfg.stage.custom_neutral else
                                     np.array([0.0, -0.3, 0.0, -2.2, 0.0, 2.0, np.pi / 4, 0.0, 0.0]))  # panda.py:67

    def _v(self, name):
        return self._env.batch.view(name)

    def get_ee_position(self):                                    # panda.py:300-302
        return self._v("observation")[0, 0:3].cpu().numpy().astype(np.float64)

    def get_ee_velocity(self):                                    # panda.py:304-306
        return self._v("observation")[0, 3:6].cpu().numpy().astype(np.float64)

    def get_joint_angles(self, joints=None):                      # core.py:146-155
        q = self._v("q")[:, 0].cpu().numpy().astype(np.float64)
        return q if joints is None else q[np.asarray(joints)]

    def get_joint_angle(self, joint):
        return float(self.get_joint_angles()[joint])

    def get_joint_velocities(self, joints=None):
        qd = self._v("qd")[:, 0].cpu().numpy().astype(np.float64)
        return qd if joints is None else qd[np.asarray(joints)]

    def get_joint_velocity(self, joint):
        return float(self.get_joint_velocities()[joint])

    def get_fingers_width(self):                                  # panda.py:309-312: the loaded URDF's finger joints are fixed
        return 0.0

    def get_manipulability(self):                                 # panda.py:314-317
        return float(self._env.batch.manipulability()[0])

    @property
    def current_joint_acceleration(self):                         # panda.py:167-172
        return self._v("joint_acc")[:, 0].cpu().numpy().astype(np.float64)

    @property
    def current_joint_jerk(self):
        return self._v("joint_jerk")[:, 0].cpu().numpy().astype(np.float64)

    def get_obs(self):                                            # panda.py:264-288
        b = self._env.batch
        w = 6 + (14 if b.cfg.obs_js else 0) + (1 if b.cfg.task == 3 else 0)
        return self._v("observation")[0, :w].cpu().numpy()


class _Task:
    """The part of ``ReachAO`` / ``Reach`` / ``Push`` (envs/tasks/*.py) that callers above the step path read."""

    def __init__(self, env):
        self._env = env
        self.scenario = getattr(env.batch.cfg, "_scenario", None)
        st = env.batch.cfg.stage
        # reach_ao.py:819-860: obstacles are named "<shape>_<index in creation order>"
        self.obstacles = {("cuboid_%d" if st.obstacle_type[j] else "sphere_%d") % j: j for j in range(st.n_obstacles)}

    def _v(self, name):
        return self._env.batch.view(name)

    @property
    def goal(self):
        return self._env.get_goal()

    def get_goal(self):
        return self._env.get_goal()

    def get_achieved_goal(self):
        return self._v("achieved_goal")[0].cpu().numpy().astype(np.float64)

    def compute_reward(self, achieved_goal, desired_goal, info=None):
        return self._env.compute_reward(achieved_goal, desired_goal, info)

    def is_success(self, achieved_goal, desired_goal):          # reach.py:80-82, reach_ao.py:1243-1261 ("reach" condition)
        a, b = np.asarray(achieved_goal), np.asarray(desired_goal)
        assert a.shape == b.shape
        d = np.round(np.linalg.norm(a - b, axis=-1), 6)          # utils.distance (panda_gym/utils.py:4-16), in the inputs' dtype
        return np.array(d < np.float32(self._env.batch.cfg.stage.ee_error_threshold), dtype=bool)

    @property
    def is_collided(self):                                        # reach_ao.py:109, 186
        return bool(int(self._v("flags")[0]) & 4)

    def is_truncated(self):                                       # reach_ao.py:1263-1264
        return self.is_collided if self._env.batch.cfg.truncate_on_collision else ()

    @property
    def distances_closest_obstacles(self):                        # reach_ao.py:897
        return self._v("d_link")[:, 0].cpu().numpy().astype(np.float64)

    @property
    def manipulability(self):                                     # reach_ao.py:1311, refreshed by every compute_reward
        return float(self._env.batch.manipulability()[0])

    def get_norm_effort(self):                                    # reach_ao.py:1299-1302
        return float(np.linalg.norm(self._env.robot.current_joint_acceleration))

    def get_norm_jerk(self):                                      # reach_ao.py:1304-1305
        return float(np.linalg.norm(self._env.robot.current_joint_jerk))

    def get_obs(self):                                            # reach_ao.py:902-959, push.py:49-63
        w = len(self._env.robot.get_obs())
        return self._v("observation")[0, w:].cpu().numpy()


class RobotTaskEnv:
    """Shared implementation; subclasses only build the pg_config."""

    metadata = {"render_modes": ["human", "rgb_array"]}

    def _setup(self, cfg, device, terminate_on_success, max_episode_steps):
        self.batch = sim.PandaBatch(cfg, device=device)
        self.terminate_on_success = terminate_on_success
        self.max_episode_steps = max_episode_steps
        self._goal_set = False
        self._saved_goal = {}
        obs, _ = self.reset()
        self.observation_space = spaces.Dict(dict(
            observation=spaces.Box(-10.0, 10.0, shape=obs["observation"].shape, dtype=np.float32),
            desired_goal=spaces.Box(-10.0, 10.0, shape=obs["achieved_goal"].shape, dtype=np.float32),
            achieved_goal=spaces.Box(-10.0, 10.0, shape=obs["achieved_goal"].shape, dtype=np.float32)))
        self.action_space = spaces.Box(-1.0, 1.0, shape=(self.batch.act_dim,), dtype=np.float32)
        # attribute surface callers poke at (evaluation/evaluate.py:131-268 uses env.task.* / env.robot.*)
        self.robot = _Robot(self)
        self.task = _Task(self)
        self.sim = _Namespace()
        self.sim.timestep = 1.0 / 500                           # pybullet.py:50
        self.sim.n_substeps = self.batch.cfg.n_substeps
        self.sim.dt = self.sim.timestep * self.sim.n_substeps   # PyBullet.dt (pybullet.py:63-66)
        self.task.sim = self.sim

    def _obs(self):
        v = self.batch.view
        return {"observation": v("observation")[0].cpu().numpy(), "achieved_goal": v("achieved_goal")[0].cpu().numpy(),
                "desired_goal": v("desired_goal")[0].cpu().numpy()}

    def get_goal(self):
        if not self._goal_set:
            raise RuntimeError("No goal yet, call reset() first")  # core.py:236-237
        return self.batch.view("goal")[:, 0].cpu().numpy().copy()

    def reset(self, seed=None, options=None):
        if seed is None:  # core.py:302: a fresh generator from OS entropy on every unseeded reset
            seed = int(np.random.SeedSequence().generate_state(1, dtype=np.uint64)[0] >> 1)
        seeds = torch.tensor([int(seed)], dtype=torch.int64, device=self.batch.device)
        self.batch.reset(seeds=seeds)
        self._goal_set = True
        flags = int(self.batch.view("flags")[0])
        return self._obs(), {"is_success": bool(flags & 1)}

    def step(self, action):
        a = torch.as_tensor(np.asarray(action, dtype=np.float32).reshape(1, -1), device=self.batch.device)
        self.batch.step(a)
        flags = int(self.batch.view("flags")[0])
        reward = float(self.batch.view("reward")[0])
        is_success = bool(flags & 1)
        terminated = bool(flags & 2)
        truncated = bool(flags & 4) or bool(flags & 8)  # task truncation (collision) or TimeLimit
        info = {"is_success": is_success, "is_truncated": bool(flags & 4)}  # core.py:365; gymnasium's TimeLimit adds no key
        return self._obs(), reward, terminated, truncated, info

    def compute_reward(self, achieved_goal, desired_goal, info=None):
        ag = np.asarray(achieved_goal, dtype=np.float32)
        dg = np.asarray(desired_goal, dtype=np.float32)
        assert ag.shape == dg.shape
        dev = self.batch.device
        out = self.batch.compute_reward(torch.from_numpy(np.ascontiguousarray(ag)).to(dev), torch.from_numpy(np.ascontiguousarray(dg)).to(dev), 0)
        out = out.cpu().numpy().reshape(ag.shape[:-1])
        return out if out.shape else np.float32(out)

    def save_state(self):
        return self.batch.save_state()

    def restore_state(self, state_id):
        self.batch.restore_state(state_id)

    def remove_state(self, state_id):
        self.batch.remove_state(state_id)

    def render(self, mode="rgb_array", width=720, height=480, target_position=None, distance=1.4, yaw=45, pitch=-30, roll=0):
        """core.py:370-389 -> PyBullet.render.  "rgb_array": the scene ray-cast on the GPU (pg_render), uint8 [height, width, 4]
        like getCameraImage's pixels; "human" (a window paced in real time) has nothing to show on a headless machine."""
        if mode != "rgb_array":
            return None
        img = self.batch.render(0, width, height, target_position, distance, yaw, pitch, roll)
        return img.cpu().numpy()

    def close(self):
        self.batch.close()


class PandaReachAOEnv(RobotTaskEnv):
    """reference panda_tasks.py:133-159"""

    def __init__(self, render=False, ee_error_threshold=0.05, speed_threshold=0.1, scenario="reachao1", config=None,
                 max_episode_steps=50, device=0):
        config = config if config is not None else TrainConfig()
        cfg = sim.make_config(1, 1, scenario=scenario, config=config, max_episode_steps=max_episode_steps,
                              ee_error_threshold=ee_error_threshold, speed_threshold=speed_threshold, auto_reset=False)
        self.config = config
        self._setup(cfg, device, config.terminate_on_success, max_episode_steps)


class PandaReachEnv(RobotTaskEnv):
    """reference panda_tasks.py:72-88"""

    def __init__(self, render=False, realtime=False, reward_type="sparse", control_type="js", goal_range=0.3, show_goal_space=False,
                 max_episode_steps=50, device=0):
        cfg = sim.make_config(0, 1, control_type=control_type, reward_type=reward_type, max_episode_steps=max_episode_steps,
                              auto_reset=False)
        if goal_range != 0.3:
            for k, (lo, hi) in enumerate(((-goal_range / 2, goal_range / 2), (-goal_range / 2, goal_range / 2), (0.0, goal_range))):
                cfg.stage.goal_lo[k], cfg.stage.goal_hi[k] = lo, hi
        self._setup(cfg, device, False, max_episode_steps)


class PandaPushEnv(RobotTaskEnv):
    """reference panda_tasks.py:55-69 + envs/tasks/push.py (block_gripper=True, cube 4 cm, mass 1)."""

    _TASK = 2

    def __init__(self, render=False, reward_type="sparse", control_type="js", max_episode_steps=50, device=0):
        cfg = sim.make_config(self._TASK, 1, control_type=control_type, reward_type=reward_type, max_episode_steps=max_episode_steps,
                              auto_reset=False)
        self._setup(cfg, device, False, max_episode_steps)
        # sim.get_base_* of the "object" body (pybullet.py:182-247), what Push.get_obs reads
        self.sim.get_base_position = lambda body="object": self.batch.view("object")[0:3, 0].cpu().numpy().astype(np.float64)
        self.sim.get_base_orientation = lambda body="object": self.batch.view("object")[3:7, 0].cpu().numpy().astype(np.float64)
        self.sim.get_base_velocity = lambda body="object": self.batch.view("object")[7:10, 0].cpu().numpy().astype(np.float64)
        self.sim.get_base_angular_velocity = lambda body="object": self.batch.view("object")[10:13, 0].cpu().numpy().astype(np.float64)


class PandaPickAndPlaceEnv(PandaPushEnv):
    """reference panda_tasks.py:37-52 + envs/tasks/pick_and_place.py.  Simulated as the reference loads it: the
    franka_panda_custom_0 URDF has FIXED finger joints, so the 4th action entry moves nothing, the finger width in the
    observation is 0 and lifted goals are out of reach (documented in DESIGN.md)."""

    _TASK = 3

    def __init__(self, render=False, reward_type="sparse", control_type="ee", max_episode_steps=50, device=0):
        super().__init__(render=render, reward_type=reward_type, control_type=control_type, max_episode_steps=max_episode_steps,
                         device=device)
