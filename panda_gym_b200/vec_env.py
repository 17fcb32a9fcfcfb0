"""``PandaVecEnv``: the batched, SB3-VecEnv-shaped front of the CUDA simulator.

Drop-in for what the reference builds with
``make_vec_env(env_id, n_envs, seed, env_kwargs, vec_env_cls=SubprocVecEnv)`` (reference
``training/utils/setup_training.py:43-47``): same ``reset / step_async / step_wait / step / env_method /
get_attr / set_attr / seed / close`` surface, same observation dict, rewards, dones and per-env ``infos`` keys
(``is_success``, ``is_truncated``, ``TimeLimit.truncated``, ``terminal_observation``, ``episode``), with
auto-reset of finished environments.  One object drives all N environments of one GPU through
``pg_step_host`` (host buffers) or ``step_device`` (zero-copy device tensors).
"""
import ctypes
import time

import numpy as np
import torch

from . import _lib, sim, spaces
from .config import TrainConfig

ENV_IDS = {}
for _rew, _rs in (("sparse", ""), ("dense", "Dense")):
    for _ct, _cs in (("ee", ""), ("joints", "Joints")):
        ENV_IDS["PandaReach%s%s-v3" % (_cs, _rs)] = dict(task=0, reward_type=_rew, control_type=_ct)
        ENV_IDS["PandaPush%s%s-v3" % (_cs, _rs)] = dict(task=2, reward_type=_rew, control_type=_ct)
        ENV_IDS["PandaPickAndPlace%s%s-v3" % (_cs, _rs)] = dict(task=3, reward_type=_rew, control_type=_ct)
ENV_IDS["PandaReachAO-v3"] = dict(task=1)


class LazyInfos:
    """``infos`` of one vector step.  SB3 indexes it per env (``infos[i]``); building 65 536 dicts per step eagerly would
    cost more than the simulation, so the per-env dict is built on ``__getitem__`` only and everything a Monitor /
    replay buffer needs is ALSO available vectorised:

      ``is_success``, ``is_truncated``, ``time_limit_truncated``  bool arrays [n]
      ``finished``                 indices of the environments whose episode ended in this step
      ``episode_returns`` / ``episode_lengths``  Monitor's ``r`` / ``l`` of those episodes (fetched once, on first use)
      ``terminal_observations()``  dict of the finished rows' last observations

    ``TimeLimit.truncated`` follows SB3's SubprocVecEnv worker on gymnasium >= 0.26: ``truncated and not terminated``
    where gymnasium's ``truncated`` is the task's ``is_truncated()`` (collision, core.py:364) OR the TimeLimit wrapper.
    So a collision-truncated ReachAO episode reports True and the learner bootstraps through it."""

    def __init__(self, desired, flags, done, fetch_terminal, t):
        self._desired, self._flags, self._done = desired, flags, done
        self._fetch, self._t = fetch_terminal, t
        self._term = None

    def __len__(self):
        return len(self._flags)

    # ---- vectorised accessors
    @property
    def is_success(self):
        return (self._flags & 1) != 0

    @property
    def is_truncated(self):
        return (self._flags & 4) != 0

    @property
    def time_limit_truncated(self):
        f = self._flags
        return ((f & 12) != 0) & ((f & 2) == 0)

    @property
    def finished(self):
        """Indices (ascending) of the environments whose episode ended in this step."""
        return self._terminal()[5]

    def _terminal(self):
        if self._term is None:
            idx, obs, ag, r, l = self._fetch()
            self._term = (None, obs, ag, r, l, idx)
        return self._term

    def _row_of(self, i):
        t = self._terminal()
        if t[0] is None:
            self._term = t = ({int(e): k for k, e in enumerate(t[5])},) + t[1:]
        return t[0][int(i)]

    @property
    def episode_returns(self):
        return self._terminal()[3]

    @property
    def episode_lengths(self):
        return self._terminal()[4]

    def terminal_observations(self):
        _, obs, ag, _, _, idx = self._terminal()
        return {"observation": obs, "achieved_goal": ag, "desired_goal": np.asarray(self._desired)[idx]}

    # ---- SB3's per-env view
    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        f = int(self._flags[i])
        info = {"is_success": bool(f & 1), "is_truncated": bool(f & 4), "TimeLimit.truncated": bool(f & 12) and not bool(f & 2)}
        if self._done[i]:
            _, obs, ag, r, l, _ = self._terminal()
            k = self._row_of(i)
            info["terminal_observation"] = {"observation": obs[k].copy(), "achieved_goal": ag[k].copy(),
                                            "desired_goal": np.array(self._desired[i], copy=True)}
            info["episode"] = {"r": float(r[k]), "l": int(l[k]), "t": self._t}
        return info

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class PandaVecEnv:
    def __init__(self, env_id="PandaReachAO-v3", n_envs=8, seed=None, scenario="reachao1", config=None, device=0,
                 max_episode_steps=50, ee_error_threshold=0.05, speed_threshold=0.1, first_env_index=0, copy_outputs=False,
                 n_host_buffers=3, numa_bind=True, pose_retry_limit=0, **env_kwargs):
        if env_id not in ENV_IDS:
            raise _lib.PandaB200Error("unknown / out-of-scope env id %r (supported: %s)" % (env_id, sorted(ENV_IDS)))
        spec = dict(ENV_IDS[env_id])
        spec.update({k: v for k, v in env_kwargs.items() if k in ("reward_type", "control_type")})
        self.env_id, self.num_envs = env_id, int(n_envs)
        self.config = config if config is not None else TrainConfig()
        self.scenario = scenario if spec["task"] == 1 else "reach_classic"
        cfg = sim.make_config(spec["task"], n_envs, scenario=self.scenario if spec["task"] == 1 else None, config=self.config,
                              control_type=spec.get("control_type"), reward_type=spec.get("reward_type"),
                              max_episode_steps=max_episode_steps, ee_error_threshold=ee_error_threshold,
                              speed_threshold=speed_threshold, auto_reset=True, pose_retry_limit=pose_retry_limit)
        # One process drives one GPU: run it (and allocate its page-locked result buffers) on the CPUs / memory node next to
        # that GPU.  On a two-socket host the per-step result copy is otherwise 2-3x slower whenever the launcher happened
        # to start the process on the far socket (profiles/r02_e2e.md).  numa_bind=False leaves the affinity alone.
        self.numa = None
        if numa_bind:
            from .sharding import bind_to_local_numa
            self.numa = bind_to_local_numa(device if isinstance(device, int) else torch.device(device).index or 0)
        self.batch = sim.PandaBatch(cfg, device=device)
        self.first_env_index = int(first_env_index)
        self.batch.seed(0 if seed is None else seed, self.first_env_index)
        self._t0 = time.time()
        self.copy_outputs = bool(copy_outputs)
        self.n_host_buffers = max(2, int(n_host_buffers))
        self._alloc_host()
        self._actions = None

    # ------------------------------------------------------------------ spaces
    def _alloc_host(self):
        """Page-locked result buffers.  ``copy_outputs=False`` (default) hands out numpy views of a ring of
        ``n_host_buffers`` (3) buffer sets: an array returned by step t stays intact until step t + 3 overwrites it --
        enough for SB3's ``_last_obs`` pattern -- and no 15 MB memcpy is paid per step.  ``copy_outputs=True`` returns
        fresh arrays like the reference (core.py:287-295)."""
        n, od, ad = self.num_envs, self.batch.obs_dim, self.batch.act_dim

        def pinned(shape, dtype):
            return torch.empty(shape, dtype=dtype).pin_memory()

        self._ring = []
        off, total = self.batch.host_layout()
        for _ in range(self.n_host_buffers):
            # one page-locked block per ring slot, laid out like the device slab: a step is ONE device-to-host copy
            blk = pinned((total,), torch.uint8)

            def carve(o, shape, dtype, blk=blk):
                nbytes = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
                return blk[o:o + nbytes].view(dtype).view(shape)

            h = dict(observation=carve(off[0], (n, od), torch.float32), achieved_goal=carve(off[1], (n, 3), torch.float32),
                     desired_goal=carve(off[2], (n, 3), torch.float32), reward=carve(off[3], (n,), torch.float32),
                     done=carve(off[4], (n,), torch.uint8), flags=carve(off[5], (n,), torch.uint8), _block=blk)
            out = _lib.HostOut(*[ctypes.c_void_p(h[k].data_ptr()) for k in
                                 ("observation", "achieved_goal", "desired_goal", "reward", "done", "flags")])
            npv = {k: v.numpy() for k, v in h.items() if not k.startswith("_")}
            npv["done_bool"] = npv["done"].view(np.bool_)
            self._ring.append((h, out, npv))
        self._ring_pos = 0
        self._h_actions = pinned((n, ad), torch.float32)
        self._np_actions = self._h_actions.numpy()
        self.observation_space = spaces.Dict(dict(
            observation=spaces.Box(-10.0, 10.0, shape=(od,), dtype=np.float32),
            desired_goal=spaces.Box(-10.0, 10.0, shape=(3,), dtype=np.float32),
            achieved_goal=spaces.Box(-10.0, 10.0, shape=(3,), dtype=np.float32)))
        self.action_space = spaces.Box(-1.0, 1.0, shape=(ad,), dtype=np.float32)
        self._last_desired = np.zeros((n, 3), np.float32)

    def _next_buffers(self):
        self._ring_pos = (self._ring_pos + 1) % len(self._ring)
        return self._ring[self._ring_pos]

    def _obs_dict(self, npv):
        if self.copy_outputs:
            return {k: npv[k].copy() for k in ("observation", "achieved_goal", "desired_goal")}
        return {k: npv[k] for k in ("observation", "achieved_goal", "desired_goal")}

    # ------------------------------------------------------------------ VecEnv API (host buffers)
    def seed(self, seed=None):
        self.batch.seed(0 if seed is None else seed, self.first_env_index)
        return [None if seed is None else seed + self.first_env_index + i for i in range(self.num_envs)]

    def reset(self):
        _, out, npv = self._next_buffers()
        self.batch.reset_host(out)
        self._last_desired = npv["desired_goal"]
        return self._obs_dict(npv)

    def step_async(self, actions):
        self._actions = actions

    def action_buffer(self):
        """The page-locked float32 [n_envs, act_dim] array the host-to-device copy of ``step`` reads from.  A policy that writes
        its actions straight into it and passes it to ``step`` saves the staging copy (1.8 MB at 65 536 environments)."""
        return self._np_actions

    def step_wait(self):
        if self._actions is not self._np_actions:
            np.copyto(self._np_actions, np.asarray(self._actions, dtype=np.float32).reshape(self.num_envs, self.batch.act_dim))
        prev_desired = self._last_desired
        _, out, npv = self._next_buffers()
        self.batch.step_host(self._np_actions, out)
        obs = self._obs_dict(npv)
        if self.copy_outputs:
            rewards, dones, flags = npv["reward"].copy(), npv["done_bool"].copy(), npv["flags"].copy()
        else:
            rewards, dones, flags = npv["reward"], npv["done_bool"], npv["flags"]
        # Finished episodes: the step kernel wrote one compact row per episode that ended (index, return, length, flags,
        # terminal achieved goal / observation) into a 3-deep device ring; LazyInfos fetches the rows of ITS step with one
        # small device-to-host copy when (and only when) somebody reads them.  Nothing is snapshotted per step.
        b = self.batch
        sid = b.step_id()

        def fetch(b=b, sid=sid):
            idx, r, l, _, ag, ob = b.fetch_finished(sid)
            return idx, ob, ag, r, l

        infos = LazyInfos(prev_desired, flags, dones, fetch, time.time() - self._t0)
        self._last_desired = obs["desired_goal"]
        return obs, rewards, dones, infos

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def start_benchmark(self):
        """Begin accumulating perform_benchmark's metrics (evaluation/evaluate.py:131-300) on the device."""
        self.batch.set_eval_metrics(True)

    def benchmark_results(self, group=None):
        """perform_benchmark's ``results`` dict over every episode finished since ``start_benchmark`` (all ranks)."""
        from .sharding import benchmark_results
        return benchmark_results(self.batch.eval_metrics_raw(), n_substeps=self.batch.cfg.n_substeps, group=group)

    def get_images(self, width=720, height=480, indices=None, **camera):
        """SB3 VecEnv.get_images: one rgb_array frame per environment in `indices` (default: the first), uint8 [h, w, 4]."""
        idx = [0] if indices is None else list(indices)
        return [self.batch.render(i, width, height, **camera).cpu().numpy() for i in idx]

    def render(self, mode="rgb_array", **kwargs):
        """SB3 VecEnv.render("rgb_array") for the first environment (a tiled view of thousands of environments is not useful)."""
        return self.get_images(**kwargs)[0] if mode == "rgb_array" else None

    def close(self):
        self.batch.close()

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False] * self.num_envs

    # per-environment state readable through get_attr (what callbacks / evaluate.py read off env.task / env.robot)
    _PER_ENV = {"goal": ("goal", 1), "joint_angles": ("q", 1), "joint_velocities": ("qd", 1), "is_collided": ("flags", 0),
                "distances_closest_obstacles": ("d_link", 1), "step_count": ("step_count", 0)}

    def get_attr(self, attr_name, indices=None):
        """SB3 ``VecEnv.get_attr``: one value per selected environment.  Per-env simulator state (``goal``,
        ``joint_angles``, ``joint_velocities``, ``distances_closest_obstacles``, ``is_collided``, ``step_count``) is read
        from the device; configuration attributes (``scenario``, ``config`` fields, spaces, ...) are shared by all
        environments of the batch."""
        idx = self._indices(indices)
        if attr_name in self._PER_ENV:
            view, soa = self._PER_ENV[attr_name]
            t = self.batch.view(view)
            sel = torch.as_tensor(idx, dtype=torch.int64, device=t.device)
            vals = (t.index_select(1, sel).t() if soa else t.index_select(0, sel)).cpu().numpy()
            if attr_name == "is_collided":
                vals = (vals & 4) != 0
            return [v.copy() if isinstance(v, np.ndarray) else v for v in vals]
        if hasattr(self, attr_name):
            return [getattr(self, attr_name) for _ in idx]
        return [getattr(self.config, attr_name) for _ in idx]

    def set_attr(self, attr_name, value, indices=None):
        """SB3 ``VecEnv.set_attr``.  The batch shares one configuration: setting an attribute for a strict subset of the
        environments cannot be honoured and raises instead of silently applying it to all."""
        if len(self._indices(indices)) != self.num_envs:
            raise _lib.PandaB200Error("set_attr on a subset of environments: the batched simulator shares its configuration")
        setattr(self, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        """``env.env_method("compute_reward", ag, dg, infos, indices=[0])`` is how SB3's HER relabels."""
        idx = self._indices(indices)
        if method_name == "compute_reward":
            return [self.compute_reward(args[0], args[1], args[2] if len(args) > 2 else None, env_index=i) for i in idx]
        if method_name == "get_manipulability":  # robot.get_manipulability (panda.py:314-317)
            m = self.batch.manipulability().cpu().numpy()
            return [float(m[i]) for i in idx]
        raise _lib.PandaB200Error("env_method(%r) is not part of the supported hot path" % method_name)

    def _indices(self, indices):
        if indices is None:
            return list(range(self.num_envs))
        if isinstance(indices, int):
            return [indices]
        return list(indices)

    def compute_reward(self, achieved_goal, desired_goal, info=None, env_index=0):
        """Task.compute_reward, vectorised over leading dims (core.py:282)."""
        ag = np.asarray(achieved_goal, dtype=np.float32)
        dg = np.asarray(desired_goal, dtype=np.float32)
        assert ag.shape == dg.shape
        dev = self.batch.device
        out = self.batch.compute_reward(torch.from_numpy(np.ascontiguousarray(ag)).to(dev),
                                        torch.from_numpy(np.ascontiguousarray(dg)).to(dev), env_index)
        return out.cpu().numpy().reshape(ag.shape[:-1])

    # ------------------------------------------------------------------ curriculum (setup_training.py:233-252)
    def set_stage(self, scenario, max_episode_steps=None, ee_error_threshold=0.05, speed_threshold=0.1):
        st = sim.stage_from_scenario(scenario, max_episode_steps or self.batch.cfg.stage.max_episode_steps,
                                     ee_error_threshold, speed_threshold)
        old = self.batch.obs_dim
        self.batch.set_stage(st)  # swaps the block and resets every environment (the reference rebuilds the VecEnv)
        self.scenario = scenario
        if self.batch.obs_dim != old:
            self._alloc_host()
        _, out, npv = self._next_buffers()
        self.batch.copy_out(out)
        self._last_desired = npv["desired_goal"]
        return self._obs_dict(npv)

    # ------------------------------------------------------------------ device-native API (no host copies)
    def step_device(self, actions):
        """actions: float32 CUDA tensor [n, act_dim]; returns zero-copy views (valid until the next step)."""
        self.batch.step(actions)
        v = self.batch.view
        return ({"observation": v("observation"), "achieved_goal": v("achieved_goal"), "desired_goal": v("desired_goal")},
                v("reward"), v("done"), v("flags"))

    def reset_device(self):
        self.batch.reset()
        v = self.batch.view
        return {"observation": v("observation"), "achieved_goal": v("achieved_goal"), "desired_goal": v("desired_goal")}

    def episode_stats(self, group=None):
        """{episodes, successes, collisions, timeouts, sum_return, sum_length, reset_failures, rates}; summed over
        ``torch.distributed`` ranks when a process group is initialised (the one collective of the path)."""
        from .sharding import allreduce_stats
        return allreduce_stats(self.batch.view("stats"), group)


def make_vec_env(env_id, n_envs=1, seed=None, env_kwargs=None, vec_env_cls=None, **kw):
    """Signature-compatible with SB3's make_vec_env as the reference calls it (setup_training.py:43-47);
    ``vec_env_cls`` is accepted and ignored (the batch lives on the GPU, not in subprocesses)."""
    env_kwargs = dict(env_kwargs or {})
    env_kwargs.update(kw)
    return PandaVecEnv(env_id, n_envs=n_envs, seed=seed, **env_kwargs)
