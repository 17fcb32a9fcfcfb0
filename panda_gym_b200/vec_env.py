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
    """``infos`` list built on demand: SB3 indexes it per env; building 65 536 dicts per step eagerly would cost
    more than the simulation.  Behaves like a read-only sequence of dicts.  Terminal observations / episode
    statistics of the finished environments are fetched from the device once, on first use, and only their rows."""

    def __init__(self, desired, flags, done, fetch_terminal, t):
        self._desired, self._flags, self._done = desired, flags, done
        self._fetch, self._t = fetch_terminal, t
        self._term = None

    def __len__(self):
        return len(self._flags)

    def _terminal(self):
        if self._term is None:
            idx = np.flatnonzero(self._done)
            obs, ag, r, l = self._fetch(idx)
            self._term = ({int(e): k for k, e in enumerate(idx)}, obs, ag, r, l)
        return self._term

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        f = int(self._flags[i])
        info = {"is_success": bool(f & 1), "is_truncated": bool(f & 4), "TimeLimit.truncated": bool(f & 8)}
        if self._done[i]:
            pos, obs, ag, r, l = self._terminal()
            k = pos[int(i)]
            info["terminal_observation"] = {"observation": obs[k].copy(), "achieved_goal": ag[k].copy(),
                                            "desired_goal": np.array(self._desired[i], copy=True)}
            info["episode"] = {"r": float(r[k]), "l": int(l[k]), "t": self._t}
        return info

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class PandaVecEnv:
    def __init__(self, env_id="PandaReachAO-v3", n_envs=8, seed=None, scenario="reachao1", config=None, device=0,
                 max_episode_steps=50, ee_error_threshold=0.05, speed_threshold=0.1, first_env_index=0, copy_outputs=False,
                 n_host_buffers=3, **env_kwargs):
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
                              speed_threshold=speed_threshold, auto_reset=True)
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
        for _ in range(self.n_host_buffers):
            h = dict(observation=pinned((n, od), torch.float32), achieved_goal=pinned((n, 3), torch.float32),
                     desired_goal=pinned((n, 3), torch.float32), reward=pinned((n,), torch.float32),
                     done=pinned((n,), torch.uint8), flags=pinned((n,), torch.uint8))
            out = _lib.HostOut(*[ctypes.c_void_p(h[k].data_ptr()) for k in
                                 ("observation", "achieved_goal", "desired_goal", "reward", "done", "flags")])
            npv = {k: v.numpy() for k, v in h.items()}
            npv["done_bool"] = npv["done"].view(np.bool_)
            self._ring.append((h, out, npv))
        self._ring_pos = 0
        v = self.batch.view
        self._term_src = tuple(v(name) for name in ("terminal_observation", "terminal_achieved", "ep_return", "ep_length"))
        self._term_ring = [tuple(torch.empty_like(x) for x in self._term_src) for _ in range(self.n_host_buffers)]
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

    def step_wait(self):
        np.copyto(self._np_actions, np.asarray(self._actions, dtype=np.float32).reshape(self.num_envs, self.batch.act_dim))
        prev_desired = self._last_desired
        _, out, npv = self._next_buffers()
        self.batch.step_host(self._np_actions, out)
        obs = self._obs_dict(npv)
        if self.copy_outputs:
            rewards, dones, flags = npv["reward"].copy(), npv["done_bool"].copy(), npv["flags"].copy()
        else:
            rewards, dones, flags = npv["reward"], npv["done_bool"], npv["flags"]
        b = self.batch
        # terminal data lives in device buffers that a later step overwrites: keep a device-side copy of them in a
        # 3-deep ring (four asynchronous device-to-device copies, ~15 MB at 65 536 envs = a few microseconds of HBM time, no
        # host synchronisation) and gather / move the finished rows to the host only if somebody asks
        snap = self._term_ring[self._ring_pos]
        for dst, src in zip(snap, self._term_src):
            dst.copy_(src, non_blocking=True)

        def fetch(idx, snap=snap):
            sel = torch.as_tensor(np.asarray(idx, dtype=np.int64), device=snap[0].device)
            return tuple(x.index_select(0, sel).cpu().numpy() for x in snap)

        infos = LazyInfos(prev_desired if self.copy_outputs else prev_desired.copy(), flags, dones, fetch, time.time() - self._t0)
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

    def close(self):
        self.batch.close()

    def env_is_wrapped(self, wrapper_class, indices=None):
        return [False] * self.num_envs

    def get_attr(self, attr_name, indices=None):
        idx = self._indices(indices)
        return [getattr(self, attr_name) if hasattr(self, attr_name) else getattr(self.config, attr_name) for _ in idx]

    def set_attr(self, attr_name, value, indices=None):
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
        self.batch.set_stage(st)
        self.scenario = scenario
        if self.batch.obs_dim != old:
            self._alloc_host()

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
