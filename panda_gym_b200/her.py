"""``HerReplayBuffer``: on-device mirror of stable_baselines3's ``HerReplayBuffer`` (SB3 >= 2.0), the replay buffer
the reference trains with (``classes/train_config.py:2,15``; default kwargs ``n_sampled_goal=4``,
``goal_selection_strategy="future"``).  Storage, episode bookkeeping, index sampling, future-goal relabelling and
the reward recompute all run in CUDA kernels (``csrc/pg_her.cuh``); ``sample`` returns device tensors.

``rng="numpy"``: index sampling consumes numpy's *global legacy* MT19937 stream exactly as SB3 does
(``np.random.choice`` + ``np.random.randint``): the state is pushed to the device before and pulled back after each
``sample`` call, so indices are bit-identical to SB3's for the same ``np.random.seed``.  ``rng="device"``: parallel
counter-based generator (same distribution, not the numpy stream) for throughput.
"""
import collections
import ctypes

import numpy as np
import torch

from . import _lib

DictReplayBufferSamples = collections.namedtuple(
    "DictReplayBufferSamples", ["observations", "actions", "next_observations", "dones", "rewards"])


class HerReplayBuffer:
    def __init__(self, buffer_size, env, n_sampled_goal=4, goal_selection_strategy="future", max_batch_size=65536, rng="numpy",
                 device_seed=0):
        """env: a ``PandaVecEnv`` (its ``batch`` provides compute_reward and the observation / action widths)."""
        if goal_selection_strategy != "future":
            raise _lib.PandaB200Error("only the reference's goal_selection_strategy='future' is built")
        self.env = env
        self.batch = env.batch
        self.lib = _lib.load()
        self.n_envs = self.batch.n
        self.obs_dim, self.act_dim = self.batch.obs_dim, self.batch.act_dim
        self.n_sampled_goal = n_sampled_goal
        self.her_ratio = 1 - (1.0 / (n_sampled_goal + 1))
        self.rng_mode = rng
        self.device_seed = int(device_seed)
        self.max_batch_size = int(max_batch_size)
        h = ctypes.c_void_p()
        _lib.check(self.lib.pg_her_create(self.batch.device_index, int(buffer_size), self.n_envs, self.obs_dim, self.act_dim,
                                          int(n_sampled_goal), self.max_batch_size, ctypes.byref(h)))
        self.h = h
        rows = ctypes.c_int()
        _lib.check(self.lib.pg_her_info(self.h, ctypes.byref(rows), None, None, None))
        self.buffer_size = rows.value  # SB3: buffer_size // n_envs rows

    def close(self):
        if getattr(self, "h", None):
            self.lib.pg_her_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_env(self, env):  # setup_training.py:176-179 on curriculum stage switch
        self.env, self.batch = env, env.batch

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.batch.device).cuda_stream)

    @staticmethod
    def _dev(x, device, dtype):
        t = x if torch.is_tensor(x) else torch.as_tensor(np.ascontiguousarray(x))
        return t.to(device=device, dtype=dtype).contiguous()

    def add(self, obs, next_obs, action, reward, done, infos=None, timeouts=None):
        """obs / next_obs: dicts {"observation", "achieved_goal", "desired_goal"} of [n_envs, ...] arrays or tensors.
        ``infos[i]["TimeLimit.truncated"]`` (or the ``timeouts`` array) marks time-limit endings."""
        dev = self.batch.device
        f = torch.float32
        if timeouts is None and infos is not None:
            if hasattr(infos, "time_limit_truncated"):  # LazyInfos: the vectorised form of the per-env lookup below
                timeouts = infos.time_limit_truncated
            else:
                timeouts = np.array([bool(infos[i].get("TimeLimit.truncated", False)) for i in range(self.n_envs)])
        t = [self._dev(obs["observation"], dev, f), self._dev(obs["achieved_goal"], dev, f), self._dev(obs["desired_goal"], dev, f),
             self._dev(next_obs["observation"], dev, f), self._dev(next_obs["achieved_goal"], dev, f),
             self._dev(next_obs["desired_goal"], dev, f), self._dev(action, dev, f), self._dev(reward, dev, f),
             self._dev(done, dev, torch.uint8)]
        to = self._dev(timeouts, dev, torch.uint8) if timeouts is not None else None
        ptrs = [ctypes.c_void_p(x.data_ptr()) for x in t] + [ctypes.c_void_p(to.data_ptr()) if to is not None else None]
        _lib.check(self.lib.pg_her_add(self.h, *ptrs, self._stream()))

    def _view(self, ptr, shape, row_stride):
        from .sim import _CudaArray
        with torch.cuda.device(self.batch.device):
            return torch.as_tensor(_CudaArray(ptr, shape, 0, self, strides=(4 * row_stride, 4)), device=self.batch.device)

    def _raise_if_empty(self):
        status = ctypes.c_int()
        _lib.check(self.lib.pg_her_last_indices(self.h, 1, None, None, None, None, ctypes.byref(status)))
        if status.value:
            raise RuntimeError("Unable to sample before the end of the first episode. We recommend choosing a value for "
                               "learning_starts that is greater than the maximum number of timesteps in the environment.")
        self._seen_valid = True

    def sample(self, batch_size, env=None):
        B = int(batch_size)
        out = _lib.HerBatch()
        exact = self.rng_mode == "numpy"
        if exact:  # hand numpy's global legacy stream to the device ...
            st = np.random.get_state()
            key = np.ascontiguousarray(st[1], dtype=np.uint32)
            _lib.check(self.lib.pg_her_set_mt_state(self.h, key.ctypes.data_as(ctypes.c_void_p), int(st[2])))
        _lib.check(self.lib.pg_her_sample(self.h, self.batch.h, B, 1 if exact else 0, self.device_seed, 0, ctypes.byref(out), self._stream()))
        # SB3 raises when no episode has finished yet.  The gather kernel leaves the batch untouched in that case; the
        # status word is read back (a 4-byte copy + sync) on every numpy-mode call, and in device mode until the first
        # successful sample (an on-device buffer that has held a finished episode practically never empties again).
        if exact or not getattr(self, "_seen_valid", False):
            self._raise_if_empty()
        if exact:  # ... and take it back, advanced exactly as np.random.choice + np.random.randint would have
            key = np.zeros(624, np.uint32); pos = ctypes.c_int()
            _lib.check(self.lib.pg_her_get_mt_state(self.h, key.ctypes.data_as(ctypes.c_void_p), ctypes.byref(pos)))
            np.random.set_state(("MT19937", key, pos.value, st[3], st[4]))  # has_gauss / cached_gaussian pass through untouched
        self.last_n_virtual = int(out.n_virtual)
        od, ad = self.obs_dim, self.act_dim
        rs = int(out.row_stride)
        self._last_packed = (out.observation, B, rs)
        v = lambda ptr, shape: self._view(ptr, shape, rs)
        observations = {"observation": v(out.observation, (B, od)), "achieved_goal": v(out.achieved_goal, (B, 3)),
                        "desired_goal": v(out.desired_goal, (B, 3))}
        next_observations = {"observation": v(out.next_observation, (B, od)), "achieved_goal": v(out.next_achieved_goal, (B, 3)),
                             "desired_goal": v(out.next_desired_goal, (B, 3))}
        return DictReplayBufferSamples(observations, v(out.action, (B, ad)), next_observations, v(out.done, (B, 1)), v(out.reward, (B, 1)))

    def last_packed(self):
        """The last sampled batch as ONE ``[B, row_stride]`` device block (record field order: observation | next observation
        | achieved | desired | next achieved | next desired | action | reward | done | ...): what
        ``sharding.allgather_samples`` exchanges with a single collective."""
        ptr, B, rs = self._last_packed
        return self._view(ptr, (B, rs), rs)

    def last_indices(self, batch_size):
        B = int(batch_size)
        row = np.zeros(B, np.int32); envi = np.zeros(B, np.int32); t = np.zeros(B, np.int32)
        nv = ctypes.c_int(); status = ctypes.c_int()
        P = ctypes.c_void_p
        _lib.check(self.lib.pg_her_last_indices(self.h, B, row.ctypes.data_as(P), envi.ctypes.data_as(P), t.ctypes.data_as(P),
                                                ctypes.byref(nv), ctypes.byref(status)))
        return row, envi, t, nv.value, status.value

    def size(self):
        pos, full = ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.pg_her_info(self.h, None, ctypes.byref(pos), ctypes.byref(full), None))
        return self.buffer_size if full.value else pos.value
