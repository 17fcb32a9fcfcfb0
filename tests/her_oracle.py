"""NumPy restatement of stable_baselines3 HerReplayBuffer (SB3 >= 2.0: add / sample / _sample_goals, strategy
"future"), SURVEY.md App. D.  TEST INFRASTRUCTURE ONLY.  [UPSTREAM-RECALL] for the bookkeeping; the index sampling
runs on numpy's own legacy RandomState, i.e. it IS the reference generator."""
import numpy as np


class HerOracle:
    def __init__(self, buffer_size, n_envs, obs_dim, act_dim, n_sampled_goal=4, compute_reward=None):
        self.buffer_size = max(buffer_size // n_envs, 1)
        self.n_envs = n_envs
        R, E = self.buffer_size, n_envs
        z = lambda *s: np.zeros(s, np.float32)
        self.obs = {"observation": z(R, E, obs_dim), "achieved_goal": z(R, E, 3), "desired_goal": z(R, E, 3)}
        self.next_obs = {"observation": z(R, E, obs_dim), "achieved_goal": z(R, E, 3), "desired_goal": z(R, E, 3)}
        self.actions = z(R, E, act_dim); self.rewards = z(R, E); self.dones = z(R, E); self.timeouts = z(R, E)
        self.ep_start = np.zeros((R, E), np.int64); self.ep_length = np.zeros((R, E), np.int64)
        self._current_ep_start = np.zeros(E, np.int64)
        self.pos = 0; self.full = False
        self.her_ratio = 1 - (1.0 / (n_sampled_goal + 1))
        self.compute_reward = compute_reward

    def add_vectorised(self, obs, next_obs, action, reward, done, timeouts):
        """Same bookkeeping as ``add`` with the two per-environment loops replaced by masked numpy updates (the 1 M-transition
        x 1 024-column test would spend minutes in the Python loops); tests/test_host_logic.py checks both forms agree."""
        R, pos = self.buffer_size, self.pos
        st, ln = self.ep_start[pos].copy(), self.ep_length[pos].copy()
        cnt = np.where(ln > 0, st + ln - pos, 0)
        for off in range(int(cnt.max()) if cnt.size else 0):
            e = np.flatnonzero(off < cnt)
            self.ep_length[(pos + off) % R, e] = 0
        self.ep_start[pos] = self._current_ep_start.copy()
        for k in self.obs:
            self.obs[k][pos] = obs[k]; self.next_obs[k][pos] = next_obs[k]
        self.actions[pos] = action; self.rewards[pos] = reward; self.dones[pos] = done; self.timeouts[pos] = timeouts
        self.pos += 1
        if self.pos == R:
            self.full = True; self.pos = 0
        d = np.asarray(done).astype(bool)
        start = self._current_ep_start.copy()
        end = np.where(self.pos < start, self.pos + R, self.pos)
        length = np.where(d, end - start, 0)
        for off in range(int(length.max()) if length.size else 0):
            e = np.flatnonzero(off < length)
            self.ep_length[(start[e] + off) % R, e] = length[e]
        self._current_ep_start[d] = self.pos

    def add(self, obs, next_obs, action, reward, done, timeouts):
        for e in range(self.n_envs):
            st, ln = self.ep_start[self.pos, e], self.ep_length[self.pos, e]
            if ln > 0:
                idx = np.arange(self.pos, st + ln) % self.buffer_size
                self.ep_length[idx, e] = 0
        self.ep_start[self.pos] = self._current_ep_start.copy()
        for k in self.obs:
            self.obs[k][self.pos] = obs[k]; self.next_obs[k][self.pos] = next_obs[k]
        self.actions[self.pos] = action; self.rewards[self.pos] = reward; self.dones[self.pos] = done; self.timeouts[self.pos] = timeouts
        self.pos += 1
        if self.pos == self.buffer_size:
            self.full = True; self.pos = 0
        for e in range(self.n_envs):
            if done[e]:
                start, end = self._current_ep_start[e], self.pos
                if end < start:
                    end += self.buffer_size
                idx = np.arange(start, end) % self.buffer_size
                self.ep_length[idx, e] = end - start
                self._current_ep_start[e] = self.pos

    def sample(self, batch_size):
        is_valid = self.ep_length > 0
        valid = np.flatnonzero(is_valid)
        sampled = np.random.choice(valid, size=batch_size, replace=True)
        rows, envs = np.unravel_index(sampled, is_valid.shape)
        nbv = int(self.her_ratio * batch_size)
        vr, rr = np.split(rows, [nbv]); ve, re_ = np.split(envs, [nbv])
        st = self.ep_start[vr, ve]; ln = self.ep_length[vr, ve]
        cur = (vr - st) % self.buffer_size
        t = np.random.randint(cur, ln)
        tr = (t + st) % self.buffer_size
        new_goal = self.next_obs["achieved_goal"][tr, ve]
        out = {}
        for name, store in (("obs", self.obs), ("next_obs", self.next_obs)):
            out[name] = {k: np.concatenate([store[k][rr, re_], store[k][vr, ve]]) for k in store}
        out["obs"]["desired_goal"][len(rr):] = new_goal
        out["next_obs"]["desired_goal"][len(rr):] = new_goal
        v_rew = self.compute_reward(self.next_obs["achieved_goal"][vr, ve], new_goal).astype(np.float32)
        out["rewards"] = np.concatenate([self.rewards[rr, re_], v_rew])
        out["actions"] = np.concatenate([self.actions[rr, re_], self.actions[vr, ve]])
        out["dones"] = np.concatenate([self.dones[rr, re_] * (1 - self.timeouts[rr, re_]), self.dones[vr, ve] * (1 - self.timeouts[vr, ve])])
        out["indices"] = (rows, envs, t)
        return out
