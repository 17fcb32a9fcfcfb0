"""Where the end-to-end step time goes (PandaVecEnv.step with host buffers, 65 536 envs): host-side segments by wall clock,
fresh random actions every step (cycling a few action sets drives the arms into their joint limits = the slow path)."""
import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from panda_gym_b200.vec_env import PandaVecEnv

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
env = PandaVecEnv("PandaReachAO-v3", n_envs=n, seed=1, scenario="reachao3", max_episode_steps=100)
env.reset()
rng = np.random.default_rng(0)
K = 40
acts = [rng.uniform(-1, 1, (n, 7)).astype(np.float32) for _ in range(3 * K + 10)]
it = iter(acts)
for k in range(10):
    env.step(next(it))
b = env.batch
# (1) the public call
t0 = time.perf_counter()
for k in range(K):
    env.step(next(it))
total = (time.perf_counter() - t0) / K
# (2) its segments
T = {"copyto": 0.0, "step_host": 0.0, "rest_of_step_wait": 0.0}
for k in range(K):
    a = next(it)
    t0 = time.perf_counter()
    np.copyto(env._np_actions, a)
    t1 = time.perf_counter()
    _, out, npv = env._next_buffers()
    b.step_host(env._np_actions, out)
    t2 = time.perf_counter()
    snap = env._term_ring[env._ring_pos]
    for dst, src in zip(snap, env._term_src):
        dst.copy_(src, non_blocking=True)
    d = npv["desired_goal"].copy()
    t3 = time.perf_counter()
    T["copyto"] += t1 - t0; T["step_host"] += t2 - t1; T["rest_of_step_wait"] += t3 - t2
# (3) device-only step with the same kind of actions, and the raw copies
dacts = [torch.from_numpy(next(it)).cuda() for _ in range(K)]
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for a in dacts: b.step(a)
e1.record(); torch.cuda.synchronize()
dev_ms = e0.elapsed_time(e1) / K
h = torch.empty((n, 56), dtype=torch.float32).pin_memory(); d = torch.empty((n, 56), dtype=torch.float32, device="cuda")
for _ in range(3): h.copy_(d, non_blocking=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): h.copy_(d, non_blocking=True); torch.cuda.synchronize()
d2h = (time.perf_counter() - t0) / 20
print({"public_step_ms": round(1e3 * total, 4), **{k: round(1e3 * v / K, 4) for k, v in T.items()}, "device_only_step_ms": round(dev_ms, 4),
       "raw_d2h_obs_ms": round(1e3 * d2h, 4), "d2h_GBps": round(h.numel() * 4 / d2h / 1e9, 1), "e2e_env_steps_per_s": n / total})
