"""Where the end-to-end step time goes (PandaVecEnv.step with host buffers, 65 536 envs): host-side segments by wall clock."""
import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from panda_gym_b200.vec_env import PandaVecEnv

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
env = PandaVecEnv("PandaReachAO-v3", n_envs=n, seed=1, scenario="reachao3", max_episode_steps=100)
env.reset()
rng = np.random.default_rng(0)
acts = [rng.uniform(-1, 1, (n, 7)).astype(np.float32) for _ in range(8)]
for k in range(10):
    env.step(acts[k % 8])
T = {"copyto": 0.0, "step_host": 0.0, "terminal_snapshot": 0.0, "infos": 0.0, "total": 0.0}
K = 50
b = env.batch
for k in range(K):
    a = acts[k % 8]
    t0 = time.perf_counter()
    np.copyto(env._np_actions, a)
    t1 = time.perf_counter()
    _, out, npv = env._next_buffers()
    b.step_host(env._np_actions, out)
    t2 = time.perf_counter()
    didx = torch.nonzero(b.view("done"), as_tuple=False).flatten()
    snap = (b.view("terminal_observation").index_select(0, didx), b.view("terminal_achieved").index_select(0, didx))
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    T["copyto"] += t1 - t0; T["step_host"] += t2 - t1; T["terminal_snapshot"] += t3 - t2
t0 = time.perf_counter()
for k in range(K):
    env.step(acts[k % 8])
T["total"] = time.perf_counter() - t0
print({k: round(1e3 * v / K, 4) for k, v in T.items()}, "ms per step; env-steps/s e2e = %.3e" % (n * K / T["total"]))
# raw copies
h = torch.empty((n, 56), dtype=torch.float32).pin_memory(); d = torch.empty((n, 56), dtype=torch.float32, device="cuda")
torch.cuda.synchronize()
for _ in range(3): h.copy_(d, non_blocking=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): h.copy_(d, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
print("raw D2H of the observation block: %.3f ms (%.1f GB/s)" % (1e3 * dt, h.numel() * 4 / dt / 1e9))
# kernel only
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
da = torch.from_numpy(acts[0]).cuda()
e0.record()
for _ in range(20): b.step(da)
e1.record(); torch.cuda.synchronize()
print("device-only step: %.3f ms" % (e0.elapsed_time(e1) / 20))
