"""Where the end-to-end step time goes (PandaVecEnv.step with host buffers, 65 536 envs, steady state): the C call alone
(pg_step_host: H2D + kernels + one D2H + synchronise), the public step, the public step + the `infos` a Monitor reads."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from panda_gym_b200.vec_env import PandaVecEnv

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
env = PandaVecEnv("PandaReachAO-v3", n_envs=n, seed=1, scenario="reachao3", max_episode_steps=100)
env.reset()
b = env.batch
for k in range(150):
    b.step(b.random_actions(7, k))
K = 60
acts = [b.random_actions(9, k).cpu().numpy() for k in range(4 * K)]
it = iter(acts)
def wall(f):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(K): f()
    torch.cuda.synchronize(); return 1e3 * (time.perf_counter() - t0) / K
def c_call():
    np.copyto(env._np_actions, next(it)); _, out, npv = env._next_buffers(); b.step_host(env._np_actions, out)
def public():
    abuf = env.action_buffer(); np.copyto(abuf, next(it)); env.step(abuf)
def public_infos():
    abuf = env.action_buffer(); np.copyto(abuf, next(it)); obs, rew, dones, infos = env.step(abuf)
    fin = infos.finished
    if len(fin):
        term = infos.terminal_observations(); float(infos.episode_returns.sum()); float(term["observation"][0, 0])
dacts = [torch.from_numpy(a).cuda() for a in acts[:K]]
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for a in dacts: b.step(a)
e1.record(); torch.cuda.synchronize()
out = {"device_only_ms": round(e0.elapsed_time(e1) / K, 4), "pg_step_host_ms": round(wall(c_call), 4), "public_step_ms": round(wall(public), 4),
       "public_step_plus_infos_ms": round(wall(public_infos), 4)}
t0 = time.perf_counter()
for _ in range(K): np.copyto(env._np_actions, acts[0])
out["copyto_actions_ms"] = round(1e3 * (time.perf_counter() - t0) / K, 4)
print(out)
