"""Where the host-API step spends its time (run on the GPU box): pg_step_host alone, + LazyInfos pieces."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from panda_gym_b200.vec_env import PandaVecEnv
N = 65536
env = PandaVecEnv("PandaReachAO-v3", n_envs=N, seed=1, scenario="reachao3", max_episode_steps=100)
env.reset()
rng = np.random.default_rng(0)
A = [rng.uniform(-1, 1, (N, 7)).astype(np.float32) for _ in range(8)]
def tm(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
k = [0]
def step_only():
    k[0] += 1; env.step(A[k[0] % 8])
def step_fin():
    k[0] += 1; o, r, d, i = env.step(A[k[0] % 8]); i.finished
def step_ret():
    k[0] += 1; o, r, d, i = env.step(A[k[0] % 8]); i.episode_returns
def step_term():
    k[0] += 1; o, r, d, i = env.step(A[k[0] % 8]); i.terminal_observations()
_, out, npv = env._next_buffers()
def host_only():
    env.batch.step_host(env._np_actions, out)
dev_a = torch.from_numpy(A[0]).cuda()
def dev_only():
    env.batch.step(dev_a)
def copy_only():
    env.batch.copy_out(out)
import ctypes
from panda_gym_b200 import _lib
lib = _lib.load()
def h2d_only():
    # the H2D leg of pg_step_host in isolation: same pinned source, same device staging size
    dev_a.copy_(env._h_actions, non_blocking=True)
for name, f in (("pg_copy_out (D2H)", copy_only), ("actions H2D", h2d_only), ("device step", dev_only), ("pg_step_host", host_only), ("env.step", step_only), ("+finished", step_fin), ("+episode_returns", step_ret), ("+terminal_obs", step_term)):
    print("%-18s %.3f ms" % (name, tm(f)))

# raw page-locked device-to-host bandwidth on this box, for reference
for mb in (16, 64, 256):
    src = torch.empty(mb << 20, dtype=torch.uint8, device="cuda"); dst = torch.empty(mb << 20, dtype=torch.uint8).pin_memory()
    def cp(): dst.copy_(src, non_blocking=True)
    print("D2H %3d MiB pinned: %.1f GB/s" % (mb, (mb << 20) / (tm(cp) * 1e-3) / 1e9))
h = torch.empty(16 << 20, dtype=torch.uint8).pin_memory(); d = torch.empty(16 << 20, dtype=torch.uint8, device="cuda")
print("H2D  16 MiB pinned: %.1f GB/s" % ((16 << 20) / (tm(lambda: d.copy_(h, non_blocking=True)) * 1e-3) / 1e9))
