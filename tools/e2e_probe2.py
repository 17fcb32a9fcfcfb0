"""Finer split of PandaVecEnv.step on the host (run on the GPU box)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from panda_gym_b200.vec_env import PandaVecEnv, LazyInfos
N = 65536
env = PandaVecEnv("PandaReachAO-v3", n_envs=N, seed=1, scenario="reachao3", max_episode_steps=100)
env.reset()
rng = np.random.default_rng(0)
A = [rng.uniform(-1, 1, (N, 7)).astype(np.float32) for _ in range(8)]
def tm(f, reps=30):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
k = [0]
def copy_actions():
    k[0] += 1; np.copyto(env._np_actions, np.asarray(A[k[0] % 8], dtype=np.float32).reshape(N, 7))
def host_same():
    env.batch.step_host(env._np_actions, env._ring[0][1])
def host_ring():
    _, out, npv = env._next_buffers(); env.batch.step_host(env._np_actions, out)
def host_ring_copy():
    k[0] += 1; np.copyto(env._np_actions, A[k[0] % 8]); _, out, npv = env._next_buffers(); env.batch.step_host(env._np_actions, out)
def full():
    k[0] += 1; env.step(A[k[0] % 8])
def full_fin():
    k[0] += 1; o, r, d, i = env.step(A[k[0] % 8]); i.finished
def fetch_only():
    env.batch.fetch_finished(env.batch.step_id())
def count_only():
    env.batch.finished_count(env.batch.step_id())
for name, f in (("copy actions", copy_actions), ("step_host same buf", host_same), ("step_host ring", host_ring), ("ring + copy", host_ring_copy),
                ("env.step", full), ("env.step + finished", full_fin), ("fetch_finished alone", fetch_only), ("finished_count", count_only)):
    print("%-22s %.3f ms" % (name, tm(f)))
print("numa", env.numa)
print("finished per step ~", env.batch.finished_count(env.batch.step_id()))
