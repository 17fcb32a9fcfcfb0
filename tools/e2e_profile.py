"""cProfile of the end-to-end host loop bench.py times (run on the GPU box)."""
import cProfile, os, pstats, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from panda_gym_b200.vec_env import PandaVecEnv
N = 65536
env = PandaVecEnv("PandaReachAO-v3", n_envs=N, seed=1234, scenario="reachao3", max_episode_steps=100)
b = env.batch
b.reset()
acts = [b.random_actions(1234, k).cpu().numpy() for k in range(60)]
def loop(K0, K):
    c = 0.0
    for k in range(K0, K0 + K):
        obs, rew, dones, infos = env.step(acts[k % 60])
        fin = infos.finished
        if len(fin):
            term = infos.terminal_observations()
            c += float(infos.episode_returns.sum()) + float(infos.episode_lengths.sum()) + float(term["observation"][0, 0]) \
                + float(infos.time_limit_truncated[fin].sum())
    return c
loop(0, 10)
torch.cuda.synchronize(); t0 = time.perf_counter(); loop(10, 40); torch.cuda.synchronize()
print("ms per e2e step: %.3f" % ((time.perf_counter() - t0) / 40 * 1e3))
pr = cProfile.Profile(); pr.enable(); loop(50, 40); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
