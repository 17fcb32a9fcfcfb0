"""BASELINE configs[4]: vectorised HER relabel+sample (k = 4 future) over a 1M-transition on-device replay buffer.
1 048 576 transitions x 1 024 env columns, synthetic episodes of length U{10..100} (seed 7), obs 56.
Reports sampled transitions/s and effective GB/s (1.08 KB algorithmic bytes per sampled transition)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from panda_gym_b200.vec_env import PandaVecEnv
from panda_gym_b200.her import HerReplayBuffer

n_envs, rows = 1024, 1024
env = PandaVecEnv("PandaReachAO-v3", n_envs=n_envs, seed=1, scenario="reachao3", max_episode_steps=100)
env.reset()
dev = env.batch.device
res = {}
for mode in ("device", "numpy"):
    her = HerReplayBuffer(n_envs * rows, env, n_sampled_goal=4, max_batch_size=1 << 20, rng=mode, device_seed=11)
    g = torch.Generator(device="cpu").manual_seed(7)
    remaining = torch.randint(10, 101, (n_envs,), generator=g)
    for r in range(rows + 200):  # fill past one wrap
        obs = {"observation": torch.randn(n_envs, 56, device=dev), "achieved_goal": torch.randn(n_envs, 3, device=dev) * 0.3,
               "desired_goal": torch.randn(n_envs, 3, device=dev) * 0.3}
        nobs = {"observation": torch.randn(n_envs, 56, device=dev), "achieved_goal": obs["achieved_goal"] + 0.01, "desired_goal": obs["desired_goal"]}
        remaining -= 1
        done = remaining <= 0
        remaining = torch.where(done, torch.randint(10, 101, (n_envs,), generator=g), remaining)
        her.add(obs, nobs, torch.rand(n_envs, 7, device=dev), torch.zeros(n_envs, device=dev), done.to(torch.uint8), timeouts=np.zeros(n_envs, bool))
    torch.cuda.synchronize()
    for B in (256, 65536, 1 << 20):
        if mode == "numpy" and B > 65536:
            continue
        np.random.seed(0)
        for _ in range(3):
            her.sample(B)
        torch.cuda.synchronize()
        K = 20 if B <= 65536 else 10
        t0 = time.perf_counter()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            her.sample(B)
        e1.record(); torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / K
        ms = e0.elapsed_time(e1) / K
        res["%s_B%d" % (mode, B)] = {"ms_per_sample_call_device": ms, "ms_per_sample_call_wall": wall * 1e3,
                                     "samples_per_s": B / (ms * 1e-3), "GBps_algorithmic": B * 1080 / (ms * 1e-3) / 1e9}
    her.close()
print(json.dumps({"workload": "HER sample, 1 048 576-transition buffer x 1024 envs, k=4 future", "results": res}))
