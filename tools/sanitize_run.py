"""Small workload touching every kernel family, meant to run under compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python tools/sanitize_run.py
k_step (block barriers + shared-memory s_q / s_tgt), k_reset<true> (32 lanes of a warp on one environment's state),
k_step_obj (contact store in dynamic shared memory), k_her_indices_exact (shared MT19937 state), k_her_gather (shared staging), k_render."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from panda_gym_b200.her import HerReplayBuffer  # noqa: E402
from panda_gym_b200.vec_env import PandaVecEnv  # noqa: E402

n = int(os.environ.get("SAN_ENVS", "512"))
rng = np.random.default_rng(0)
for scenario in ("reachao_rand", "wangexp-3"):
    env = PandaVecEnv("PandaReachAO-v3", n_envs=n, seed=1, scenario=scenario, max_episode_steps=4)
    obs = env.reset()
    her = HerReplayBuffer(n * 8, env, max_batch_size=1024, rng="numpy")
    for k in range(6):
        a = rng.uniform(-1, 1, (n, 7)).astype(np.float32)
        nobs, rew, dones, infos = env.step(a)
        her.add(obs, nobs, a, rew, dones, infos=infos)
        obs = nobs
    np.random.seed(0)
    s = her.sample(1000)
    her.rng_mode = "device"
    s = her.sample(1024)
    torch.cuda.synchronize()
    print(scenario, "ok", float(s.rewards.sum()), env.episode_stats()["episodes"])
    her.close(); env.close()
env = PandaVecEnv("PandaPush-v3", n_envs=128, seed=2, max_episode_steps=3)
env.reset()
for k in range(4):
    env.step(rng.uniform(-1, 1, (128, 3)).astype(np.float32))
torch.cuda.synchronize()
img = env.render(width=64, height=48)
torch.cuda.synchronize()
print("push ok", env.episode_stats()["episodes"], "render", img.shape)
env.close()
