"""Short run of the numpy-exact HER index kernel for ncu: a wrapped 256-row x 1024-env buffer, a few sample(65536) calls."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from panda_gym_b200.vec_env import PandaVecEnv
from panda_gym_b200.her import HerReplayBuffer
n_envs, rows = 1024, 256
env = PandaVecEnv("PandaReachAO-v3", n_envs=n_envs, seed=1, scenario="reachao3", max_episode_steps=100)
env.reset()
dev = env.batch.device
her = HerReplayBuffer(n_envs * rows, env, n_sampled_goal=4, max_batch_size=1 << 16, rng="numpy")
g = torch.Generator(device="cpu").manual_seed(7)
remaining = torch.randint(10, 101, (n_envs,), generator=g)
obs = {"observation": torch.randn(n_envs, 56, device=dev), "achieved_goal": torch.randn(n_envs, 3, device=dev) * 0.3,
       "desired_goal": torch.randn(n_envs, 3, device=dev) * 0.3}
for r in range(rows + 60):
    remaining -= 1
    done = remaining <= 0
    remaining = torch.where(done, torch.randint(10, 101, (n_envs,), generator=g), remaining)
    her.add(obs, obs, torch.rand(n_envs, 7, device=dev), torch.zeros(n_envs, device=dev), done.to(torch.uint8), timeouts=np.zeros(n_envs, bool))
np.random.seed(0)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    her.sample(65536)
torch.cuda.synchronize()
print("ok")
