import sys
sys.path.insert(0, '/root/repo')
import torch
from panda_gym_b200 import sim
cfg = sim.make_config(1, 65536, scenario="reachao3", max_episode_steps=100, auto_reset=True)
b = sim.PandaBatch(cfg); b.seed(1); b.reset()
a = b.random_actions(5, 0)
for k in range(45): b.step(a)
torch.cuda.synchronize(); print("ok")
