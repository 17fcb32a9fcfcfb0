"""Short run of the Push step path for ncu."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from panda_gym_b200 import sim
K = int(sys.argv[1]) if len(sys.argv) > 1 else 8
cfg = sim.make_config(2, 16384, max_episode_steps=50, auto_reset=True)
b = sim.PandaBatch(cfg); b.seed(1234); b.reset()
acts = [b.random_actions(1234, k) for k in range(K)]
for k in range(K):
    b.step(acts[k])
torch.cuda.synchronize()
print("ok", b.view("stats").cpu().numpy())
