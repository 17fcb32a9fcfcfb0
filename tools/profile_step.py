"""Short single-GPU run of the step path for ncu: W warm-up + K timed steps of BASELINE configs[1]."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from panda_gym_b200 import sim

scenario = sys.argv[1] if len(sys.argv) > 1 else "reachao3"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 6
cfg = sim.make_config(1, 65536, scenario=scenario, max_episode_steps=100, auto_reset=True)
b = sim.PandaBatch(cfg); b.seed(1234); b.reset()
acts = [b.random_actions(1234, k) for k in range(K)]
for k in range(K):
    b.step(acts[k])
torch.cuda.synchronize()
print("ok", b.view("stats").cpu().numpy())
