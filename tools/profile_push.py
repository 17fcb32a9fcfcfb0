"""Short run of the Push step path for ncu: K steps from reset (the gripper has sunk onto the table by step ~15)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from panda_gym_b200 import sim
K = int(sys.argv[1]) if len(sys.argv) > 1 else 8
cfg = sim.make_config(2, 16384, max_episode_steps=50, auto_reset=True)
b = sim.PandaBatch(cfg); b.seed(1234); b.reset()
acts = [b.random_actions(1234, k) for k in range(K)]
ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
ev[0].record()
for k in range(K):
    b.step(acts[k]); ev[k + 1].record()
torch.cuda.synchronize()
print("ms per step:", " ".join("%.2f" % ev[k].elapsed_time(ev[k + 1]) for k in range(K)))
print("ok", b.view("stats").cpu().numpy())
