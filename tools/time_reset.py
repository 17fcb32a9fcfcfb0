"""Per-launch device time of the auto-reset kernel at steady state: run under
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_reset -s 130 -c 40 --csv ... python tools/time_reset.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from panda_gym_b200 import sim
sc = sys.argv[1] if len(sys.argv) > 1 else "reachao3"
b = sim.PandaBatch(sim.make_config(1, 65536, scenario=sc, max_episode_steps=100, auto_reset=True)); b.seed(1234); b.reset()
for k in range(190):
    b.step(b.random_actions(1234, k))
torch.cuda.synchronize()
print("ok")
