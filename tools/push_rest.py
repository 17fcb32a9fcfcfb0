import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
from panda_gym_b200 import sim
np.set_printoptions(precision=6, suppress=True, linewidth=220)
n = 64
cfg = sim.make_config(2, n, max_episode_steps=1000, auto_reset=False)
b = sim.PandaBatch(cfg, device=0); b.seed(7); b.reset(); torch.cuda.synchronize()
o0 = b.view("object").cpu().numpy().T.copy()
a = torch.zeros((n, 3), device='cuda')
for s in range(50):
    b.step(a)
    if s % 10 == 9:
        torch.cuda.synchronize()
        o = b.view("object").cpu().numpy().T
        d = np.abs(o[:, :3] - o0[:, :3]).max(axis=1)
        w = int(np.argmax(d))
        print(s, "max drift", d.max(), "median", np.median(d), "worst env", w, o[w])
