import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
import oracle_lib as O
from panda_gym_b200 import sim
dev = torch.device('cuda', 0)
n = 4
nsub = int(sys.argv[1]) if len(sys.argv) > 1 else 1
cfg = sim.make_config(2, n, max_episode_steps=100000, auto_reset=False)
cfg.n_substeps = nsub
b = sim.PandaBatch(cfg, device=0); b.seed(11); b.reset()
envs = [O.Env(task=2, n_substeps=nsub, max_episode_steps=100000) for _ in range(n)]
for i, e in enumerate(envs): e.reset(11 + i)
np.set_printoptions(precision=5, suppress=True, linewidth=200)
for s in range(int(sys.argv[2]) if len(sys.argv) > 2 else 400):
    a = np.tile(np.array([0.3, 0.0, -1.0], np.float32), (n, 1))
    b.step(torch.from_numpy(a).to(dev)); torch.cuda.synchronize()
    og = b.view("observation").cpu().numpy(); qg = b.view("q").cpu().numpy().T
    bad = False
    for i, e in enumerate(envs):
        o, r, f = e.step(a[i])
        err = np.abs(o["observation"] - og[i])
        if err.max() > 3e-3 and not bad:
            bad = True
            print("step", s, "env", i, "err", err.max(), "at", err.argmax())
            print(" oracle", o["observation"]); print(" gpu   ", og[i])
            print(" q oracle", e.arm.q()[0], "\n q gpu   ", qg[i])
            print(" contacts", [(c['kind'], c['link'], round(c['d'], 5)) for c in e.contacts()])
    if bad: break
else:
    print("no mismatch")
