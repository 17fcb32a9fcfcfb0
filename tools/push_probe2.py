import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
import oracle_lib as O
from panda_gym_b200 import sim
dev = torch.device('cuda', 0)
n = 48
for mode in ("random", "down", "chase"):
    cfg = sim.make_config(2, n, max_episode_steps=1000, auto_reset=False)
    b = sim.PandaBatch(cfg, device=0); b.seed(11); b.reset()
    envs = [O.Env(task=2, max_episode_steps=1000) for _ in range(n)]
    for i, e in enumerate(envs): e.reset(11 + i)
    rng = np.random.default_rng(5)
    ee_err = []; ob_err = []; q_err = []
    for s in range(50):
        a = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
        if mode == "down": a[:, 2] = -np.abs(a[:, 2])
        if mode == "chase":
            obj = b.view("object").cpu().numpy().T[:, :3]; ee = b.view("observation").cpu().numpy()[:, :3]
            d = obj - ee; d[:, 2] += 0.02
            a = np.clip(0.6 * a + 8.0 * d, -1, 1).astype(np.float32)
        b.step(torch.from_numpy(a).to(dev)); torch.cuda.synchronize()
        og = b.view("observation").cpu().numpy(); qg = b.view("q").cpu().numpy().T
        e1 = np.zeros(n); e2 = np.zeros(n); e3 = np.zeros(n)
        for i, e in enumerate(envs):
            o, r, f = e.step(a[i])
            e1[i] = np.abs(o["observation"][:3] - og[i][:3]).max(); e2[i] = np.abs(o["observation"][6:9] - og[i][6:9]).max()
            e3[i] = np.abs(e.arm.q()[0] - qg[i]).max()
        ee_err.append(e1); ob_err.append(e2); q_err.append(e3)
    ee_err = np.array(ee_err); ob_err = np.array(ob_err); q_err = np.array(q_err)
    qs = [0.5, 0.9, 1.0]
    print(mode, "ee err @10/25/50:", [np.round(np.quantile(ee_err[k], qs), 5) for k in (9, 24, 49)])
    print(mode, "q  err @10/25/50:", [np.round(np.quantile(q_err[k], qs), 5) for k in (9, 24, 49)])
    print(mode, "obj err @10/25/50:", [np.round(np.quantile(ob_err[k], qs), 5) for k in (9, 24, 49)])
    w = int(np.argmax(q_err[-1])); print("  worst env", w, "q err by step", np.round(q_err[::5, w], 5), "ee z", og[w][2], "contacts", [(c['kind'], c['link'], round(c['d'], 4)) for c in envs[w].contacts() if c['kind']])
