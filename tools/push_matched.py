"""Matched-state probe: every step starts from the ORACLE's state (rounded to fp32) on both sides; one env step; compare."""
import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
import oracle_lib as O
from panda_gym_b200 import sim
dev = torch.device('cuda', 0)
np.set_printoptions(precision=5, suppress=True, linewidth=220)
n = 48
nsub = int(sys.argv[1]) if len(sys.argv) > 1 else 20
cfg = sim.make_config(2, n, max_episode_steps=1000, auto_reset=False); cfg.n_substeps = nsub
b = sim.PandaBatch(cfg, device=0); b.seed(11); b.reset()
envs = [O.Env(task=2, max_episode_steps=1000, n_substeps=nsub) for _ in range(n)]
for i, e in enumerate(envs): e.reset(11 + i)
rng = np.random.default_rng(5)
worst = (0, None)
all_q = []; all_o = []
for s in range(60 * 20 // nsub):
    Q = np.zeros((n, 7), np.float32); QD = np.zeros((n, 7), np.float32); OB = np.zeros((n, 13), np.float32)
    for i, e in enumerate(envs):
        q, qd = e.arm.q(); p, qt, v, w = e.object_state()
        Q[i] = q; QD[i] = qd; OB[i] = np.concatenate([p, qt, v, w])
        e.arm.set_q(Q[i].astype(np.float64), QD[i].astype(np.float64))
        e.set_object_state(OB[i, :3].astype(np.float64), OB[i, 3:7].astype(np.float64), OB[i, 7:10].astype(np.float64), OB[i, 10:13].astype(np.float64))
    b.set_joint_state(torch.tensor(Q.T.copy()), torch.tensor(QD.T.copy())); b.set_object_state(torch.tensor(OB.T.copy())); b.observe()
    ee = np.array([e.get_obs()["observation"][:3] for e in envs])
    a = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    d = OB[:, :3] - ee; d[:, 2] += 0.02
    a = np.clip(0.6 * a + 8.0 * d, -1, 1).astype(np.float32)
    pre = [e.contacts() for e in envs]
    b.step(torch.from_numpy(a).to(dev)); torch.cuda.synchronize()
    og = b.view("observation").cpu().numpy(); qg = b.view("q").cpu().numpy().T; obg = b.view("object").cpu().numpy().T
    for i, e in enumerate(envs):
        o, r, f = e.step(a[i])
        p, qt, v, w = e.object_state()
        err = max(np.abs(e.arm.q()[0] - qg[i]).max(), np.abs(p - obg[i, :3]).max())
        all_q.append(np.abs(e.arm.q()[0] - qg[i]).max()); all_o.append(np.abs(p - obg[i, :3]).max())
        err = np.abs(e.arm.q()[0] - qg[i]).max()
        if err > worst[0]:
            worst = (err, dict(step=s, env=i, q0=Q[i], qd0=QD[i], ob0=OB[i], a=a[i], q_or=e.arm.q()[0], q_gpu=qg[i], ob_or=np.concatenate([p, qt, v, w]), ob_gpu=obg[i],
                               pre=[(c['kind'], c['link'], round(c['d'], 5)) for c in pre[i]]))
print("q err quantiles 50/90/99/99.9/100", np.quantile(all_q, [.5,.9,.99,.999,1]))
print("obj err quantiles 50/90/99/99.9/100", np.quantile(all_o, [.5,.9,.99,.999,1]))
print("worst single-step err", worst[0])
for k, v in worst[1].items(): print(" ", k, v)
