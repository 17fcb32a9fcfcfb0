import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
import oracle_lib as O
from panda_gym_b200 import sim
import test_gpu_push as T
np.set_printoptions(precision=5, suppress=True, linewidth=220)
n = 16
b, envs, _ = T.make_pair(sim, O, 2, n, seed=3, max_episode_steps=1000)
down = np.tile(np.array([0.0, 0.0, -1.0], np.float32), (n, 1))
for s in range(12):
    b.step(torch.from_numpy(down).to(T.dev()))
    for i, e in enumerate(envs): e.step(down[i])
rng = np.random.default_rng(4)
for i, e in enumerate(envs):
    ee = e.get_obs()["observation"][:3].astype(float)
    e.set_object_state([ee[0] + 0.055 + 0.01 * rng.random(), ee[1] - 0.01 + 0.01 * (rng.random() - 0.5), 0.02])
T.push_state(b, envs)
a = np.tile(np.array([0.35, 0.0, -0.2], np.float32), (n, 1))
for s in range(10):
    b.step(torch.from_numpy(a).to(T.dev())); torch.cuda.synchronize()
    og = b.view("observation").cpu().numpy(); qg = b.view("q").cpu().numpy().T
    eq = []; eo = []
    for i, e in enumerate(envs):
        o, r, f = e.step(a[i])
        eq.append(np.abs(e.arm.q()[0] - qg[i]).max()); eo.append(np.abs(o["observation"][6:9] - og[i][6:9]).max())
    print(s, "q err max %.5f  obj err max %.5f" % (max(eq), max(eo)), "env", int(np.argmax(eq)), "obj v", og[int(np.argmax(eq))][12:15], [(c['kind'], c['link'], round(c['d'], 4)) for c in envs[int(np.argmax(eq))].contacts() if c['kind']])
