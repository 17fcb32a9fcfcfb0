import sys
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from panda_gym_b200 import sim
n = 65536
F = [87.0, 87.0, 87.0, 87.0, 12.0, 120.0, 120.0]
neutral = np.array([0, -0.3, 0, -2.2, 0, 2.0, np.pi / 4], np.float32)
def run(label, frac_at_limit):
    Q = np.tile(neutral, (n, 1)); T = Q.copy()
    k = int(frac_at_limit * n)
    idx = np.random.default_rng(0).permutation(n)[:k]
    Q[idx, 3] = 0.0005; T[idx, 3] = 0.3     # joint 4 upper limit is 0.0
    q = torch.tensor(Q.T.copy(), device="cuda").contiguous(); qd = torch.zeros_like(q); t = torch.tensor(T.T.copy(), device="cuda").contiguous()
    sim.debug_substeps(q, qd, t, F, 20); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); sim.debug_substeps(q, qd, t, F, 20); e1.record(); torch.cuda.synchronize()
    print(label, "20 substeps: %.3f ms" % e0.elapsed_time(e1))
run("no env at a limit    ", 0.0)
run("1% of envs at a limit", 0.01)
run("50% at a limit       ", 0.5)
run("all at a limit       ", 1.0)
