import sys, time, ctypes
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from panda_gym_b200 import sim
def run(label, constant):
    cfg = sim.make_config(1, 65536, scenario="reachao3", max_episode_steps=100, auto_reset=True)
    b = sim.PandaBatch(cfg); b.seed(1); b.reset()
    acts = [b.random_actions(5, k) for k in range(64)]
    ev = []
    for k in range(60):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); b.step(acts[0] if constant else acts[k]); e1.record(); ev.append((e0, e1))
    torch.cuda.synchronize()
    ms = [a.elapsed_time(c) for a, c in ev]
    st = b.view("stats").cpu().numpy()
    q = b.view("q").cpu().numpy()
    lo = np.array([-2.9671, -1.8326, -2.9671, -3.1416, -2.9671, -0.0873, -2.9671])[:, None]; hi = np.array([2.9671, 1.8326, 2.9671, 0.0, 2.9671, 3.8223, 2.9671])[:, None]
    at_limit = ((q <= lo) | (q >= hi)).any(0).mean()
    print(label, "ms/step first10 %.2f mid %.2f last10 %.2f | episodes %d collisions %d | envs at a joint limit now: %.1f%%" % (
        np.mean(ms[:10]), np.mean(ms[25:35]), np.mean(ms[-10:]), st[0], st[2], 100 * at_limit))
    print("   per-step ms:", " ".join("%.1f" % m for m in ms[::3]))
    b.close()
run("random  ", False)
run("constant", True)
