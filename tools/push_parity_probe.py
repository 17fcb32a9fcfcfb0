"""Probe: CUDA Push / PickAndPlace path vs the CPU oracle (reset, matched-state steps, rollouts).  Prints error statistics."""
import sys, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
import oracle_lib as O
from panda_gym_b200 import sim

dev = torch.device('cuda', 0)


def soa(x, dt=torch.float32):
    return torch.tensor(np.ascontiguousarray(np.asarray(x).T), dtype=dt, device=dev).contiguous()


for task in (2, 3):
    n = 64
    cfg = sim.make_config(task, n, max_episode_steps=50, auto_reset=False)
    b = sim.PandaBatch(cfg, device=0)
    b.seed(11)
    b.reset()
    torch.cuda.synchronize()
    og = b.view("observation").cpu().numpy(); dg = b.view("desired_goal").cpu().numpy(); ag = b.view("achieved_goal").cpu().numpy()
    envs = [O.Env(task=task) for _ in range(n)]
    w = 0.0
    for i, e in enumerate(envs):
        o, _ = e.reset(11 + i)
        w = max(w, np.abs(o["observation"] - og[i]).max(), np.abs(o["desired_goal"] - dg[i]).max(), np.abs(o["achieved_goal"] - ag[i]).max())
    print("task", task, "obs_dim", b.obs_dim, "act_dim", b.act_dim, "reset max err", w)
    # rollout: ee actions biased downward and toward the object
    rng = np.random.default_rng(5)
    errs = []
    t0 = time.time()
    for s in range(50):
        a = rng.uniform(-1, 1, (n, b.act_dim)).astype(np.float32)
        obj = b.view("object").cpu().numpy().T[:, :3]
        ee = b.view("observation").cpu().numpy()[:, :3]
        d = obj - ee
        d[:, 2] += 0.02
        a[:, :3] = np.clip(0.6 * a[:, :3] + 8.0 * d, -1, 1)
        b.step(torch.from_numpy(a).to(dev))
        torch.cuda.synchronize()
        og = b.view("observation").cpu().numpy(); rg = b.view("reward").cpu().numpy(); fg = b.view("flags").cpu().numpy()
        step_err = np.zeros(n)
        for i, e in enumerate(envs):
            o, r, f = e.step(a[i])
            k = b.obs_dim - 12
            step_err[i] = max(np.abs(o["observation"][:3] - og[i][:3]).max(), np.abs(o["observation"][k:k + 3] - og[i][k:k + 3]).max())
        errs.append(step_err)
    errs = np.array(errs)
    print("  rollout: per-step max POSITION err (ee, object) quantiles (50, 90, 99, 100%%): %s  (%.1fs)" % (np.quantile(errs, [0.5, 0.9, 0.99, 1.0]), time.time() - t0))
    print("  envs within 1e-3 after 50 steps: %d / %d; moved objects: %d" % ((errs[-1] < 1e-3).sum(), n,
          (np.abs(b.view("object").cpu().numpy().T[:, 7:10]).max(axis=1) > 1e-3).sum()))
    worst = np.argmax(errs[-1])
    print("  worst env", worst, "err by step", np.round(errs[:, worst], 5)[::5])
