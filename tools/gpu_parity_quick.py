import sys, time, ctypes
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch
import oracle_lib as O
from panda_gym_b200 import sim
torch.manual_seed(0)
dev = torch.device('cuda', 0)
# ---- 1. minv parity
n = 256
rng = np.random.default_rng(0)
lo = np.array([-2.7,-1.7,-2.7,-3.0,-2.7,0.1,-2.7]); hi = np.array([2.7,1.7,2.7,-0.1,2.7,3.7,2.7])
Q = rng.uniform(lo, hi, (n, 7))
q = torch.tensor(Q.T.copy(), dtype=torch.float32, device=dev).contiguous()
mi = sim.debug_minv(q).cpu().numpy()
arm = O.Arm(model=0)
worst = 0
for e in range(n):
    arm.set_q(Q[e].astype(np.float32).astype(np.float64), np.zeros(7)); M = arm.minv()
    for i in range(7):
        for j in range(i+1):
            ref = M[i, j]; got = mi[i*(i+1)//2+j, e]
            worst = max(worst, abs(ref-got)/ (abs(M).max()))
print("minv rel err (vs max entry):", worst)
# ---- 2. substeps parity: 20 substeps from random state
QD = rng.uniform(-1, 1, (n, 7))
T = Q + rng.uniform(-0.05, 0.05, (n, 7))
q = torch.tensor(Q.T.copy(), dtype=torch.float32, device=dev).contiguous()
qd = torch.tensor(QD.T.copy(), dtype=torch.float32, device=dev).contiguous()
t = torch.tensor(T.T.copy(), dtype=torch.float32, device=dev).contiguous()
F = [87,87,87,87,12,120,120]
sim.debug_substeps(q, qd, t, F, 20)
torch.cuda.synchronize()
qg = q.cpu().numpy().T; qdg = qd.cpu().numpy().T
eq = 0; ev = 0
for e in range(n):
    arm.set_q(Q[e].astype(np.float32).astype(np.float64), QD[e].astype(np.float32).astype(np.float64))
    for d in range(7): arm.set_motor(d, 1, float(np.float32(T[e, d])), F[d])
    for k in range(20): arm.substep()
    qq, qqd = arm.q()
    eq = max(eq, np.abs(qq - qg[e]).max()); ev = max(ev, np.abs(qqd - qdg[e]).max())
print("20 substeps: max |dq| %.3e rad, max |dqd| %.3e rad/s" % (eq, ev))
# ---- 3. env rollout parity, reachao3
N = 64
cfg = sim.make_config(1, N, scenario="reachao3", max_episode_steps=100, auto_reset=False)
b = sim.PandaBatch(cfg)
b.seed(1000)
b.reset()
torch.cuda.synchronize()
obs0 = b.view("observation").cpu().numpy()
envs = [O.Env(task=1, scenario="reachao3", max_episode_steps=100) for _ in range(N)]
r0 = [envs[i].reset(1000 + i) for i in range(N)]
o_ref = np.stack([r[0]["observation"] for r in r0])
print("reset obs max err", np.abs(o_ref - obs0).max(), "goal err", np.abs(np.stack([e.goal() for e in envs]).T - b.view("goal").cpu().numpy()).max())
act_rng = np.random.default_rng(5)
alive = np.ones(N, bool); maxerr_q = 0; maxerr_ee = 0; flag_mismatch = 0; rew_mismatch = 0; steps_cmp = 0
for step in range(50):
    A = act_rng.uniform(-1, 1, (N, 7)).astype(np.float32)
    b.step(torch.tensor(A, device=dev))
    torch.cuda.synchronize()
    og = b.view("observation").cpu().numpy(); rg = b.view("reward").cpu().numpy(); fg = b.view("flags").cpu().numpy(); dg_ = b.view("done").cpu().numpy()
    for i in range(N):
        if not alive[i]: continue
        o, r, f = envs[i].step(A[i])
        fl = (1 if f["is_success"] else 0) | (2 if f["terminated"] else 0) | (4 if f["truncated"] else 0) | (8 if (f["time_limit"] and not f["terminated"]) else 0)
        steps_cmp += 1
        maxerr_q = max(maxerr_q, np.abs(o["observation"][6:13] - og[i, 6:13]).max())
        maxerr_ee = max(maxerr_ee, np.abs(o["observation"][0:3] - og[i, 0:3]).max())
        if fl != fg[i]: flag_mismatch += 1; print("flag mismatch env", i, "step", step, fl, fg[i], o["observation"][20:29], og[i,20:29])
        if abs(r - rg[i]) > 1e-6: rew_mismatch += 1
        if fl & 14: alive[i] = False
print("rollout: compared", steps_cmp, "env-steps; max |dq| %.3e, max |dee| %.3e, flag mismatches %d, reward mismatches %d, alive at end %d" % (maxerr_q, maxerr_ee, flag_mismatch, rew_mismatch, alive.sum()))
# ---- 4. throughput quick look
for N in [65536]:
    cfg = sim.make_config(1, N, scenario="reachao3", max_episode_steps=100, auto_reset=True)
    bb = sim.PandaBatch(cfg); bb.seed(0); bb.reset()
    acts = bb.random_actions(1, 0)
    for i in range(3): bb.step(acts)
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    K = 10
    for i in range(K): bb.random_actions(1, i+1, acts); bb.step(acts)
    t1.record(); torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / K
    print("N=%d: %.3f ms/step -> %.3e env-steps/s; stats %s" % (N, ms, N / ms * 1e3, bb.view("stats").cpu().numpy()))
