"""How well can a key known at the START of a step predict an environment's PGS work in that step?  Variant built with
-DPG_SWEEP_PROBE counts the sweep pairs every environment runs per step; this script compares, for several keys, the lane
efficiency of the solver loop if the 443 environments of a block were sorted by the key before being dealt to warps
(sum of per-warp maxima against the sum of the counts; the step total is a proxy for the per-substep maxima).
PANDA_B200_LIB=panda_gym_b200/variants/libpanda_b200_probe.so python tools/sweep_probe.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from panda_gym_b200 import sim, _lib
lib = _lib.load()
N = 65536
b = sim.PandaBatch(sim.make_config(1, N, scenario="reachao3", max_episode_steps=100, auto_reset=True)); b.seed(1234); b.reset()
for k in range(160):
    b.step(b.random_actions(1234, k))
torch.cuda.synchronize()
buf = np.zeros(N, np.int32)
lib.pg_debug_sweeps_read(buf.ctypes.data_as(ctypes.c_void_p), N)
prev = None
def eff(counts, key):
    tot = 0.0; ideal = 0.0
    for blk in range(0, N, 448):
        c = counts[blk:blk + 448]; k = key[blk:blk + 448]
        c = c[np.argsort(k, kind="stable")]
        pad = (-len(c)) % 32
        c = np.concatenate([c, np.zeros(pad, c.dtype)]).reshape(-1, 32)
        tot += c.max(1).sum() * 32; ideal += c.sum()
    return ideal / tot
for k in range(160, 166):
    a = b.random_actions(1234, k)
    q = b.view("q").clone(); qd = b.view("qd").clone()
    b.step(a); torch.cuda.synchronize()
    lib.pg_debug_sweeps_read(buf.ctypes.data_as(ctypes.c_void_p), N)
    c = buf.astype(np.float64).copy()
    an = a.cpu().numpy(); qn = q.cpu().numpy(); qdn = qd.cpu().numpy()
    if qn.shape[0] == 7: qn = qn.T; qdn = qdn.T
    keys = {"none (env order)": np.arange(N), "oracle (the count itself)": c, "|a|_inf": np.abs(an).max(1), "|a|_1": np.abs(an).sum(1),
            "|qd|_inf at start": np.abs(qdn).max(1), "|0.05 a - qd dt|_1": np.abs(0.05 * an - qdn * 0.04).sum(1)}
    if prev is not None: keys["previous step's count"] = prev
    print("step %d: pairs per env-step mean %.1f min %d max %d (of %d); corr with prev %s" % (
        k, c.mean(), c.min(), c.max(), 25 * 20, "%.2f" % np.corrcoef(c, prev)[0, 1] if prev is not None else "-"))
    for name, key in keys.items():
        print("   sorted by %-28s lane efficiency %.3f   corr %.2f" % (name, eff(c, key), np.corrcoef(c, key)[0, 1] if name != "none (env order)" else 0))
    prev = c
