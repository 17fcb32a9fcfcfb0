"""How much of the idle-lane loss of k_step's PGS loop could a re-ordering of a block's environments recover?  A variant built
with -DPG_SWEEP_PROBE counts the sweep pairs every environment runs in every substep; this script computes the lane efficiency
of the loop (sum of counts / 32 x sum over substeps of per-warp maxima) for several orderings of the 448 threads of a block:
keys known at the start of the step, the previous step's work, and the work of the step's own first substeps (what a
re-binning inside the step could use).
PANDA_B200_LIB=panda_gym_b200/variants/libpanda_b200_probe.so python tools/sweep_probe.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from panda_gym_b200 import sim, _lib
lib = _lib.load()
N, T = 65536, 148 * 512
b = sim.PandaBatch(sim.make_config(1, N, scenario="reachao3", max_episode_steps=100, auto_reset=True)); b.seed(1234); b.reset()
for k in range(160):
    b.step(b.random_actions(1234, k))
torch.cuda.synchronize()
buf = np.zeros((20, T), np.int32)
lib.pg_debug_sweeps_read(buf.ctypes.data_as(ctypes.c_void_p))
def eff(c, key, first=0):
    """c [20][N]; order by key, efficiency over substeps first..19"""
    tot = 0.0; ideal = 0.0
    for blk in range(0, N, 448):
        o = np.argsort(key[blk:blk + 448], kind="stable")
        cc = c[first:, blk:blk + 448][:, o]
        pad = (-cc.shape[1]) % 32
        cc = np.concatenate([cc, np.zeros((cc.shape[0], pad), cc.dtype)], 1).reshape(cc.shape[0], -1, 32)
        tot += cc.max(2).sum() * 32; ideal += cc.sum()
    return ideal / tot
prev = None
for k in range(160, 164):
    a = b.random_actions(1234, k)
    b.step(a); torch.cuda.synchronize()
    lib.pg_debug_sweeps_read(buf.ctypes.data_as(ctypes.c_void_p))
    c = buf[:, :N].astype(np.float64).copy()
    tot = c.sum(0)
    print("step %d: pairs per env-substep mean %.1f of 25; per substep: %s" % (k, c.mean(), " ".join("%.0f" % x for x in c.mean(1))))
    print("   share of env-substeps at the cap (25 pairs): %.2f; below 5 pairs: %.2f" % ((c >= 25).mean(), (c < 5).mean()))
    keys = [("environment index (as built)", np.arange(N), 0), ("the step's total (unattainable)", tot, 0), ("|a|_inf", np.abs(a.cpu().numpy()).max(1), 0)]
    if prev is not None: keys.append(("previous step's total", prev, 0))
    for s in (1, 2, 4):
        keys.append(("work of substeps 0..%d, applied from substep %d on" % (s - 1, s), c[:s].sum(0), s))
        keys.append(("   (env order over the same substeps)", np.arange(N), s))
    for name, key, first in keys:
        print("   %-52s lane efficiency %.3f" % (name, eff(c, key, first)))
    prev = tot
