"""Phase stamps of one steady-state auto-reset launch (variant built with -DPG_RESET_TRACE, see pg_reset_body.cuh):
PANDA_B200_LIB=panda_gym_b200/variants/libpanda_b200_trace.so python tools/trace_reset.py [scenario]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from panda_gym_b200 import sim, _lib
sc = sys.argv[1] if len(sys.argv) > 1 else "reachao3"
lib = _lib.load()
b = sim.PandaBatch(sim.make_config(1, 65536, scenario=sc, max_episode_steps=100, auto_reset=True)); b.seed(1234); b.reset()
for k in range(189):
    b.step(b.random_actions(1234, k))
torch.cuda.synchronize()
for rep in range(3):
    lib.pg_debug_reset_trace_clear()
    b.step(b.random_actions(1234, 189 + rep))
    torch.cuda.synchronize()
    buf = np.zeros((65536, 8), np.uint64)
    n = lib.pg_debug_reset_trace_read(buf.ctypes.data_as(ctypes.c_void_p), 65536)
    t = buf[:n].astype(np.int64)
    span = (t[:, 6].max() - t[:, 0].min()) / 1e3
    dur = t[:, 5] - t[:, 1]
    ph = np.stack([t[:, 2] - t[:, 1], t[:, 3] - t[:, 2], t[:, 4] - t[:, 3], t[:, 5] - t[:, 4]], 1)
    sm = (t[:, 7] >> 8) & 0xffffff
    print("launch %d: %d resets, span %.1f us (globaltimer), start spread %.1f us; cycles per reset: median %d, p90 %d, max %d" % (
        rep, n, span, (t[:, 0].max() - t[:, 0].min()) / 1e3, np.median(dur), np.percentile(dur, 90), dur.max()))
    print("  phases median (goal, obstacles, pose+stores, observation):", np.median(ph, 0).astype(int), " max:", ph.max(0))
    cnt = np.bincount(sm, minlength=148)
    print("  resets per SM: min %d max %d; SMs with none: %d" % (cnt.min(), cnt.max(), (cnt == 0).sum()))
    order = np.argsort(-dur)[:6]
    for i in order:
        print("   slow: env %d sm %d cycles %d phases %s  start +%.1f us end +%.1f us" % (t[i, 7] >> 32, sm[i], dur[i], ph[i], (t[i, 0] - t[:, 0].min()) / 1e3, (t[i, 6] - t[:, 0].min()) / 1e3))
    # when does each SM finish
    end_by_sm = np.array([(t[sm == s, 6].max() - t[:, 0].min()) / 1e3 if (sm == s).any() else 0 for s in range(148)])
    print("  SM finish time us: median %.1f p90 %.1f max %.1f" % (np.median(end_by_sm), np.percentile(end_by_sm, 90), end_by_sm.max()))
np.save("gpurun_out/reset_trace.npy", t)
