"""BASELINE configs[3]: PandaPush-v3 / PandaPickAndPlace-v3 with ee control via IK, 16 384 envs per GPU, TimeLimit 50,
random actions, auto-reset.  Prints one JSON line per task (env-steps/s on one GPU, CUDA-event timed, L2 flushed
between steps) and the oracle's CPU rate on the host threads next to it."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from panda_gym_b200 import sim  # noqa: E402


def cpu_rate(task, seconds=6.0):
    import ctypes
    import threading
    import oracle_lib as O
    L = O.lib()
    cfg = O.EnvConfig()
    L.pgo_env_default_config(ctypes.byref(cfg), task)
    cfg.max_episode_steps = 50
    nt = os.cpu_count() or 1
    steps = 400
    t0 = time.perf_counter()
    th = [threading.Thread(target=lambda i=i: L.pgo_bench_rollout(ctypes.byref(cfg), 1000 + i, steps, None)) for i in range(nt)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    return steps * nt / dt, nt, dt


def main():
    n = int(os.environ.get("PUSH_ENVS", 16384))
    steps, warmup = int(os.environ.get("PUSH_STEPS", 60)), 10
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for task, name in ((2, "PandaPush-v3"), (3, "PandaPickAndPlace-v3")):
        cfg = sim.make_config(task, n, max_episode_steps=50, auto_reset=True)
        b = sim.PandaBatch(cfg, device=0)
        b.seed(1234)
        b.reset()
        acts = [b.random_actions(99, k) for k in range(steps + warmup)]
        for k in range(warmup):
            b.step(acts[k])
        torch.cuda.synchronize()
        ms = []
        for k in range(steps):
            flush.fill_(k & 255)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            b.step(acts[warmup + k])
            e1.record()
            e1.synchronize()
            ms.append(e0.elapsed_time(e1))
        st = b.view("stats").cpu().numpy()
        ob = b.view("object").cpu().numpy().T
        cpu, nt, dt = cpu_rate(task)
        print(json.dumps({
            "metric": "env-steps/sec (one GPU)", "workload": "BASELINE configs[3]: %s, ee control via IK, %d envs, TimeLimit 50, U(-1,1) actions, auto-reset" % (name, n),
            "value": n / (np.mean(ms) * 1e-3), "ms_per_step": float(np.mean(ms)), "ms_min": float(np.min(ms)), "ms_max": float(np.max(ms)),
            "steps": steps, "episodes": float(st[0]), "successes": float(st[1]), "objects_off_table": int((ob[:, 2] < -0.1).sum()),
            "cpu_baseline": {"value": cpu, "unit": "env-steps/s", "cores": nt, "kind": "port", "sample": "%d envs x 400 steps, %.1f s" % (nt, dt)}}))
        b.close()


if __name__ == "__main__":
    main()
