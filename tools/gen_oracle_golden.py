#!/usr/bin/env python3
"""Freeze seeded rollouts of the CPU oracle as a committed fixture (tests/golden/oracle_rollouts.json).

NOT reference output: PyBullet cannot be installed here, so these vectors come from oracle/ (itself pinned
on the reference's known-answer tests).  They (1) guard the oracle against silent drift between rounds -- a changed oracle
changes what "parity" means -- and (2) let the GPU path be checked against committed numbers without building the oracle.

    python tools/gen_oracle_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402

CASES = [
    dict(name="reach_joints", task=0, kw=dict(control=0, max_episode_steps=50), seeds=[0, 1, 2], steps=20),
    dict(name="reach_ee", task=0, kw=dict(control=1, max_episode_steps=50), seeds=[0, 1], steps=20),
    dict(name="reachao3", task=1, kw=dict(scenario="reachao3", max_episode_steps=100), seeds=[10, 11, 12, 13], steps=25),
    dict(name="reachao_rand", task=1, kw=dict(scenario="reachao_rand", max_episode_steps=100), seeds=[20, 21], steps=25),
    dict(name="wangexp3_random_base", task=1, kw=dict(scenario="wangexp-3", randomize_robot_pose=1, max_episode_steps=100), seeds=[30, 31], steps=15),
    dict(name="push", task=2, kw=dict(max_episode_steps=50), seeds=[40, 41], steps=15),
]


def run_case(c):
    out = []
    for seed in c["seeds"]:
        e = O.Env(task=c["task"], **c["kw"])
        o, info = e.reset(seed)
        rng = np.random.default_rng(1000 + seed)
        rec = dict(seed=seed, reset_obs=[float(x) for x in o["observation"]], goal=[float(x) for x in o["desired_goal"]],
                   actions=[], rewards=[], flags=[], obs=[], achieved=[])
        for _ in range(c["steps"]):
            a = rng.uniform(-1, 1, e.act_dim).astype(np.float32)
            o, r, f = e.step(a)
            rec["actions"].append([float(x) for x in a]); rec["rewards"].append(float(r))
            rec["flags"].append([int(f["is_success"]), int(f["terminated"]), int(f["truncated"]), int(f["time_limit"])])
            rec["obs"].append([float(x) for x in o["observation"]]); rec["achieved"].append([float(x) for x in o["achieved_goal"]])
            if f["terminated"] or f["truncated"] or f["time_limit"]:
                break
        out.append(rec)
    return out


if __name__ == "__main__":
    O.build()
    data = {"source": "oracle/ (CPU restatement), NOT the reference; see tools/gen_oracle_golden.py",
            "cases": [dict(name=c["name"], task=c["task"], kw=c["kw"], steps=c["steps"], rollouts=run_case(c)) for c in CASES]}
    path = os.path.join(ROOT, "tests", "golden", "oracle_rollouts.json")
    with open(path, "w") as f:
        json.dump(data, f)
    print("wrote", path, os.path.getsize(path), "bytes")
