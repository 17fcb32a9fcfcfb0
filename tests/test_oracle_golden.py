"""Oracle vs golden vectors generated from the importable part of the reference (tools/gen_golden.py ->
tests/golden/utils_vectors.json) and vs numpy itself (the RNG the reference's resets use)."""
import ctypes
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "utils_vectors.json")


def test_distance_bit_exact(oracle):
    g = json.load(open(GOLDEN))
    a = np.array(g["a"], np.float32); b = np.array(g["b"], np.float32)
    out = np.zeros(len(a), np.float32)
    oracle.lib().pgo_test_distance(oracle.dp(a), oracle.dp(b), len(a), oracle.dp(out))
    assert np.array_equal(out.view(np.uint32), np.array(g["distance_bits"], np.uint32))
    assert np.array_equal((out < np.float32(0.05)).astype(int), np.array(g["success_reach_ao"]))


def test_compute_reward_matches_reference_formula(oracle):
    g = json.load(open(GOLDEN))
    a = np.array(g["a"], np.float32); b = np.array(g["b"], np.float32)
    classic = oracle.Env(task=0)
    assert np.array_equal(classic.compute_reward(a, b), np.array(g["sparse_reward_classic"], np.float32))
    ao = oracle.Env(task=1, scenario="reachao3"); ao.reset(0)
    # reach_ao.py:1317-1320 with is_collided False: -1 + (d < thr)
    assert np.array_equal(ao.compute_reward(a, b), -1.0 + np.array(g["success_reach_ao"], np.float32))


def test_pcg64_stream_matches_numpy(oracle):
    L = oracle.lib()
    for seed in [0, 1, 12345, 2**32 + 5, 2**63 + 11]:
        g = oracle.Pcg64()
        L.pgo_pcg64_seed(ctypes.byref(g), seed)
        mine = np.array([L.pgo_pcg64_double(ctypes.byref(g)) for _ in range(16)])
        assert np.array_equal(mine, np.random.default_rng(seed).random(16))
        g2 = oracle.Pcg64(); L.pgo_pcg64_seed(ctypes.byref(g2), seed)
        u = np.array([L.pgo_pcg64_uniform(ctypes.byref(g2), -0.3, 0.7) for _ in range(8)])
        assert np.array_equal(u, np.random.default_rng(seed).uniform(-0.3, 0.7, 8))


def test_integers_and_shuffle_match_numpy(oracle):
    L = oracle.lib()
    for seed in range(100):
        for skip in (0, 3):
            perm = np.zeros(6, np.int32)
            k = L.pgo_test_integers_shuffle(seed, skip, 4, 6, 6, oracle.dp(perm))
            g = np.random.default_rng(seed)
            if skip:
                g.random(skip)
            k2 = g.integers(4, 6); lst = list(range(6)); g.shuffle(lst)
            assert k == k2 and list(perm) == lst


def test_reach_goal_is_numpy_uniform(oracle):
    # reach.py:75-78: goal = np_random.uniform(goal_range_low, goal_range_high) with np_random = default_rng(seed)
    for seed in (0, 7, 99):
        e = oracle.Env(task=0)
        e.reset(seed)
        ref = np.random.default_rng(seed).uniform([-0.15, -0.15, 0.0], [0.15, 0.15, 0.3])
        assert np.array_equal(e.goal(), ref)


def test_hollow_sphere_sampler_matches_numpy_restatement(oracle):
    # reach_ao.py:1188-1211 restated with numpy on the same stream (reach2: upper half, three-quarter front)
    def sample(g, rmin, rmax, upper, front, q34):
        phi = g.uniform(0, 2 * np.pi)
        theta = g.uniform(0, 0.5 * np.pi) if upper else g.uniform(0, np.pi)
        if front:
            phi = g.uniform(-0.5 * np.pi, 0.5 * np.pi)
        if q34:
            phi = g.uniform(-0.75 * np.pi, 0.75 * np.pi)
        r = np.cbrt(g.uniform(rmin ** 3, rmax ** 3))
        return np.array([r * np.sin(theta) * np.cos(phi), r * np.sin(theta) * np.sin(phi), r * np.cos(theta)])

    hits = 0
    for seed in range(40):
        e = oracle.Env(task=1, scenario="reach2")
        e.reset(seed)
        g = np.random.default_rng(seed)
        first = sample(g, 0.5, 0.85, True, False, True)
        # reach2 has no obstacle randomisation: the goal is drawn twice (margin 0.1 then 0.03, reach_ao.py:981-992);
        # when both first draws are accepted the final goal is the second draw
        second = sample(g, 0.5, 0.85, True, False, True)
        if np.allclose(e.goal(), second, atol=1e-12):
            hits += 1
        else:  # a rejection happened: the goal must still lie in the shell
            r = np.linalg.norm(e.goal()); assert 0.5 - 1e-9 <= r <= 0.85 + 1e-9 and e.goal()[2] >= 0
        assert first.shape == (3,)
    assert hits > 10


def test_oracle_reproduces_committed_rollouts(oracle):
    """Drift guard: tests/golden/oracle_rollouts.json (tools/gen_oracle_golden.py) froze seeded oracle rollouts; a changed
    oracle would silently change what GPU 'parity' means."""
    import json
    import os
    import numpy as np
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "oracle_rollouts.json")
    data = json.load(open(path))
    n = 0
    for c in data["cases"]:
        for rec in c["rollouts"]:
            e = oracle.Env(task=c["task"], **c["kw"])
            o, _ = e.reset(rec["seed"])
            assert np.allclose(o["observation"], rec["reset_obs"], rtol=0, atol=1e-6), c["name"]
            assert np.allclose(o["desired_goal"], rec["goal"], rtol=0, atol=1e-7)
            for a, r, f, ob in zip(rec["actions"], rec["rewards"], rec["flags"], rec["obs"]):
                o, rr, ff = e.step(np.asarray(a, np.float32))
                assert rr == r and [int(ff["is_success"]), int(ff["terminated"]), int(ff["truncated"]), int(ff["time_limit"])] == f
                assert np.allclose(o["observation"], ob, rtol=0, atol=1e-6), c["name"]
                n += 1
    assert n > 200
