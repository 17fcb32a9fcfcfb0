"""Oracle env layer: the reference's own smoke / determinism tests restated (test/envs_test.py, test/seed_test.py)
plus the ReachAO rules of reach_ao.py."""
import numpy as np
import pytest

NEUTRAL = np.array([0, -0.3, 0, -2.2, 0, 2.0, np.pi / 4])


@pytest.mark.parametrize("scenario,obs_dim,nobs", [("reachao1", 56, 1), ("reachao2", 56, 2), ("reachao3", 56, 3),
                                                   ("wangexp-3", 56, 3), ("wang-2", 56, 2), ("reachao_rand", 56, 6),
                                                   ("reach1", 47, 0), ("reach2", 47, 0), ("reach3", 47, 0)])
def test_scenarios_reset(oracle, scenario, obs_dim, nobs):
    e = oracle.Env(task=1, scenario=scenario)
    o, info = e.reset(3)
    assert info["status"] == 0 and not info["is_success"]
    assert e.obs_dim == obs_dim and o["observation"].shape == (obs_dim,) and e.act_dim == 7
    assert len(e.obstacles()) == nobs
    assert np.allclose(o["observation"][6:13], NEUTRAL, atol=1e-7) and np.all(o["observation"][13:20] == 0)
    assert np.allclose(o["achieved_goal"], [0.4816, 0.0100, 0.4375], atol=1e-4)   # COM of link 11
    g = e.goal()
    assert g[2] - 0.05 > 0.1 - 1e-9, "goal probe (r 0.05) keeps 0.1 clearance above the table (reach_ao.py:1101)"
    if nobs == 0:
        assert np.all(o["observation"][20:] == 1.0)  # reach_ao.py:904
    else:
        assert not e.check_collided()[0]


def test_seed_determinism(oracle):
    # reference test/seed_test.py: same seed and actions -> identical observations
    acts = np.random.default_rng(0).uniform(-1, 1, (6, 7)).astype(np.float32)
    outs = []
    for _ in range(2):
        e = oracle.Env(task=1, scenario="reachao3")
        e.reset(12345)
        outs.append(np.stack([e.step(a)[0]["observation"] for a in acts]))
    assert np.array_equal(outs[0], outs[1])


def test_random_rollout_smoke(oracle):
    # reference test/envs_test.py: random steps, nothing numeric
    for task, kw in ((0, {}), (1, dict(scenario="reachao2"))):
        e = oracle.Env(task=task, **kw)
        e.reset(0)
        rng = np.random.default_rng(1)
        for i in range(150):
            o, r, f = e.step(rng.uniform(-1, 1, e.act_dim).astype(np.float32))
            assert np.isfinite(o["observation"]).all()
            if f["terminated"] or f["truncated"] or f["time_limit"]:
                e.reset(i)


def test_reward_and_flags_rules(oracle):
    e = oracle.Env(task=1, scenario="reachao3", max_episode_steps=5)
    e.reset(0)
    # success: place the goal on the end-effector
    ee = e.get_obs()["achieved_goal"].astype(np.float64)
    e.set_goal(ee)
    o, r, f = e.step(np.zeros(7, np.float32))
    assert f["is_success"] and f["terminated"] and r == 0.0
    # collision: put an obstacle into link 5 -> truncated, reward -1 - 100 (reach_ao.py:1317-1320, 1376-1377)
    e.reset(1)
    p = e.arm.link_state(4)[0]
    e.set_obstacles([oracle.make_obstacle("sphere", p, 0.05)])
    o, r, f = e.step(np.zeros(7, np.float32))
    assert f["truncated"] and not f["is_success"] and r == -101.0
    # time limit
    e.reset(2)
    flags = [e.step(np.zeros(7, np.float32))[2] for _ in range(5)]
    assert not flags[3]["time_limit"] and flags[4]["time_limit"]


def test_collision_stops_substeps(oracle):
    """reach_ao.py:186-188: the substep loop breaks at the first colliding substep."""
    e1 = oracle.Env(task=1, scenario="reachao1"); e1.reset(5)
    e2 = oracle.Env(task=1, scenario="reachao1"); e2.reset(5)
    p = e1.arm.link_state(6)[0] + np.array([0, 0, 0.12])
    e1.set_obstacles([oracle.make_obstacle("sphere", p, 0.04)])
    e2.set_obstacles([oracle.make_obstacle("sphere", p + 5.0, 0.04)])
    a = np.array([0, -1, 0, 1, 0, -1, 0], np.float32)
    moved1 = moved2 = 0
    for _ in range(6):
        q0 = e1.arm.q()[0].copy(); o1, r1, f1 = e1.step(a); moved1 = np.abs(e1.arm.q()[0] - q0).max()
        q0 = e2.arm.q()[0].copy(); o2, r2, f2 = e2.step(a); moved2 = np.abs(e2.arm.q()[0] - q0).max()
        if f1["truncated"]:
            break
    assert f1["truncated"] and not f2["truncated"]
    assert moved1 < moved2


def test_ee_control_uses_ik(oracle):
    e = oracle.Env(task=0, control=1)
    assert e.act_dim == 3
    o, _ = e.reset(0)
    z0 = o["achieved_goal"][2]
    for _ in range(10):
        o, r, f = e.step(np.array([0, 0, 1.0], np.float32))
    assert o["achieved_goal"][2] > z0 + 0.1


def test_random_robot_base_reset(oracle):
    """N2 (reach_ao.py:1238-1241, :1035-1060): with randomize_robot_pose the wangexp scenario draws joint 1 first from
    the env's numpy stream, keeps the other joints neutral and ends collision-free with 0.05 margin to every obstacle."""
    import numpy as np
    e = oracle.Env(task=1, scenario="wangexp-3", randomize_robot_pose=1, max_episode_steps=100)
    moved = 0
    for seed in range(40):
        o, info = e.reset(seed)
        q = e.arm.q()[0]
        first = np.random.default_rng(seed).uniform(-2.7437, 2.7437)
        assert np.allclose(q[1:7], [-0.3, 0.0, -2.2, 0.0, 2.0, np.pi / 4])
        assert -2.7437 <= q[0] <= 2.7437
        moved += int(q[0] != first)         # set_coll_free_robot re-drew the pose
        if q[0] == first:
            assert info["status"] == 0
        for _, p, size in e.obstacles():    # no obstacle within 0.05 of the robot
            for link in range(0, 12):
                d, _, _ = e.arm.link_obstacle_distance(link, oracle.make_obstacle(0, p, size))
                assert d > 0.05 - 1e-12
        # the observation reports the turned pose: q0 sits right after ee position / velocity (panda.py:264-288)
        assert abs(o["observation"][6] - np.float32(q[0])) < 1e-6
    assert moved < 40  # most seeds keep their first draw
    plain = oracle.Env(task=1, scenario="wangexp-3", max_episode_steps=100)
    plain.reset(3)
    assert plain.arm.q()[0][0] == 0.0      # flag off: Panda.reset to the neutral pose


# ------------------------------------------------------------------------------------------ N2 / a16: the numpy and libc pieces
def test_crand_is_glibc_rand(oracle):
    """roboticstoolbox's ik_lm_chan draws restart configurations from the C library's rand(), never seeded: the oracle's
    pgo_crand must be glibc's srand(1) stream (checked against the libc of this machine)."""
    import ctypes
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(1)
    want = [libc.rand() for _ in range(2000)]

    class CRand(ctypes.Structure):
        _fields_ = [("r", ctypes.c_uint32 * 31), ("f", ctypes.c_int), ("b", ctypes.c_int)]
    L = oracle.lib()
    L.pgo_crand_seed.argtypes = [ctypes.c_void_p, ctypes.c_uint32]; L.pgo_crand_next.argtypes = [ctypes.c_void_p]
    c = CRand()
    L.pgo_crand_seed(ctypes.byref(c), 1)
    assert [L.pgo_crand_next(ctypes.byref(c)) for _ in range(2000)] == want


def test_dirichlet_is_numpys(oracle):
    """Generator.dirichlet([1, 1, 1]) and standard_exponential (256-layer ziggurat, tables of tools/gen_ziggurat.py) bit for bit."""
    import ctypes
    L = oracle.lib()
    for seed in (0, 77, 2 ** 40 + 5):
        g = oracle.Pcg64(); L.pgo_pcg64_seed(ctypes.byref(g), seed)
        ref = np.random.default_rng(seed)
        for _ in range(3000):
            out = np.zeros(3)
            L.pgo_pcg64_dirichlet111(ctypes.byref(g), oracle.dp(out))
            assert np.array_equal(out, ref.dirichlet(alpha=[1, 1, 1], size=1)[0])
        want = ref.standard_exponential(20000)
        got = np.array([L.pgo_pcg64_standard_exponential(ctypes.byref(g)) for _ in range(20000)])
        assert np.array_equal(got, want)


def test_rtb_ik_reaches_its_target(oracle):
    """Panda.rtb_ik (panda.py:220-224): the Levenberg-Marquardt solution satisfies ik_lm_chan's own stopping rule (E < 0.01 with
    the neutral orientation as target), lies inside the toolbox model's joint limits, and most targets need no restart."""
    import ctypes

    class CRand(ctypes.Structure):
        _fields_ = [("r", ctypes.c_uint32 * 31), ("f", ctypes.c_int), ("b", ctypes.c_int)]
    L = oracle.lib()
    L.pgo_crand_seed.argtypes = [ctypes.c_void_p, ctypes.c_uint32]
    c = CRand(); L.pgo_crand_seed(ctypes.byref(c), 1)
    arm = oracle.Arm(model=0)
    neutral = np.array([0.0, -0.3, 0.0, -2.2, 0.0, 2.0, np.pi / 4])
    arm.set_q(np.concatenate([neutral, np.zeros(0)])[:7], np.zeros(7))
    R0, _ = arm.link_frame(7)
    lo = np.array([-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973]); hi = np.array([2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973])
    rng = np.random.default_rng(0)
    codes = []
    for _ in range(300):
        phi = rng.uniform(0, 2 * np.pi); th = rng.uniform(0, 0.5 * np.pi); r = np.cbrt(rng.uniform(0.5 ** 3, 0.8 ** 3))
        t = np.array([r * np.sin(th) * np.cos(phi), r * np.sin(th) * np.sin(phi), r * np.cos(th)])
        q = np.zeros(7); its = ctypes.c_int(0)
        rc = L.pgo_arm_ik_lm_chan(arm.h, oracle.dp(t), oracle.dp(neutral), 100, 100, 0.01, 1.0, ctypes.byref(c), oracle.dp(q), ctypes.byref(its))
        codes.append(rc)
        if rc == 2:
            continue
        assert (q >= lo).all() and (q <= hi).all()
        arm.set_q(q, np.zeros(7))
        R, p = arm.link_frame(7)
        Rerr = R0 @ R.T
        ang = np.arccos(np.clip((np.trace(Rerr) - 1) / 2, -1, 1))
        E = 0.5 * (np.sum((t - p) ** 2) + ang ** 2)
        assert E < 0.01 + 1e-9
    assert codes.count(0) > 200 and codes.count(2) < 10


def test_pose_randomizers_and_random_sizes(oracle):
    """Each scenario's robot_pose_randomizer moves the arm; set_robot_random_pose keeps the EE height inside [0.4, 0.6];
    reachao_rand_shape re-draws cuboid half extents 0.2 * Dirichlet(1, 1, 1) = numpy's own rows."""
    for scn in ("reach1", "reach3", "wang-2"):
        e = oracle.Env(task=1, scenario=scn, randomize_robot_pose=1, max_episode_steps=100)
        qs = []
        for s in range(20):
            o, info = e.reset(50 + s)
            assert info["status"] == 0 and 0.4 <= o["achieved_goal"][2] <= 0.6
            qs.append(e.arm.q()[0][:7].copy())
        assert np.std(qs, axis=0).max() > 0.1
    e = oracle.Env(task=1, scenario="reachao_rand_shape", max_episode_steps=100)
    for s in range(10):
        e.reset(s)
        g = np.random.default_rng(s)
        for t, p, sz in e.obstacles():
            if t == 1:
                assert np.array_equal(sz, 0.2 * g.dirichlet(alpha=[1, 1, 1], size=1)[0])
