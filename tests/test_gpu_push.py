"""GPU parity tests for Push / PickAndPlace: k_step_obj through the C ABI against oracle/pgo_world.c.

Contact dynamics amplify rounding differences (fp32 kernel vs fp64 oracle): a violent event -- the arm slamming the
box into the table -- is chaotic in the oracle itself (a 1e-6 rad perturbation changes the outcome by 10 %).  Parity
is therefore asserted (a) exactly for the seeded resets, the flags and the rewards at matched states, (b) within the
north-star tolerance (1e-3 rad / 1 mm) for contact-free motion, rest, sliding and a well-conditioned push, and (c) as
quantile bounds over single steps started from matched states for aggressive random pushing.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sim():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from panda_gym_b200 import sim as s
    return s


def dev():
    return torch.device("cuda", 0)


def soa(x):
    return torch.tensor(np.ascontiguousarray(np.asarray(x, dtype=np.float32).T), dtype=torch.float32, device=dev()).contiguous()


def make_pair(sim, oracle, task, n, seed, **kw):
    cfg = sim.make_config(task, n, max_episode_steps=kw.get("max_episode_steps", 50), auto_reset=False,
                          control_type=kw.get("control_type"), reward_type=kw.get("reward_type"))
    b = sim.PandaBatch(cfg, device=0)
    b.seed(seed)
    b.reset()
    okw = dict(max_episode_steps=kw.get("max_episode_steps", 50))
    if kw.get("control_type") in ("joints", "js"):
        okw["control"] = 0
    if kw.get("reward_type") == "dense":
        okw["reward_type"] = 1
    envs = [oracle.Env(task=task, **okw) for _ in range(n)]
    first = [e.reset(seed + i)[0] for i, e in enumerate(envs)]
    torch.cuda.synchronize()
    return b, envs, first


def push_state(b, envs):
    """Copy every oracle state (rounded to fp32) to both sides: the matched state a step starts from."""
    n = len(envs)
    Q = np.zeros((n, 7), np.float32); QD = np.zeros((n, 7), np.float32); OB = np.zeros((n, 13), np.float32)
    for i, e in enumerate(envs):
        q, qd = e.arm.q()
        p, qt, v, w = e.object_state()
        Q[i] = q; QD[i] = qd; OB[i] = np.concatenate([p, qt, v, w])
        e.arm.set_q(Q[i].astype(np.float64), QD[i].astype(np.float64))
        e.set_object_state(OB[i, :3].astype(np.float64), OB[i, 3:7].astype(np.float64), OB[i, 7:10].astype(np.float64),
                           OB[i, 10:13].astype(np.float64))
    b.set_joint_state(soa(Q), soa(QD))
    b.set_object_state(soa(OB))
    b.observe()
    return Q, QD, OB


@pytest.mark.parametrize("task", [2, 3])
def test_seeded_reset_is_identical(sim, oracle, task):
    n = 256
    b, envs, first = make_pair(sim, oracle, task, n, seed=123)
    og = b.view("observation").cpu().numpy(); ag = b.view("achieved_goal").cpu().numpy(); dg = b.view("desired_goal").cpu().numpy()
    assert b.obs_dim == (18 if task == 2 else 19) and b.act_dim == (3 if task == 2 else 4)
    for i in range(n):
        assert np.array_equal(first[i]["desired_goal"], dg[i])
        assert np.array_equal(first[i]["achieved_goal"], ag[i])
        assert np.abs(first[i]["observation"] - og[i]).max() < 2e-6
    if task == 3:
        assert np.all(og[:, 6] == 0.0)              # fingers width of the fixed-finger URDF
        assert (dg[:, 2] > 0.021).mean() > 0.5      # lifted goals (pick_and_place.py:72-77)


@pytest.mark.parametrize("task,control,steps", [(2, "ee", 15), (2, "joints", 50), (3, "ee", 15), (3, "joints", 50)])
def test_contact_free_rollout_within_tolerance(sim, oracle, task, control, steps):
    """Small random actions: the arm never touches anything, the box rests.  1e-3 rad / 1 mm.
    (ee control sinks ~2 cm per step by itself -- the IK drives the link-11 FRAME to where its COM was, a quirk of
    the reference's calculateInverseKinematics call restated as is -- so those runs are kept short and biased upward.)"""
    n = 64
    b, envs, _ = make_pair(sim, oracle, task, n, seed=7, control_type=control)
    assert b.act_dim == (3 if control == "ee" else 7) + (1 if task == 3 else 0)
    rng = np.random.default_rng(2)
    for s in range(steps):
        a = rng.uniform(-1, 1, (n, b.act_dim)).astype(np.float32)
        a *= 0.3   # a random walk around the neutral pose: stays clear of the table and the box
        if control == "ee":
            a[:, 2] = 0.4
        b.step(torch.from_numpy(a).to(dev()))
        torch.cuda.synchronize()
        og = b.view("observation").cpu().numpy(); qg = b.view("q").cpu().numpy().T
        rg = b.view("reward").cpu().numpy(); fg = b.view("flags").cpu().numpy()
        for i, e in enumerate(envs):
            o, r, f = e.step(a[i])
            assert np.abs(e.arm.q()[0] - qg[i]).max() < 1e-3
            assert np.abs(o["observation"][:3] - og[i][:3]).max() < 1e-3
            k = b.obs_dim - 12
            assert np.abs(o["observation"][k:k + 3] - og[i][k:k + 3]).max() < 1e-3
            assert r == rg[i] and bool(fg[i] & 1) == f["is_success"] and bool(fg[i] & 8) == f["time_limit"]


def test_sliding_and_falling_object(sim, oracle):
    """No robot interaction: sliding with friction, dropping onto the table, leaving the table for the plane."""
    n = 6
    b, envs, _ = make_pair(sim, oracle, 2, n, seed=1, max_episode_steps=1000)
    states = [
        ([-0.2, 0.0, 0.02], [0, 0, 0, 1], [0.5, 0.0, 0.0], [0, 0, 0]),
        ([-0.2, 0.1, 0.02], [0, 0, 0, 1], [0.3, -0.4, 0.0], [0, 0, 2.0]),
        ([0.0, 0.0, 0.30], [0, 0, 0, 1], [0, 0, 0], [0, 0, 0]),
        ([0.0, -0.1, 0.15], [0.1825742, 0.3651484, 0.5477226, 0.7302967], [0, 0, 0], [1.0, -2.0, 0.5]),
        ([0.30, 0.0, 0.05], [0, 0, 0, 1], [0.2, 0, 0], [0, 0, 0]),
        ([0.1, 0.2, 0.02], [0, 0, 0.3826834, 0.9238795], [0, 0, 0], [0, 0, 0]),
    ]
    for e, st in zip(envs, states):
        e.set_object_state(*st)
    push_state(b, envs)
    worst = 0.0
    for s in range(12):
        a = np.zeros((n, 3), np.float32); a[:, 2] = 0.2
        b.step(torch.from_numpy(a).to(dev()))
        torch.cuda.synchronize()
        obg = b.view("object").cpu().numpy().T
        for i, e in enumerate(envs):
            e.step(a[i])
            p, q, v, w = e.object_state()
            if i in (0, 1, 5) or s < 4:          # smooth regimes: slide / rest / the free-fall phase
                worst = max(worst, np.abs(p - obg[i, :3]).max())
                assert np.abs(p - obg[i, :3]).max() < 1e-3, (s, i)
                assert np.abs(v - obg[i, 7:10]).max() < 2e-2, (s, i)
    # everything has come to rest on the table or on the plane on both sides
    obg = b.view("object").cpu().numpy().T
    for i, e in enumerate(envs):
        p = e.object_state()[0]
        assert abs(p[2] - obg[i, 2]) < 2e-3
    assert abs(obg[4, 2] + 0.38) < 3e-3 and abs(obg[2, 2] - 0.02) < 2e-3


def test_well_conditioned_push(sim, oracle):
    """The lowered gripper bar pushes the box along the table, ten steps free-running (no re-matching).
    Frictional pushing amplifies perturbations: the ORACLE run twice, the second time with its start state perturbed by 1e-7
    / 1e-6 (one / ten fp32 ulps of a joint angle), drifts apart by 0.13 mm / 3.2 mm over these ten steps (measured, DESIGN.md
    "Push parity").  fp32 rounding over 200 substeps is a perturbation of that size, so the free-running bound is 3 mm; the
    1 mm north-star tolerance is asserted where it is meaningful, from matched states
    (test_matched_state_single_steps_under_aggressive_pushing)."""
    n = 16
    b, envs, _ = make_pair(sim, oracle, 2, n, seed=3, max_episode_steps=1000)
    down = np.tile(np.array([0.0, 0.0, -1.0], np.float32), (n, 1))
    for s in range(12):
        b.step(torch.from_numpy(down).to(dev()))
        for i, e in enumerate(envs):
            e.step(down[i])
    torch.cuda.synchronize()
    rng = np.random.default_rng(4)
    for i, e in enumerate(envs):
        ee = e.get_obs()["observation"][:3].astype(float)
        e.set_object_state([ee[0] + 0.055 + 0.01 * rng.random(), ee[1] - 0.01 + 0.01 * (rng.random() - 0.5), 0.02])
    push_state(b, envs)
    start = [e.object_state()[0].copy() for e in envs]
    a = np.tile(np.array([0.35, 0.0, -0.2], np.float32), (n, 1))
    moved = 0
    for s in range(10):
        b.step(torch.from_numpy(a).to(dev()))
        torch.cuda.synchronize()
        og = b.view("observation").cpu().numpy(); qg = b.view("q").cpu().numpy().T
        for i, e in enumerate(envs):
            o, r, f = e.step(a[i])
            # the arm presses on the table while it pushes: its redundant (elbow) self-motion is weakly determined,
            # task-space quantities stay within the north-star tolerance
            assert np.abs(e.arm.q()[0] - qg[i]).max() < 1e-2, (s, i)
            assert np.abs(o["observation"][6:9] - og[i][6:9]).max() < 3e-3, (s, i)
            assert np.abs(o["observation"][:3] - og[i][:3]).max() < 3e-3, (s, i)
    for i, e in enumerate(envs):
        moved += e.object_state()[0][0] - start[i][0] > 0.01
    assert moved >= n // 2


def test_matched_state_single_steps_under_aggressive_pushing(sim, oracle):
    """Every step starts from the oracle's state on both sides; actions chase the box.  Quantile bounds on the
    one-step error, exact rewards / flags for the state each side reached."""
    n = 64
    b, envs, _ = make_pair(sim, oracle, 2, n, seed=11, max_episode_steps=1000)
    rng = np.random.default_rng(5)
    eq, eo = [], []
    hits = 0
    for s in range(40):
        Q, QD, OB = push_state(b, envs)
        ee = np.array([e.get_obs()["observation"][:3] for e in envs])
        a = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
        d = OB[:, :3] - ee
        d[:, 2] += 0.02
        a = np.clip(0.6 * a + 8.0 * d, -1, 1).astype(np.float32)
        b.step(torch.from_numpy(a).to(dev()))
        torch.cuda.synchronize()
        qg = b.view("q").cpu().numpy().T; obg = b.view("object").cpu().numpy().T
        ag = b.view("achieved_goal").cpu().numpy(); dg = b.view("desired_goal").cpu().numpy()
        rg = b.view("reward").cpu().numpy(); fg = b.view("flags").cpu().numpy()
        for i, e in enumerate(envs):
            e.step(a[i])
            p, qt, v, w = e.object_state()
            eq.append(np.abs(e.arm.q()[0] - qg[i]).max()); eo.append(np.abs(p - obg[i, :3]).max())
            hits += np.abs(v).max() > 0.05
            # flags / rewards are exact functions of the state the kernel reached
            assert np.array_equal(ag[i], obg[i, :3])
            r_ref = e.compute_reward(ag[i], dg[i])[0]
            assert r_ref == rg[i]
            assert bool(fg[i] & 1) == (r_ref == 0.0)
    eq, eo = np.array(eq), np.array(eo)
    assert hits > 100                                    # the box really gets pushed around
    assert np.quantile(eq, 0.5) < 1e-5 and np.quantile(eo, 0.5) < 1e-5
    assert np.quantile(eq, 0.99) < 1e-3 and np.quantile(eo, 0.99) < 1e-3
    assert eq.max() < 3e-2 and eo.max() < 3e-2


def test_auto_reset_and_episode_stats(sim):
    """TimeLimit 50 + SB3 auto-reset for the object tasks; the object is re-sampled on the table."""
    n = 512
    cfg = sim.make_config(2, n, max_episode_steps=5, auto_reset=True)
    b = sim.PandaBatch(cfg, device=0)
    b.seed(0)
    b.reset()
    for s in range(5):
        a = b.random_actions(9, s)
        b.step(a)
    torch.cuda.synchronize()
    done = b.view("done").cpu().numpy()
    assert done.all()
    ob = b.view("object").cpu().numpy().T
    assert np.all(np.abs(ob[:, :2]) <= 0.15 + 1e-6) and np.allclose(ob[:, 2], 0.02) and np.allclose(ob[:, 6], 1.0)
    assert b.view("stats").cpu().numpy()[0] == n
    assert np.all(b.view("step_count").cpu().numpy() == 0)
