"""Benchmark scenes (N1): the static-box scene pass of the CUDA path against the oracle, scene by scene.

Both sides get the same boxes (panda_gym_b200/data/benchmark_scenes.json, generated from the reference assets by
tools/gen_scene_data.py; tests/test_host_logic.py checks the loader against the assets where they exist)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

Q_LO = np.array([-2.7, -1.7, -2.7, -3.0, -2.7, 0.1, -2.7])
Q_HI = np.array([2.7, 1.7, 2.7, -0.1, 2.7, 3.7, 2.7])


@pytest.fixture(scope="module")
def sim():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from panda_gym_b200 import sim as s
    return s


def oracle_env(oracle, scenes, name, stage=None, **kw):
    sc = scenes.SCENES[name]
    extra = dict(custom_neutral=1, neutral_q=sc["neutral"]) if sc["neutral"] is not None else {}
    if stage is not None and stage.force_randomize_pose:   # narrow_tunnel / workshop2 / industrial install a start-pose randomizer
        extra.update(scene_pose_randomizer=stage.pose_randomizer, scene_force_randomize=1,
                     scene_pose_lo=list(stage.pose_lo), scene_pose_hi=list(stage.pose_hi))
    e = oracle.Env(task=1, scenario="static", scene_goal_lo=sc["goal_lo"], scene_goal_hi=sc["goal_hi"], **extra, **kw)
    kind = "box" if name == "wall" else "urdf_box"   # wall: a create_obstacle_cuboid body (safe margin), not a URDF import
    e.set_obstacles([oracle.make_obstacle(kind, b[:3], b[3:6]) for b in scenes.builtin_scene_boxes(name)])
    return e


ALL = ["bookshelves", "warehouse", "countertop", "kitchen", "raised_shelves", "tabletop", "tabletop2", "industrial", "tunnel",
       "narrow_tunnel", "workshop", "workshop2", "workshop3", "library", "library1", "library2", "wall"]


@pytest.mark.parametrize("name", ALL)
def test_scene_reset_distances_and_flags(sim, oracle, name):
    from panda_gym_b200 import scenes
    n = 48
    cfg = sim.make_config(1, n, scenario=name, max_episode_steps=50, auto_reset=False)
    assert cfg.stage.needs_scene == 1
    b = sim.PandaBatch(cfg); b.seed(321); b.reset()
    torch.cuda.synchronize()
    goal = b.view("goal").cpu().numpy(); obs = b.view("observation").cpu().numpy(); q0 = b.view("q").cpu().numpy()
    e = oracle_env(oracle, scenes, name, stage=cfg.stage, max_episode_steps=50)
    # ---- seeded reset: goal from the scene's box, collision-free against robot, table and every box of the scene
    fails = 0
    for i in range(12):
        if cfg.stage.force_randomize_pose:   # the start pose depends on the environment's own history (C rand() stream, Bullet IK
            e = oracle_env(oracle, scenes, name, stage=cfg.stage, max_episode_steps=50)   # seed): device env i is a fresh one
        o, info = e.reset(321 + i)
        # status 1 = no collision-free goal in 10 000 draws, the goal falls back to the EE position (reach_ao.py:1108-1115):
        # what happens in the scenes whose goal box lies inside the furniture and that the reference drives with
        # config.fixed_target (industrial :410, workshop2 :381) or whose builder loads another scene's URDF (warehouse :473)
        if info["status"] == 0:
            assert np.array_equal(e.goal(), goal[:, i]), (name, i, e.goal(), goal[:, i])
        else:
            assert np.abs(e.goal() - goal[:, i]).max() < 1e-6, (name, i, e.goal(), goal[:, i])
            fails += 1
        assert np.abs(e.arm.q()[0] - q0[:, i]).max() < 1e-6
        assert np.abs(o["observation"][:29] - obs[i][:29]).max() < 2e-5, (name, i, np.abs(o["observation"] - obs[i]).argmax())
        assert np.abs(o["observation"][29:] - obs[i][29:]).max() < 5e-4, (name, i)   # unit vectors (fp32 closest-point direction)
    assert int(b.view("stats").cpu().numpy()[6]) * 12 == fails * n or fails in (0, 12), "reset-failure counter"
    # ---- matched random poses: per-link distances (observation), collision flag
    rng = np.random.default_rng(5)
    Q = (Q_LO + rng.random((n, 7)) * (Q_HI - Q_LO)).astype(np.float32)
    Q[: n // 3] = q0[:, : n // 3].T + 0.15 * rng.standard_normal((n // 3, 7)).astype(np.float32)   # near the reset pose too
    b.set_joint_state(torch.tensor(np.ascontiguousarray(Q.T), device=b.device), None)
    b.observe()
    torch.cuda.synchronize()
    og = b.view("observation").cpu().numpy(); fg = b.view("flags").cpu().numpy(); dl = b.view("d_link").cpu().numpy()
    n_coll = 0
    for i in range(n):
        e.arm.set_q(Q[i].astype(np.float64), np.zeros(7))
        coll, d, dt = e.check_collided()
        assert np.abs(np.asarray(d) - dl[:, i]).max() < 3e-5, (name, i, d, dl[:, i])
        margin = min(np.abs(d).min(), abs(dt))
        if margin > 1e-4:
            assert bool(fg[i] & 4) == bool(coll), (name, i, d, dt)
        n_coll += bool(coll)
        o = e.get_obs()["observation"]
        k = np.abs(o - og[i])
        # unit vectors flip between equally close features; distances and the robot part must agree
        assert k[:6].max() < 2e-5
    if name in ("warehouse", "tabletop2", "wall"):   # tabletop2.urdf as placed by its json puts a table top through the robot's
        assert n_coll == n                           # base link, and so does the cuboid of create_stage_wall
    else:
        assert 0 < n_coll < n, "poses must cover both outcomes"


@pytest.mark.parametrize("name", ["kitchen", "industrial", "library"])
def test_scene_rollout_flags(sim, oracle, name):
    """Free-running 12 steps with truncate-on-collision: same termination step and flags on both sides."""
    from panda_gym_b200 import scenes
    n = 16
    cfg = sim.make_config(1, n, scenario=name, max_episode_steps=30, auto_reset=False)
    b = sim.PandaBatch(cfg); b.seed(77); b.reset()
    envs = []
    for i in range(n):
        e = oracle_env(oracle, scenes, name, stage=cfg.stage, max_episode_steps=30)
        e.reset(77 + i)
        envs.append(e)
    alive = np.ones(n, bool)
    for k in range(12):
        a = b.random_actions(9, k)
        b.step(a)
        torch.cuda.synchronize()
        an = a.cpu().numpy(); fg = b.view("flags").cpu().numpy(); og = b.view("observation").cpu().numpy()
        for i, e in enumerate(envs):
            if not alive[i]:
                continue
            o, r, f = e.step(an[i])
            d = np.abs(np.asarray(e.check_collided()[1])).min()
            if d > 1e-4:
                assert bool(fg[i] & 4) == f["truncated"], (name, k, i)
            assert np.abs(o["observation"][:6] - og[i][:6]).max() < 1e-4
            if f["truncated"] or f["terminated"]:
                alive[i] = False


def test_scene_swap_and_removal(sim):
    """set_stage with a benchmark scene brings its boxes; a training stage removes them again."""
    cfg = sim.make_config(1, 32, scenario="reachao1", max_episode_steps=50, auto_reset=False)
    b = sim.PandaBatch(cfg); b.seed(1); b.reset()
    b.set_stage(sim.stage_from_scenario("bookshelves"))
    torch.cuda.synchronize()
    d1 = b.view("d_link").cpu().numpy().copy()
    assert d1.min() < 0.6            # shelves are within reach of the arm
    b.set_stage(sim.stage_from_scenario("reachao1"))
    torch.cuda.synchronize()
    assert b._scene_boxes is None
    img = b.render(0, 96, 64).cpu().numpy()
    assert img.shape == (64, 96, 4)


def test_fixed_target(sim, oracle):
    """TrainConfig.fixed_target (train_config.py:29): with re-sampled obstacles the goal IS the fixed target and the first
    goal-sampling pass draws nothing (the obstacle draws that follow start at the same stream position on both sides);
    with fixed obstacles the second pass replaces it (reach_ao.py:981-988)."""
    from panda_gym_b200.config import TrainConfig
    tgt = [0.55, 0.05, 0.4]
    for scenario, kept in (("reachao3", True), ("reach1", False)):
        c = TrainConfig(); c.fixed_target = tgt
        cfg = sim.make_config(1, 16, scenario=scenario, config=c, max_episode_steps=50, auto_reset=False)
        b = sim.PandaBatch(cfg); b.seed(40); b.reset()
        torch.cuda.synchronize()
        goal = b.view("goal").cpu().numpy(); opos = b.view("obstacle_pos").cpu().numpy()
        for i in range(16):
            e = oracle.Env(task=1, scenario=scenario, has_fixed_target=1, fixed_target=tgt, max_episode_steps=50)
            e.reset(40 + i)
            assert np.allclose(e.goal(), goal[:, i], rtol=0, atol=1e-12)
            assert np.array_equal(e.goal(), np.array(tgt)) == kept
            for j, (_, p, _) in enumerate(e.obstacles()):
                assert np.allclose(p, opos[j, :, i], rtol=0, atol=1e-9), (scenario, i, j)
