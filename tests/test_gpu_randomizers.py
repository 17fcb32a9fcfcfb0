"""Start-pose randomisers and per-episode cuboid sizes (row a16) of the CUDA reset against the oracle.

robot_pose_randomizer kinds: reach_ao.py:62-63, 521-604, 669 (set_robot_random_pose / ..._ik_goal_space / ..._ik_sphere, all
through Panda.rtb_ik = roboticstoolbox ik_lm_chan, panda.py:220-224); reachao_rand_shape: reach_ao.py:606-608, 969-979, 1084-1089.
Both sides walk the same numpy PCG64 stream per episode and the same C rand() stream per environment (ik_lm_chan restarts)."""
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


def _cfg(sim, scenario, n, randomize=True, **kw):
    from panda_gym_b200.config import TrainConfig
    tc = TrainConfig()
    tc.randomize_robot_pose = randomize
    return sim.make_config(1, n, scenario=scenario, config=tc, max_episode_steps=100, auto_reset=False, **kw)


def _compare_reset(b, envs, seeds, tag):
    goal = b.view("goal").cpu().numpy(); opos = b.view("obstacle_pos").cpu().numpy()
    obs = b.view("observation").cpu().numpy(); q = b.view("q").cpu().numpy()
    failed = 0
    for i, (e, s) in enumerate(zip(envs, seeds)):
        o, info = e.reset(int(s))
        failed += info["status"] != 0
        qo = e.arm.q()[0][:7]
        # joint values come out of ~10 Levenberg-Marquardt iterations in fp64 on both sides (device FMA contraction, libm vs CUDA
        # sin / cos): 1e-9 apart at most, then cast to the simulator's fp32 state
        assert np.abs(qo - q[:, i]).max() < 1e-6, (tag, i, qo, q[:, i])
        assert np.abs(e.goal() - goal[:, i]).max() < 1e-9, (tag, i)
        for j, (_, p, _) in enumerate(e.obstacles()):
            assert np.abs(p - opos[j, :, i]).max() < 1e-9, (tag, i, j)
        assert np.abs(o["observation"][:20] - obs[i, :20]).max() < 2e-5, (tag, i)
        if obs.shape[1] >= 29:
            assert np.abs(o["observation"][20:29] - obs[i, 20:29]).max() < 5e-5, (tag, i)
    return failed


@pytest.mark.parametrize("scenario", ["reach1", "reach3", "reachao1", "reachao2", "reachao3", "wang-2", "reachao_rand"])
def test_pose_randomizer_reset_matches_oracle(sim, oracle, scenario):
    n = 40
    b = sim.PandaBatch(_cfg(sim, scenario, n))
    envs = [oracle.Env(task=1, scenario=scenario, randomize_robot_pose=1, max_episode_steps=100) for _ in range(n)]
    total_failed = 0
    for episode in range(3):   # the C rand() stream of an environment carries over from episode to episode
        seeds = 5000 + 1000 * episode + np.arange(n)
        b.reset(seeds=torch.tensor(seeds, dtype=torch.int64, device=b.device))
        torch.cuda.synchronize()
        total_failed += _compare_reset(b, envs, seeds, (scenario, episode))
    q = b.view("q").cpu().numpy()
    assert np.std(q[0]) > 0.05 and np.std(q[3]) > 0.05, "start poses are not randomised"
    assert int(b.view("stats").cpu().numpy()[6]) == total_failed    # reset_failures counts what the oracle flags
    b.close()


def test_ik_sphere_randomizer(sim, oracle):
    """reachao_rand_start (reach_ao.py:601-604): set_robot_random_joint_position_ik_sphere draws poses until check_collided
    is false -- against the obstacles where they ARE, i.e. at create_obstacle_*'s (0.1, 0, 0.1) in the very first reset, which
    touches panda_link1 whatever the pose: the reference never leaves that loop; here it stops after 1000 tries (flagged)."""
    n = 12
    cfg = sim.make_config(1, n, scenario="reachao_rand_start", max_episode_steps=100, auto_reset=False)
    assert cfg.stage.force_randomize_pose == 1
    b = sim.PandaBatch(cfg)
    envs = [oracle.Env(task=1, scenario="reachao_rand_start", max_episode_steps=100) for _ in range(n)]
    seeds = 40 + np.arange(n)
    b.reset(seeds=torch.tensor(seeds, dtype=torch.int64, device=b.device)); torch.cuda.synchronize()
    failed = _compare_reset(b, envs, seeds, "first")
    assert failed == n and int(b.view("stats").cpu().numpy()[6]) == n
    for episode in range(2):   # afterwards the obstacles lie where the previous episode put them and the loop ends quickly
        seeds = 900 + 50 * episode + np.arange(n)
        b.reset(seeds=torch.tensor(seeds, dtype=torch.int64, device=b.device)); torch.cuda.synchronize()
        _compare_reset(b, envs, seeds, ("later", episode))
    for e in envs:
        assert not e.check_collided()[0]
    b.close()


def test_random_size_cuboids(sim, oracle):
    """reachao_rand_shape: cuboid half extents 0.2 * Dirichlet(1, 1, 1) from the episode's stream (bit-equal to numpy's own
    Generator.dirichlet), then layouts, observations, a rollout and matched-pose distances with per-environment box shapes."""
    n = 48
    b = sim.PandaBatch(_cfg(sim, "reachao_rand_shape", n, randomize=False))
    b.seed(777); b.reset(); torch.cuda.synchronize()
    size = b.view("obstacle_size").cpu().numpy(); opos = b.view("obstacle_pos").cpu().numpy()
    goal = b.view("goal").cpu().numpy(); obs = b.view("observation").cpu().numpy()
    envs = []
    for i in range(n):
        e = oracle.Env(task=1, scenario="reachao_rand_shape", max_episode_steps=100)
        o, info = e.reset(777 + i)
        assert info["status"] == 0
        envs.append(e)
        g = np.random.default_rng(777 + i)   # the Dirichlet rows are the first draws of the episode
        want = np.array([[0.2 * f for f in g.dirichlet(alpha=[1, 1, 1], size=1)[0]] for _ in range(3)])
        for j, (t, p, sz) in enumerate(e.obstacles()):
            if t == 1:
                assert np.array_equal(sz, want[j - 3]), "oracle cuboid sizes differ from numpy's dirichlet"
                assert np.abs(size[j, :, i] - sz).max() < 1e-15, (i, j, size[j, :, i], sz)
            assert np.abs(p - opos[j, :, i]).max() < 1e-9
        assert np.abs(e.goal() - goal[:, i]).max() < 1e-12
        assert np.abs(o["observation"][:29] - obs[i, :29]).max() < 3e-5
    rng = np.random.default_rng(3)
    alive = np.ones(n, bool)
    n_trunc = 0
    for k in range(12):
        A = rng.uniform(-1, 1, (n, 7)).astype(np.float32)
        b.step(torch.from_numpy(A).to(b.device)); torch.cuda.synchronize()
        og = b.view("observation").cpu().numpy(); fg = b.view("flags").cpu().numpy()
        for i, e in enumerate(envs):
            if not alive[i]:
                continue
            o, r, f = e.step(A[i])
            d = np.abs(np.asarray(e.check_collided()[1])).min()
            if d > 1e-4:
                assert bool(fg[i] & 4) == f["truncated"], (k, i)
            assert np.abs(o["observation"][:29] - og[i, :29]).max() < 1e-3
            if f["truncated"] or f["terminated"]:
                alive[i] = False; n_trunc += f["truncated"]
    assert n_trunc > 0
    # a second seeded reset re-draws the sizes
    b.reset(); torch.cuda.synchronize()
    size2 = b.view("obstacle_size").cpu().numpy()
    assert np.abs(size2[3:6] - size[3:6]).max() > 1e-3 and np.allclose(size2[3:6].sum(axis=1), 0.2)
    img = b.render(0, 96, 64).cpu().numpy()
    assert img.shape == (64, 96, 4)
    b.close()


def test_pose_randomizer_under_auto_reset(sim):
    """Warp-cooperative auto-reset (k_reset<true>) with an rtb_ik randomizer: poses differ between environments, observations
    report the pose of the reset, the per-environment C rand() stream keeps advancing, nothing NaN."""
    from panda_gym_b200.config import TrainConfig
    from panda_gym_b200.vec_env import PandaVecEnv
    tc = TrainConfig(); tc.randomize_robot_pose = True
    n = 256
    env = PandaVecEnv("PandaReachAO-v3", n_envs=n, seed=3, scenario="reachao2", config=tc, max_episode_steps=5)
    env.reset()
    rng = np.random.default_rng(1)
    seen = []
    for t in range(11):
        obs, rew, dones, infos = env.step(rng.uniform(-1, 1, (n, 7)).astype(np.float32))
        assert np.isfinite(obs["observation"]).all()
        if dones.any():
            q = env.batch.view("q").cpu().numpy()[:, dones]
            assert np.abs(obs["observation"][dones][:, 6:13] - q.T).max() < 1e-6
            seen.append(q.T.copy())
    seen = np.concatenate(seen)
    assert len(seen) >= 2 * n and seen.std(axis=0).min() > 0.02
    env.close()


def test_futile_robot_loop_short_circuit_is_exact_for_random_base(sim, oracle):
    """set_coll_free_robot with set_random_robot_base (wangexp-N): an obstacle within 0.05 of the robot's fixed core fails all
    10 000 re-draws.  The device (and the oracle by default) jump the numpy stream instead of looping; the oracle with
    no_shortcuts=1 runs the loop as the reference writes it -- joint 1, goal, obstacles and the draws that follow must be the same."""
    from panda_gym_b200.config import TrainConfig
    n = 600
    tc = TrainConfig(); tc.randomize_robot_pose = True
    b = sim.PandaBatch(sim.make_config(1, n, scenario="wangexp-3", config=tc, max_episode_steps=100, auto_reset=False))
    seeds = 7000 + np.arange(n)
    b.reset(seeds=torch.tensor(seeds, dtype=torch.int64, device=b.device)); torch.cuda.synchronize()
    q = b.view("q").cpu().numpy(); goal = b.view("goal").cpu().numpy(); opos = b.view("obstacle_pos").cpu().numpy()
    futile = 0
    for i in range(n):
        # a fresh oracle environment per seed, like device environment i: the obstacles not yet re-sampled stand at their creation pose
        e = oracle.Env(task=1, scenario="wangexp-3", randomize_robot_pose=1, max_episode_steps=100, no_shortcuts=1)
        o, info = e.reset(int(seeds[i]))
        futile += info["status"] != 0
        assert np.float32(e.arm.q()[0][0]) == q[0, i], (i, info)
        assert np.abs(e.goal() - goal[:, i]).max() < 1e-13
        for j, (_, p, _) in enumerate(e.obstacles()):
            assert np.abs(p - opos[j, :, i]).max() < 1e-12
    assert futile >= 3 and int(b.view("stats").cpu().numpy()[6]) == futile
    b.close()


def test_pose_retry_limit(sim, oracle):
    """pg_config.pose_retry_limit caps set_coll_free_robot's tries on both sides the same way."""
    from panda_gym_b200.config import TrainConfig
    n = 64
    tc = TrainConfig(); tc.randomize_robot_pose = True
    b = sim.PandaBatch(sim.make_config(1, n, scenario="reachao2", config=tc, max_episode_steps=100, auto_reset=False, pose_retry_limit=5))
    envs = [oracle.Env(task=1, scenario="reachao2", randomize_robot_pose=1, max_episode_steps=100, pose_retry_limit=5) for _ in range(n)]
    for episode in range(2):
        seeds = 300 + 100 * episode + np.arange(n)
        b.reset(seeds=torch.tensor(seeds, dtype=torch.int64, device=b.device)); torch.cuda.synchronize()
        _compare_reset(b, envs, seeds, ("limit", episode))
    b.close()


def _episode_seed(base, g, k):
    """pg_kernels.cuh episode_seed: seed of episode k of global environment g (auto-resets after the first episode)."""
    M = (1 << 64) - 1
    if k == 0:
        return (base + g) & M
    x = (base + g + 0x9E3779B97F4A7C15 * k) & M
    x = (x + 0x9E3779B97F4A7C15) & M
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M
    return x ^ (x >> 31)


@pytest.mark.parametrize("scenario", ["reachao2", "reachao3", "wangexp-3", "reachao_rand_start", "reachao_rand_shape", "wang-3", "reachao1"])
def test_warp_cooperative_auto_reset_matches_oracle(sim, oracle, scenario):
    """The auto-reset path (k_reset list mode, one warp per environment: warp-parallel ik_lm_chan restarts and the speculative
    32-tries-at-once form of set_coll_free_robot) against the oracle, which does one try at a time: every environment is reset at
    every step (TimeLimit 1) with the library's per-episode seeds, over enough episodes to meet re-draws and restarts."""
    from panda_gym_b200.config import TrainConfig
    n, base = 384, 4100
    tc = TrainConfig(); tc.randomize_robot_pose = True
    if scenario == "reachao_rand_start":
        n = 96   # its first reset runs the 1 000-try loop in every environment (test_ik_sphere_randomizer)
    b = sim.PandaBatch(sim.make_config(1, n, scenario=scenario, config=tc, max_episode_steps=1, auto_reset=True))
    b.seed(base); b.reset(); torch.cuda.synchronize()
    envs = [oracle.Env(task=1, scenario=scenario, randomize_robot_pose=1, max_episode_steps=1) for _ in range(n)]
    for i, e in enumerate(envs):
        e.reset(_episode_seed(base, i, 0))
    moved = 0
    for k in range(1, 5):
        b.step(torch.zeros((n, 7), device=b.device)); torch.cuda.synchronize()   # every episode ends, every environment auto-resets
        q = b.view("q").cpu().numpy(); goal = b.view("goal").cpu().numpy(); opos = b.view("obstacle_pos").cpu().numpy()
        for i, e in enumerate(envs):
            o, info = e.reset(_episode_seed(base, i, k))
            assert np.abs(e.arm.q()[0][:7] - q[:, i]).max() < 1e-6, (scenario, k, i, info)
            assert np.abs(e.goal() - goal[:, i]).max() < 1e-9
            for j, (_, p, _) in enumerate(e.obstacles()):
                assert np.abs(p - opos[j, :, i]).max() < 1e-9, (scenario, k, i, j)
            moved += info["status"] != 0
        if scenario == "reachao_rand_shape":   # per-environment cuboids together with a start-pose randomizer (k_reset_envbox<., true>)
            size = b.view("obstacle_size").cpu().numpy()
            for i, e in enumerate(envs):
                for j, (t, _, sz) in enumerate(e.obstacles()):
                    if t == 1:
                        assert np.abs(size[j, :, i] - sz).max() < 1e-15
    b.close()


@pytest.mark.parametrize("scenario", ["wangexp-3", "reachao3", "wang-3"])
def test_one_thread_and_warp_resets_agree(sim, scenario):
    """The auto-reset gives every environment a warp while the finished list is short and a thread each when it is long (with the
    long rejection loops deferred to a warp pass): the same environments, seeded alike, must come out identical either way.
    One batch of 4 096 (every step resets all of them: one-thread mode) against four shards of 1 024 (warp mode)."""
    from panda_gym_b200.config import TrainConfig
    tc = TrainConfig(); tc.randomize_robot_pose = True
    N, S = 4096, 1024
    big = sim.PandaBatch(sim.make_config(1, N, scenario=scenario, config=tc, max_episode_steps=1, auto_reset=True))
    big.seed(90, 0); big.reset()
    shards = []
    for r in range(N // S):
        b = sim.PandaBatch(sim.make_config(1, S, scenario=scenario, config=tc, max_episode_steps=1, auto_reset=True))
        b.seed(90, r * S); b.reset()
        shards.append(b)
    for k in range(4):
        a = torch.zeros((N, 7), device=big.device)
        big.step(a)
        for b in shards:
            b.step(a[:S])
        torch.cuda.synchronize()
        for name in ("q", "goal", "obstacle_pos"):
            full = big.view(name).cpu().numpy()
            for r, b in enumerate(shards):
                part = b.view(name).cpu().numpy()
                assert np.array_equal(full[..., r * S:(r + 1) * S], part), (scenario, k, name, r)
        # the observation is fp32 arithmetic compiled twice (warp / one-thread instantiation): equal to rounding, not bit for bit
        d = np.abs(big.view("observation").cpu().numpy() - np.concatenate([b.view("observation").cpu().numpy() for b in shards]))
        assert d[:, :29].max() < 2e-6 and d.max() < 1e-3, (scenario, k, d[:, :29].max(), d.max())   # [29:] = unit vectors
    assert sum(int(b.view("stats").cpu().numpy()[6]) for b in shards) == int(big.view("stats").cpu().numpy()[6])
    big.close()
    for b in shards:
        b.close()
