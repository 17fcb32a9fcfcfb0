"""On-device HER buffer vs the numpy restatement of SB3's HerReplayBuffer: bookkeeping, bit-exact index sampling for the
same np.random stream, relabelled goals / rewards, output order, ring-buffer overwrite."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from panda_gym_b200.vec_env import PandaVecEnv
    e = PandaVecEnv("PandaReachAO-v3", n_envs=16, seed=3, scenario="reachao3", max_episode_steps=12)
    yield e
    e.close()


def _fill(env, her, ora, steps, seed):
    rng = np.random.default_rng(seed)
    obs = env.reset()
    for _ in range(steps):
        a = rng.uniform(-1, 1, (env.num_envs, 7)).astype(np.float32)
        nobs, rew, dones, infos = env.step(a)
        real_next = {k: v.copy() for k, v in nobs.items()}
        timeouts = np.zeros(env.num_envs, bool)
        for i in np.flatnonzero(dones):  # SB3 stores the terminal observation as next_obs
            info = infos[i]
            for k in real_next:
                real_next[k][i] = info["terminal_observation"][k]
            timeouts[i] = info["TimeLimit.truncated"]
        her.add(obs, real_next, a, rew, dones, timeouts=timeouts)
        ora.add(obs, real_next, a, rew, dones.astype(np.float32), timeouts.astype(np.float32))
        obs = nobs


@pytest.mark.parametrize("buffer_size,steps", [(16 * 64, 40), (16 * 24, 70)])  # second case wraps and overwrites episodes
def test_her_matches_sb3_restatement(env, oracle, buffer_size, steps):
    from panda_gym_b200.her import HerReplayBuffer
    from her_oracle import HerOracle
    her = HerReplayBuffer(buffer_size, env, n_sampled_goal=4, max_batch_size=4096, rng="numpy")
    ref_env = oracle.Env(task=1, scenario="reachao3"); ref_env.reset(0)
    ora = HerOracle(buffer_size, env.num_envs, 56, 7, 4, compute_reward=lambda ag, dg: ref_env.compute_reward(ag, dg))
    _fill(env, her, ora, steps, seed=buffer_size)
    assert her.size() == (ora.buffer_size if ora.full else ora.pos)
    for B, seed in ((256, 0), (256, 1), (1000, 7), (4096, 123)):
        np.random.seed(seed)
        ref = ora.sample(B)
        state_after_ref = np.random.get_state()
        np.random.seed(seed)
        got = her.sample(B)
        state_after_dev = np.random.get_state()
        row, envi, t, nv, status = her.last_indices(B)
        assert status == 0 and nv == int((ora.ep_length > 0).sum())
        nbv = int(0.8 * B)
        assert np.array_equal(row, ref["indices"][0]) and np.array_equal(envi, ref["indices"][1]), "np.random.choice stream"
        assert np.array_equal(t[:nbv], ref["indices"][2]), "np.random.randint(cur, ep_len) stream"
        assert np.array_equal(state_after_ref[1], state_after_dev[1]) and state_after_ref[2] == state_after_dev[2], \
            "the MT19937 state must advance exactly as numpy's"
        g = lambda x: x.cpu().numpy()
        for k in ("observation", "achieved_goal", "desired_goal"):
            assert np.array_equal(g(got.observations[k]), ref["obs"][k])
            assert np.array_equal(g(got.next_observations[k]), ref["next_obs"][k])
        assert np.array_equal(g(got.actions), ref["actions"])
        assert np.array_equal(g(got.rewards)[:, 0], ref["rewards"])
        assert np.array_equal(g(got.dones)[:, 0], ref["dones"])
        # relabelled part: desired goal == a future achieved goal of the same episode, reward in {-1, 0}
        assert set(np.unique(g(got.rewards)[B - nbv:, 0])) <= {-1.0, 0.0}
    her.close()


def test_her_device_rng_and_empty(env):
    from panda_gym_b200.her import HerReplayBuffer
    her = HerReplayBuffer(16 * 64, env, rng="numpy", max_batch_size=512)
    with pytest.raises(RuntimeError):
        np.random.seed(0); her.sample(32)          # nothing stored yet (SB3 raises too)
    her.close()
    her = HerReplayBuffer(16 * 64, env, rng="device", max_batch_size=65536, device_seed=5)
    obs = env.reset()
    rng = np.random.default_rng(0)
    for _ in range(30):
        a = rng.uniform(-1, 1, (16, 7)).astype(np.float32)
        nobs, rew, dones, infos = env.step(a)
        her.add(obs, nobs, a, rew, dones, timeouts=np.zeros(16, bool))
        obs = nobs
    s = her.sample(65536)
    row, envi, t, nv, status = her.last_indices(65536)
    assert status == 0 and nv > 0
    assert s.observations["observation"].shape == (65536, 56) and torch.isfinite(s.rewards).all()
    # every virtual t lies inside its episode and not before the sampled transition
    nbv = int(0.8 * 65536)
    assert (t[:nbv] >= 0).all()
    # all valid transitions are reachable and roughly uniformly hit
    flat = row.astype(np.int64) * 16 + envi
    counts = np.bincount(flat, minlength=64 * 16)
    assert (counts > 0).sum() == nv and counts.max() < 6 * 65536 / nv
    her.close()


def test_her_at_baseline_size(oracle):
    """BASELINE configs[4] / SURVEY 8(d) config 5: 1 048 576-transition buffer x 1 024 environment columns (1 024 rows),
    synthetic episodes of length U{10..100} (seed 7), the ring wrapped once; (row, env, t) and the post-call MT19937 state
    bit-exact against numpy's own RandomState at B = 256, 4 096 and 65 536, gathered fields equal at B = 4 096."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from panda_gym_b200.vec_env import PandaVecEnv
    from panda_gym_b200.her import HerReplayBuffer
    from her_oracle import HerOracle
    E, cap, od, ad = 1024, 1 << 20, 56, 7
    env = PandaVecEnv("PandaReachAO-v3", n_envs=E, seed=3, scenario="reachao3", max_episode_steps=100)
    env.reset()
    her = HerReplayBuffer(cap, env, n_sampled_goal=4, max_batch_size=65536, rng="numpy")
    ref_env = oracle.Env(task=1, scenario="reachao3"); ref_env.reset(0)
    ora = HerOracle(cap, E, od, ad, 4, compute_reward=lambda ag, dg: ref_env.compute_reward(ag, dg))
    assert her.buffer_size == ora.buffer_size == 1024
    rng = np.random.default_rng(7)
    remaining = rng.integers(10, 101, E)
    dev = env.batch.device
    adds = 1024 + 160  # wraps: the oldest 160 rows are overwritten, episodes straddling the cursor get invalidated
    for step in range(adds):
        blk = rng.random((E, 2 * od + 12 + ad + 1), dtype=np.float32)
        obs = {"observation": blk[:, :od], "achieved_goal": blk[:, 2 * od:2 * od + 3], "desired_goal": blk[:, 2 * od + 3:2 * od + 6]}
        nobs = {"observation": blk[:, od:2 * od], "achieved_goal": blk[:, 2 * od + 6:2 * od + 9], "desired_goal": blk[:, 2 * od + 9:2 * od + 12]}
        act, rew = blk[:, 2 * od + 12:2 * od + 12 + ad], blk[:, -1]
        remaining -= 1
        done = remaining == 0
        timeouts = done & (rng.random(E) < 0.3)
        remaining[done] = rng.integers(10, 101, int(done.sum()))
        ora.add_vectorised(obs, nobs, act, rew, done.astype(np.float32), timeouts.astype(np.float32))
        t = torch.from_numpy(blk).to(dev)
        her.add({"observation": t[:, :od], "achieved_goal": t[:, 2 * od:2 * od + 3], "desired_goal": t[:, 2 * od + 3:2 * od + 6]},
                {"observation": t[:, od:2 * od], "achieved_goal": t[:, 2 * od + 6:2 * od + 9], "desired_goal": t[:, 2 * od + 9:2 * od + 12]},
                t[:, 2 * od + 12:2 * od + 12 + ad], t[:, -1], torch.from_numpy(done).to(dev), timeouts=torch.from_numpy(timeouts).to(dev))
    assert ora.full and her.size() == 1024
    g = lambda x: x.cpu().numpy()
    for B, seed in ((256, 0), (4096, 11), (65536, 5), (65536, 6)):
        np.random.seed(seed)
        ref = ora.sample(B)
        state_ref = np.random.get_state()
        np.random.seed(seed)
        got = her.sample(B)
        state_dev = np.random.get_state()
        row, envi, t, nv, status = her.last_indices(B)
        nbv = int(0.8 * B)
        assert status == 0 and nv == int((ora.ep_length > 0).sum()) and nv > 900000
        assert np.array_equal(row, ref["indices"][0]) and np.array_equal(envi, ref["indices"][1])
        assert np.array_equal(t[:nbv], ref["indices"][2])
        assert np.array_equal(state_ref[1], state_dev[1]) and state_ref[2] == state_dev[2]
        if B <= 4096:
            for k in ("observation", "achieved_goal", "desired_goal"):
                assert np.array_equal(g(got.observations[k]), ref["obs"][k])
                assert np.array_equal(g(got.next_observations[k]), ref["next_obs"][k])
            assert np.array_equal(g(got.actions), ref["actions"]) and np.array_equal(g(got.rewards)[:, 0], ref["rewards"])
            assert np.array_equal(g(got.dones)[:, 0], ref["dones"])
    her.close(); env.close()


def test_her_numpy_stream_details(env):
    """Two places where float / state handling could drift from SB3: nb_virtual = int(her_ratio * B) is a DOUBLE product
    (n_sampled_goal = 2, B = 3 -> 2, where float32 gives 1), and a Gaussian cached in numpy's legacy state (has_gauss)
    must survive a sample() call."""
    from panda_gym_b200.her import HerReplayBuffer
    her = HerReplayBuffer(16 * 64, env, n_sampled_goal=2, rng="numpy", max_batch_size=512)
    obs = env.reset()
    rng = np.random.default_rng(1)
    for _ in range(30):
        a = rng.uniform(-1, 1, (16, 7)).astype(np.float32)
        nobs, rew, dones, infos = env.step(a)
        her.add(obs, nobs, a, rew, dones, infos=infos)
        obs = nobs
    for B in (3, 6, 300, 333):
        np.random.seed(B)
        np.random.normal()                      # leaves a cached second Gaussian in the legacy state
        before = np.random.get_state()
        assert before[3] == 1
        s = her.sample(B)
        after = np.random.get_state()
        assert after[3] == 1 and after[4] == before[4], "cached Gaussian lost"
        assert her.last_n_virtual == int((1 - (1.0 / 3)) * B), "nb_virtual must be the double product SB3 computes"
        assert s.observations["observation"].shape == (B, 56)
    assert int((1 - (1.0 / 3)) * 3) == 2  # the product that a float32 her_ratio (0.6666667 -> 1.9999999) truncates to 1
    her.close()
