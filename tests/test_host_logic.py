"""Host-side logic that needs no GPU: the C ABI loads and exports every declared symbol, config / scenario blocks,
the reference-shaped Python surface, loud failure without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    ge.build_cuda()
    from panda_gym_b200 import _lib
    return _lib


def test_abi_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "panda_b200.h")).read()
    declared = set(re.findall(r"^\s*(?:int64_t|int|const char\*)\s+(pg_\w+)\s*\(", header, re.M))
    assert len(declared) >= 28
    L = ctypes.CDLL(lib.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), "libpanda_b200.so does not export %s" % name
    bound = {s[0] for s in lib.SYMBOLS}
    assert declared == bound, "ctypes table and header differ: %s" % (declared ^ bound)
    assert b"sm_100a" in lib.load().pg_version()


def test_scenario_blocks(lib):
    L = lib.load()
    st = lib.Stage()
    assert L.pg_stage_from_scenario(ctypes.byref(st), b"reachao3") == 0
    assert st.n_obstacles == 3 and st.goal_sampler == 1 and (st.goal_rmin, st.goal_rmax) == (0.5, 0.8)
    assert st.obstacle_sampler == 1 and st.randomize_obstacle_position == 1 and st.goal_upper_half == 1
    assert L.pg_stage_from_scenario(ctypes.byref(st), b"reachao1") == 0
    assert st.n_obstacles == 1 and st.obstacle_size[0][0] == 0.04 and st.goal_sampler == 0
    assert np.allclose(st.goal_lo[:], [0.4, -0.2, 0.0]) and np.allclose(st.goal_hi[:], [0.8, 0.2, 0.4])
    assert L.pg_stage_from_scenario(ctypes.byref(st), b"wangexp-5") == 0 and st.n_obstacles == 5
    assert L.pg_stage_from_scenario(ctypes.byref(st), b"reachao_rand") == 0
    assert st.n_obstacles == 6 and list(st.obstacle_type[:6]) == [0, 0, 0, 1, 1, 1] and st.allow_overlapping_obstacles == 1
    assert (st.sample_size_lo, st.sample_size_hi) == (4, 6)
    # reach_ao.py:262-264 / SURVEY B.2: "wangexp_3" is not a key
    assert L.pg_stage_from_scenario(ctypes.byref(st), b"wangexp_3") != 0
    assert b"not found" in L.pg_last_error()


def test_default_configs(lib):
    L = lib.load()
    c = lib.Config()
    assert L.pg_default_config(ctypes.byref(c), 1, None) == 0
    assert (c.obs_js, c.obstacle_obs, c.truncate_on_collision, c.terminate_on_success, c.n_substeps) == (1, 1, 1, 1, 20)
    assert c.collision_reward == -100.0 and list(c.base_position[:]) == [0.0, 0.0, 0.0]
    assert L.pg_default_config(ctypes.byref(c), 0, None) == 0
    assert c.obs_js == 0 and c.terminate_on_success == 0 and c.base_position[0] == -0.6


def test_make_config_maps_trainconfig(lib):
    from panda_gym_b200 import sim
    from panda_gym_b200.config import TrainConfig
    cfg = sim.make_config(1, 16, scenario="reachao2", config=TrainConfig(reward_type="kumar", goal_condition="halt",
                                                                          action_limiter="scale", collision_reward=-50,
                                                                          task_observations={"obstacles": "closest_per_link", "prior": None}),
                          max_episode_steps=75, speed_threshold=0.01)
    assert (cfg.reward_type, cfg.goal_condition, cfg.action_limiter, cfg.obstacle_obs) == (4, 1, 1, 3)
    assert cfg.collision_reward == -50 and cfg.stage.max_episode_steps == 75 and cfg.stage.speed_threshold == 0.01
    # robot_pose_randomizer of each scenario (reach_ao.py:62-63, 521-604, 669, 711): 0 torus, 1 random base, 2 torus front half,
    # 3 ik goal space, 4 ik sphere, 5 ik box, 6 Bullet ik box
    for scn, kind in [("reachao1", 2), ("reach1", 2), ("reach2", 2), ("reach3", 0), ("reachao2", 3), ("reachao3", 3), ("reachao_rand", 3),
                      ("reachao_rand_shape", 3), ("wang-2", 0), ("wangexp-3", 1)]:
        cfg = sim.make_config(1, 1, scenario=scn, config=TrainConfig(randomize_robot_pose=True))
        assert cfg.randomize_robot_pose == 1 and cfg.stage.pose_randomizer == kind and cfg.stage.force_randomize_pose == 0, scn
    cfg = sim.make_config(1, 1, scenario="wangexp-3", config=TrainConfig(randomize_robot_pose=True))  # set_random_robot_base
    assert cfg.stage.n_obstacles == 3
    cfg = sim.make_config(1, 1, scenario="reachao_rand_start")   # switches randomize_robot_pose on itself (reach_ao.py:601-604)
    assert cfg.stage.force_randomize_pose == 1 and cfg.stage.pose_randomizer == 4 and (cfg.stage.pose_lo[0], cfg.stage.pose_hi[0]) == (0.45, 0.7)
    cfg = sim.make_config(1, 1, scenario="reachao_rand_shape")
    assert cfg.stage.random_size_cuboids == 1 and cfg.stage.n_obstacles == 6
    for scn, kind in [("narrow_tunnel", 5), ("workshop2", 5), ("industrial", 6)]:   # reach_ao.py:310-320, 363-375, 395-403
        cfg = sim.make_config(1, 1, scenario=scn)
        assert cfg.stage.force_randomize_pose == 1 and cfg.stage.pose_randomizer == kind, scn
    with pytest.raises(lib.PandaB200Error):
        sim.make_config(1, 1, scenario="no_such_scenario")


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from panda_gym_b200 import sim
    cfg = sim.make_config(1, 4, scenario="reachao1")
    with pytest.raises(lib.PandaB200Error):
        sim.PandaBatch(cfg)
    h = ctypes.c_void_p()
    assert lib.load().pg_create(ctypes.byref(cfg), 0, ctypes.byref(h)) != 0
    assert b"no CUDA device" in lib.load().pg_last_error() or b"CUDA" in lib.load().pg_last_error()


def test_lazy_infos():
    from panda_gym_b200.vec_env import LazyInfos
    flags = np.array([0, 1 | 2, 4, 8], np.uint8); done = np.array([0, 1, 1, 1], bool)
    term = np.arange(3 * 5, dtype=np.float32).reshape(3, 5); ag = np.zeros((3, 3), np.float32)
    calls = []

    def fetch():  # what PandaBatch.fetch_finished delivers: the finished rows of this step, sorted by environment index
        calls.append(1)
        return np.array([1, 2, 3]), term, ag, np.array([-3, -101, -50], np.float32), np.array([4, 7, 50])

    infos = LazyInfos(np.ones((4, 3), np.float32), flags, done, fetch, 1.5)
    assert len(infos) == 4 and "terminal_observation" not in infos[0] and calls == []
    assert infos[1]["is_success"] and infos[1]["episode"] == {"r": -3.0, "l": 4, "t": 1.5}
    # SB3's SubprocVecEnv worker on gymnasium >= 0.26: TimeLimit.truncated = truncated and not terminated, and gymnasium's
    # truncated is the task's is_truncated() (collision) OR the TimeLimit wrapper
    assert infos[2]["is_truncated"] and infos[2]["TimeLimit.truncated"]
    assert not infos[1]["TimeLimit.truncated"] and not infos[0]["TimeLimit.truncated"]
    assert np.array_equal(infos.time_limit_truncated, [False, False, True, True])
    assert np.array_equal(infos.finished, [1, 2, 3]) and np.array_equal(infos.episode_lengths, [4, 7, 50])
    assert infos.terminal_observations()["observation"].shape == (3, 5)
    assert infos[3]["TimeLimit.truncated"] and np.array_equal(infos[3]["terminal_observation"]["observation"], term[2])
    assert [i["is_success"] for i in infos] == [False, True, False, False]
    assert calls == [1], "the finished rows are fetched once, on first use"


def test_spaces_and_ids():
    from panda_gym_b200 import spaces
    from panda_gym_b200.vec_env import ENV_IDS
    b = spaces.Box(-1.0, 1.0, shape=(7,), dtype=np.float32)
    assert b.sample().shape == (7,) and b.sample().dtype == np.float32
    assert set(ENV_IDS) >= {"PandaReachAO-v3", "PandaReach-v3", "PandaReachJoints-v3", "PandaReachDense-v3", "PandaReachJointsDense-v3"}


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "panda_gym_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle_lib" not in src and "liboracle" not in src, f
                assert not re.search(r'#include\s+"[^"]*pgo', src) and not re.search(r"\bpgo_\w+\s*\(", src), f


def test_benchmark_results_from_accumulators():
    """sharding.benchmark_results: perform_benchmark's results dict (evaluation/evaluate.py:286-300) from the 12 device sums."""
    from panda_gym_b200.sharding import benchmark_results
    #      succ  len_s  eff   man  jerk  speed  eps  len   coll  ret    timeouts
    raw = [4.0, 40.0, 2.0, 0.4, 8.0, 12.0, 10.0, 130.0, 5.0, -420.0, 1.0, 0.0]
    r = benchmark_results(raw, n_substeps=20)
    assert r["success_rate"] == 0.4 and r["collision_rate"] == 0.5 and r["timeout_rate"] == 0.1 and r["num_episodes"] == 10.0
    assert r["mean_reward"] == -42.0 and r["mean_ep_length"] == 13.0 and r["mean_ep_length_success"] == 10.0
    assert r["mean_num_sim_steps"] == 260.0 and r["mean_num_sim_steps_success"] == 200.0
    assert r["mean_effort"] == 0.5 and r["mean_manipulability"] == 0.1 and r["mean_norm_jerk"] == 0.2 and r["mean_ee_speed"] == 0.3
    empty = benchmark_results([0.0] * 12)
    assert empty["num_episodes"] == 0.0 and empty["mean_effort"] != empty["mean_effort"]  # nan, like np.mean([])


def test_her_oracle_vectorised_add_equals_loop():
    """The numpy restatement of SB3's HerReplayBuffer has a loop form (reads like SB3) and a vectorised form (used at the
    1 M-transition size): identical bookkeeping on a wrapping buffer with ragged episodes."""
    from her_oracle import HerOracle
    rng = np.random.default_rng(3)
    E, R, od, ad = 13, 37, 5, 2
    a, b = HerOracle(R * E, E, od, ad), HerOracle(R * E, E, od, ad)
    remaining = rng.integers(3, 60, E)  # some episodes are longer than the buffer has rows
    for step in range(150):
        obs = {"observation": rng.random((E, od), np.float32), "achieved_goal": rng.random((E, 3), np.float32),
               "desired_goal": rng.random((E, 3), np.float32)}
        nobs = {k: v + 1 for k, v in obs.items()}
        remaining -= 1
        done = (remaining == 0)
        remaining[done] = rng.integers(3, 60, int(done.sum()))
        args = (obs, nobs, rng.random((E, ad), np.float32), rng.random(E, np.float32), done.astype(np.float32), np.zeros(E, np.float32))
        a.add(*args); b.add_vectorised(*args)
        assert np.array_equal(a.ep_start, b.ep_start) and np.array_equal(a.ep_length, b.ep_length), step
        assert np.array_equal(a._current_ep_start, b._current_ep_start) and a.pos == b.pos and a.full == b.full
    assert a.full and (a.ep_length > 0).any()


def test_benchmark_scene_loader():
    """scenes.py: shipped numbers (data/benchmark_scenes.json) = what the loader extracts from the reference assets, where
    those exist (not on the GPU box); counts and extents of the shipped table always."""
    import os
    import importlib.util
    import numpy as np
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("pg_scenes", os.path.join(root, "panda_gym_b200", "scenes.py"))
    scenes = importlib.util.module_from_spec(spec); spec.loader.exec_module(scenes)
    counts = {"bookshelves": 6, "warehouse": 50, "countertop": 52, "kitchen": 50, "raised_shelves": 10, "tabletop": 47, "tabletop2": 50,
              "industrial": 16, "tunnel": 3, "narrow_tunnel": 3, "workshop": 6, "library": 7}
    for name, k in counts.items():
        b = scenes.builtin_scene_boxes(name)
        assert b.shape == (k, 7) and (b[:, 3:6] > b[:, 6:7]).all() and (b[:, 6] == 0.001).all()
        assert k <= 64
    # the URDFs give every link's joint AND collision origin: a box sits at their sum times globalScaling (tabletop: 1.4)
    t = scenes.builtin_scene_boxes("bookshelves")
    assert np.allclose(t[0], [0.765, 0.0, 0.14, 0.21, 0.385, 0.007, 0.001])
    assets = "/root/reference/panda_gym/assets/scenarios"
    if os.path.isdir(assets):
        for name in scenes.SCENES:
            assert np.allclose(scenes.load_scene_boxes(assets, name), scenes.builtin_scene_boxes(name), atol=1e-9), name
