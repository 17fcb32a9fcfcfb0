"""Oracle checks for the Push / PickAndPlace world (oracle/pgo_world.c): the free rigid box, its contacts and the
task layer (push.py, pick_and_place.py).  Pinned by the reference's free-body KATs (test/pybullet_test.py:38-100)
and by numpy's own Generator stream for the reset sampling; contact behaviour is checked against closed-form
physics (Coulomb stopping distance, rest), since the reference holds no vector for it (parity unpinned)."""
import numpy as np
import pytest


def make(oracle, task=2, **kw):
    return oracle.Env(task=task, **kw)


def test_free_fall_kat(oracle):
    """test/pybullet_test.py:56-64: a free box after one step() (20 substeps) has v = [0, 0, -0.392], atol 1e-3."""
    e = make(oracle)
    e.reset(0)
    e.set_object_state([0.0, 0.0, 1.0])
    for _ in range(20):
        e.world_substep()
    p, q, v, w = e.object_state()
    assert np.allclose(v, [0.0, 0.0, -0.392], atol=1e-3)
    assert np.allclose(q, [0, 0, 0, 1], atol=1e-3) and np.allclose(w, 0, atol=1e-3)   # :67-100
    assert abs(p[2] - (1.0 - 0.5 * 9.81 * 0.04 * 0.04 - 0.5 * 9.81 * 0.04 * 0.002)) < 1e-4  # semi-implicit Euler


def test_base_getters_at_creation(oracle):
    """test/pybullet_test.py:47-54, 67-100: position, orientation, euler rotation, angular velocity of a fresh box."""
    e = make(oracle)
    o, _ = e.reset(3)
    ob = o["observation"]
    p, q, v, w = e.object_state()
    assert np.allclose(ob[6:9], p, atol=1e-7) and np.allclose(ob[9:12], 0, atol=1e-7)
    assert np.allclose(q, [0, 0, 0, 1]) and np.allclose(ob[12:18], 0)


def test_euler_from_quaternion(oracle):
    """pybullet.getEulerFromQuaternion == scipy 'xyz' extrinsic angles (roll, pitch, yaw)."""
    from scipy.spatial.transform import Rotation
    L = oracle.lib()
    rng = np.random.default_rng(0)
    for _ in range(200):
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        out = np.zeros(3)
        L.pgo_quat_to_euler(oracle.dp(q), oracle.dp(out))
        ref = Rotation.from_quat(q).as_euler("xyz")
        assert np.allclose(Rotation.from_euler("xyz", out).as_matrix(), Rotation.from_euler("xyz", ref).as_matrix(), atol=1e-9)


def test_rest_is_stable(oracle):
    e = make(oracle)
    e.reset(1)
    p0 = e.object_state()[0].copy()
    for _ in range(500):  # one second
        e.world_substep()
    p, q, v, w = e.object_state()
    assert np.abs(p - p0).max() < 1e-4 and np.abs(v).max() < 1e-4 and np.abs(w).max() < 1e-2


def test_coulomb_stopping_distance(oracle):
    """A box sliding on the table stops after ~ v0^2 / (2 mu g) with mu = 0.5 * 0.5 (Bullet's default lateral friction,
    combined by product)."""
    e = make(oracle)
    e.reset(1)
    v0 = 0.5
    e.set_object_state([-0.2, 0.0, 0.02], lin=[v0, 0.0, 0.0])
    for _ in range(500):
        e.world_substep()
    p, q, v, w = e.object_state()
    expect = v0 * v0 / (2 * 0.25 * 9.81)
    assert np.abs(v).max() < 1e-3
    assert abs((p[0] + 0.2) - expect) < 0.1 * expect   # linear damping 0.04 shortens it slightly
    assert abs(p[2] - 0.02) < 1e-4 and abs(p[1]) < 1e-4


def test_falls_off_the_table_onto_the_plane(oracle):
    e = make(oracle)
    e.reset(1)
    e.set_object_state([0.30, 0.0, 0.05])   # beyond the table edge (x = 0.25)
    for _ in range(1000):
        e.world_substep()
    p = e.object_state()[0]
    assert abs(p[2] - (-0.4 + 0.02)) < 2e-3


@pytest.mark.parametrize("task", [2, 3])
def test_reset_sampling_matches_numpy(oracle, task):
    """Push.reset / PickAndPlace.reset draw order (push.py:69-87, pick_and_place.py:65-85) on numpy's own Generator."""
    e = make(oracle, task=task)
    for seed in range(40):
        o, _ = e.reset(seed)
        rng = np.random.default_rng(seed)
        hi_z = 0.2 if task == 3 else 0.0
        goal = np.array([0.0, 0.0, 0.02]) + 0
        noise = rng.uniform(np.array([-0.15, -0.15, 0.0]), np.array([0.15, 0.15, hi_z]))
        if task == 3 and rng.random() < 0.3:
            noise[2] = 0.0
        goal = goal + noise
        obj = np.array([0.0, 0.0, 0.02]) + rng.uniform(np.array([-0.15, -0.15, 0.0]), np.array([0.15, 0.15, 0.0]))
        assert np.array_equal(o["desired_goal"], goal.astype(np.float32))
        assert np.array_equal(o["achieved_goal"], obj.astype(np.float32))
        k = e.obs_dim - 12
        assert np.array_equal(o["observation"][k:k + 3], obj.astype(np.float32))
    assert e.obs_dim == (18 if task == 2 else 19) and e.act_dim == (3 if task == 2 else 4)


def test_push_moves_the_object_and_reward_follows(oracle):
    e = make(oracle)
    e.reset(3)
    for _ in range(12):
        o, r, f = e.step(np.array([0, 0, -1.0], np.float32))
    ee = o["observation"][:3].astype(float)
    start = np.array([ee[0] + 0.06, ee[1] - 0.01, 0.02])
    e.set_object_state(start)
    e.set_goal(start + np.array([0.03, 0.0, 0.0]))
    succ = []
    for _ in range(4):
        o, r, f = e.step(np.array([0.4, 0, -0.2], np.float32))
        d = np.linalg.norm(o["achieved_goal"] - o["desired_goal"])
        assert f["is_success"] == bool(np.float32(round(float(d), 6)) < np.float32(0.05))
        assert r == (0.0 if d <= 0.05 else -1.0)
        succ.append(f["is_success"])
    assert o["achieved_goal"][0] > start[0] + 0.02  # pushed along +x
    assert abs(o["achieved_goal"][2] - 0.02) < 2e-3  # still flat on the table


def test_render_oracle_camera_and_scene():
    """pgo_render (the checker of k_render): Bullet's yaw / pitch / roll camera and the scene content.
    The camera target projects to the image centre; yaw 0 looks along +y, yaw 90 along -x (pybullet.py:152-180 defaults)."""
    import oracle_lib as OL
    e = OL.Env(task=1, scenario="reachao3")
    e.reset(3)
    g = np.array([0.35, 0.1, 0.3])
    e.set_goal(g)
    w, h = 121, 81
    img = e.render(width=w, height=h, target_position=tuple(g), distance=0.5, yaw=30.0, pitch=-20.0, roll=0.0)
    assert img.shape == (h, w, 4) and (img[..., 3] == 255).all()
    c = img[h // 2, w // 2, :3].astype(int)
    assert c[1] > c[0] + 40 and c[1] > c[2] + 40          # the green goal marker sits at the centre pixel
    assert (img[..., :3] == (179, 179, 204)).all(axis=2).any()   # background (0.7, 0.7, 0.8) above the horizon
    # deterministic
    assert np.array_equal(img, e.render(width=w, height=h, target_position=tuple(g), distance=0.5, yaw=30.0, pitch=-20.0, roll=0.0))
    # yaw turns the camera about the world z axis through the target: an obstacle-free Reach scene seen from yaw and
    # yaw + 180 shows the goal at the centre both times, the table edge mirrored
    e0 = OL.Env(task=0)
    e0.reset(1)
    e0.set_goal(np.array([0.4, 0.3, 0.5]))
    for yaw in (0.0, 90.0, 180.0):
        im = e0.render(width=w, height=h, target_position=(0.4, 0.3, 0.5), distance=0.4, yaw=yaw, pitch=-10.0)
        c = im[h // 2, w // 2, :3].astype(int)
        assert c[1] > c[0] + 15                        # 30 % green over the bluish background
    # looking straight down from above the table: the floor / table fill the frame, no background
    top = e0.render(width=w, height=h, target_position=(0.0, 0.0, 0.0), distance=2.0, yaw=0.0, pitch=-89.9)
    assert not (top[..., :3] == (179, 179, 204)).all(axis=2).any()
