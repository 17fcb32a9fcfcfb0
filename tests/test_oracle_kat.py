"""Pin the CPU oracle against every known-answer vector the reference's own tests hold for the path
(reference test/pybullet_test.py, all atol 1e-3) plus the golden vectors recovered from shipped artefacts
(SURVEY.md section 4 / 8c)."""
import numpy as np


def test_dt(oracle):
    # test/pybullet_test.py:30-35: dt == timestep * n_substeps == 0.04.  Measured on the oracle: a free box (no contact) under
    # gravity gains v_z = -g * t over one env step (n_substeps substeps of the oracle's own timestep), so t = -v_z / g.
    e = oracle.Env(task=2)
    e.reset(0)
    e.set_object_state([0.0, 0.0, 1.0])
    for _ in range(e.cfg.n_substeps):
        e.world_substep()
    v = e.object_state()[2]
    # per-substep damping m v (k + k |v|) with k = 0.04 shaves 4e-5 relative: well inside the reference's atol
    assert abs(-v[2] / 9.81 - 0.04) < 1e-4 and e.cfg.n_substeps == 20


def test_link_position_stock(oracle):
    # test/pybullet_test.py:124-136: stock URDF, q = 0, getLinkState(1)[0]
    arm = oracle.Arm(model=1)
    assert np.allclose(arm.link_state(1)[0], [0.000, 0.060, 0.373], atol=1e-3)


def _motor_kat(oracle):
    arm = oracle.Arm(model=1)
    arm.set_motor(5, 1, 0.3, 5.0)  # control_joints("panda", [5], [0.3], [5.0]) -> POSITION_CONTROL
    for _ in range(20):
        arm.substep()
    return arm


def test_joint_angle_kat(oracle):
    # test/pybullet_test.py:190-204
    arm = _motor_kat(oracle)
    assert np.allclose(arm.q()[0][5], 0.063, atol=1e-3)


def test_link_orientation_kat(oracle):
    # test/pybullet_test.py:139-153
    arm = _motor_kat(oracle)
    assert np.allclose(arm.link_state(5)[1], [0.707, -0.02, 0.02, 0.707], atol=1e-3)


def test_link_velocity_kat(oracle):
    # test/pybullet_test.py:156-170
    arm = _motor_kat(oracle)
    assert np.allclose(arm.link_state(5)[2], [-0.0068, 0.0000, 0.1186], atol=1e-3)


def test_link_angular_velocity_kat(oracle):
    # test/pybullet_test.py:173-187 -- this one pins the inertia-from-AABB rule (mesh + 4 margins) and the
    # btMultiBody damping 0.04: 3 margins give 2.9835, 5 give 2.9545, damping 0 gives 2.9727.
    arm = _motor_kat(oracle)
    w = arm.link_state(5)[3]
    assert np.allclose(w, [0.000, -2.969, 0.000], atol=1e-3)
    assert abs(w[1] + 2.969) < 2e-4


def test_set_joint_angles(oracle):
    # test/pybullet_test.py:220-251
    arm = oracle.Arm(model=1)
    arm.reset_joint(3, 0.4)
    arm.reset_joint(4, 0.5)
    q, qd = arm.q()
    assert np.allclose(q[3:5], [0.4, 0.5], atol=1e-3) and np.all(qd == 0)


def test_inverse_kinematics_kat(oracle):
    # test/pybullet_test.py:254-266 -- pins: DLS with joint damping 0.5, 20 iterations, URDF link frame
    # (not COM), normalised target quaternion.
    arm = oracle.Arm(model=1)
    out = arm.ik(6, [0.4, 0.5, 0.6], [0.707, -0.02, 0.02, 0.707])
    assert np.allclose(out, [1.000, 1.223, -1.113, -0.021, -0.917, 0.666, -0.499, 0.0, 0.0], atol=1e-3)


def test_free_fall(oracle):
    # test/pybullet_test.py:56-64: a box created at z = 1 has get_base_velocity == [0, 0, -0.392] after one step()
    # (20 substeps of semi-implicit Euler at g = 9.81); the oracle's free body is the Push / PickAndPlace object
    e = oracle.Env(task=2)
    e.reset(0)
    e.set_object_state([0.0, 0.0, 1.0])
    for _ in range(20):
        e.world_substep()
    p, q, v, w = e.object_state()
    assert np.allclose(v, [0.0, 0.0, -0.392], atol=1e-3)
    assert np.allclose(q, [0.0, 0.0, 0.0, 1.0], atol=1e-3) and np.allclose(w, 0.0, atol=1e-3)   # :67-100


def test_neutral_pose_golden(oracle):
    # shipped checkpoints (training/run_data/wandb/deep-frog-298 ... best_model.zip::_last_obs.achieved_goal):
    # real PyBullet FK of panda_grasptarget (link 12 of custom_0) at the neutral pose, float32
    arm = oracle.Arm(model=0)
    for d, v in enumerate([0, -0.3, 0, -2.2, 0, 2.0, np.pi / 4]):
        arm.reset_joint(d, v)
    assert np.allclose(arm.link_state(12)[0], [0.48420656, -5.19e-12, 0.41103777], atol=2e-6)
    # the current code's end-effector = COM of link 11 (rightfinger)
    assert np.allclose(arm.link_state(11)[0], [0.4816, 0.0100, 0.4375], atol=1e-4)


def test_minv_symmetric_positive(oracle):
    arm = oracle.Arm(model=0)
    rng = np.random.default_rng(0)
    for _ in range(5):
        arm.set_q(rng.uniform(-1, 1, 7), np.zeros(7))
        M = arm.minv()
        assert np.allclose(M, M.T, atol=1e-9)
        assert np.all(np.linalg.eigvalsh(M) > 0)


def test_manipulability_matches_franka_dh_table():
    """a15: the oracle's Yoshikawa index (panda.py:314-317) against an independent statement of the kinematics --
    Franka's published modified-DH table (what roboticstoolbox's DH Panda uses) with a numerical Jacobian."""
    import oracle_lib as O

    def fk(q):
        aa = [0, 0, 0, 0.0825, -0.0825, 0, 0.088, 0]
        dd = [0.333, 0, 0.316, 0, 0.384, 0, 0, 0.107]
        al = [0, -np.pi / 2, np.pi / 2, np.pi / 2, -np.pi / 2, np.pi / 2, np.pi / 2, 0]
        T = np.eye(4)
        for i in range(8):
            th = q[i] if i < 7 else 0.0
            ca, sa, ct, st = np.cos(al[i]), np.sin(al[i]), np.cos(th), np.sin(th)
            T = T @ np.array([[ct, -st, 0, aa[i]], [st * ca, ct * ca, -sa, -dd[i] * sa], [st * sa, ct * sa, ca, dd[i] * ca], [0, 0, 0, 1]])
        return T[:3, 3]

    def manip(q):
        J = np.zeros((3, 7))
        for i in range(7):
            e = np.zeros(7); e[i] = 1e-6
            J[:, i] = (fk(q + e) - fk(q - e)) / 2e-6
        return np.sqrt(max(np.linalg.det(J @ J.T), 0.0))

    arm = O.Arm(model=0, base=(-0.6, 0.0, 0.0))
    rng = np.random.default_rng(4)
    qs = [np.zeros(7), np.array([0.0, 0.41, 0.0, -1.85, 0.0, 2.26, 0.79])] + [rng.uniform(-2.5, 2.5, 7) for _ in range(50)]
    for q in qs:
        arm.set_q(q, np.zeros(7))
        assert abs(arm.manipulability() - manip(q)) < 1e-7
