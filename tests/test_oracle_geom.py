"""Oracle geometry: GJK against brute force, analytic pairs against first principles, table fast path vs GJK."""
import numpy as np


def _brute_cyl_box(rng, c, a, r, h, b, e, n=60000):
    u = np.cross(a, [1.0, 0, 0])
    if np.linalg.norm(u) < 1e-3:
        u = np.cross(a, [0, 1.0, 0])
    u /= np.linalg.norm(u); w = np.cross(a, u)
    t = rng.uniform(-h, h, n); th = rng.uniform(0, 2 * np.pi, n); rr = r * np.sqrt(rng.uniform(0, 1, n))
    rr[: n // 2] = r; t[: n // 4] = h; t[n // 4: n // 2] = -h
    P = c + np.outer(t, a) + np.outer(rr * np.cos(th), u) + np.outer(rr * np.sin(th), w)
    d = np.maximum(np.abs(P - b) - e, 0)
    return np.sqrt((d ** 2).sum(1)).min()


def test_gjk_cylinder_box_vs_brute_force(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(0)
    for i in range(120):
        c = rng.uniform(-.5, .5, 3); a = rng.normal(size=3); a /= np.linalg.norm(a)
        r = rng.uniform(.02, .08); h = rng.uniform(.005, .15)
        b = rng.uniform(-.5, .5, 3); e = rng.uniform(.02, .3, 3)
        if i % 10 == 0:  # the table: a huge box (this case exposed a degenerate-tetrahedron bug in round 1)
            b = np.array([0, 0, -0.2]); e = np.array([0.98, 0.63, 0.18]); c = np.array([rng.uniform(-.8, .8), rng.uniform(-.5, .5), rng.uniform(0.05, .6)])
        pa = np.zeros(3); pb = np.zeros(3)
        d = L.pgo_gjk_cyl_box(oracle.dp(c), oracle.dp(a), r, h, oracle.dp(b), oracle.dp(e), oracle.dp(pa), oracle.dp(pb))
        bf = _brute_cyl_box(rng, c, a, r, h, b, e)
        assert d <= bf + 1e-9, "GJK must be a lower bound of any sampled distance"
        assert bf - d < 4e-3
        if d > 0:
            assert abs(np.linalg.norm(pa - pb) - d) < 1e-9
            assert np.all(np.abs(pb - b) <= e + 1e-9)


def test_cylinder_over_table_closed_form(oracle):
    """A cylinder whose footprint lies inside the table top: distance = lowest point height (minus margins)."""
    rng = np.random.default_rng(1)
    arm = oracle.Arm(model=0)
    table = oracle.make_obstacle("box", (0, 0, -0.2), (1.0, 0.65, 0.2))
    L = oracle.lib()
    checked = 0
    for _ in range(200):
        q = np.array([0, -0.3, 0, -2.2, 0, 2.0, np.pi / 4]) + rng.uniform(-0.6, 0.6, 7)
        arm.set_q(q, np.zeros(7))
        for prim, link in [(6, 2), (9, 3), (12, 4), (18, 5), (21, 6), (30, 9)]:  # cylinders of links 3..7, ee
            pa = np.zeros(3); pb = np.zeros(3)
            import ctypes
            d = L.pgo_prim_obstacle_distance(arm.h, prim, ctypes.byref(table), oracle.dp(pa), oracle.dp(pb))
            if d > 0.01 and abs(pa[0]) < 0.7 and abs(pa[1]) < 0.4:
                # closest point on the table is straight below the closest point on the cylinder
                assert abs(pb[2]) < 1e-9 and abs(pa[2] - d) < 1e-9
                assert np.allclose(pa[:2], pb[:2], atol=1e-6)
                checked += 1
    assert checked > 300


def test_sphere_pairs(oracle):
    arm = oracle.Arm(model=0)
    for d, v in enumerate([0, -0.3, 0, -2.2, 0, 2.0, np.pi / 4]):
        arm.reset_joint(d, v)
    # link index 1 (panda_link2) sphere primitives: brute-force the link distance to a sphere obstacle
    rng = np.random.default_rng(2)
    for _ in range(50):
        o = rng.uniform([-0.3, -0.5, 0.0], [0.8, 0.5, 0.9])
        ob = oracle.make_obstacle("sphere", o, 0.05)
        for link in (0, 3, 6):
            dd, pa, pb = arm.link_obstacle_distance(link, ob)
            if dd > 1e-6:
                assert abs(np.linalg.norm(pb - pa) - dd) < 1e-9          # points sit on the two surfaces
                assert abs(np.linalg.norm(pb - o) - 0.05) < 1e-9
    assert abs(oracle.lib().pgo_box_margin(oracle.dp(np.array([0.05, 0.05, 0.05]))) - 0.005) < 1e-15
    assert abs(oracle.lib().pgo_box_margin(oracle.dp(np.array([1.0, 0.65, 0.2]))) - 0.02) < 1e-15
