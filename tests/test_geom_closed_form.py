"""The closed-form cylinder-box bounds that replaced GJK on the per-substep collision path (csrc/pg_geom.cuh: seg_box,
cyl_box_bound), compiled for the HOST and checked against the oracle's independent fp64 GJK and against brute force.
Properties: the segment-box distance is exact; lb <= dist(cylinder, box) <= ub always; lb == dist when `exact`."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "native", "geom_host.cu")
OUT = os.path.join(ROOT, "tests", "native", "_build", "libgeomhost.so")


@pytest.fixture(scope="module")
def geom():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    hdr = os.path.join(ROOT, "panda_gym_b200", "csrc", "pg_geom.cuh")
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < max(os.path.getmtime(SRC), os.path.getmtime(hdr)):
        subprocess.run(["nvcc", "-std=c++17", "-O2", "-Xcompiler", "-fPIC", "-shared", "-o", OUT, SRC], check=True)
    L = ctypes.CDLL(OUT)
    P, D = ctypes.c_void_p, ctypes.c_double
    L.geom_cyl_box_bound_f64.argtypes = [P, P, D, D, P, P]
    L.geom_cyl_box_bound_f32.argtypes = [P, P, D, D, P, P]
    L.geom_seg_box_f64.argtypes = [P, P, D, P, P]
    for f in (L.geom_gjk_decide_f64, L.geom_gjk_decide_f32):
        f.argtypes = [P, P, D, D, P, D]; f.restype = D
    return L


def _dp(x):
    return x.ctypes.data_as(ctypes.c_void_p)


def _cases(rng, n):
    for _ in range(n):
        a = rng.normal(size=3); a /= np.linalg.norm(a)
        if rng.random() < 0.25:   # axis-parallel / in-plane axes: the degenerate (flat minimum) configurations
            a = np.eye(3)[rng.integers(3)] * rng.choice([-1.0, 1.0])
        e = rng.uniform(0.02, 0.3, 3)
        c = rng.uniform(-0.5, 0.5, 3)
        yield c, a, rng.choice([0.025, 0.04, 0.06]), rng.uniform(0.005, 0.15), e


def test_segment_box_distance_is_exact(geom):
    rng = np.random.default_rng(0)
    out = np.zeros(2)
    for c, a, r, h, e in _cases(rng, 400):
        geom.geom_seg_box_f64(_dp(c), _dp(a), h, _dp(e), _dp(out))
        t = np.linspace(-h, h, 20001)
        p = c[None, :] + t[:, None] * a[None, :]
        d = np.linalg.norm(p - np.clip(p, -e, e), axis=1)
        assert out[0] <= d.min() + 1e-12 and abs(out[0] - d.min()) < 1e-8 + 1e-4 * h, (out, d.min())


def test_bounds_bracket_the_oracle_gjk(geom, oracle):
    L = oracle.lib()
    rng = np.random.default_rng(1)
    out = np.zeros(7); out32 = np.zeros(7)
    pa = np.zeros(3); pb = np.zeros(3); zero = np.zeros(3)
    n_exact = n_slanted = 0
    for c, a, r, h, e in _cases(rng, 3000):
        d = L.pgo_gjk_cyl_box(_dp(c), _dp(a), float(r), float(h), _dp(zero), _dp(e), _dp(pa), _dp(pb))
        geom.geom_cyl_box_bound_f64(_dp(c), _dp(a), float(r), float(h), _dp(e), _dp(out))
        lb, exact, s, ub = out[0], out[1] > 0, out[2], out[6]
        assert lb <= d + 1e-9 and d <= max(ub, 0.0) + 1e-9, (lb, d, ub)
        if exact:
            n_exact += 1
            assert abs(lb - d) < 1e-9, (lb, d)
            if d > 1e-3:   # direction robot -> box equals the GJK pair's (GJK pins its closest POINTS only to ~sqrt(tolerance))
                u = (pb - pa) / np.linalg.norm(pb - pa)
                assert np.abs(u - out[3:6]).max() < 2e-7 / d + 1e-6, (d, u, out[3:6])
        else:
            n_slanted += 1
            assert lb >= s - r - 1e-12   # never weaker than the capsule bound
        geom.geom_cyl_box_bound_f32(_dp(c), _dp(a), float(r), float(h), _dp(e), _dp(out32))
        if exact and out32[1] > 0:
            assert abs(out32[0] - d) < 2e-6
    assert n_exact > 500 and n_slanted > 500


def test_thresholded_gjk_decides_like_the_converged_one(geom, oracle):
    """gjk_cyl_box_decide (the reset's accept / reject tests in fp64, the flag pass's sign test in fp32) stops as soon as one of
    GJK's two bounds settles `distance <= thr`: whatever it returns must lie on the same side of thr as the oracle's converged
    fp64 GJK distance, for thresholds well inside, at and well outside the distance (ties within the arithmetic's resolution
    excluded), and must itself be a valid bound (<= thr and >= d when within; > thr and <= d when beyond)."""
    L = oracle.lib()
    rng = np.random.default_rng(5)
    pa = np.zeros(3); pb = np.zeros(3); zero = np.zeros(3)
    n_in = n_out = 0
    for c, a, r, h, e in _cases(rng, 2500):
        d = L.pgo_gjk_cyl_box(_dp(c), _dp(a), float(r), float(h), _dp(zero), _dp(e), _dp(pa), _dp(pb))
        for thr in (0.5 * d, d - 1e-6, d + 1e-6, 1.5 * d + 1e-3, 0.031, 0.051):
            if thr <= 0:
                continue
            v = geom.geom_gjk_decide_f64(_dp(c), _dp(a), float(r), float(h), _dp(e), float(thr))
            if abs(d - thr) > 1e-9:
                assert (v <= thr) == (d <= thr), (d, thr, v)
                if d <= thr:
                    n_in += 1
                    assert v >= d - 1e-9
                else:
                    n_out += 1
                    assert v <= d + 1e-9
            v32 = geom.geom_gjk_decide_f32(_dp(c), _dp(a), float(r), float(h), _dp(e), float(thr))
            if abs(d - thr) > 2e-5 + 1e-4 * d:   # fp32 resolution of the inputs and of GJK's own stop rule
                assert (v32 <= np.float32(thr)) == (d <= thr), (d, thr, v32)
    assert n_in > 2000 and n_out > 2000
