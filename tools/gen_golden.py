#!/usr/bin/env python3
"""Generate golden vectors from the parts of the reference that ARE importable in the build container.

Only `panda_gym/utils.py` (numpy only) loads: every other module of the reference imports pybullet / gymnasium at
module top.  Its `distance` (float32 norm + np.round(., 6)) decides every success flag and sparse reward of the
path, so its exact float32 behaviour is pinned here.  Output: tests/golden/utils_vectors.json (committed).

    python tools/gen_golden.py        # needs /root/reference
"""
import importlib.util
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("ref_utils", "/root/reference/panda_gym/utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

rng = np.random.default_rng(20240611)
n = 4000
a = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
b = (a + rng.normal(0, 0.05, (n, 3))).astype(np.float32)
# rows engineered to sit on the 0.05 threshold and on rounding boundaries
for i in range(400):
    d = rng.normal(size=3); d /= np.linalg.norm(d)
    b[i] = (a[i].astype(np.float64) + d * (0.05 + rng.uniform(-2e-6, 2e-6))).astype(np.float32)
dist = ref.distance(a, b)
assert dist.dtype == np.float32
thr32 = np.float32(0.05)
pa = rng.uniform(-1, 1, (200, 3)); pb = rng.uniform(-1, 1, (200, 3)); pb[:5] = pa[:5]
with np.errstate(all="ignore"):
    unit = np.array([ref.unit_vector(x.copy(), y.copy()) for x, y in zip(pa, pb)])
out = {
    "source": "reference panda_gym/utils.py distance()/unit_vector(), numpy %s" % np.__version__,
    "a": a.tolist(), "b": b.tolist(), "distance_bits": dist.view(np.uint32).tolist(),
    "success_reach_ao": (dist < thr32).astype(int).tolist(),          # reach_ao.py:1260 (threshold np.float32(0.05))
    "sparse_reward_classic": (-(dist > 0.05).astype(np.float32)).tolist(),  # reach.py:84-87
    "unit_a": pa.tolist(), "unit_b": pb.tolist(), "unit": unit.tolist(),
}
os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
with open(os.path.join(ROOT, "tests", "golden", "utils_vectors.json"), "w") as f:
    json.dump(out, f)
print("wrote tests/golden/utils_vectors.json", n, "rows; on-threshold successes:", int((dist[:400] < thr32).sum()), "of 400")
