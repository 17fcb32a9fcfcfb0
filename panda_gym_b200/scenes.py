"""Benchmark scenes of ReachAO as static axis-aligned box sets.

The reference builds them with ``ReachAO.setup_benchmark_scenario`` (reach_ao.py:292-306): a json file per scene lists
URDF bodies with ``basePosition`` / ``baseOrientation`` / ``globalScaling``; every URDF is a tree of fixed joints whose
links carry one ``<box>`` collision each (bookshelves, countertop, industrial, kitchen, raised_shelves, tabletop,
tabletop2, tunnel, warehouse) or a single 8-vertex cuboid OBJ mesh (library, narrow_tunnel, workshop).  Bullet keeps
every link as its own collision object (reach_ao.py:136-148), so the robot-to-scene distances are minima over boxes.
No transform in these assets rotates a box off the world axes (the only rotations are half turns about z), hence
"axis-aligned box set"; a rotated asset is refused instead of approximated.

``load_scene_boxes`` parses a panda_gym checkout's ``assets/scenarios``; ``builtin_scene_boxes`` returns the same numbers
from ``data/benchmark_scenes.json`` (generated from the reference assets by tools/gen_scene_data.py, numbers only) for
users without the asset tree.  ``SCENES`` restates the per-scenario settings of the ``create_scenario_*`` builders
(reach_ao.py:308-516).
"""
import json
import os
import xml.etree.ElementTree as ET

import numpy as np

URDF_MARGIN = 0.001  # gUrdfDefaultCollisionMargin: btBoxShape / btConvexHullShape created by the URDF importer

# scenario name -> asset directory, goal box (reach_ao.py line of the builder), neutral joint values when the builder
# replaces Panda.neutral_joint_values (None = panda.py:55 default), whether the builder installs a pose randomizer
SCENES = {
    "bookshelves":    dict(asset="bookshelves", goal_lo=(0.6, -0.35, 0.2), goal_hi=(0.7, 0.35, 0.8), neutral=None),        # :449-455
    "warehouse":      dict(asset="tabletop2", goal_lo=(0.5, -0.3, 0.0), goal_hi=(0.85, 0.3, 0.3), neutral=None),            # :470-476 (loads tabletop2)
    "countertop":     dict(asset="countertop", goal_lo=(0.5, -0.3, 0.0), goal_hi=(0.85, 0.3, 0.3), neutral=None),           # :478-484
    "kitchen":        dict(asset="kitchen", goal_lo=(0.5, -0.3, 0.0), goal_hi=(0.85, 0.3, 0.3), neutral=None),              # :486-492
    "raised_shelves": dict(asset="raised_shelves", goal_lo=(0.5, -0.3, 0.0), goal_hi=(0.85, 0.3, 0.3), neutral=None),       # :494-500
    "tabletop":       dict(asset="tabletop", goal_lo=(0.5, -0.3, 0.0), goal_hi=(0.85, 0.3, 0.3), neutral=None),             # :502-508
    "tabletop2":      dict(asset="tabletop2", goal_lo=(0.5, -0.3, 0.0), goal_hi=(0.85, 0.3, 0.3), neutral=None),            # :510-516
    "industrial":     dict(asset="industrial", goal_lo=(0.5, -0.1, 0.55), goal_hi=(0.6, 0.1, 0.75), neutral=None),          # :393-411
    "tunnel":         dict(asset="tunnel", goal_lo=(0.55, 0.2, 0.2), goal_hi=(0.75, 0.4, 0.75),
                           neutral=(-1.0, -0.3, 0.0, -2.2, 0.0, 2.0, 0.7853981633974483)),                                  # :336-349
    "narrow_tunnel":  dict(asset="narrow_tunnel", goal_lo=(0.55, 0.2, 0.2), goal_hi=(0.75, 0.4, 0.75),
                           neutral=(-1.0, -0.3, 0.0, -2.2, 0.0, 2.0, 0.7853981633974483)),                                  # :308-334
    "workshop":       dict(asset="workshop", goal_lo=(-0.7, -0.7, 0.4), goal_hi=(0.1, -0.4, 0.7),
                           neutral=(0.00887326, -0.05377409, -0.03621967, -1.9094068, 0.08791409, 2.00265486, 0.76681184)),  # :351-391
    "workshop2":      dict(asset="workshop", goal_lo=(0.5, -0.15, 0.4), goal_hi=(0.6, 0.15, 0.5),
                           neutral=(0.00887326, -0.05377409, -0.03621967, -1.9094068, 0.08791409, 2.00265486, 0.76681184)),
    "workshop3":      dict(asset="workshop", goal_lo=(-0.7, -0.7, 0.4), goal_hi=(0.1, -0.4, 0.7),
                           neutral=(0.00887326, -0.05377409, -0.03621967, -1.9094068, 0.08791409, 2.00265486, 0.76681184)),
    "library":        dict(asset="library", goal_lo=(0.2, -0.3, 0.0), goal_hi=(0.7, 0.3, 0.6),
                           neutral=(0.0, 0.12001979, 0.0, -1.64029458, 0.02081271, 3.1, 0.77979846)),                       # :422-447
    "library1":       dict(asset="library", goal_lo=(0.5, -0.3, 0.0), goal_hi=(0.85, 0.3, 0.3),
                           neutral=(0.0, 0.12001979, 0.0, -1.64029458, 0.02081271, 3.1, 0.77979846)),   # set_joint_neutral ran before the reassignment
    "library2":       dict(asset="library", goal_lo=(-0.7, -0.4, 0.4), goal_hi=(-0.55, 0.4, 0.85),
                           neutral=(0.0, 0.12001979, 0.0, -1.64029458, 0.02081271, 3.1, 0.77979846)),
}
# create_stage_wall (reach_ao.py:457-468): one cuboid from create_obstacle_cuboid (half extents = `size`, btBoxShape safe margin),
# never re-sampled -- a one-box static scene
SCENES["wall"] = dict(asset=None, goal_lo=(0.45, -0.6, 0.1), goal_hi=(0.7, -0.1, 0.3),
                      neutral=(0.94551719, 0.65262327, 0.12742699, -1.74347465, -0.16996126, 1.97424632, 0.88058222),
                      boxes=((0.0, 0.0, 0.1, 0.2, 0.05, 0.3, 0.005),))
# kasys (reach_ao.py:413-420) references ../meshes/trolley.obj and wall.obj, which the reference repository does not ship:
# the reference cannot load it either.


def _rpy(r, p, y):
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                     [-sp, cp * sr, cp * cr]])


def _quat(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def _origin(el):
    o = el.find("origin") if el is not None else None
    xyz = np.array([float(v) for v in (o.get("xyz") if o is not None and o.get("xyz") else "0 0 0").split()])
    rpy = [float(v) for v in (o.get("rpy") if o is not None and o.get("rpy") else "0 0 0").split()]
    return _rpy(*rpy), xyz


def _axis_aligned(R, what):
    """R must map axes onto axes (up to the 3e-6 rad by which the assets' 3.14159 misses pi)."""
    Rr = np.round(R)
    if np.abs(R - Rr).max() > 1e-5 or not np.allclose(np.abs(Rr).sum(axis=0), 1) or not np.allclose(np.abs(Rr).sum(axis=1), 1):
        raise ValueError("%s: a box rotated off the world axes (oriented boxes are not supported)" % what)
    return Rr


def _obj_cuboid(path, scale):
    V = np.array([[float(x) for x in l.split()[1:4]] for l in open(path) if l.startswith("v ")]) * scale
    lo, hi = V.min(axis=0), V.max(axis=0)
    corners = np.array([[(lo, hi)[i][0], (lo, hi)[j][1], (lo, hi)[k][2]] for i in (0, 1) for j in (0, 1) for k in (0, 1)])
    for c in corners:  # all 8 corners of the bounding box are vertices: the convex hull Bullet builds IS that box
        if np.abs(V - c).sum(axis=1).min() > 1e-9 * max(1.0, np.abs(V).max()):
            raise ValueError("%s: the convex hull of this mesh is not an axis-aligned cuboid" % path)
    return 0.5 * (lo + hi), 0.5 * (hi - lo)


def urdf_boxes(path, base_position=(0, 0, 0), base_orientation=(0, 0, 0, 1), global_scaling=1.0):
    """World-frame (centre[3], half[3], margin) of every collision box of one URDF body, as Bullet places them."""
    root = ET.parse(path).getroot()
    parent = {}
    for j in root.findall("joint"):
        if j.get("type") != "fixed":
            raise ValueError("%s: joint %s is not fixed" % (path, j.get("name")))
        parent[j.find("child").get("link")] = (j.find("parent").get("link"),) + _origin(j)
    Rb, pb = _quat(base_orientation), np.asarray(base_position, float)
    out = []
    for link in root.findall("link"):
        # link frame in the body frame: the chain of fixed joints up to the root
        R, p, name = np.eye(3), np.zeros(3), link.get("name")
        while name in parent:
            name, Rj, pj = parent[name]
            R, p = Rj @ R, Rj @ p + pj
        for col in link.findall("collision"):
            Rc, pc = _origin(col)
            geom = col.find("geometry")
            box, mesh = geom.find("box"), geom.find("mesh")
            if box is not None:
                half, local, grow = 0.5 * np.array([float(v) for v in box.get("size").split()]), np.zeros(3), 0.0
            elif mesh is not None:
                sc = np.array([float(v) for v in (mesh.get("scale") or "1 1 1").split()])
                local, half = _obj_cuboid(os.path.normpath(os.path.join(os.path.dirname(path), mesh.get("filename"))), sc)
                grow = URDF_MARGIN  # a convex hull carries its margin outside the vertices, a btBoxShape inside its faces
            else:
                raise ValueError("%s: unsupported collision geometry in link %s" % (path, link.get("name")))
            Rw = Rb @ R @ Rc
            A = _axis_aligned(Rw, path)
            centre = pb + Rb @ (global_scaling * (R @ (Rc @ local + pc) + p))
            out.append((centre, np.abs(A) @ (global_scaling * half) + grow, URDF_MARGIN))
    return out


def load_scene_boxes(assets_dir, scenario):
    """All collision boxes of benchmark scenario ``scenario`` from a panda_gym ``assets/scenarios`` directory:
    float64 array [n, 7] = centre, half extents, collision margin."""
    if SCENES[scenario].get("boxes") is not None:
        return np.array(SCENES[scenario]["boxes"], dtype=np.float64)
    asset = SCENES[scenario]["asset"]
    d = os.path.join(assets_dir, asset)
    with open(os.path.join(d, asset + ".json")) as f:
        bodies = json.load(f)
    rows = []
    for kw in bodies.values():
        for c, h, m in urdf_boxes(os.path.join(d, "urdf", kw["fileName"]), kw.get("basePosition", (0, 0, 0)),
                                  kw.get("baseOrientation", (0, 0, 0, 1)), kw.get("globalScaling", 1.0)):
            rows.append(np.concatenate([c, h, [m]]))
    return np.array(rows, dtype=np.float64)


_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "benchmark_scenes.json")


def builtin_scene_boxes(scenario):
    if SCENES[scenario].get("boxes") is not None:
        return np.array(SCENES[scenario]["boxes"], dtype=np.float64)
    with open(_DATA) as f:
        return np.array(json.load(f)[SCENES[scenario]["asset"]], dtype=np.float64)


def scene_boxes(scenario, assets_dir=None):
    return load_scene_boxes(assets_dir, scenario) if assets_dir else builtin_scene_boxes(scenario)
