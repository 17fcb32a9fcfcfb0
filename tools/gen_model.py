#!/usr/bin/env python3
"""Generate the numeric robot-model tables used by the oracle and the CUDA kernels.

Run in the build container (where /root/reference exists):

    python tools/gen_model.py

Reads the two Panda URDFs of the reference
  * panda_gym/assets/robots/franka_panda_custom_0/panda.urdf  (loaded by Panda, panda.py:54)
  * panda_gym/assets/robots/franka_panda/panda.urdf           (stock; used by test/pybullet_test.py KATs)
and writes
  * oracle/panda_model_data.h            generic link/primitive tables (double) for both models

Only NUMBERS leave the reference (link frames, masses, primitive sizes, mesh AABBs); no source.

Rules restated from Bullet / PyBullet (all [UPSTREAM-RECALL]):
  * link index = pre-order DFS over joints in file order, base = -1
  * multibody link frame sits at the inertial origin (COM)
  * link inertia is NOT read from the file (no URDF_USE_INERTIA_FROM_FILE): it is the solid-box
    inertia of the AABB of the link's collision compound, expressed in the inertial frame:
      - sphere child: AABB = centre +- r
      - cylinder child: 32-gon convex hull (vertices hit +-r on x,y), margin 0.001 baked into the
        cached local AABB and added once more by getAabb  -> half = (r, r, L/2) + 0.002, rotated by |R|
      - mesh child: compound of convex hulls -> half = vertex half-extent + MESH_PAD
      - the per-link compound has margin 0.001 (added to every half extent of the union AABB);
        I = m/12 * (ly^2+lz^2, lx^2+lz^2, lx^2+ly^2) from the AABB EXTENTS only
      - a link without collision elements gets an empty compound: AABB extents 0 -> I = 0
"""
import math
import os
import sys
import xml.etree.ElementTree as ET

REF = "/root/reference/panda_gym/assets/robots"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "panda_model_data.h")

HULL_MARGIN = 0.001  # PyBullet default collision margin for URDF convex hulls
CYL_PAD = 2 * HULL_MARGIN  # cached local AABB (+margin) and getAabb (+margin)
MESH_PAD = 3 * HULL_MARGIN  # as above, + margin of the per-mesh compound (createConvexHullFromShapes)
LINK_COMPOUND_MARGIN = HULL_MARGIN  # convertLinkCollisionShapes: compoundShape->setMargin(default margin)
# The sum "mesh vertices + 4 margins" is PINNED by the reference KAT test/pybullet_test.py:173-187:
# it reproduces w = -2.969 rad/s to 4 digits (2.9690); 3 margins give 2.9835, 5 give 2.9545.


def rpy_to_R(r, p, y):
    """URDF fixed-axis rpy -> rotation matrix (child->parent), R = Rz(y) Ry(p) Rx(r)."""
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return [
        [cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
        [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
        [-sp, cp * sr, cp * cr],
    ]


def floats(s, n=3):
    v = [float(x) for x in s.split()]
    assert len(v) == n, s
    return v


def origin_of(el):
    o = el.find("origin") if el is not None else None
    if o is None:
        return [0.0, 0.0, 0.0], [0.0, 0.0, 0.0]
    return floats(o.get("xyz", "0 0 0")), floats(o.get("rpy", "0 0 0"))


def obj_aabb(path):
    lo = [1e30] * 3
    hi = [-1e30] * 3
    with open(path) as f:
        for line in f:
            if line.startswith("v "):
                v = [float(x) for x in line.split()[1:4]]
                for k in range(3):
                    lo[k] = min(lo[k], v[k])
                    hi[k] = max(hi[k], v[k])
    return lo, hi


def parse(urdf_path):
    root = ET.parse(urdf_path).getroot()
    base_dir = os.path.dirname(urdf_path)
    links = {}
    for l in root.findall("link"):
        name = l.get("name")
        inert = l.find("inertial")
        com, com_rpy = origin_of(inert)
        assert all(abs(x) < 1e-12 for x in com_rpy), "inertial rpy != 0 unsupported"
        mass = float(inert.find("mass").get("value")) if inert is not None else 0.0
        prims = []
        for c in l.findall("collision"):
            xyz, rpy = origin_of(c)
            g = c.find("geometry")
            if g.find("sphere") is not None:
                prims.append(dict(type="sphere", xyz=xyz, rpy=rpy, r=float(g.find("sphere").get("radius")), L=0.0))
            elif g.find("cylinder") is not None:
                cy = g.find("cylinder")
                prims.append(dict(type="cylinder", xyz=xyz, rpy=rpy, r=float(cy.get("radius")), L=float(cy.get("length"))))
            elif g.find("mesh") is not None:
                fn = g.find("mesh").get("filename").replace("package://", "")
                lo, hi = obj_aabb(os.path.join(base_dir, fn))
                prims.append(dict(type="mesh", xyz=xyz, rpy=rpy, lo=lo, hi=hi, r=0.0, L=0.0))
            else:
                raise ValueError("unsupported geometry")
        links[name] = dict(name=name, com=com, mass=mass, prims=prims)
    joints = []
    for j in root.findall("joint"):
        xyz, rpy = origin_of(j)
        ax = j.find("axis")
        axis = floats(ax.get("xyz")) if ax is not None else [1.0, 0.0, 0.0]
        lim = j.find("limit")
        jt = j.get("type")
        lower = upper = 0.0
        has_limit = 0
        if lim is not None and jt in ("revolute", "prismatic"):
            lower = float(lim.get("lower", "0"))
            upper = float(lim.get("upper", "0"))
            has_limit = 1 if lower <= upper else 0
        joints.append(dict(name=j.get("name"), type=jt, parent=j.find("parent").get("link"),
                           child=j.find("child").get("link"), xyz=xyz, rpy=rpy, axis=axis,
                           lower=lower, upper=upper, has_limit=has_limit))
    children = {j["child"] for j in joints}
    roots = [n for n in links if n not in children]
    assert len(roots) == 1
    order = []  # (link name, joint, parent index)

    def dfs(name, parent_idx):
        for j in joints:  # file order
            if j["parent"] == name:
                idx = len(order)
                order.append((j["child"], j, parent_idx))
                dfs(j["child"], idx)

    dfs(roots[0], -1)
    return links, roots[0], order


def prim_aabb_in_inertial(p, com):
    """AABB (lo, hi) of one collision child in the link's inertial frame, Bullet-style."""
    R = rpy_to_R(*p["rpy"])
    c = [p["xyz"][k] - com[k] for k in range(3)]
    if p["type"] == "sphere":
        h = [p["r"]] * 3
        ctr = c
    elif p["type"] == "cylinder":
        hl = [p["r"] + CYL_PAD, p["r"] + CYL_PAD, p["L"] / 2 + CYL_PAD]
        h = [sum(abs(R[i][k]) * hl[k] for k in range(3)) for i in range(3)]
        ctr = c
    else:  # mesh
        lc = [(p["lo"][k] + p["hi"][k]) / 2 for k in range(3)]
        hl = [(p["hi"][k] - p["lo"][k]) / 2 + MESH_PAD for k in range(3)]
        h = [sum(abs(R[i][k]) * hl[k] for k in range(3)) for i in range(3)]
        ctr = [c[i] + sum(R[i][k] * lc[k] for k in range(3)) for i in range(3)]
    return [ctr[k] - h[k] for k in range(3)], [ctr[k] + h[k] for k in range(3)]


def bullet_inertia(link):
    if link["mass"] == 0.0 or not link["prims"]:
        return [0.0, 0.0, 0.0]
    lo = [1e30] * 3
    hi = [-1e30] * 3
    for p in link["prims"]:
        l, h = prim_aabb_in_inertial(p, link["com"])
        for k in range(3):
            lo[k] = min(lo[k], l[k])
            hi[k] = max(hi[k], h[k])
    lx, ly, lz = (hi[k] - lo[k] + 2 * LINK_COMPOUND_MARGIN for k in range(3))
    m = link["mass"]
    return [m / 12.0 * (ly * ly + lz * lz), m / 12.0 * (lx * lx + lz * lz), m / 12.0 * (lx * lx + ly * ly)]


def fmt(v):
    return "{" + ", ".join(repr(float(x)) for x in v) + "}"


def emit_model(tag, urdf_path, out):
    links, root, order = parse(urdf_path)
    JT = {"fixed": 0, "revolute": 1, "continuous": 1, "prismatic": 2}
    prim_rows = []
    link_rows = []
    def add_prims(L, idx):
        for p in L["prims"]:
            if p["type"] == "mesh":
                continue  # meshes only feed the inertia rule (stock model); no distance queries on them
            Rp = rpy_to_R(*p["rpy"])
            axis = [Rp[0][2], Rp[1][2], Rp[2][2]]  # cylinder z-axis in link frame
            prim_rows.append("  {%d, %d, %s, %s, %r, %r}" % (
                idx, 0 if p["type"] == "sphere" else 1, fmt(p["xyz"]), fmt(axis), p["r"], p["L"] / 2))

    for idx, (name, j, parent) in enumerate(order):
        L = links[name]
        R = rpy_to_R(*j["rpy"])
        first = len(prim_rows)
        add_prims(L, idx)
        inertia = bullet_inertia(L)
        link_rows.append("  {/*%2d %s*/ %d, %d, %s, %s, %s, %d, %r, %r, %s, %r, %s, %d, %d}" % (
            idx, name, parent, JT[j["type"]], fmt(j["xyz"]), fmt([x for row in R for x in row]), fmt(j["axis"]),
            j["has_limit"], j["lower"], j["upper"], fmt(L["com"]), L["mass"], fmt(inertia), first,
            len(prim_rows) - first))
    add_prims(links[root], -1)  # base link (fixed): only body-level closest-point queries see it
    out.append("/* %s : %s */" % (tag, urdf_path.replace("/root/reference/", "")))
    out.append("static const pgo_link_t %s_LINKS[%d] = {" % (tag, len(link_rows)))
    out.append(",\n".join(link_rows))
    out.append("};")
    out.append("static const pgo_prim_t %s_PRIMS[%d] = {" % (tag, max(1, len(prim_rows))))
    out.append(",\n".join(prim_rows) if prim_rows else "  {0, 0, {0,0,0}, {0,0,1}, 0, 0}")
    out.append("};")
    out.append("#define %s_NLINKS %d" % (tag, len(link_rows)))
    out.append("#define %s_NPRIMS %d" % (tag, len(prim_rows)))
    out.append("")
    return links, order


def main():
    out = [
        "/* GENERATED by tools/gen_model.py -- do not edit. Numbers only (link frames, masses, Bullet-rule",
        " * inertias, collision primitives) derived from the reference URDFs; see the generator's docstring. */",
        "#ifndef PGO_MODEL_DATA_H",
        "#define PGO_MODEL_DATA_H",
        "typedef struct {",
        "  int parent;        /* multibody parent link index, -1 = base */",
        "  int jtype;         /* 0 fixed, 1 revolute, 2 prismatic */",
        "  double org_xyz[3]; /* joint origin in the parent's URDF link frame */",
        "  double org_R[9];   /* row-major rotation child->parent at q = 0 */",
        "  double axis[3];    /* joint axis in the child frame */",
        "  int has_limit; double lower, upper;",
        "  double com[3];     /* inertial origin in the link frame */",
        "  double mass; double inertia[3]; /* Bullet AABB-box rule, diag in the inertial frame */",
        "  int prim0, nprim;  /* collision primitives of this link */",
        "} pgo_link_t;",
        "typedef struct {",
        "  int link; int type; /* 0 sphere, 1 cylinder */",
        "  double xyz[3];      /* centre in the link frame */",
        "  double axis[3];     /* cylinder axis in the link frame */",
        "  double radius, half_len;",
        "} pgo_prim_t;",
        "",
    ]
    emit_model("PGO_CUSTOM0", os.path.join(REF, "franka_panda_custom_0/panda.urdf"), out)
    emit_model("PGO_STOCK", os.path.join(REF, "franka_panda/panda.urdf"), out)
    out.append("#endif")
    with open(OUT, "w") as f:
        f.write("\n".join(out) + "\n")
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
