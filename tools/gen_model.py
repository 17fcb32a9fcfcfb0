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
    links["__root__"] = root
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


def mat_vec(R, v):
    return [sum(R[i][k] * v[k] for k in range(3)) for i in range(3)]


def mat_mul(A, B):
    return [[sum(A[i][k] * B[k][j] for k in range(3)) for j in range(3)] for i in range(3)]


def snap(x):
    """Joint-origin rotations of the Panda are multiples of 90 deg written with 11 digits; snap |x| < 1e-9."""
    for t in (0.0, 1.0, -1.0):
        if abs(x - t) < 1e-9:
            return t
    return x


def ff(x):
    return repr(float(x)) + "f"


def emit_cuda(links, order, path):
    """Specialised constants of franka_panda_custom_0 for the sm_100a kernels: 7 revolute z-axis joints whose
    origins are Rx(0 / +-90 deg); everything rigidly attached to link7 (link8, hand, ee, fingers) is expressed
    in the link7 frame."""
    o = [
        "// GENERATED by tools/gen_model.py from the reference's franka_panda_custom_0/panda.urdf (numbers only).",
        "// Link i (1..7) = child of revolute joint i; frames are the URDF link frames (origin on the joint axis).",
        "#pragma once",
        "namespace pgm {",
        "template <int I> struct Link;   // kinematic + inertial constants of moving link I (1..7)",
        "template <int K> struct Sub;    // rigid bodies fixed to link7 (own mass / COM / inertia; Bullet damps each)",
        "template <int K> struct Prim;   // collision primitives of the 9 ReachAO collision links",
    ]
    names = [n for n, _, _ in order]
    # transforms of links 7.. relative to link7 (index 6)
    rel = {6: ([[1, 0, 0], [0, 1, 0], [0, 0, 1]], [0.0, 0.0, 0.0])}
    for idx in range(7, len(order)):
        name, j, parent = order[idx]
        assert j["type"] == "fixed"
        Rp, pp = rel[parent]
        R = rpy_to_R(*j["rpy"])
        rel[idx] = (mat_mul(Rp, R), [pp[k] + mat_vec(Rp, j["xyz"])[k] for k in range(3)])
    for i in range(7):
        name, j, parent = order[i]
        L = links[name]
        R = [[snap(x) for x in row] for row in rpy_to_R(*j["rpy"])]
        if R == [[1, 0, 0], [0, 1, 0], [0, 0, 1]]:
            rot = 0
        elif R == [[1, 0, 0], [0, 0, -1], [0, 1, 0]]:
            rot = 1   # Rx(+90): (x, y, z) -> (x, -z, y)
        elif R == [[1, 0, 0], [0, 0, 1], [0, -1, 0]]:
            rot = -1  # Rx(-90): (x, y, z) -> (x, z, -y)
        else:
            raise ValueError("unexpected joint origin rotation %r" % R)
        assert j["axis"] == [0.0, 0.0, 1.0] and j["type"] == "revolute"
        I = bullet_inertia(L)
        o.append("template <> struct Link<%d> {  // %s" % (i + 1, name))
        o.append("  static constexpr int rot = %d;  // joint origin = Rx(rot * 90 deg)" % rot)
        o.append("  static constexpr float px = %s, py = %s, pz = %s;  // joint origin in the parent frame" % tuple(ff(x) for x in j["xyz"]))
        o.append("  static constexpr float cx = %s, cy = %s, cz = %s;  // COM" % tuple(ff(x) for x in L["com"]))
        o.append("  static constexpr float m = %s, ixx = %s, iyy = %s, izz = %s;  // Bullet AABB-box inertia" % (ff(L["mass"]), ff(I[0]), ff(I[1]), ff(I[2])))
        o.append("  static constexpr float lo = %s, hi = %s;  // joint limits" % (ff(j["lower"]), ff(j["upper"])))
        o.append("};")
    # sub-bodies on link7 with mass
    k = 0
    for idx in range(7, len(order)):
        L = links[names[idx]]
        if L["mass"] == 0.0:
            continue
        R, p = rel[idx]
        c = [p[t] + mat_vec(R, L["com"])[t] for t in range(3)]
        I = bullet_inertia(L)
        D = [[I[0], 0, 0], [0, I[1], 0], [0, 0, I[2]]]
        Rt = [[R[j2][i2] for j2 in range(3)] for i2 in range(3)]
        IR = mat_mul(mat_mul(R, D), Rt)
        o.append("template <> struct Sub<%d> {  // %s, in the link7 frame" % (k, names[idx]))
        o.append("  static constexpr float cx = %s, cy = %s, cz = %s, m = %s;" % (ff(c[0]), ff(c[1]), ff(c[2]), ff(L["mass"])))
        o.append("  static constexpr float ixx = %s, iyy = %s, izz = %s, ixy = %s, ixz = %s, iyz = %s;" % (
            ff(IR[0][0]), ff(IR[1][1]), ff(IR[2][2]), ff(IR[0][1]), ff(IR[0][2]), ff(IR[1][2])))
        o.append("};")
        k += 1
    o.append("constexpr int kNumSub = %d;" % k)
    # end-effector = COM of link 11 (panda.py:68), in the link7 frame, + its orientation
    R, p = rel[11]
    c = [p[t] + mat_vec(R, links[names[11]]["com"])[t] for t in range(3)]
    o.append("// end-effector = COM of link 11 (%s), link7 frame; link-11 frame origin (IK drives this point)" % names[11])
    o.append("constexpr float kEEx = %s, kEEy = %s, kEEz = %s;" % tuple(ff(x) for x in c))
    o.append("constexpr float kEEox = %s, kEEoy = %s, kEEoz = %s;" % tuple(ff(x) for x in p))
    o.append("constexpr float kEErot_c = %s, kEErot_s = %s;  // link11 axes = Rz(angle) * link7 axes: cos, sin" % (ff(R[0][0]), ff(R[1][0])))
    # COM of link index 0 (sample_obstacle_wang uses get_link_position(0))
    # collision primitives: 9 collision links, link8 / ee expressed in the link7 frame
    coll = [0, 1, 2, 3, 4, 5, 6, 7, 9, 8]  # slot 9 = panda_hand: not a ReachAO collision link, used by the contact pass of Push / PickAndPlace
    k = 0
    o.append("// collision primitives: slot = index into the 9 ReachAO collision links (link1..link8, panda_ee);")
    o.append("// frame = moving link whose frame the numbers are in (1..7); type 0 sphere, 1 cylinder")
    for slot, idx in enumerate(coll):
        L = links[names[idx]]
        frame = min(idx, 6)
        R, p = rel[idx] if idx >= 6 else ([[1, 0, 0], [0, 1, 0], [0, 0, 1]], [0, 0, 0])
        for pr in L["prims"]:
            Rp = rpy_to_R(*pr["rpy"])
            axis = mat_vec(R, [Rp[0][2], Rp[1][2], Rp[2][2]])
            ctr = [p[t] + mat_vec(R, pr["xyz"])[t] for t in range(3)]
            axis = [0.0 if abs(x) < 1e-12 else x for x in axis]
            o.append("template <> struct Prim<%d> { static constexpr int slot = %d, frame = %d, type = %d; "
                     "static constexpr float x = %s, y = %s, z = %s, ax = %s, ay = %s, az = %s, r = %s, h = %s; };" % (
                         k, slot, frame + 1, 0 if pr["type"] == "sphere" else 1, ff(ctr[0]), ff(ctr[1]), ff(ctr[2]),
                         ff(axis[0]), ff(axis[1]), ff(axis[2]), ff(pr["r"]), ff(pr["L"] / 2)))
            k += 1
    o.append("constexpr int kNumPrim = %d;" % k)
    o.append("}  // namespace pgm")
    with open(path, "w") as f:
        f.write("\n".join(o) + "\n")
    print("wrote", path)


def emit_neutral(links, root, order, path):
    """World placement (base at the origin) of all 36 collision primitives with the arm at Panda's neutral pose
    (panda.py:67), in double: the reset path's rejection sampling (reach_ao.py:1101-1181) only ever queries the
    robot at this pose (randomize_robot_pose=False)."""
    neutral = [0.0, -0.3, 0.0, -2.2, 0.0, 2.0, math.pi / 4]
    I3 = [[1.0, 0, 0], [0, 1.0, 0], [0, 0, 1.0]]
    Rw = {-1: I3}
    ow = {-1: [0.0, 0.0, 0.0]}
    rows = []
    names = [n for n, _, _ in order]

    def add(L, R, o):
        for pr in L["prims"]:
            Rp = rpy_to_R(*pr["rpy"])
            ax = mat_vec(R, [Rp[0][2], Rp[1][2], Rp[2][2]])
            c = [o[k] + mat_vec(R, pr["xyz"])[k] for k in range(3)]
            rows.append("  {%d, {%r, %r, %r}, {%r, %r, %r}, %r, %r}" % (
                0 if pr["type"] == "sphere" else 1, c[0], c[1], c[2], ax[0], ax[1], ax[2], pr["r"], pr["L"] / 2))

    for idx, (name, j, parent) in enumerate(order):
        R = rpy_to_R(*j["rpy"])
        if j["type"] == "revolute":
            q = neutral[idx]
            c, s = math.cos(q), math.sin(q)
            R = mat_mul(R, [[c, -s, 0], [s, c, 0], [0, 0, 1]])
        Rw[idx] = mat_mul(Rw[parent], R)
        ow[idx] = [ow[parent][k] + mat_vec(Rw[parent], j["xyz"])[k] for k in range(3)]
        add(links[name], Rw[idx], ow[idx])
    add(links[root], I3, [0.0, 0.0, 0.0])
    # the same primitives in the frame of the moving link that carries them (frame 0 = base, i = child of revolute joint i),
    # for the start-pose randomisers that put the arm anywhere (reach_ao.py:782-817): fp64 forward kinematics on the device
    Rl = {-1: I3}; ol = {-1: [0.0, 0.0, 0.0]}; fr = {-1: 0}
    lrows = []; jrows = []

    def add_local(L, R, o, frame):
        for pr in L["prims"]:
            Rp = rpy_to_R(*pr["rpy"])
            ax = mat_vec(R, [Rp[0][2], Rp[1][2], Rp[2][2]])
            c = [o[k] + mat_vec(R, pr["xyz"])[k] for k in range(3)]
            lrows.append("  {%d, %d, {%r, %r, %r}, {%r, %r, %r}, %r, %r}" % (
                frame, 0 if pr["type"] == "sphere" else 1, c[0], c[1], c[2], ax[0], ax[1], ax[2], pr["r"], pr["L"] / 2))

    nrev = 0
    for idx, (name, j, parent) in enumerate(order):
        R = rpy_to_R(*j["rpy"])
        if j["type"] == "revolute":
            nrev += 1
            assert fr[parent] == nrev - 1 and Rl[parent] == I3 and ol[parent] == [0.0, 0.0, 0.0], "serial chain of revolute joints expected"
            jrows.append("  {%s}" % ", ".join(repr(float(x)) for x in list(j["xyz"]) + [R[r][c] for r in range(3) for c in range(3)]))
            Rl[idx] = I3; ol[idx] = [0.0, 0.0, 0.0]; fr[idx] = nrev
        else:
            Rl[idx] = mat_mul(Rl[parent], R)
            ol[idx] = [ol[parent][k] + mat_vec(Rl[parent], j["xyz"])[k] for k in range(3)]
            fr[idx] = fr[parent]
        add_local(links[name], Rl[idx], ol[idx], fr[idx])
    add_local(links[root], I3, [0.0, 0.0, 0.0], 0)
    assert len(lrows) == len(rows) and nrev == 7
    ee_l = [ol[11][k] + mat_vec(Rl[11], links[names[11]]["com"])[k] for k in range(3)]
    l0_l = [ol[0][k] + mat_vec(Rl[0], links[names[0]]["com"])[k] for k in range(3)]
    fl_l = ol[7]  # panda_link8 (the flange: roboticstoolbox's end-effector link)
    assert fr[11] == 7 and fr[0] == 1 and fr[7] == 7
    ee = [ow[11][k] + mat_vec(Rw[11], links[names[11]]["com"])[k] for k in range(3)]
    l0 = [ow[0][k] + mat_vec(Rw[0], links[names[0]]["com"])[k] for k in range(3)]
    o = [
        "// GENERATED by tools/gen_model.py: robot collision primitives at the neutral pose, base at the origin (double).",
        "#pragma once",
        "namespace pgm {",
        "struct NeutralPrim { int type; double c[3]; double a[3]; double r, h; };  // type 0 sphere, 1 cylinder",
        "constexpr int kNumNeutralPrim = %d;" % len(rows),
        "static __device__ __constant__ NeutralPrim kNeutralPrims[%d] = {" % len(rows),
        ",\n".join(rows),
        "};",
        "static __device__ __constant__ double kNeutralEE[3] = {%r, %r, %r};     // COM of link 11 (get_ee_position)" % tuple(ee),
        "static __device__ __constant__ double kNeutralLink0[3] = {%r, %r, %r};  // COM of pybullet link 0 = panda_link1 (get_link_position(0)); turns with joint 1" % tuple(l0),
        "static __device__ __constant__ double kNeutralQ[7] = {%s};" % ", ".join(repr(x) for x in neutral),
        "// link-local form of the same table, for forward kinematics at an arbitrary pose: frame 0 = base, i = moving link i",
        "struct LocalPrim { int frame, type; double c[3]; double a[3]; double r, h; };",
        "static __device__ __constant__ LocalPrim kLocalPrims[%d] = {" % len(lrows),
        ",\n".join(lrows),
        "};",
        "// joint origins: xyz in the parent frame | rotation child -> parent at q = 0 (row-major), joint axis = child z",
        "static __device__ __constant__ double kJointOrigin[7][12] = {",
        ",\n".join(jrows),
        "};",
        "static __device__ __constant__ double kLocalEE[3] = {%r, %r, %r};      // COM of link 11, link-7 frame" % tuple(ee_l),
        "static __device__ __constant__ double kLocalLink0[3] = {%r, %r, %r};   // COM of panda_link1, link-1 frame" % tuple(l0_l),
        "static __device__ __constant__ double kLocalFlange[3] = {%r, %r, %r};  // origin of panda_link8, link-7 frame" % tuple(fl_l),
        "static __device__ __constant__ double kLocalEEOrigin[3] = {%r, %r, %r};  // origin of the link-11 frame (what PyBullet's IK drives), link-7 frame" % tuple(ol[11]),
        "}  // namespace pgm",
    ]
    with open(path, "w") as f:
        f.write("\n".join(o) + "\n")
    print("wrote", path)


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
    links0, order0 = emit_model("PGO_CUSTOM0", os.path.join(REF, "franka_panda_custom_0/panda.urdf"), out)
    csrc = os.path.join(os.path.dirname(os.path.dirname(OUT)), "panda_gym_b200", "csrc")
    emit_cuda(links0, order0, os.path.join(csrc, "pg_model.cuh"))
    emit_neutral(links0, links0["__root__"], order0, os.path.join(csrc, "pg_neutral.h"))
    emit_model("PGO_STOCK", os.path.join(REF, "franka_panda/panda.urdf"), out)
    out.append("#endif")
    with open(OUT, "w") as f:
        f.write("\n".join(out) + "\n")
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
