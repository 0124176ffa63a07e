"""indep.py -- TEST INFRASTRUCTURE: a second, independent restatement of the path in numpy / scipy.

It shares NOTHING with the product (no import from `flash`, not its model builders, not its packing, not its URDF /
STL reader) and nothing with the C++ oracle: the models are restated here straight from the reference sources
(src/models.jl, examples/irb_and_squishable.ipynb, examples/data/IRB140 -- the URDF text and STL bytes of the fixture
copies under tests/golden/IRB140), the kinematics is plain numpy, the RBF skin is numpy.linalg.solve, the convex
signed distance is scipy.spatial.ConvexHull planes + brute-force point-triangle distances.  tests/test_indep.py holds
the C++ oracle and the CUDA path against it, so that a wrong joint order, a dropped visual origin or a welding bug
cannot hide behind a builder both sides share.  Slow (seconds for 10^4 points): small cases only.
"""
import ctypes as C
import math
import os
import struct
import xml.etree.ElementTree as ET

import numpy as np
from scipy.spatial import ConvexHull

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD_URDF = os.path.join(HERE, "golden", "IRB140", "urdf", "irb_140_convhull.urdf")
FIXED, REVOLUTE, QUAT = 0, 1, 2


def _I34():
    return np.hstack([np.eye(3), np.zeros((3, 1))])


# ---------------------------------------------------------------------------------------------- models (src/models.jl)
def two_link_arm():
    """src/models.jl:19-71: two revolute links about z, link 2 attached at (1, 0, 0) of link 1; ONE rigid skin whose
    points are pushed link by link (surface ring points, then the link's extras), skeleton points likewise."""
    L, r = 1.0, 0.1
    bodies = [dict(parent=-1, jtype=REVOLUTE, axis=(0, 0, 1), j2p=_I34()),
              dict(parent=0, jtype=REVOLUTE, axis=(0, 0, 1), j2p=np.hstack([np.eye(3), [[L], [0.0], [0.0]]]))]
    surface, skeleton = [], []
    for i in (0, 1):
        for x in np.linspace(0.3 * L, 0.7 * L, 3):
            for y in (-r, r):
                for z in (-r, r):
                    surface.append((i, (x, y, z)))
            for z in (-math.sqrt(2) * r, math.sqrt(2) * r):
                surface.append((i, (x, 0.0, z)))
        if i == 0:
            for z in (-math.sqrt(2) * r, math.sqrt(2) * r):
                surface.append((i, (L, 0.0, z)))
            surface.append((i, (0.0, 0.0, 0.0)))
        else:
            surface.append((i, (L, 0.0, 0.0)))
        for x in np.linspace(0.2 * L, 0.8 * L, 3):
            skeleton.append((i, (x, 0.0, 0.0)))
    return dict(bodies=bodies, surfaces=[dict(kind="rigid", surface=surface, skeleton=skeleton)])


def _floating(name):
    return dict(parent=-1, jtype=QUAT, axis=(0, 0, 1), j2p=_I34(), name=name)


def beanbag():
    """src/models.jl:73-99: -e then +e along x, y, z; skeleton at the origin; a DEFORMABLE skin."""
    surface = []
    for axis in range(3):
        for s in (-1.0, 1.0):
            x = [0.0, 0.0, 0.0]; x[axis] = s
            surface.append((0, tuple(x)))
    return dict(bodies=[_floating("body1")], surfaces=[dict(kind="deformable", surface=surface, skeleton=[(0, (0.0, 0.0, 0.0))])])


def squishable():
    """src/models.jl:101-136"""
    radii = (0.44 / 2, 0.40 / 2, 0.30 / 2)
    surface = []
    for axis in (1, 2, 3):
        for i_sign in (-1, 1):
            for j_sign in (-1, 1):
                theta = math.pi / 4
                x = [0.0, 0.0, 0.0]
                i = axis % 3 + 1
                j = i % 3 + 1
                a, b = radii[i - 1] * 1.25, radii[j - 1] * 1.25
                x[i - 1] = i_sign * math.sqrt(a ** 2 * b ** 2 / (a ** 2 * math.tan(theta) ** 2 + b ** 2))
                x[j - 1] = j_sign * math.sqrt(b ** 2 * (1 - b ** 2 / (a ** 2 * math.tan(theta) ** 2 + b ** 2)))
                surface.append((0, tuple(x)))
    return dict(bodies=[_floating("squishable_body")], surfaces=[dict(kind="deformable", surface=surface, skeleton=[(0, (0.0, 0.0, 0.0))])])


def table():
    """examples/irb_and_squishable.ipynb cell 3: an SVector{8} of box corners handed to ConvexGeometry"""
    w, t = 0.25, 0.05
    verts = [(x, y, z) for z in (-t, t) for x in (-w, w) for y in (-w, w)]
    return dict(bodies=[_floating("table_body")], surfaces=[dict(kind="convex", body=0, frame=_I34(), verts=np.array(verts, float))])


def _rpy(r, p, y):
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return (np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]]) @ np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
            @ np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]]))


def _origin(elem):
    T = _I34()
    o = None if elem is None else elem.find("origin")
    if o is not None:
        T[:, :3] = _rpy(*[float(v) for v in o.get("rpy", "0 0 0").split()])
        T[:, 3] = [float(v) for v in o.get("xyz", "0 0 0").split()]
    return T


def read_stl_vertices(path):
    """binary STL: 80-byte header, u32 triangle count, 50 bytes per triangle -> welded vertices in order of first occurrence"""
    data = open(path, "rb").read()
    ntri = struct.unpack_from("<I", data, 80)[0]
    assert len(data) == 84 + 50 * ntri, "not a binary STL"
    out, seen = [], set()
    for t in range(ntri):
        v = struct.unpack_from("<9f", data, 84 + 50 * t + 12)
        for k in range(3):
            p = tuple(float(c) for c in v[3 * k:3 * k + 3])
            if p not in seen:
                seen.add(p); out.append(p)
    return np.array(out, float)


def load_urdf(path=GOLD_URDF, package_dir=os.path.join(HERE, "golden")):
    """src/models.jl:154-171 on the fixture copy of examples/data/IRB140/urdf/irb_140_convhull.urdf: breadth-first tree from
    the root link (welded to the world), one convex surface per <visual> mesh, frame = the visual origin."""
    root = ET.parse(path).getroot()
    links = {l.get("name"): l for l in root.findall("link")}
    joints = root.findall("joint")
    children = {j.find("child").get("link") for j in joints}
    (root_name,) = [n for n in links if n not in children]
    bodies, index, queue = [dict(parent=-1, jtype=FIXED, axis=(0, 0, 1), j2p=_I34(), name=root_name)], {root_name: 0}, [root_name]
    while queue:
        name = queue.pop(0)
        for j in joints:
            if j.find("parent").get("link") != name:
                continue
            child = j.find("child").get("link")
            jt = {"revolute": REVOLUTE, "continuous": REVOLUTE, "fixed": FIXED, "floating": QUAT}[j.get("type")]
            ax = j.find("axis")
            bodies.append(dict(parent=index[name], jtype=jt, axis=tuple(float(v) for v in ax.get("xyz").split()) if ax is not None else (1, 0, 0),
                               j2p=_origin(j), name=child))
            index[child] = len(bodies) - 1
            queue.append(child)
    surfaces = []
    for b, body in enumerate(bodies):
        for vis in links[body["name"]].findall("visual"):
            mesh = vis.find("geometry/mesh")
            fn = mesh.get("filename")
            assert fn.startswith("package://")
            surfaces.append(dict(kind="convex", body=b, frame=_origin(vis), verts=read_stl_vertices(os.path.join(package_dir, fn[len("package://"):]))))
    return dict(bodies=bodies, surfaces=surfaces)


def merge(m1, m2):
    """merge!, src/models.jl:173-177: m2's tree hangs off m1's root, surfaces are concatenated"""
    off = len(m1["bodies"])
    sh = lambda b: b + off if b >= 0 else -1
    for b in m2["bodies"]:
        m1["bodies"].append(dict(b, parent=sh(b["parent"])))
    for s in m2["surfaces"]:
        if s["kind"] == "convex":
            m1["surfaces"].append(dict(s, body=sh(s["body"])))
        else:
            m1["surfaces"].append(dict(s, surface=[(sh(b), p) for b, p in s["surface"]], skeleton=[(sh(b), p) for b, p in s["skeleton"]]))
    return m1


def irb140(floating_base=False):
    m = load_urdf()
    if floating_base:   # change_joint_type!(..., QuaternionFloating), examples/irb_and_squishable.ipynb cell 4
        m["bodies"][0]["jtype"] = QUAT
    return m


def irb_and_squishable():
    return merge(merge(irb140(True), squishable()), table())


MODELS = {"two_link_arm": two_link_arm, "beanbag": beanbag, "squishable": squishable, "table": table, "irb140": irb140, "irb_and_squishable": irb_and_squishable}


# ---------------------------------------------------------------------------------------------- state layout
def nq_of(m):
    return sum({FIXED: 0, REVOLUTE: 1, QUAT: 7}[b["jtype"]] for b in m["bodies"])


def nx_of(m):
    return nq_of(m) + 3 * sum(len(s["surface"]) for s in m["surfaces"] if s["kind"] == "deformable")


def zero_state(m):
    x = np.zeros(nx_of(m)); o = 0
    for b in m["bodies"]:
        if b["jtype"] == QUAT:
            x[o] = 1.0
        o += {FIXED: 0, REVOLUTE: 1, QUAT: 7}[b["jtype"]]
    return x


# ---------------------------------------------------------------------------------------------- kinematics, fields
def body_poses(m, x):
    """body -> world as 4x4, quaternions normalised inside (normalize!, src/gradientdescent.jl:19-30)"""
    T, o = [], 0
    for b in m["bodies"]:
        J = np.eye(4)
        if b["jtype"] == REVOLUTE:
            a = np.array(b["axis"], float); a /= np.linalg.norm(a)
            K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
            th = x[o]
            J[:3, :3] = np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * (K @ K)
            o += 1
        elif b["jtype"] == QUAT:
            w, qx, qy, qz = x[o:o + 4] / np.linalg.norm(x[o:o + 4])
            J[:3, :3] = [[1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - w * qz), 2 * (qx * qz + w * qy)],
                         [2 * (qx * qy + w * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - w * qx)],
                         [2 * (qx * qz - w * qy), 2 * (qy * qz + w * qx), 1 - 2 * (qx * qx + qy * qy)]]
            J[:3, 3] = x[o + 4:o + 7]
            o += 7
        P = np.vstack([np.asarray(b["j2p"], float).reshape(3, 4), [0, 0, 0, 1]])
        parent = T[b["parent"]] if b["parent"] >= 0 else np.eye(4)
        T.append(parent @ P @ J)
    return T


def _tri_dist(p, a, b, c):
    """distance from points p (n,3) to triangle abc (Ericson's closest-point regions, vectorised)"""
    ab, ac, ap = b - a, c - a, p - a
    d1, d2 = ap @ ab, ap @ ac
    bp = p - b; d3, d4 = bp @ ab, bp @ ac
    cp = p - c; d5, d6 = cp @ ab, cp @ ac
    vc = d1 * d4 - d3 * d2; vb = d5 * d2 - d1 * d6; va = d3 * d6 - d5 * d4
    out = np.empty(len(p)); done = np.zeros(len(p), bool)
    def put(mask, q):
        mask = mask & ~done
        out[mask] = np.linalg.norm(p[mask] - q[mask] if q.ndim == 2 else p[mask] - q, axis=1); done[mask] = True
    put((d1 <= 0) & (d2 <= 0), a)
    put((d3 >= 0) & (d4 <= d3), b)
    with np.errstate(divide="ignore", invalid="ignore"):
        put((vc <= 0) & (d1 >= 0) & (d3 <= 0), a + np.outer(d1 / (d1 - d3), ab))
        put((d6 >= 0) & (d5 <= d6), c)
        put((vb <= 0) & (d2 >= 0) & (d6 <= 0), a + np.outer(d2 / (d2 - d6), ac))
        put((va <= 0) & (d4 - d3 >= 0) & (d5 - d6 >= 0), b + np.outer((d4 - d3) / ((d4 - d3) + (d5 - d6)), c - b))
        den = va + vb + vc
        put(np.ones(len(p), bool), a + np.outer(vb / den, ab) + np.outer(vc / den, ac))
    return out


def hull_signed_distance(verts, p):
    """signed distance of points p (hull frame) to conv(verts): outside the Euclidean distance to the nearest facet triangle,
    inside the largest (least negative) plane distance"""
    h = ConvexHull(verts)
    plane = (h.equations[:, :3] @ p.T + h.equations[:, 3:4]).max(0)
    d = np.full(len(p), np.inf)
    for tri in h.simplices:
        d = np.minimum(d, _tri_dist(p, verts[tri[0]], verts[tri[1]], verts[tri[2]]))
    return np.where(plane <= 0, plane, d)


def skin_value(m, s, x, T, p):
    """InterpolatingSurface(points, values, XCubed(), true) (src/Flash.jl:207-213) restated: cubic RBF + linear polynomial
    through the control points (0 on the surface, -1 on the skeleton), value f / |grad f|"""
    off = nq_of(m)
    for t in m["surfaces"]:
        if t is s:
            break
        if t["kind"] == "deformable":
            off += 3 * len(t["surface"])
    cs = []
    for k, (b, q) in enumerate(list(s["surface"]) + list(s["skeleton"])):
        q = np.array(q, float)
        if s["kind"] == "deformable" and k < len(s["surface"]):
            q = q + x[off + 3 * k:off + 3 * k + 3]
        cs.append((T[b] @ np.r_[q, 1.0])[:3] if b >= 0 else q)
    c = np.array(cs); n = len(c)
    A = np.zeros((n + 4, n + 4))
    A[:n, :n] = np.linalg.norm(c[:, None] - c[None], axis=2) ** 3
    A[:n, n] = 1; A[n, :n] = 1; A[:n, n + 1:] = c; A[n + 1:, :n] = c.T
    rhs = np.r_[np.zeros(len(s["surface"])), -np.ones(len(s["skeleton"])), np.zeros(4)]
    w = np.linalg.solve(A, rhs)
    u = p[:, None, :] - c[None]
    r = np.linalg.norm(u, axis=2)
    f = (r ** 3) @ w[:n] + w[n] + p @ w[n + 1:]
    g = np.einsum("ij,ijk->ik", 3 * w[:n] * r, u) + w[n + 1:]
    return f / np.linalg.norm(g, axis=1)


def surfaces(m, x, pts):
    """(n, S): every surface's value at every world point under state x"""
    x = np.asarray(x, float); pts = np.asarray(pts, float)
    T = body_poses(m, x)
    out = np.empty((len(pts), len(m["surfaces"])))
    for k, s in enumerate(m["surfaces"]):
        if s["kind"] == "convex":
            W = (T[s["body"]] if s["body"] >= 0 else np.eye(4)) @ np.vstack([np.asarray(s["frame"], float).reshape(3, 4), [0, 0, 0, 1]])
            local = (pts - W[:3, 3]) @ W[:3, :3]       # R^T (p - t)
            out[:, k] = hull_signed_distance(np.asarray(s["verts"], float), local)
        else:
            out[:, k] = skin_value(m, s, x, T, pts)
    return out


def cost(m, x, pts, w_def=10.0):
    """src/gradientdescent.jl:28-39: sum of squared skin values (min over surfaces, first minimum) + w * |deformations|^2"""
    v = surfaces(m, x, pts)
    return float((v.min(1) ** 2).sum() + w_def * (np.asarray(x, float)[nq_of(m):] ** 2).sum()), v.argmin(1)


# ---------------------------------------------------------------------------------------------- flash_model_desc for the C++ oracle
class ModelDesc(C.Structure):    # include/flashsdf.h flash_model_desc, declared here so that nothing of the product's Python is imported
    _fields_ = [("abi_version", C.c_int32), ("n_bodies", C.c_int32), ("body_parent", C.POINTER(C.c_int32)), ("joint_type", C.POINTER(C.c_int32)),
                ("joint_axis", C.POINTER(C.c_double)), ("joint_to_parent", C.POINTER(C.c_double)), ("n_surfaces", C.c_int32),
                ("surface_kind", C.POINTER(C.c_int32)), ("surface_body", C.POINTER(C.c_int32)), ("surface_frame", C.POINTER(C.c_double)),
                ("point_offset", C.POINTER(C.c_int32)), ("n_surface_points", C.POINTER(C.c_int32)), ("points", C.POINTER(C.c_double)),
                ("point_body", C.POINTER(C.c_int32))]


class Packed:
    """the arrays of a flash_model_desc built from one of the dict models above (keeps them alive)"""

    def __init__(self, m, abi_version=2):
        i32 = lambda a: np.ascontiguousarray(a, np.int32)
        f64 = lambda a: np.ascontiguousarray(a, np.float64)
        B = m["bodies"]
        self.a = dict(parent=i32([b["parent"] for b in B]), jtype=i32([b["jtype"] for b in B]), axis=f64([b["axis"] for b in B]).reshape(-1),
                      j2p=f64([np.asarray(b["j2p"], float).reshape(12) for b in B]).reshape(-1))
        kind, sbody, frame, offs, nsurf, pts, pbody = [], [], [], [0], [], [], []
        for s in m["surfaces"]:
            if s["kind"] == "convex":
                kind.append(0); sbody.append(s["body"]); frame.append(np.asarray(s["frame"], float).reshape(12))
                pts += [list(v) for v in np.asarray(s["verts"], float)]; pbody += [-1] * len(s["verts"]); nsurf.append(len(s["verts"]))
            else:
                kind.append(2 if s["kind"] == "deformable" else 1); sbody.append(-1); frame.append(_I34().reshape(12))
                for b, p in list(s["surface"]) + list(s["skeleton"]):
                    pts.append([float(v) for v in p]); pbody.append(b)
                nsurf.append(len(s["surface"]))
            offs.append(len(pts))
        self.a.update(kind=i32(kind), sbody=i32(sbody), frame=f64(frame).reshape(-1), offs=i32(offs), nsurf=i32(nsurf), pts=f64(pts).reshape(-1), pbody=i32(pbody))
        d = ModelDesc()
        pi = lambda k: self.a[k].ctypes.data_as(C.POINTER(C.c_int32))
        pd = lambda k: self.a[k].ctypes.data_as(C.POINTER(C.c_double))
        d.abi_version = abi_version; d.n_bodies = len(B); d.body_parent = pi("parent"); d.joint_type = pi("jtype"); d.joint_axis = pd("axis"); d.joint_to_parent = pd("j2p")
        d.n_surfaces = len(kind); d.surface_kind = pi("kind"); d.surface_body = pi("sbody"); d.surface_frame = pd("frame"); d.point_offset = pi("offs")
        d.n_surface_points = pi("nsurf"); d.points = pd("pts"); d.point_body = pi("pbody")
        self.desc = d
        self.n_surfaces = len(kind)
