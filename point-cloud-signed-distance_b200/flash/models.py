"""Mirror of src/models.jl: model builders, URDF ingestion, merge!."""
import json
import math
import os
import struct
import xml.etree.ElementTree as ET

import numpy as np

from .core import (Body, ConvexGeometry, DeformableInterpolatingSkin, Manipulator, Mechanism,
                   RigidInterpolatingSkin, JOINT_FIXED, JOINT_QUAT_FLOATING, JOINT_REVOLUTE, _identity34)

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def _translation(x, y, z):
    T = _identity34()
    T[:, 3] = (x, y, z)
    return T


def two_link_arm(deformable=False):
    """src/models.jl:19-71.  (`deformable` is ignored by the reference too: line 68 always builds a
    RigidInterpolatingSkin.)"""
    link_length, radius = 1.0, 0.1
    bodies, surface_points, skeleton_points = [], [], []
    parent = -1
    for i in (1, 2):
        j2p = _translation(link_length, 0.0, 0.0) if i > 1 else _identity34()
        bodies.append(Body("body%d" % i, parent, JOINT_REVOLUTE, (0.0, 0.0, 1.0), j2p))
        b = len(bodies) - 1
        parent = b
        for x in np.linspace(0.3 * link_length, 0.7 * link_length, 3):
            for y in (-radius, radius):
                for z in (-radius, radius):
                    surface_points.append((b, (x, y, z)))
            for z in (-math.sqrt(2) * radius, math.sqrt(2) * radius):
                surface_points.append((b, (x, 0.0, z)))
        if i == 1:
            for z in (-math.sqrt(2) * radius, math.sqrt(2) * radius):
                surface_points.append((b, (link_length, 0.0, z)))
            surface_points.append((b, (0.0, 0.0, 0.0)))
        else:
            surface_points.append((b, (link_length, 0.0, 0.0)))
        for x in np.linspace(0.2 * link_length, 0.8 * link_length, 3):
            skeleton_points.append((b, (x, 0.0, 0.0)))
    return Manipulator(Mechanism(bodies), [RigidInterpolatingSkin(surface_points, skeleton_points)])


def beanbag():
    """src/models.jl:73-98"""
    bodies = [Body("body1", -1, JOINT_QUAT_FLOATING)]
    surface_points = []
    for axis in range(3):
        for s in (-1.0, 1.0):
            x = [0.0, 0.0, 0.0]
            x[axis] = s
            surface_points.append((0, tuple(x)))
    skeleton_points = [(0, (0.0, 0.0, 0.0))]
    return Manipulator(Mechanism(bodies), [DeformableInterpolatingSkin(surface_points, skeleton_points)])


def squishable():
    """src/models.jl:100-136"""
    bodies = [Body("squishable_body", -1, JOINT_QUAT_FLOATING)]
    radii = (0.44 / 2, 0.40 / 2, 0.30 / 2)
    surface_points = []
    theta = math.pi / 4
    for axis in (1, 2, 3):
        for i_sign in (-1, 1):
            for j_sign in (-1, 1):
                x = [0.0, 0.0, 0.0]
                i = axis % 3 + 1
                j = i % 3 + 1
                a = radii[i - 1] * 1.25
                b = radii[j - 1] * 1.25
                x[i - 1] = i_sign * math.sqrt(a ** 2 * b ** 2 / (a ** 2 * math.tan(theta) ** 2 + b ** 2))
                x[j - 1] = j_sign * math.sqrt(b ** 2 * (1 - b ** 2 / (a ** 2 * math.tan(theta) ** 2 + b ** 2)))
                surface_points.append((0, tuple(x)))
    skeleton_points = [(0, (0.0, 0.0, 0.0))]
    return Manipulator(Mechanism(bodies), [DeformableInterpolatingSkin(surface_points, skeleton_points)])


def table():
    """The 8-point convex box of examples/irb_and_squishable.ipynb cell 3 (an SVector{8} of points
    handed to ConvexGeometry) on a quaternion-floating body."""
    bodies = [Body("table_body", -1, JOINT_QUAT_FLOATING)]
    width, thickness = 0.25, 0.05
    pts = [(x, y, z) for z in (-thickness, thickness) for x in (-width, width) for y in (-width, width)]
    return Manipulator(Mechanism(bodies), [ConvexGeometry(pts, 0)])


def merge_(manip1, manip2):
    """merge!, src/models.jl:173-177: attach manip2's mechanism to manip1's root, concatenate surfaces."""
    off = len(manip1.mechanism.bodies)
    for b in manip2.mechanism.bodies:
        manip1.mechanism.bodies.append(Body(b.name, b.parent + off if b.parent >= 0 else -1, b.joint_type, b.axis, b.joint_to_parent))
    for s in manip2.surfaces:
        if isinstance(s, ConvexGeometry):
            manip1.surfaces.append(ConvexGeometry(s.vertices, s.body + off if s.body >= 0 else -1, s.frame))
        else:
            sp = [(b + off if b >= 0 else -1, p) for b, p in s.surface_points]
            sk = [(b + off if b >= 0 else -1, p) for b, p in s.skeleton_points]
            manip1.surfaces.append(type(s)(sp, sk))
    manip1.invalidate()
    return manip1


def change_joint_type_(manip, body_index, joint_type):
    """RigidBodyDynamics.change_joint_type! as used at examples/irb_and_squishable.ipynb cell 4."""
    manip.mechanism.bodies[body_index].joint_type = int(joint_type)
    manip.invalidate()
    return manip


# ------------------------------------------------------------------ URDF + STL ingestion
def _rpy_to_R(r, p, y):
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _origin(elem):
    T = _identity34()
    if elem is None:
        return T
    o = elem.find("origin")
    if o is None:
        return T
    xyz = [float(v) for v in o.get("xyz", "0 0 0").split()]
    rpy = [float(v) for v in o.get("rpy", "0 0 0").split()]
    T[:, :3] = _rpy_to_R(*rpy)
    T[:, 3] = xyz
    return T


def read_stl(path):
    """Binary or ASCII STL -> welded unique vertices (n, 3) float64."""
    data = open(path, "rb").read()
    ntri = struct.unpack("<I", data[80:84])[0] if len(data) >= 84 else -1
    if ntri >= 0 and len(data) == 84 + 50 * ntri:
        rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=ntri, offset=84)
        verts = rec["v"].reshape(-1, 3).astype(np.float64)
    else:
        verts = []
        for line in data.decode("ascii", "replace").splitlines():
            t = line.split()
            if len(t) == 4 and t[0] == "vertex":
                verts.append([float(t[1]), float(t[2]), float(t[3])])
        verts = np.array(verts, float)
    if len(verts) == 0:
        raise IOError("no vertices in STL %s" % path)
    _, idx = np.unique(verts, axis=0, return_index=True)
    return verts[np.sort(idx)]


def read_obj_vertices(path):
    verts = [[float(v) for v in line.split()[1:4]] for line in open(path) if line.startswith("v ")]
    verts = np.array(verts, float)
    _, idx = np.unique(verts, axis=0, return_index=True)
    return verts[np.sort(idx)]


def _resolve(filename, urdf_dir, package_path):
    if filename.startswith("package://"):
        rel = filename[len("package://"):]
        for root in list(package_path) + [urdf_dir]:
            cand = os.path.join(root, rel)
            if os.path.exists(cand):
                return cand
            # also accept a package_path entry that IS the package directory
            parts = rel.split("/", 1)
            if len(parts) == 2 and os.path.exists(os.path.join(root, parts[1])):
                return os.path.join(root, parts[1])
        raise IOError("cannot resolve %s in package_path %r" % (filename, list(package_path)))
    return filename if os.path.isabs(filename) else os.path.join(urdf_dir, filename)


def load_urdf(filename, package_path=()):
    """src/models.jl:166-171: parse_urdf + parse_urdf_visuals + extract_convex_surfaces.

    Each <visual> mesh becomes a ConvexGeometry (the convex hull of its vertices, the
    EnhancedGJK.NeighborMesh semantics of src/models.jl:152) whose frame is the visual origin.
    The root link is welded to the world (RigidBodyDynamics.parse_urdf)."""
    root = ET.parse(filename).getroot()
    urdf_dir = os.path.dirname(os.path.abspath(filename))
    links = {l.get("name"): l for l in root.findall("link")}
    joints = root.findall("joint")
    children = {j.find("child").get("link") for j in joints}
    roots = [n for n in links if n not in children]
    if len(roots) != 1:
        raise ValueError("URDF must have exactly one root link, found %r" % roots)
    by_parent = {}
    for j in joints:
        by_parent.setdefault(j.find("parent").get("link"), []).append(j)
    bodies, index = [Body(roots[0], -1, JOINT_FIXED)], {roots[0]: 0}
    queue = [roots[0]]
    while queue:
        name = queue.pop(0)
        for j in by_parent.get(name, []):
            child = j.find("child").get("link")
            jt = j.get("type")
            if jt in ("revolute", "continuous"):
                kind = JOINT_REVOLUTE
            elif jt == "fixed":
                kind = JOINT_FIXED
            elif jt == "floating":
                kind = JOINT_QUAT_FLOATING
            else:
                raise ValueError("unsupported URDF joint type %r (joint %s)" % (jt, j.get("name")))
            ax = j.find("axis")
            axis = [float(v) for v in ax.get("xyz").split()] if ax is not None else [1.0, 0.0, 0.0]
            bodies.append(Body(child, index[name], kind, axis, _origin(j)))
            index[child] = len(bodies) - 1
            queue.append(child)
    surfaces = []
    for body in bodies:  # vis_data is keyed by link frame in mechanism order
        for vis in links[body.name].findall("visual"):
            mesh = vis.find("geometry/mesh")
            if mesh is None:
                continue
            path = _resolve(mesh.get("filename"), urdf_dir, package_path)
            verts = read_obj_vertices(path) if path.lower().endswith(".obj") else read_stl(path)
            sc = mesh.get("scale")
            if sc:
                verts = verts * np.array([float(v) for v in sc.split()])
            surfaces.append(ConvexGeometry(verts, index[body.name], _origin(vis)))
    return Manipulator(Mechanism(bodies), surfaces)


# ------------------------------------------------------------------ committed fixture (IRB140 convex hulls)
def manipulator_to_dict(m):
    out = {"bodies": [], "surfaces": []}
    for b in m.mechanism.bodies:
        out["bodies"].append({"name": b.name, "parent": b.parent, "joint_type": b.joint_type,
                              "axis": [float(v) for v in b.axis], "joint_to_parent": [float(v) for v in b.joint_to_parent.reshape(-1)]})
    for s in m.surfaces:
        if isinstance(s, ConvexGeometry):
            out["surfaces"].append({"kind": "convex", "body": s.body, "frame": [float(v) for v in s.frame.reshape(-1)],
                                    "vertices": [[float(np.float32(v)) for v in p] for p in s.vertices]})
        else:
            out["surfaces"].append({"kind": "deformable" if isinstance(s, DeformableInterpolatingSkin) else "rigid",
                                    "surface_points": [[b, list(p)] for b, p in s.surface_points],
                                    "skeleton_points": [[b, list(p)] for b, p in s.skeleton_points]})
    return out


def manipulator_from_dict(d):
    bodies = [Body(b["name"], b["parent"], b["joint_type"], b["axis"], np.array(b["joint_to_parent"]).reshape(3, 4)) for b in d["bodies"]]
    surfaces = []
    for s in d["surfaces"]:
        if s["kind"] == "convex":
            surfaces.append(ConvexGeometry(np.array(s["vertices"], float), s["body"], np.array(s["frame"]).reshape(3, 4)))
        else:
            cls = DeformableInterpolatingSkin if s["kind"] == "deformable" else RigidInterpolatingSkin
            surfaces.append(cls([(b, tuple(p)) for b, p in s["surface_points"]], [(b, tuple(p)) for b, p in s["skeleton_points"]]))
    return Manipulator(Mechanism(bodies), surfaces)


def irb140(floating_base=False):
    """The IRB140 convex-hull manipulator of examples/irb140.ipynb cell 3
    (load_urdf("irb_140_convhull.urdf")), from the committed fixture data/irb140_convhull.json
    (made by tools/make_irb140_fixture.py with load_urdf above)."""
    with open(os.path.join(_DATA, "irb140_convhull.json")) as f:
        m = manipulator_from_dict(json.load(f))
    if floating_base:
        change_joint_type_(m, 0, JOINT_QUAT_FLOATING)
    return m


def irb_and_squishable():
    """The combined scene of examples/irb_and_squishable.ipynb cell 4:
    IRB140 with a quaternion-floating base + squishable + table -> 10 links, 9 surfaces, 63 states."""
    m = irb140(floating_base=True)
    merge_(m, squishable())
    merge_(m, table())
    return m
