"""Mirror of src/models.jl: model builders, URDF ingestion, merge!."""
import json
import math
import os

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
def load_urdf(filename, package_path=()):
    """src/models.jl:166-171: parse_urdf + parse_urdf_visuals + extract_convex_surfaces.

    The parsing lives in the library (flash_urdf_load, csrc/flash_urdf.h: XML, binary / ASCII STL, OBJ, welding,
    `package://` resolution); this wrapper only turns the returned flash_model_desc into a Manipulator.
    Each <visual> mesh becomes a ConvexGeometry (the convex hull of its vertices, the EnhancedGJK.NeighborMesh
    semantics of src/models.jl:152) whose frame is the visual origin; the root link is welded to the world
    (RigidBodyDynamics.parse_urdf)."""
    import ctypes as C
    from . import _lib
    L = _lib.lib()
    dirs = [os.fsencode(os.path.abspath(p)) for p in package_path]
    arr = (C.c_char_p * max(len(dirs), 1))(*dirs)
    h = C.c_void_p()
    rc = L.flash_urdf_load(os.fsencode(os.path.abspath(filename)), arr, len(dirs), C.byref(h))
    if rc != 0:
        msg = L.flash_last_error(None)
        raise IOError("flash_urdf_load failed (%d): %s" % (rc, msg.decode() if msg else "?"))
    try:
        d = L.flash_urdf_desc(h).contents
        bodies = []
        for b in range(d.n_bodies):
            bodies.append(Body(L.flash_urdf_body_name(h, b).decode(), d.body_parent[b], d.joint_type[b],
                               [d.joint_axis[3 * b + k] for k in range(3)], np.array([d.joint_to_parent[12 * b + k] for k in range(12)]).reshape(3, 4)))
        surfaces = []
        for s_ in range(d.n_surfaces):
            p0, p1 = d.point_offset[s_], d.point_offset[s_ + 1]
            verts = np.array([d.points[k] for k in range(3 * p0, 3 * p1)]).reshape(-1, 3)
            surfaces.append(ConvexGeometry(verts, d.surface_body[s_], np.array([d.surface_frame[12 * s_ + k] for k in range(12)]).reshape(3, 4)))
    finally:
        L.flash_urdf_free(h)
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
