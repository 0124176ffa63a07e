"""Mirror of src/Flash.jl: geometry types, Manipulator, ManipulatorState, skin."""
import ctypes as C

import numpy as np

from ._desc import ABI_VERSION, Options, PackedModel, Timing, c_double_p, c_int32_p
from . import _lib

JOINT_FIXED, JOINT_REVOLUTE, JOINT_QUAT_FLOATING = 0, 1, 2
SDF_EXACT, SDF_GRID = 0, 1
_NQ = {JOINT_FIXED: 0, JOINT_REVOLUTE: 1, JOINT_QUAT_FLOATING: 7}


class FlashError(RuntimeError):
    pass


def _identity34():
    return np.hstack([np.eye(3), np.zeros((3, 1))])


class Body:
    """A moving body and the joint that attaches it to its parent (RigidBodyDynamics attach!)."""

    def __init__(self, name, parent, joint_type, axis=(0.0, 0.0, 1.0), joint_to_parent=None):
        self.name = name
        self.parent = int(parent)
        self.joint_type = int(joint_type)
        self.axis = np.asarray(axis, float)
        self.joint_to_parent = _identity34() if joint_to_parent is None else np.asarray(joint_to_parent, float).reshape(3, 4)


class Mechanism:
    def __init__(self, bodies=None):
        self.bodies = list(bodies or [])

    def num_positions(self):
        return sum(_NQ[b.joint_type] for b in self.bodies)

    def q_offsets(self):
        offs, o = [], 0
        for b in self.bodies:
            offs.append(o)
            o += _NQ[b.joint_type]
        return offs

    def zero_configuration(self):
        q = np.zeros(self.num_positions())
        for b, o in zip(self.bodies, self.q_offsets()):
            if b.joint_type == JOINT_QUAT_FLOATING:
                q[o] = 1.0
        return q


class BodyGeometry:  # src/Flash.jl:30
    pass


class InterpolatingGeometry(BodyGeometry):  # src/Flash.jl:33
    def __init__(self, surface_points, skeleton_points):
        # each point is (body_index, (x, y, z)) -- a Point3D in that body's frame
        self.surface_points = [(int(b), tuple(map(float, p))) for b, p in surface_points]
        self.skeleton_points = [(int(b), tuple(map(float, p))) for b, p in skeleton_points]


class DeformableInterpolatingSkin(InterpolatingGeometry):  # src/Flash.jl:35-38
    pass


class RigidInterpolatingSkin(InterpolatingGeometry):  # src/Flash.jl:40-43
    pass


class ConvexGeometry(BodyGeometry):  # src/Flash.jl:45-48
    def __init__(self, vertices, body, frame=None):
        self.vertices = np.asarray(vertices, float).reshape(-1, 3)
        self.body = int(body)
        self.frame = _identity34() if frame is None else np.asarray(frame, float).reshape(3, 4)


def num_deformations(x):  # src/Flash.jl:79-88
    if isinstance(x, DeformableInterpolatingSkin):
        return len(x.surface_points)
    if isinstance(x, BodyGeometry):
        return 0
    return sum(num_deformations(s) for s in x.surfaces)


class Manipulator:  # src/Flash.jl:62-65
    def __init__(self, mechanism, surfaces):
        self.mechanism = mechanism
        self.surfaces = list(surfaces)
        self._ctx = None
        self._packed = None
        self._points_owner = None   # the CostFunctor whose cloud the ctx currently holds
        self._options = dict(sdf_mode=SDF_EXACT, grid_spacing=0.0, grid_margin=0.0,
                             rbf_normalize_by_gradient=1, rbf_polynomial_degree=1, device=-1, deterministic=0, devices=())

    def __repr__(self):  # src/Flash.jl:67
        return "Manipulator with %d links and %d surfaces" % (len(self.mechanism.bodies) + 1, len(self.surfaces))

    # ---- C-ABI context management
    def set_options(self, **kw):
        for k in kw:
            if k not in self._options:
                raise KeyError(k)
        self._options.update(kw)
        self.invalidate()

    def invalidate(self):
        if self._ctx is not None:
            _lib.lib().flash_destroy(self._ctx)
        self._ctx = None
        self._packed = None
        self._points_owner = None

    def packed(self):
        if self._packed is None:
            self._packed = PackedModel(self)
        return self._packed

    def ctx(self):
        if self._ctx is None:
            L = _lib.lib()
            opt = Options()
            L.flash_options_default(C.byref(opt))
            for k, v in self._options.items():
                if k in ("grid_spacing", "grid_margin"):
                    if v > 0:
                        setattr(opt, k, v)
                elif k == "devices":
                    opt.n_devices = len(v)
                    for i, dv in enumerate(v):
                        opt.devices[i] = int(dv)
                else:
                    setattr(opt, k, v)
            out = C.c_void_p()
            rc = L.flash_model_create(C.byref(self.packed().desc), C.byref(opt), C.byref(out))
            if rc != 0:
                msg = L.flash_last_error(None)
                raise FlashError("flash_model_create failed (%d): %s" % (rc, msg.decode() if msg else "?"))
            self._ctx = out
        return self._ctx

    def check(self, rc):
        if rc != 0:
            msg = _lib.lib().flash_last_error(self._ctx)
            raise FlashError("libflashsdf error %d: %s" % (rc, msg.decode() if msg else "?"))

    def last_timing(self):
        t = Timing()
        self.check(_lib.lib().flash_last_timing(self.ctx(), C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timing._fields_}

    def grid_report(self, surface):
        """grid mode: (dims, spacing, max |trilinear - exact| sampled at cell centres) of a convex surface."""
        dims = (C.c_int32 * 3)(); sp = C.c_double(); err = C.c_double()
        self.check(_lib.lib().flash_grid_report(self.ctx(), surface, dims, C.byref(sp), C.byref(err)))
        return tuple(dims), sp.value, err.value

    def grid_download(self, surface):
        dims, sp, _ = self.grid_report(surface)
        org = np.zeros(3); g2b = np.zeros(12); vals = np.empty(dims[2] * dims[1] * dims[0], np.float32)
        self.check(_lib.lib().flash_grid_download(self.ctx(), surface, org.ctypes.data_as(c_double_p), g2b.ctypes.data_as(c_double_p),
                                                  vals.ctypes.data_as(C.POINTER(C.c_float))))
        return vals.reshape(dims[2], dims[1], dims[0]), org, sp, g2b.reshape(3, 4)

    def __del__(self):
        try:
            self.invalidate()
        except Exception:
            pass


def num_positions(manip):
    return manip.mechanism.num_positions()


def num_states(manip):  # src/Flash.jl:90
    return num_positions(manip) + 3 * num_deformations(manip)


class _MechanismState:
    def __init__(self, mechanism):
        self.mechanism = mechanism
        self.q = mechanism.zero_configuration()


class ManipulatorState:  # src/Flash.jl:71-125
    def __init__(self, manipulator):
        self.manipulator = manipulator
        self.mechanism_state = _MechanismState(manipulator.mechanism)
        self.deformation_data = np.zeros(3 * num_deformations(manipulator))
        self.deformations = {}
        off = 0
        for s in manipulator.surfaces:
            views = []
            for _ in range(num_deformations(s)):
                views.append(self.deformation_data[off:off + 3])
                off += 3
            self.deformations[s] = views


def set_configuration_(state, joint_angles):  # set_configuration!, src/Flash.jl:127-129
    q = np.asarray(joint_angles, float)
    if q.shape != state.mechanism_state.q.shape:
        raise ValueError("configuration has %d entries, expected %d" % (q.size, state.mechanism_state.q.size))
    state.mechanism_state.q[:] = q


def _flatten(state):
    return np.concatenate([state.mechanism_state.q, state.deformation_data])


class Skin:
    """The closure returned by Flash.skin(state): x -> minimum(s(x) for s in all_surfaces)."""

    def __init__(self, manipulator, x):
        self.manipulator = manipulator
        self.x = np.ascontiguousarray(x, dtype=np.float64)

    def eval(self, points, return_argmin=False):
        pts = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
        n = len(pts)
        dist = np.empty(n)
        arg = np.empty(n, dtype=np.int32)
        m = self.manipulator
        m.check(_lib.lib().flash_skin_eval(m.ctx(), self.x.ctypes.data_as(c_double_p), pts.ctypes.data_as(c_double_p), n,
                                           dist.ctypes.data_as(c_double_p), arg.ctypes.data_as(c_int32_p)))
        return (dist, arg) if return_argmin else dist

    def __call__(self, point):
        p = np.asarray(point, float)
        if p.ndim == 1:
            return float(self.eval(p.reshape(1, 3))[0])
        return self.eval(p)


def skin(state):  # src/Flash.jl:265-268
    return Skin(state.manipulator, _flatten(state))


def surface_points(state, surface):
    """Body-frame control points of a skin including deformations (world transform is on device)."""
    pts = np.array([p for _, p in surface.surface_points], float)
    if isinstance(surface, DeformableInterpolatingSkin):
        pts = pts + np.array(state.deformations[surface]).reshape(-1, 3)
    return pts


def _body_poses(mechanism, q):
    """4x4 body->world transforms (host-side, only to place a drawing region around a few points)."""
    X = []
    for b, o in zip(mechanism.bodies, mechanism.q_offsets()):
        J = np.eye(4)
        if b.joint_type == JOINT_REVOLUTE:
            a = b.axis / np.linalg.norm(b.axis)
            K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
            J[:3, :3] = np.eye(3) + np.sin(q[o]) * K + (1 - np.cos(q[o])) * (K @ K)
        elif b.joint_type == JOINT_QUAT_FLOATING:
            w, x, y, z = q[o:o + 4] / np.linalg.norm(q[o:o + 4])
            J[:3, :3] = [[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                         [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                         [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]
            J[:3, 3] = q[o + 4:o + 7]
        P = np.eye(4); P[:3, :] = b.joint_to_parent
        T = (X[b.parent] if b.parent >= 0 else np.eye(4)) @ P @ J
        X.append(T)
    return X


def drawing_region(state, surface):  # src/Flash.jl:270-295
    """World-frame box the reference contours a surface in: the control points' AABB grown by half its
    widths (skins), the hull's AABB grown by a tenth (convex).  Control points / hull vertices are posed
    by one skin_eval-free FK on the host (a handful of points)."""
    m = state.manipulator
    X = _body_poses(m.mechanism, state.mechanism_state.q)
    if isinstance(surface, ConvexGeometry):
        T = X[surface.body] if surface.body >= 0 else np.eye(4)
        F = np.eye(4); F[:3, :] = np.asarray(surface.frame, float).reshape(3, 4)
        P = (T @ F @ np.c_[np.asarray(surface.vertices, float), np.ones(len(surface.vertices))].T).T[:, :3]
        grow = 0.1
    else:
        pts = []
        sp = surface_points(state, surface)
        for (b, _), p in zip(surface.surface_points, sp):
            pts.append((X[b] @ np.r_[p, 1.0])[:3] if b >= 0 else p)
        for b, p in surface.skeleton_points:
            pts.append((X[b] @ np.r_[np.asarray(p, float), 1.0])[:3] if b >= 0 else np.asarray(p, float))
        P = np.array(pts)
        grow = 0.5
    lb, ub = P.min(0), P.max(0)
    w = ub - lb
    return lb - grow * w, ub + grow * w


def contour_mesh(state, lower, upper, isosurface_value=0.0, resolution=0.1, surface=None):
    """Iso-surface of one surface's field (or of skin(state) when surface is None) inside [lower, upper],
    extracted on the device (flash_contour_mesh).  Returns triangles [T][3][3] (world frame)."""
    m = state.manipulator
    x = _flatten(state)
    idx = -1 if surface is None else (surface if isinstance(surface, int) else m.surfaces.index(surface))
    lo = np.ascontiguousarray(lower, float); hi = np.ascontiguousarray(upper, float)
    n = C.c_int64(0)
    L = _lib.lib()
    fp = C.POINTER(C.c_float)
    m.check(L.flash_contour_mesh(m.ctx(), x.ctypes.data_as(c_double_p), idx, lo.ctypes.data_as(c_double_p), hi.ctypes.data_as(c_double_p),
                                 float(resolution), float(isosurface_value), None, 0, C.byref(n)))
    tris = np.empty((max(n.value, 1), 3, 3), np.float32)
    m.check(L.flash_contour_mesh(m.ctx(), x.ctypes.data_as(c_double_p), idx, lo.ctypes.data_as(c_double_p), hi.ctypes.data_as(c_double_p),
                                 float(resolution), float(isosurface_value), tris.ctypes.data_as(fp), n.value, C.byref(n)))
    return tris[: n.value]


def draw(state, draw_skin=True, path=None, resolution=None):  # src/Flash.jl:297-329, headless
    """The reference sends control points and one contour mesh per surface to a remote viewer (LCMGL /
    DrakeVisualizer: unavailable on this box).  Here the same meshes are returned -- and written as a
    Wavefront OBJ when `path` is given -- one group per surface."""
    meshes = []
    if draw_skin:
        for s_ in state.manipulator.surfaces:
            lb, ub = drawing_region(state, s_)
            h = resolution if resolution else float(np.max(ub - lb)) / 40.0
            meshes.append(contour_mesh(state, lb, ub, 0.0, h, surface=s_))
    if path is not None:
        with open(path, "w") as f:
            base = 1
            for k, t in enumerate(meshes):
                f.write("g surface_%d\n" % k)
                for v in t.reshape(-1, 3):
                    f.write("v %.7g %.7g %.7g\n" % tuple(v))
                for i in range(len(t)):
                    f.write("f %d %d %d\n" % (base + 3 * i, base + 3 * i + 1, base + 3 * i + 2))
                base += 3 * len(t)
    return meshes
