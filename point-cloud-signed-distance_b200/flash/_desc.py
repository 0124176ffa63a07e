"""ctypes image of include/flashsdf.h's structs + marshalling of a Manipulator into flash_model_desc."""
import ctypes as C
import numpy as np

ABI_VERSION = 2
c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)


class ModelDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("n_bodies", C.c_int32),
        ("body_parent", c_int32_p), ("joint_type", c_int32_p),
        ("joint_axis", c_double_p), ("joint_to_parent", c_double_p),
        ("n_surfaces", C.c_int32),
        ("surface_kind", c_int32_p), ("surface_body", c_int32_p), ("surface_frame", c_double_p),
        ("point_offset", c_int32_p), ("n_surface_points", c_int32_p),
        ("points", c_double_p), ("point_body", c_int32_p),
    ]


class Options(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("device", C.c_int32),
        ("rbf_normalize_by_gradient", C.c_int32), ("rbf_polynomial_degree", C.c_int32),
        ("sdf_mode", C.c_int32), ("grid_spacing", C.c_double), ("grid_margin", C.c_double),
        ("deterministic", C.c_int32), ("n_devices", C.c_int32), ("devices", C.c_int32 * 16),
        ("reserved", C.c_int32 * 6),
    ]


class SolverParams(C.Structure):
    _fields_ = [
        ("rate", C.c_double), ("max_step", C.c_double), ("iteration_limit", C.c_int32),
        ("gradient_convergence_tolerance", C.c_double), ("deformation_cost_weight", C.c_double),
    ]


class Timing(C.Structure):
    _fields_ = [
        ("ms_prepare", C.c_float), ("ms_main", C.c_float), ("ms_allreduce", C.c_float),
        ("ms_epilogue", C.c_float), ("ms_total", C.c_float), ("launches", C.c_int32), ("ms_group", C.c_float),
    ]


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class PackedModel:
    """Owns the numpy buffers a flash_model_desc points into."""

    def __init__(self, manipulator):
        from .core import ConvexGeometry, DeformableInterpolatingSkin
        mech = manipulator.mechanism
        nb = len(mech.bodies)
        self.body_parent = _i32([b.parent for b in mech.bodies])
        self.joint_type = _i32([b.joint_type for b in mech.bodies])
        self.joint_axis = _f64([b.axis for b in mech.bodies]).reshape(nb, 3)
        self.joint_to_parent = _f64([np.asarray(b.joint_to_parent).reshape(12) for b in mech.bodies]).reshape(nb, 12)
        kinds, sbody, sframe, offs, nsurf, pts, pbody = [], [], [], [0], [], [], []
        ident = np.hstack([np.eye(3), np.zeros((3, 1))]).reshape(12)
        for s in manipulator.surfaces:
            if isinstance(s, ConvexGeometry):
                kinds.append(0); sbody.append(s.body); sframe.append(np.asarray(s.frame, float).reshape(12))
                v = np.asarray(s.vertices, float).reshape(-1, 3)
                pts.extend(v.tolist()); pbody.extend([-1] * len(v)); nsurf.append(len(v))
            else:
                kinds.append(2 if isinstance(s, DeformableInterpolatingSkin) else 1)
                sbody.append(-1); sframe.append(ident)
                for (b, p) in list(s.surface_points) + list(s.skeleton_points):
                    pts.append(list(map(float, p))); pbody.append(b)
                nsurf.append(len(s.surface_points))
            offs.append(len(pts))
        ns = len(kinds)
        self.surface_kind = _i32(kinds); self.surface_body = _i32(sbody)
        self.surface_frame = _f64(sframe).reshape(ns, 12)
        self.point_offset = _i32(offs); self.n_surface_points = _i32(nsurf)
        self.points = _f64(pts).reshape(-1, 3); self.point_body = _i32(pbody)
        d = ModelDesc()
        d.abi_version = ABI_VERSION
        d.n_bodies = nb
        d.body_parent = self.body_parent.ctypes.data_as(c_int32_p)
        d.joint_type = self.joint_type.ctypes.data_as(c_int32_p)
        d.joint_axis = self.joint_axis.ctypes.data_as(c_double_p)
        d.joint_to_parent = self.joint_to_parent.ctypes.data_as(c_double_p)
        d.n_surfaces = ns
        d.surface_kind = self.surface_kind.ctypes.data_as(c_int32_p)
        d.surface_body = self.surface_body.ctypes.data_as(c_int32_p)
        d.surface_frame = self.surface_frame.ctypes.data_as(c_double_p)
        d.point_offset = self.point_offset.ctypes.data_as(c_int32_p)
        d.n_surface_points = self.n_surface_points.ctypes.data_as(c_int32_p)
        d.points = self.points.ctypes.data_as(c_double_p)
        d.point_body = self.point_body.ctypes.data_as(c_int32_p)
        self.desc = d
