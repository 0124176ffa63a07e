"""ctypes binding of oracle/liboracle.so (CPU fp64 restatement of the reference path).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's CPU legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")

import sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import indep  # noqa: E402  (the oracle's model description is packed by the independent restatement, not by the product's flash._desc)
from indep import ModelDesc  # noqa: E402

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)


class SolverParams(C.Structure):   # flash_solver_params of include/flashsdf.h
    _fields_ = [("rate", C.c_double), ("max_step", C.c_double), ("iteration_limit", C.c_int32),
                ("gradient_convergence_tolerance", C.c_double), ("deformation_cost_weight", C.c_double)]


def model_dict(manipulator):
    """A product-side Manipulator (duck-typed: nothing of `flash` is imported here) as the dict form of tests/indep.py.
    Tests that want the oracle independent of the product's model BUILDERS too pass indep.MODELS[...]() instead."""
    if isinstance(manipulator, dict):
        return manipulator
    bodies = [dict(parent=b.parent, jtype=b.joint_type, axis=tuple(float(v) for v in b.axis), j2p=np.asarray(b.joint_to_parent, float).reshape(3, 4), name=getattr(b, "name", ""))
              for b in manipulator.mechanism.bodies]
    surfaces = []
    for s_ in manipulator.surfaces:
        if hasattr(s_, "vertices"):
            surfaces.append(dict(kind="convex", body=s_.body, frame=np.asarray(s_.frame, float).reshape(3, 4), verts=np.asarray(s_.vertices, float).reshape(-1, 3)))
        else:
            surfaces.append(dict(kind="deformable" if type(s_).__name__.startswith("Deformable") else "rigid",
                                 surface=[(int(b), tuple(p)) for b, p in s_.surface_points], skeleton=[(int(b), tuple(p)) for b, p in s_.skeleton_points]))
    return dict(bodies=bodies, surfaces=surfaces)

BRUTE, GJK = 0, 1
_lib = None


def build():
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("flash_oracle_capi.cpp", "flash_oracle.hpp")] + [os.path.join(ROOT, "include", "flashsdf.h")]
    if os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(s) for s in srcs):
        return
    subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"])


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        vp = C.c_void_p
        L.fo_last_error.restype = C.c_char_p
        L.fo_model_create.argtypes = [C.POINTER(ModelDesc), C.c_int, C.c_int]; L.fo_model_create.restype = vp
        L.fo_model_destroy.argtypes = [vp]; L.fo_model_destroy.restype = None
        L.fo_num_states.argtypes = [vp]; L.fo_num_positions.argtypes = [vp]
        L.fo_hull_info.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.fo_cost.argtypes = [vp, c_double_p, c_double_p, C.c_int64, C.c_double, C.c_int, C.c_int, c_double_p, c_int32_p]
        L.fo_cost_grad_dual.argtypes = [vp, c_double_p, c_double_p, C.c_int64, C.c_double, C.c_int, C.c_int, c_double_p, c_double_p, C.POINTER(C.c_int)]
        L.fo_cost_grad_adjoint.argtypes = [vp, c_double_p, c_double_p, C.c_int64, C.c_double, C.c_int, C.c_int, c_double_p, c_double_p]
        L.fo_skin_eval.argtypes = [vp, c_double_p, c_double_p, C.c_int64, C.c_int, C.c_int, c_double_p, c_int32_p]
        L.fo_surface_eval.argtypes = [vp, c_double_p, c_double_p, C.c_int64, C.c_int, c_double_p]
        L.fo_kinect_rays.argtypes = [C.c_int, C.c_int, C.c_double, C.c_double, c_double_p]
        L.fo_raycast.argtypes = [vp, c_double_p, c_double_p, C.c_int64, c_double_p, C.c_int, C.c_int, c_double_p, c_double_p, C.POINTER(C.c_int64)]
        L.fo_estimate_state.argtypes = [vp, c_double_p, c_double_p, C.c_int64, C.POINTER(SolverParams), c_double_p, C.c_int, C.c_int, c_double_p, C.POINTER(C.c_int), c_double_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(c_double_p)


class Oracle:
    def __init__(self, manipulator, normalize_by_gradient=1, poly_degree=1, threads=None):
        L = lib()
        self.packed = indep.Packed(model_dict(manipulator))
        self.h = L.fo_model_create(C.byref(self.packed.desc), normalize_by_gradient, poly_degree)
        if not self.h:
            raise RuntimeError("oracle: " + L.fo_last_error().decode())
        self.nx = L.fo_num_states(self.h)
        self.nq = L.fo_num_positions(self.h)
        self.n_surfaces = self.packed.n_surfaces
        self.threads = threads or L.fo_max_threads()

    def __del__(self):
        try:
            lib().fo_model_destroy(self.h)
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise RuntimeError("oracle: " + lib().fo_last_error().decode())

    def hull_info(self, s):
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        self._chk(lib().fo_hull_info(self.h, s, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def cost(self, x, pts, w_def=10.0, mode=BRUTE, argmin=False):
        x = np.ascontiguousarray(x, float); pts = np.ascontiguousarray(pts, float).reshape(-1, 3)
        c = C.c_double()
        arg = np.empty(len(pts), np.int32) if argmin else None
        self._chk(lib().fo_cost(self.h, _p(x), _p(pts), len(pts), w_def, mode, self.threads, C.byref(c), arg.ctypes.data_as(c_int32_p) if argmin else None))
        return (c.value, arg) if argmin else c.value

    def cost_grad(self, x, pts, w_def=10.0, mode=BRUTE, method="dual"):
        x = np.ascontiguousarray(x, float); pts = np.ascontiguousarray(pts, float).reshape(-1, 3)
        c = C.c_double(); g = np.zeros(self.nx); passes = C.c_int()
        if method == "dual":
            self._chk(lib().fo_cost_grad_dual(self.h, _p(x), _p(pts), len(pts), w_def, mode, self.threads, C.byref(c), _p(g), C.byref(passes)))
        else:
            self._chk(lib().fo_cost_grad_adjoint(self.h, _p(x), _p(pts), len(pts), w_def, mode, self.threads, C.byref(c), _p(g)))
        return c.value, g

    def skin(self, x, pts, mode=BRUTE, argmin=False):
        x = np.ascontiguousarray(x, float); pts = np.ascontiguousarray(pts, float).reshape(-1, 3)
        d = np.empty(len(pts)); arg = np.empty(len(pts), np.int32)
        self._chk(lib().fo_skin_eval(self.h, _p(x), _p(pts), len(pts), mode, self.threads, _p(d), arg.ctypes.data_as(c_int32_p)))
        return (d, arg) if argmin else d

    def surfaces(self, x, pts, mode=BRUTE):
        x = np.ascontiguousarray(x, float); pts = np.ascontiguousarray(pts, float).reshape(-1, 3)
        out = np.empty((len(pts), self.n_surfaces))
        self._chk(lib().fo_surface_eval(self.h, _p(x), _p(pts), len(pts), mode, _p(out)))
        return out

    def raycast(self, x, rays, cam, mode=BRUTE):
        x = np.ascontiguousarray(x, float); rays = np.ascontiguousarray(rays, float).reshape(-1, 3)
        cam = np.ascontiguousarray(cam, float).reshape(12)
        depths = np.empty(len(rays)); hits = np.empty((len(rays), 3)); nh = C.c_int64()
        self._chk(lib().fo_raycast(self.h, _p(x), _p(rays), len(rays), _p(cam), mode, self.threads, _p(depths), _p(hits), C.byref(nh)))
        return depths, hits[:nh.value].copy()

    def estimate_state(self, x0, pts, rate=0.1, max_step=0.5, iteration_limit=30, tol=0.0, divisors=None, w_def=10.0, mode=BRUTE, trace=False):
        x = np.array(x0, float); pts = np.ascontiguousarray(pts, float).reshape(-1, 3)
        sp = SolverParams(rate, max_step, iteration_limit, tol, w_def)
        c = C.c_double(); it = C.c_int()
        tr = np.full((iteration_limit, 1 + self.nx), np.nan) if trace else None
        div = None if divisors is None else np.ascontiguousarray(divisors, float)
        self._chk(lib().fo_estimate_state(self.h, _p(x), _p(pts), len(pts), C.byref(sp), None if div is None else _p(div), mode, self.threads,
                                          C.byref(c), C.byref(it), None if tr is None else _p(tr)))
        return (x, c.value, it.value, tr) if trace else (x, c.value, it.value)


def kinect_rays(rows, cols, vfov=0.4682, hfov=0.5449):
    r = np.empty((rows, cols, 3))
    lib().fo_kinect_rays(rows, cols, vfov, hfov, _p(r))
    return r
