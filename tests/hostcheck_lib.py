"""ctypes binding of tests/hostcheck/libhostcheck.so: the product's __host__ __device__ math compiled
for the CPU and replayed serially (TEST INFRASTRUCTURE; lets CPU-only CI check the kernel formulas)."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HC_DIR = os.path.join(ROOT, "tests", "hostcheck")
LIB = os.path.join(HC_DIR, "libhostcheck.so")
CSRC = os.path.join(ROOT, "point-cloud-signed-distance_b200", "csrc")

import sys
sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
from flash._desc import ModelDesc, Options, PackedModel, c_double_p, c_int32_p  # noqa: E402

_lib = None
c_float_p = C.POINTER(C.c_float)


def build():
    srcs = [os.path.join(HC_DIR, "hostcheck.cpp"), os.path.join(CSRC, "flash_dev.h"), os.path.join(CSRC, "flash_host.h"), os.path.join(CSRC, "flash_bsp.h"),
            os.path.join(ROOT, "include", "flashsdf.h")]
    if os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(s) for s in srcs):
        return
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-Wno-unused-function", "-Wno-maybe-uninitialized", "-Wno-unknown-pragmas",
                           "-ffp-contract=off", "-DFLASH_COUNTERS", "-o", LIB, os.path.join(HC_DIR, "hostcheck.cpp")])


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        vp = C.c_void_p
        L.hc_last_error.restype = C.c_char_p
        L.hc_model_create.argtypes = [C.POINTER(ModelDesc), C.POINTER(Options)]; L.hc_model_create.restype = vp
        L.hc_model_destroy.argtypes = [vp]; L.hc_model_destroy.restype = None
        L.hc_info.argtypes = [vp, c_int32_p]; L.hc_info.restype = None
        L.hc_convex_info.argtypes = [vp, C.c_int, c_int32_p, c_int32_p, c_float_p, c_float_p, c_float_p]; L.hc_convex_info.restype = None
        L.hc_prepare.argtypes = [vp, c_double_p, c_double_p, c_double_p]; L.hc_prepare.restype = None
        L.hc_accumulate.argtypes = [vp, c_double_p, c_float_p, C.c_longlong, C.c_int, c_double_p, c_double_p, c_int32_p]; L.hc_accumulate.restype = None
        L.hc_epilogue.argtypes = [vp, c_double_p, c_double_p, c_double_p, c_double_p, C.c_double, c_double_p, c_double_p]; L.hc_epilogue.restype = None
        L.hc_counters.argtypes = [C.POINTER(C.c_longlong), C.c_int]; L.hc_counters.restype = None
        L.hc_hull_sd.argtypes = [vp, C.c_int, c_float_p, C.c_longlong, c_float_p, c_float_p]; L.hc_hull_sd.restype = None
        L.hc_hull_sd_bsp.argtypes = [vp, C.c_int, c_float_p, C.c_longlong, c_float_p, c_float_p, c_int32_p]; L.hc_hull_sd_bsp.restype = C.c_int
        L.hc_bsp_stats.argtypes = [vp, C.c_int, c_double_p]; L.hc_bsp_stats.restype = None
        L.hc_bsp_depth.argtypes = [vp, C.c_int, c_float_p, C.c_longlong, c_int32_p]; L.hc_bsp_depth.restype = None
        L.hc_skin_bound.argtypes = [vp, C.c_int, c_double_p, c_float_p, C.c_longlong, c_float_p, c_float_p, c_double_p]; L.hc_skin_bound.restype = None
        L.hc_set_group_cull.argtypes = [C.c_int]; L.hc_set_group_cull.restype = None
        L.hc_group_stats.argtypes = [C.POINTER(C.c_longlong), C.c_int]; L.hc_group_stats.restype = None
        L.hc_set_skin_cull.argtypes = [C.c_int]; L.hc_set_skin_cull.restype = None
        L.hc_set_exact.argtypes = [C.c_int]; L.hc_set_exact.restype = None
        L.hc_skin_groups_skipped.argtypes = [C.c_int]; L.hc_skin_groups_skipped.restype = C.c_longlong
        L.hc_skin_ball_lower.argtypes = [vp, c_double_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float]; L.hc_skin_ball_lower.restype = C.c_float
        _lib = L
    return _lib


class HostCheck:
    def __init__(self, manipulator):
        L = lib()
        self.packed = PackedModel(manipulator)
        self.h = L.hc_model_create(C.byref(self.packed.desc), None)
        if not self.h:
            raise RuntimeError("hostcheck: " + L.hc_last_error().decode())
        info = np.zeros(8, np.int32)
        L.hc_info(self.h, info.ctypes.data_as(c_int32_p))
        self.nx, self.nq, self.tab_len, self.acc_len, self.aux_len, self.n_convex, self.n_skins = [int(v) for v in info[:7]]

    def __del__(self):
        try:
            lib().hc_model_destroy(self.h)
        except Exception:
            pass

    def prepare(self, x):
        x = np.ascontiguousarray(x, float)
        tab = np.zeros(self.tab_len); aux = np.zeros(self.aux_len)
        lib().hc_prepare(self.h, x.ctypes.data_as(c_double_p), tab.ctypes.data_as(c_double_p), aux.ctypes.data_as(c_double_p))
        return tab, aux

    def accumulate(self, tab, pts, use_grid=0):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        acc = np.zeros(self.acc_len); dist = np.empty(len(pts)); arg = np.empty(len(pts), np.int32)
        lib().hc_accumulate(self.h, tab.ctypes.data_as(c_double_p), pts.ctypes.data_as(c_float_p), len(pts), use_grid,
                            acc.ctypes.data_as(c_double_p), dist.ctypes.data_as(c_double_p), arg.ctypes.data_as(c_int32_p))
        return acc, dist, arg

    def epilogue(self, x, tab, aux, acc, w_def=10.0):
        x = np.ascontiguousarray(x, float)
        cost = C.c_double(); grad = np.zeros(self.nx)
        lib().hc_epilogue(self.h, x.ctypes.data_as(c_double_p), tab.ctypes.data_as(c_double_p), aux.ctypes.data_as(c_double_p),
                          acc.ctypes.data_as(c_double_p), w_def, C.byref(cost), grad.ctypes.data_as(c_double_p))
        return cost.value, grad

    def cost_grad(self, x, pts, w_def=10.0):
        tab, aux = self.prepare(x)
        acc, dist, arg = self.accumulate(tab, pts)
        c, g = self.epilogue(x, tab, aux, acc, w_def)
        return c, g, dist, arg

    def convex_info(self, j):
        nv, npl = C.c_int32(), C.c_int32(); lo = np.zeros(3, np.float32); hi = np.zeros(3, np.float32); db = C.c_float()
        lib().hc_convex_info(self.h, j, C.byref(nv), C.byref(npl), lo.ctypes.data_as(c_float_p), hi.ctypes.data_as(c_float_p), C.byref(db))
        return nv.value, npl.value, lo, hi, db.value

    def skin_bound(self, k, tab, pts):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 3)
        lo = np.empty(len(pts), np.float32); hi = np.empty(len(pts), np.float32); s = np.empty(len(pts))
        lib().hc_skin_bound(self.h, k, tab.ctypes.data_as(c_double_p), pts.ctypes.data_as(c_float_p), len(pts), lo.ctypes.data_as(c_float_p),
                            hi.ctypes.data_as(c_float_p), s.ctypes.data_as(c_double_p))
        return lo, hi, s

    def skin_ball_lower(self, k, tab, centre, rho):
        c = np.asarray(centre, np.float32)
        return float(lib().hc_skin_ball_lower(self.h, tab.ctypes.data_as(c_double_p), k, float(c[0]), float(c[1]), float(c[2]), float(np.float32(rho))))

    def hull_sd(self, j, y):
        y = np.ascontiguousarray(y, np.float32).reshape(-1, 3)
        d = np.empty(len(y), np.float32); n = np.empty((len(y), 3), np.float32)
        lib().hc_hull_sd(self.h, j, y.ctypes.data_as(c_float_p), len(y), d.ctypes.data_as(c_float_p), n.ctypes.data_as(c_float_p))
        return d, n

    def hull_sd_bsp(self, j, y):
        """signed distance through the closest-feature BSP tree; returns d, n, leaf words (None if the hull has no tree)"""
        y = np.ascontiguousarray(y, np.float32).reshape(-1, 3)
        d = np.empty(len(y), np.float32); n = np.empty((len(y), 3), np.float32); leaf = np.empty(len(y), np.int32)
        ok = lib().hc_hull_sd_bsp(self.h, j, y.ctypes.data_as(c_float_p), len(y), d.ctypes.data_as(c_float_p), n.ctypes.data_as(c_float_p), leaf.ctypes.data_as(c_int32_p))
        return (d, n, leaf) if ok else None

    def bsp_stats(self, j):
        out = np.zeros(5)
        lib().hc_bsp_stats(self.h, j, out.ctypes.data_as(c_double_p))
        return dict(zip(("nodes", "max_depth", "mean_leaf_depth", "cells", "dropped"), out.tolist()))

    def bsp_depth(self, j, y):
        y = np.ascontiguousarray(y, np.float32).reshape(-1, 3)
        dep = np.empty(len(y), np.int32)
        lib().hc_bsp_depth(self.h, j, y.ctypes.data_as(c_float_p), len(y), dep.ctypes.data_as(c_int32_p))
        return dep


def counters(reset=True):
    out = (C.c_longlong * 5)()
    lib().hc_counters(out, 1 if reset else 0)
    return dict(zip(("points", "narrow", "lb_tests", "skin64", "bsp_levels"), list(out)))


def set_group_cull(on):
    lib().hc_set_group_cull(1 if on else 0)


def set_skin_cull(on):
    lib().hc_set_skin_cull(1 if on else 0)


def set_exact(on):
    """exact centre distances + BSP start nodes in the group pre-pass (group_prepass) on / off"""
    lib().hc_set_exact(1 if on else 0)


def skin_groups_skipped(reset=True):
    return int(lib().hc_skin_groups_skipped(1 if reset else 0))


def group_stats(reset=True):
    out = (C.c_longlong * 2)()
    lib().hc_group_stats(out, 1 if reset else 0)
    return {"candidates": out[0], "groups": out[1]}
