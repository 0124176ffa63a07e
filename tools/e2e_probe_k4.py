"""Host-side time of each call of the pipelined K4 batch step (8 frames x 16 configs)."""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bench, flash
from flash import _lib
from flash._desc import c_double_p
L = _lib.lib()
m = bench.scene_model(flash); ctx = m.ctx()
X4 = bench.scene_states(flash, m, 16)
F, Cn, nx = 8, 16, X4.shape[1]
X = np.ascontiguousarray(np.repeat(X4[None], F, 0))
rng = np.random.default_rng(0)
keep, ptrs = [], []
for f in range(F):
    t = torch.empty((307200, 3), dtype=torch.float64, pin_memory=True); a = t.numpy()
    a[:] = bench.scene_points(307200, seed=50 + f); keep.append(t); ptrs.append(a.ctypes.data_as(c_double_p))
P_ = (c_double_p * F)(*ptrs); cnt = (C.c_int64 * F)(*[307200] * F)
cost = np.empty((F, Cn)); grad = np.empty((F, Cn, nx))
pX, pc, pg = X.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p)
m.check(L.flash_frames_set(ctx, P_, cnt, F))
for _ in range(3): m.check(L.flash_cost_grad_frames(ctx, pX, Cn, 10.0, pc, pg))
t0 = time.perf_counter(); m.check(L.flash_cost_grad_frames(ctx, pX, Cn, 10.0, pc, pg)); print("cost_grad_frames resident %.3f ms" % (1e3 * (time.perf_counter() - t0)))
t0 = time.perf_counter(); m.check(L.flash_frames_set(ctx, P_, cnt, F)); print("frames_set blocking %.3f ms" % (1e3 * (time.perf_counter() - t0)))
t0 = time.perf_counter(); m.check(L.flash_frames_set_async(ctx, P_, cnt, F)); t1 = time.perf_counter(); m.check(L.flash_points_sync(ctx)); print("frames_set_async call %.3f ms, until done %.3f ms" % (1e3 * (t1 - t0), 1e3 * (time.perf_counter() - t0)))
for k in range(5):
    t0 = time.perf_counter(); m.check(L.flash_points_swap(ctx)); t1 = time.perf_counter()
    m.check(L.flash_frames_set_async(ctx, P_, cnt, F)); t2 = time.perf_counter()
    m.check(L.flash_cost_grad_frames(ctx, pX, Cn, 10.0, pc, pg)); t3 = time.perf_counter()
    m.check(L.flash_points_sync(ctx)); t4 = time.perf_counter()
    print("step %d: swap %.3f set_async %.3f cost_grad %.3f copy-stream tail %.3f ms" % (k, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3)))
