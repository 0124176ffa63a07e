"""Host-side time of each call of the pipelined end-to-end step (K5): where does the overlap go?"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bench, flash
from flash import _lib
from flash._desc import c_double_p
L = _lib.lib()
m = bench.scene_model(flash); ctx = m.ctx()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
Cn = 64
X = np.ascontiguousarray(bench.scene_states(flash, m, Cn)); nx = X.shape[1]
pin = torch.empty((N, 3), dtype=torch.float64, pin_memory=True); pts = pin.numpy(); pts[:] = bench.scene_points(N)
cost = np.empty(Cn); grad = np.empty((Cn, nx))
pX, pc, pg, pp = X.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p), pts.ctypes.data_as(c_double_p)
m.check(L.flash_points_set(ctx, pp, N))
for _ in range(3): m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))
t0 = time.perf_counter(); m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None)); print("cost_grad resident %.2f ms" % (1e3 * (time.perf_counter() - t0)))
t0 = time.perf_counter(); m.check(L.flash_points_set(ctx, pp, N)); print("points_set blocking %.2f ms" % (1e3 * (time.perf_counter() - t0)))
t0 = time.perf_counter(); m.check(L.flash_points_set_async(ctx, pp, N)); t1 = time.perf_counter(); m.check(L.flash_points_sync(ctx)); print("set_async call %.2f ms, until done %.2f ms" % (1e3 * (t1 - t0), 1e3 * (time.perf_counter() - t0)))
for k in range(5):
    t0 = time.perf_counter(); m.check(L.flash_points_swap(ctx)); t1 = time.perf_counter()
    m.check(L.flash_points_set_async(ctx, pp, N)); t2 = time.perf_counter()
    m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None)); t3 = time.perf_counter()
    m.check(L.flash_points_sync(ctx)); t4 = time.perf_counter()
    print("step %d: swap %.2f set_async %.2f cost_grad %.2f copy-stream tail %.2f ms" % (k, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3)))
