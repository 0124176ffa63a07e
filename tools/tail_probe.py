"""Runs ON THE GPU BOX: cost_kernel time on the full K5 cloud and on the shard of rank 0 of 8 (one GPU), for A/B of
scheduling changes.  usage: [ENV...] python tools/tail_probe.py"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
import numpy as np
import bench, flash
from flash import _lib
from flash._desc import Timing, c_double_p
L = _lib.lib()
m = bench.scene_model(flash); ctx = m.ctx()
Cn = 64
X = np.ascontiguousarray(bench.scene_states(flash, m, Cn)); nx = X.shape[1]
cost = np.empty(Cn); grad = np.empty((Cn, nx))
pX, pc, pg = X.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p)
clouds = {"rank 0 of 8 (blobs)": bench.scene_points(10_000_000, 0, 8), "rank 3 of 8 (blobs)": bench.scene_points(10_000_000, 3, 8)}
if "--full" in sys.argv:
    clouds["full 10M"] = bench.scene_points(10_000_000)
m.check(L.flash_profiling(ctx, 1))
tm = Timing()
for name, p in clouds.items():
    p = np.ascontiguousarray(p)
    m.check(L.flash_points_set(ctx, p.ctypes.data_as(c_double_p), len(p)))
    ts = []
    for k in range(8):
        m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))
        L.flash_last_timing(ctx, C.byref(tm))
        if k >= 3: ts.append((tm.ms_main - tm.ms_group, tm.ms_group, tm.ms_total))
    a = np.mean(ts, 0)
    print("%-22s n %8d  cost %.3f ms  group %.3f ms  total %.3f ms -> %.2f ps per evaluation  checksum %.10g" % (name, len(p), a[0], a[1], a[2], 1e9 * a[0] / (len(p) * Cn), cost.sum()))
