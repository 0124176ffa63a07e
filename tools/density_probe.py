"""Per-evaluation cost of cost_kernel against cloud size and density (one GPU): the 10 M-point K5 cloud, every 8th point of it
(thin), one compact eighth (dense), the interleaved-slab shard of rank 0 of 8 (dense), and 2.5 M / 5 M thin clouds."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
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
full = bench.scene_points(10_000_000)
order = np.argsort(full[:, 0], kind="stable")
clouds = {"full 10M": full, "thin 5M (every 2nd)": full[::2], "thin 2.5M (every 4th)": full[::4], "thin 1.25M (every 8th)": full[::8],
          "dense compact eighth (x-sorted, middle)": full[order[4_375_000:5_625_000]], "dense slabs of rank 0 of 8": bench.scene_points(10_000_000, 0, 8),
          "dense compact half": full[order[2_500_000:7_500_000]]}
m.check(L.flash_profiling(ctx, 1))
tm = Timing()
for name, p in clouds.items():
    p = np.ascontiguousarray(p)
    m.check(L.flash_points_set(ctx, p.ctypes.data_as(c_double_p), len(p)))
    ts = []
    for k in range(6):
        m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))
        L.flash_last_timing(ctx, C.byref(tm))
        if k >= 2: ts.append((tm.ms_main - tm.ms_group, tm.ms_group))
    a = np.mean(ts, 0)
    print("%-42s n %8d  cost %.3f ms  group %.3f ms  -> %.2f ps per evaluation (cost kernel)" % (name, len(p), a[0], a[1], 1e9 * a[0] / (len(p) * Cn)))
