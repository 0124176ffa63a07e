"""Runs ON THE GPU BOX: the on-device solver loop on models WITH hulls (IRB140: 6 joints, one 640 x 480 frame;
the K4 / K5 scene: IRB140 floating + squishable + table), 30 iterations per frame.  usage: [ENV...] python tools/k2_track_probe.py"""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
import numpy as np  # noqa: E402
import bench  # noqa: E402
import flash  # noqa: E402
from flash import _lib  # noqa: E402
from flash._desc import SolverParams, c_double_p  # noqa: E402

L = _lib.lib()
out = {}
rng = np.random.default_rng(2)
cases = {}
m = flash.models.irb140()
cam = flash.depthsensors.affine((0, 1.5, 0.5), (1, 0, 0), np.pi / 2)
pts, _ = bench.frame_cloud(flash, flash.ManipulatorState(m), cam)
cases["irb140_307200_points"] = (m, pts, np.zeros(6) + 0.1 * rng.standard_normal(6), SolverParams(0.1, 0.5, 30, 0.0, 10.0))
m4 = bench.scene_model(flash)
st4 = flash.ManipulatorState(m4)
x4 = bench.scene_states(flash, m4, 1)[0]
flash.gradientdescent.unflatten_(st4, x4)
cam4 = flash.depthsensors.affine((0.3, 1.8, 1.0), (1, 0, 0), np.pi / 2 + 0.25)
p4, _ = bench.frame_cloud(flash, st4, cam4)
x4p = x4.copy(); x4p[:flash.num_positions(m4)] += 0.01 * rng.standard_normal(flash.num_positions(m4))
cases["scene_307200_points"] = (m4, p4, x4p, SolverParams(0.01, 0.1, 30, 0.0, 10.0))
for name, (mm, cloud, x0, sp) in cases.items():
    f = flash.gradientdescent.CostFunctor(mm, cloud)
    c0 = f(x0)
    ts, ends = [], []
    for _ in range(5):
        X = x0.reshape(1, -1).copy(); cost = np.zeros(1)
        t0 = time.perf_counter()
        mm.check(L.flash_estimate_state(mm.ctx(), X.ctypes.data_as(c_double_p), 1, C.byref(sp), None, cost.ctypes.data_as(c_double_p), None))
        ts.append(time.perf_counter() - t0); ends.append(float(cost[0]))
    out[name] = {"ms_per_frame": 1e3 * min(ts[1:]), "cost_start": float(c0), "cost_end": ends[-1], "x_end_head": X[0, :4].tolist()}
print(json.dumps(out))
