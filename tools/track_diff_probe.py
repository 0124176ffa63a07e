"""Runs ON THE GPU BOX: host loop vs device loop of estimate_state on random clouds (squishable), per iteration count."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
import numpy as np
import flash
rng = np.random.default_rng(14)
m = flash.models.squishable()
nx, nq = flash.num_states(m), flash.num_positions(m)
x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m))
x0 = x0 + np.r_[0.05 * rng.standard_normal(nq), 0.005 * rng.standard_normal(nx - nq)]
big = rng.uniform([-0.35, -0.35, -0.3], [0.35, 0.35, 0.3], (307200, 3)).astype(np.float32).astype(np.float64)
div = np.r_[np.ones(7), 1e-3 * np.ones(36)]
for name, cloud in (("big", big), ("small", big[:1000]), ("tiny", big[:20])):
    for it in (1, 2, 4, 8, 12):
        solver = flash.tracking.NaiveSolver(43, rate=0.01, max_step=0.1, iteration_limit=it, precondition_divisors=div)
        xh = flash.tracking.estimate_state(m, cloud, x0, solver=solver)
        xd = flash.tracking.estimate_state(m, cloud, x0, solver=solver, on_device=True)
        print(name, it, "max |host - device| %.3e" % np.abs(xh - xd).max(), "moved %.3e" % np.abs(xd - x0).max())
