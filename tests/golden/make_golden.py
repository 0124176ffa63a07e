"""Writes tests/golden/golden_v1.npz: seeded inputs and the fp64 oracle's outputs for every model
builder of src/models.jl plus the IRB140 scenes.  Julia cannot run here, so these are ORACLE outputs
(regression pins + GPU parity targets), not reference outputs; the only reference-held known answer
is test/runtests.jl:17 and it is asserted directly in tests/test_oracle.py.

Also writes squishable_unsquished_xyz.npy: the xyz columns (as float32) of the reference's real Kinect
cloud examples/data/squishable_unsquished_xyzrgb.txt, read with flash.depthdata.read_point_cloud
(only when /root/reference is present).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import flash  # noqa: E402
import oracle_lib as ol  # noqa: E402

CASES = {
    "two_link_arm": (flash.models.two_link_arm, [-2.2, -2.2, -0.5], [2.2, 2.2, 0.5]),
    "beanbag": (flash.models.beanbag, [-3, -3, -3], [3, 3, 3]),
    "squishable": (flash.models.squishable, [-0.5, -0.5, -0.5], [0.5, 0.5, 0.5]),
    "irb140": (flash.models.irb140, [-0.6, -0.7, -0.2], [0.9, 0.7, 1.1]),
    "irb_and_squishable": (flash.models.irb_and_squishable, [-0.8, -0.8, -0.3], [1.0, 0.8, 1.2]),
}


def main():
    out = {}
    for i, (name, (builder, lo, hi)) in enumerate(CASES.items()):
        rng = np.random.default_rng(100 + i)
        m = builder()
        o = ol.Oracle(m, threads=1)
        nx, nq = o.nx, o.nq
        x = flash.gradientdescent.flatten(flash.ManipulatorState(m))
        x = x + np.r_[0.25 * rng.standard_normal(nq), 0.02 * rng.standard_normal(nx - nq)]
        pts = rng.uniform(lo, hi, (256, 3)).astype(np.float32).astype(np.float64)
        cost, grad = o.cost_grad(x, pts, mode=ol.BRUTE, method="dual")
        per = o.surfaces(x, pts, mode=ol.BRUTE)
        out[name + "/x"] = x; out[name + "/pts"] = pts; out[name + "/cost"] = np.array(cost)
        out[name + "/grad"] = grad; out[name + "/per_surface"] = per; out[name + "/argmin"] = per.argmin(1).astype(np.int32)
        print(name, "nx", nx, "cost", cost)
    np.savez_compressed(os.path.join(HERE, "golden_v1.npz"), **out)
    ref = "/root/reference/examples/data/squishable_unsquished_xyzrgb.txt"
    if os.path.exists(ref):
        pc = flash.depthdata.read_point_cloud(ref)
        np.save(os.path.join(HERE, "squishable_unsquished_xyz.npy"), pc.positions.astype(np.float32))
        print(pc)


if __name__ == "__main__":
    main()
