"""CPU tests of the oracle (oracle/, the fp64 restatement of the reference path): the reference's
only known-answer test, the committed golden vectors, and internal cross-checks for every boundary
the reference leaves unpinned."""
import os

import numpy as np
import pytest

import flash
import oracle_lib as ol

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_v1.npz"))
BUILDERS = {"two_link_arm": flash.models.two_link_arm, "beanbag": flash.models.beanbag, "squishable": flash.models.squishable,
            "irb140": flash.models.irb140, "irb_and_squishable": flash.models.irb_and_squishable}


def test_reference_known_answer_beanbag_skin():
    """test/runtests.jl:13-18: skin(SVector(100.0, 0, 0)) ~ 99.0 with rtol 2e-2 on the default beanbag."""
    m = flash.models.beanbag()
    x = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    v = ol.Oracle(m).skin(x, [[100.0, 0.0, 0.0]])[0]
    assert np.isclose(v, 99.0, rtol=2e-2)
    assert np.isclose(v, 98.8930103718821, rtol=1e-12)  # the value of the restated definition


def test_rbf_switches_select_the_pinned_definition():
    """only r^3 + polynomial + f/|grad f| satisfies the reference test."""
    m = flash.models.beanbag()
    x = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    raw = ol.Oracle(m, normalize_by_gradient=0).skin(x, [[100.0, 0.0, 0.0]])[0]
    assert np.isclose(raw, 162.25712187121368, rtol=1e-10) and not np.isclose(raw, 99.0, rtol=2e-2)
    const_poly = ol.Oracle(m, poly_degree=0).skin(x, [[100.0, 0.0, 0.0]])[0]
    assert np.isclose(const_poly, 98.8930103718821, rtol=1e-9)  # symmetric beanbag cannot tell const from linear


@pytest.mark.parametrize("name", sorted(BUILDERS))
def test_golden_vectors(name):
    o = ol.Oracle(BUILDERS[name](), threads=2)
    x, pts = GOLD[name + "/x"], GOLD[name + "/pts"]
    c, g = o.cost_grad(x, pts, mode=ol.BRUTE, method="dual")
    assert np.isclose(c, float(GOLD[name + "/cost"]), rtol=1e-12)
    assert np.allclose(g, GOLD[name + "/grad"], rtol=1e-9, atol=1e-10 * np.abs(g).max())
    per = o.surfaces(x, pts, mode=ol.BRUTE)
    assert np.allclose(per, GOLD[name + "/per_surface"], rtol=1e-12, atol=1e-14)
    _, arg = o.cost(x, pts, argmin=True)
    assert (arg == GOLD[name + "/argmin"]).all()


@pytest.mark.parametrize("name", sorted(BUILDERS))
def test_gradient_three_ways(name):
    """ForwardDiff-style chunked duals == single-pass adjoint == central differences."""
    o = ol.Oracle(BUILDERS[name](), threads=2)
    x, pts = GOLD[name + "/x"], GOLD[name + "/pts"][:64]
    c1, g1 = o.cost_grad(x, pts, method="dual")
    c2, g2 = o.cost_grad(x, pts, method="adjoint")
    assert np.isclose(c1, c2, rtol=1e-12)
    assert np.abs(g1 - g2).max() <= 1e-9 * np.abs(g1).max()
    h = 1e-6
    gf = np.array([(o.cost(x + h * e, pts) - o.cost(x - h * e, pts)) / (2 * h) for e in np.eye(o.nx)])
    assert np.abs(g1 - gf).max() <= 2e-6 * np.abs(g1).max()


def test_forwarddiff_chunk_rule():
    """63 states -> Dual{9} in 7 passes (examples/irb_and_squishable.ipynb:479-483); 43 -> 9 in 5."""
    import ctypes as C
    o = ol.Oracle(flash.models.irb_and_squishable(), threads=2)
    c = C.c_double(); g = np.zeros(o.nx); passes = C.c_int()
    x, pts = GOLD["irb_and_squishable/x"], np.ascontiguousarray(GOLD["irb_and_squishable/pts"][:8])
    ol.lib().fo_cost_grad_dual(o.h, ol._p(x), ol._p(pts), 8, 10.0, 0, 1, C.byref(c), ol._p(g), C.byref(passes))
    assert passes.value == 7
    o2 = ol.Oracle(flash.models.squishable())
    ol.lib().fo_cost_grad_dual(o2.h, ol._p(GOLD["squishable/x"]), ol._p(pts), 8, 10.0, 0, 1, C.byref(c), ol._p(np.zeros(43)), C.byref(passes))
    assert passes.value == 5


def test_gjk_equals_brute_force_point_polytope_distance():
    rng = np.random.default_rng(0)
    m = flash.models.irb_and_squishable()
    o = ol.Oracle(m)
    x = GOLD["irb_and_squishable/x"]
    pts = rng.uniform([-0.8, -0.8, -0.3], [1.0, 0.8, 1.2], (4000, 3))
    a, b = o.surfaces(x, pts, mode=ol.BRUTE), o.surfaces(x, pts, mode=ol.GJK)
    assert np.abs(a - b).max() < 1e-12
    assert (a[:, :7] < 0).any()  # some points are inside hulls: the exact-depth branch is exercised


def test_convex_distance_against_analytic_box():
    """The 8-point table of examples/irb_and_squishable.ipynb cell 3 is a box: closed-form distance."""
    m = flash.models.table()
    o = ol.Oracle(m)
    x = np.array([1.0, 0, 0, 0, 0, 0, 0])
    rng = np.random.default_rng(1)
    pts = rng.uniform(-0.6, 0.6, (2000, 3))
    hw = np.array([0.25, 0.25, 0.05])
    q = np.abs(pts) - hw
    ref = np.linalg.norm(np.maximum(q, 0), axis=1) + np.minimum(q.max(1), 0)
    assert np.abs(o.skin(x, pts) - ref).max() < 1e-12
    assert o.hull_info(0) == (8, 6, 12)


def test_kinematics_irb140_joint_chain():
    """FK self-check against the URDF joint table: with q = 0 a vertex of link_6's hull sits
    at base * joint origins * visual origin; rotating joint 1 by 90 deg about (0,-1,0) in the
    rpy(-1.57079,0,0) frame rotates the arm about world z."""
    m = flash.models.irb140()
    o = ol.Oracle(m)
    probe = np.array([[0.515, 0.0, 0.712]])   # on link_6's flange axis in the world at q = 0 (derived from SURVEY 9)
    d0 = o.surfaces(np.zeros(6), probe)[0]
    assert d0[6] < 0.03                        # close to / inside link_6's hull
    q = np.zeros(6); q[0] = np.pi / 2
    c, s = np.cos(np.pi / 2), np.sin(np.pi / 2)
    for sign in (1, -1):
        Rz = np.array([[c, -sign * s, 0], [sign * s, c, 0], [0, 0, 1]])
        if abs(o.surfaces(q, probe @ Rz.T)[0][6] - d0[6]) < 1e-4:
            break
    else:
        raise AssertionError("joint 1 does not rotate the arm about world z")


def test_quaternion_normalisation_inside_cost():
    """src/gradientdescent.jl:30: cost(x) is invariant to the scale of each quaternion, so the gradient
    is orthogonal to it."""
    m = flash.models.squishable()
    o = ol.Oracle(m)
    x, pts = GOLD["squishable/x"].copy(), GOLD["squishable/pts"][:64]
    c0, g = o.cost_grad(x, pts)
    x2 = x.copy(); x2[:4] *= 3.7
    assert np.isclose(o.cost(x2, pts), c0, rtol=1e-12)
    assert abs(g[:4] @ x[:4]) < 1e-9 * np.abs(g).max()


def test_deformation_regulariser():
    """src/gradientdescent.jl:33-37 with default_deformation_cost_weight = 10 (:7)."""
    m = flash.models.beanbag()
    o = ol.Oracle(m)
    x = GOLD["beanbag/x"]
    pts = np.zeros((0, 3))
    assert np.isclose(o.cost(x, pts, w_def=10.0), 10.0 * (x[7:] ** 2).sum())
    assert np.isclose(o.cost(x, pts, w_def=3.0), 3.0 * (x[7:] ** 2).sum())


def test_raycast_two_link_arm_k1():
    """K1 (examples/manipulator.ipynb cells 2, 6): 41x41 Kinect at (0,0,4) looking down, q = [pi, 1.3].
    Hits lie on the zero level set; misses are NaN; hit order is row-major."""
    m = flash.models.two_link_arm()
    o = ol.Oracle(m)
    rays = ol.kinect_rays(41, 41)
    assert np.allclose(np.linalg.norm(rays, axis=2), 1.0)
    cx = 21.0
    assert np.allclose(rays[20, 20], [0, 0, 1])          # centre pixel, camera_cx = (cols + 1) / 2
    assert np.isclose(rays[0, 40, 0] / rays[0, 40, 2], (41 - cx) * np.tan(0.4682) / cx)   # x uses vertical_fov (sic)
    cam = flash.depthsensors.affine((0, 0, 4), (1, 0, 0), np.pi)
    x = np.array([np.pi, 1.3])
    depths, hits = o.raycast(x, rays.reshape(-1, 3), cam)
    n = np.isfinite(depths).sum()
    assert 20 < n < 400 and len(hits) == n
    assert np.abs(o.skin(x, hits)).max() < 1e-2           # 1000 * EPS acceptance band (src/depthsensors.jl:76)
    idx = np.nonzero(np.isfinite(depths))[0]
    w = (cam[:, :3] @ (rays.reshape(-1, 3)[idx] * depths[idx, None]).T).T + cam[:, 3]
    assert np.allclose(w, hits, atol=1e-12)


def test_estimate_state_converges_on_k1():
    """src/tracking.jl:8-27 on the two-link arm started within 0.5 rad (examples/manipulator.ipynb cell 13)."""
    m = flash.models.two_link_arm()
    o = ol.Oracle(m)
    cam = flash.depthsensors.affine((0, 0, 4), (1, 0, 0), np.pi)
    x_true = np.array([np.pi, 1.3])
    _, pts = o.raycast(x_true, ol.kinect_rays(41, 41).reshape(-1, 3), cam)
    x0 = x_true + np.array([0.3, -0.25])
    xs, cost, iters, tr = o.estimate_state(x0, pts, rate=1.0, max_step=0.5, iteration_limit=30, trace=True)
    assert iters == 30
    assert (np.diff(tr[:, 0]) < 1e-9).all()          # monotone descent with this step size
    assert cost < 0.1 * tr[0, 0]
    assert np.abs(xs - x_true).max() < 0.5 * np.abs(x0 - x_true).max()
    # the reference default (rate 0.1, max_step 0.5, 30 iterations, src/tracking.jl:12-15) also descends
    _, c_def, _ = o.estimate_state(x0, pts)
    assert c_def < tr[0, 0]
