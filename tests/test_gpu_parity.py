"""GPU parity tests proper: the CUDA path (through the C ABI) against the fp64 oracle on the same
seeded inputs.  Tolerances (BASELINE.json north_star): cost and gradient within 1e-5 relative
(gradient relative to its max-norm), argmin identical except on ties within fp32 rounding."""
import numpy as np
import pytest

import flash
import oracle_lib as ol

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _state(m, rng, qs=0.2, ds=0.02):
    nx, nq = flash.num_states(m), flash.num_positions(m)
    x = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    return x + np.r_[qs * rng.standard_normal(nq), ds * rng.standard_normal(nx - nq)]


def _points(rng, n, lo, hi):
    return rng.uniform(lo, hi, (n, 3)).astype(np.float32).astype(np.float64)


def _compare(m, X, pts, tie_band=2e-6, w_def=10.0):
    f = flash.gradientdescent.CostFunctor(m, pts)
    cost, grad, arg = f.batch(X, gradient=True, argmin=True, deformation_cost_weight=w_def)
    o = ol.Oracle(m)
    for c in range(len(X)):
        co, go = o.cost_grad(X[c], pts, w_def=w_def, method="adjoint", mode=ol.GJK)
        assert abs(cost[c] - co) <= RTOL * abs(co), (c, cost[c], co)
        assert np.abs(grad[c] - go).max() <= RTOL * np.abs(go).max(), (c, np.abs(grad[c] - go).max(), np.abs(go).max())
        # argmin: identical except where the two best surfaces are within fp32 rounding of each other
        per = o.surfaces(X[c], pts, mode=ol.GJK)
        ao = per.argmin(1)
        bad = np.nonzero(arg[c] != ao)[0]
        for i in bad:
            srt = np.sort(per[i])
            assert srt[1] - srt[0] <= tie_band * max(1.0, abs(srt[0])), (i, per[i], arg[c][i], ao[i])
        assert len(bad) <= max(2, len(pts) // 2000)
    return cost, grad, arg


def test_k2_irb140_cost_grad_argmin():
    rng = np.random.default_rng(2)
    m = flash.models.irb140()
    X = rng.uniform(-np.pi / 2, np.pi / 2, (2, 6))  # examples/irb140.ipynb cell 9
    pts = _points(rng, 30000, [-0.6, -0.7, -0.2], [0.9, 0.7, 1.1])
    _compare(m, X, pts)


def test_k3_squishable():
    rng = np.random.default_rng(3)
    m = flash.models.squishable()
    X = np.stack([_state(m, rng) for _ in range(3)])
    pts = _points(rng, 20000, [-0.5, -0.5, -0.5], [0.5, 0.5, 0.5])
    _compare(m, X, pts)


def test_beanbag_known_answer_and_grad():
    m = flash.models.beanbag()
    st = flash.ManipulatorState(m)
    # the reference's only known-answer test, test/runtests.jl:17
    assert np.isclose(flash.skin(st)([100.0, 0.0, 0.0]), 99.0, rtol=2e-2)
    rng = np.random.default_rng(4)
    X = np.stack([_state(m, rng, ds=0.1) for _ in range(2)])
    _compare(m, X, _points(rng, 5000, [-3, -3, -3], [3, 3, 3]))


def test_k1_two_link_arm():
    rng = np.random.default_rng(1)
    m = flash.models.two_link_arm()
    X = np.array([[np.pi, 1.3], [0.3, -0.7], [2.0, 0.1]]) + rng.uniform(-0.5, 0.5, (3, 2))
    _compare(m, X, _points(rng, 4000, [-2.2, -2.2, -0.5], [2.2, 2.2, 0.5]))


def test_k4_combined_scene_batched():
    rng = np.random.default_rng(5)
    m = flash.models.irb_and_squishable()
    assert flash.num_states(m) == 63
    x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    # placement of examples/irb_and_squishable.ipynb cell 6
    x0[4:7] = [0.0, 0.0, 0.75]
    nq = flash.num_positions(m)
    x0[13 + 4:13 + 7] = [0.55, 0.45, 0.8]
    x0[20 + 4:20 + 7] = [0.4, 0.0, 0.6]
    X = np.stack([x0 + np.r_[0.05 * rng.standard_normal(nq), 0.005 * rng.standard_normal(63 - nq)] for _ in range(4)])
    pts = _points(rng, 20000, [-0.5, -0.5, 0.5], [1.0, 0.5, 2.0])  # scene AABB, cell 6
    _compare(m, X, pts)


def test_skin_eval_matches_oracle():
    rng = np.random.default_rng(6)
    m = flash.models.irb_and_squishable()
    x = _state(m, rng, qs=0.1)
    pts = _points(rng, 10000, [-0.8, -0.8, -0.3], [1.0, 0.8, 1.2])
    st = flash.ManipulatorState(m)
    flash.gradientdescent.unflatten_(st, x)
    d, a = flash.skin(st).eval(pts, return_argmin=True)
    do, ao = ol.Oracle(m).skin(x, pts, mode=ol.GJK, argmin=True)
    assert np.abs(d - do).max() < 2e-6
    assert (a != ao).sum() <= 5


def test_linearity_in_points_full_size():
    """Size-independent property at K2's full size (307,200 points): the cost and gradient of a
    cloud are the sums over any partition of it (what the multi-GPU sharding relies on)."""
    rng = np.random.default_rng(8)
    m = flash.models.irb140()
    X = rng.uniform(-1, 1, (2, 6))
    pts = _points(rng, 307200, [-0.6, -0.7, -0.2], [0.9, 0.7, 1.1])
    full = flash.gradientdescent.CostFunctor(m, pts).batch(X)
    parts = [flash.gradientdescent.CostFunctor(m, pts[s]).batch(X) for s in (slice(0, 100000), slice(100000, 307200))]
    # per-point values do not depend on the partition; only the fp32 warp-level partial sums of the
    # convex wrench terms (rounded ~6e-8 each before the fp64 accumulation) regroup
    assert np.allclose(full[0], parts[0][0] + parts[1][0], rtol=1e-9)
    gsum = parts[0][1] + parts[1][1]
    assert np.abs(full[1] - gsum).max() <= 1e-6 * np.abs(full[1]).max(), np.abs(full[1] - gsum).max() / np.abs(full[1]).max()


def test_empty_and_tiny_inputs():
    m = flash.models.squishable()
    x = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    x[7:] = 0.01
    f = flash.gradientdescent.CostFunctor(m, np.zeros((0, 3)))
    c, g, _ = f.batch(x)
    assert np.isclose(c[0], 10 * 36 * 1e-4)          # only the regulariser, src/gradientdescent.jl:33-37
    assert np.allclose(g[0][7:], 2 * 10 * 0.01) and np.allclose(g[0][:7], 0)
    f1 = flash.gradientdescent.CostFunctor(m, np.array([[0.3, 0.1, 0.2]]))
    c1, g1, a1 = f1.batch(x, argmin=True)
    co, go = ol.Oracle(m).cost_grad(x, [[0.3, 0.1, 0.2]], method="dual")
    assert np.isclose(c1[0], co, rtol=1e-6) and np.allclose(g1[0], go, rtol=1e-5, atol=1e-6 * np.abs(go).max()) and a1[0, 0] == 0


def test_errors_are_reported_not_thrown_across_abi():
    m = flash.models.squishable()
    f = flash.gradientdescent.CostFunctor(m, np.zeros((4, 3)))
    with pytest.raises(ValueError):
        f.batch(np.zeros(5))
    with pytest.raises(flash.FlashError):
        bad = flash.models.beanbag()
        bad.surfaces[0].surface_points[0] = bad.surfaces[0].surface_points[1]  # coincident control points
        flash.gradientdescent.CostFunctor(bad, np.ones((4, 3)))(flash.gradientdescent.flatten(flash.ManipulatorState(bad)))


# ------------------------------------------------------------------ golden vectors through the CUDA path
import os
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_v1.npz"))
BUILDERS = {"two_link_arm": flash.models.two_link_arm, "beanbag": flash.models.beanbag, "squishable": flash.models.squishable,
            "irb140": flash.models.irb140, "irb_and_squishable": flash.models.irb_and_squishable}


@pytest.mark.parametrize("name", sorted(BUILDERS))
def test_golden_vectors_on_gpu(name):
    m = BUILDERS[name]()
    x, pts = GOLD[name + "/x"], GOLD[name + "/pts"]
    c, g, a = flash.gradientdescent.CostFunctor(m, pts).batch(x, argmin=True)
    assert abs(c[0] - float(GOLD[name + "/cost"])) <= RTOL * float(GOLD[name + "/cost"])
    gg = GOLD[name + "/grad"]
    assert np.abs(g[0] - gg).max() <= RTOL * np.abs(gg).max()
    per = GOLD[name + "/per_surface"]
    for i in np.nonzero(a[0] != GOLD[name + "/argmin"])[0]:
        s = np.sort(per[i]); assert s[1] - s[0] < 2e-6


def test_points_f32_and_device_entry_points():
    import ctypes as C
    import torch
    from flash import _lib
    rng = np.random.default_rng(9)
    m = flash.models.irb140()
    X = rng.uniform(-1, 1, (2, 6))
    pts = _points(rng, 50000, [-0.6, -0.7, -0.2], [0.9, 0.7, 1.1])
    f = flash.gradientdescent.CostFunctor(m, pts)
    ref = f.batch(X, argmin=True)
    L = _lib.lib()
    p32 = np.ascontiguousarray(pts, np.float32)
    m.check(L.flash_points_set_f32(m.ctx(), p32.ctypes.data_as(C.POINTER(C.c_float)), len(p32)))
    got = f.batch(X, argmin=True)
    assert np.allclose(got[0], ref[0], rtol=1e-9) and np.abs(got[1] - ref[1]).max() <= 1e-6 * np.abs(ref[1]).max() and (got[2] == ref[2]).all()
    d4 = torch.zeros((len(pts), 4), dtype=torch.float32, device="cuda")
    d4[:, :3] = torch.from_numpy(p32).cuda()
    torch.cuda.synchronize()
    m.check(L.flash_points_set_device(m.ctx(), C.c_void_p(d4.data_ptr()), len(pts)))
    got = f.batch(X, argmin=True)          # caller's order, no binning
    assert np.allclose(got[0], ref[0], rtol=1e-9) and np.abs(got[1] - ref[1]).max() <= 1e-6 * np.abs(ref[1]).max() and (got[2] == ref[2]).all()


def test_k3_real_kinect_cloud():
    """examples/squishable.ipynb: the 25,571-point real cloud (fixture tests/golden/squishable_unsquished_xyz.npy)."""
    pts = np.load(os.path.join(os.path.dirname(__file__), "golden", "squishable_unsquished_xyz.npy")).astype(np.float64)
    m = flash.models.squishable()
    x = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    x[4:7] = pts.mean(0)
    X = np.stack([x, x + 0.01])
    _compare(m, X, pts)


# ------------------------------------------------------------------ depth rendering (row A9)
def _raycast_case(m, x, sensor, cam):
    """GPU ray marcher vs the oracle's.  doRaycast (src/depthsensors.jl:56-81) is a secant iteration
    that either converges onto the surface (|field| <= 1e-5) or, for rays that pass within ~1 cm of
    a silhouette, wanders for 60 iterations and is accepted iff its last sample has |field| <= 1e-2.
    The first kind is well defined and must match; the second kind is chaotic (a 1e-7 difference in
    one field value changes where it ends), so hit/miss flips are accepted only there."""
    st = flash.ManipulatorState(m)
    flash.gradientdescent.unflatten_(st, x)
    skin = flash.skin(st)
    depths = flash.depthsensors.raycast_depths(skin, sensor, cam)
    hits = flash.depthsensors.raycast_points(skin, sensor, cam)
    o = ol.Oracle(m)
    rays = sensor.rays.reshape(-1, 3)
    do, ho = o.raycast(x, rays, cam, mode=ol.GJK)
    d = depths.reshape(-1)
    hit_g, hit_o = np.isfinite(d), np.isfinite(do)
    assert len(hits) == hit_g.sum()
    # hits are in ray order and lie on the rays
    idx = np.nonzero(hit_g)[0]
    w = (cam[:, :3] @ (rays[idx] * d[idx, None]).T).T + cam[:, 3]
    assert np.allclose(w, hits, atol=1e-9)
    f_g = np.full(len(d), np.inf); f_g[hit_g] = np.abs(o.skin(x, hits, mode=ol.GJK))
    f_o = np.full(len(d), np.inf); f_o[hit_o] = np.abs(o.skin(x, ho, mode=ol.GJK))
    assert f_g[hit_g].max() < 1.1e-2                      # every hit passes the 1000*EPS acceptance test
    conv_o, conv_g = f_o <= 1.5e-5, f_g <= 3e-5           # converged marches
    assert hit_g[conv_o].all() and hit_o[conv_g].all()     # no flips on well-defined rays
    both = conv_o & conv_g
    diff = np.abs(d[both] - do[both])
    # converged marches stop anywhere inside the 1e-5 band: depths agree to band/slope; a ray with two
    # crossings a few cm apart can settle on the other one (<= 1%)
    assert (diff > 2e-4).sum() <= max(1, int(0.01 * both.sum())), (diff > 2e-4).sum()
    assert np.median(diff) < 2e-5
    flips = (hit_g != hit_o).sum()
    assert flips <= 0.05 * hit_o.sum()
    return both.sum()


def test_raycast_two_link_arm_k1():
    m = flash.models.two_link_arm()
    sensor = flash.depthsensors.Kinect(41, 41)
    cam = flash.depthsensors.affine((0, 0, 4), (1, 0, 0), np.pi)          # examples/manipulator.ipynb cell 2
    assert _raycast_case(m, np.array([np.pi, 1.3]), sensor, cam) > 20


def test_raycast_irb140():
    m = flash.models.irb140()
    sensor = flash.depthsensors.Kinect(120, 160)
    cam = flash.depthsensors.affine((0, 1.5, 0.5), (1, 0, 0), np.pi / 2)  # examples/irb140.ipynb cell 4
    assert _raycast_case(m, np.zeros(6), sensor, cam) > 500


def test_raycast_then_track_squishable():
    """examples/deformable_manipulator.ipynb flow on the squishable: render a deformed state, start
    from the identity pose, run estimate_state; host loop and on-device loop agree."""
    rng = np.random.default_rng(11)
    m = flash.models.squishable()
    true = flash.ManipulatorState(m)
    true.mechanism_state.q[4:7] = [0.1, -0.05, 0.08]
    true.deformation_data[:] = 0.03 * (rng.random(36) - 0.5)
    sensor = flash.depthsensors.Kinect(80, 80)
    cam = flash.depthsensors.affine((0, 0, 2), (1, 0, 0), np.pi)
    pts = flash.depthsensors.raycast(true, sensor, cam)
    assert len(pts) > 100
    x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    div = np.r_[np.ones(7), 1e-3 * np.ones(36)]                            # examples/squishable.ipynb cell 9
    solver = flash.tracking.NaiveSolver(43, rate=0.01, max_step=0.1, iteration_limit=20, precondition_divisors=div)
    trace = []
    xh = flash.tracking.estimate_state(m, pts, x0, callback=lambda x, c: trace.append(c), solver=solver)
    xd = flash.tracking.estimate_state(m, pts, x0, solver=solver, on_device=True)
    assert len(trace) == 20 and trace[-1] < trace[0]
    assert np.abs(xh - xd).max() < 1e-6 * max(1.0, np.abs(xh).max())
    xo, co, it = ol.Oracle(m).estimate_state(x0, pts, rate=0.01, max_step=0.1, iteration_limit=20, divisors=div, mode=ol.GJK)
    assert np.abs(xh - xo).max() < 1e-4                                    # 20 iterations of a 1e-5-accurate gradient


def test_on_device_estimate_state_two_link_k1():
    """K1 (BASELINE configs[0]) through the on-device loop: the two-link arm's 50 x 50 RBF system takes the
    generic (multi-element-per-thread) inversion path of iterate_kernel; host loop, device loop and the oracle's
    loop must land on the same state (examples/manipulator.ipynb cell 13)."""
    m = flash.models.two_link_arm()
    cam = flash.depthsensors.affine((0, 0, 4), (1, 0, 0), np.pi)
    x_true = np.array([np.pi, 1.3])
    o = ol.Oracle(m)
    _, pts = o.raycast(x_true, ol.kinect_rays(41, 41).reshape(-1, 3), cam)
    x0 = x_true + np.array([0.3, -0.25])
    solver = flash.tracking.NaiveSolver(2, rate=1.0, max_step=0.5, iteration_limit=30)
    xh = flash.tracking.estimate_state(m, pts, x0, solver=solver)
    xd = flash.tracking.estimate_state(m, pts, x0, solver=solver, on_device=True)
    xo, co, it = o.estimate_state(x0, pts, rate=1.0, max_step=0.5, iteration_limit=30, mode=ol.GJK)
    assert np.abs(xh - xd).max() < 1e-8
    assert np.abs(xd - xo).max() < 1e-5
    assert np.abs(xd - x_true).max() < 0.5 * np.abs(x0 - x_true).max()


def test_on_device_estimate_state_batched_with_tolerance():
    rng = np.random.default_rng(12)
    m = flash.models.irb140()
    true = np.zeros(6)
    st = flash.ManipulatorState(m)
    sensor = flash.depthsensors.Kinect(60, 80)
    cam = flash.depthsensors.affine((0, 1.5, 0.5), (1, 0, 0), np.pi / 2)
    pts = flash.depthsensors.raycast(st, sensor, cam)
    import ctypes as C
    from flash import _lib
    from flash._desc import SolverParams, c_double_p, c_int32_p
    f = flash.gradientdescent.CostFunctor(m, pts)
    X = true + 0.2 * rng.standard_normal((4, 6))
    X0 = X.copy()
    sp = SolverParams(20.0, 0.5, 40, 1e-4, 10.0)                            # examples/irb140.ipynb cell 9
    cost = np.zeros(4); iters = np.zeros(4, np.int32)
    m.check(_lib.lib().flash_estimate_state(m.ctx(), X.ctypes.data_as(c_double_p), 4, C.byref(sp), None, cost.ctypes.data_as(c_double_p), iters.ctypes.data_as(c_int32_p)))
    c0, _, _ = f.batch(X0, gradient=False)
    assert (cost < c0).all() and (iters > 0).all() and (iters <= 40).all()
    c1, _, _ = f.batch(X, gradient=False)
    assert np.allclose(c1, cost, rtol=1e-9)


def test_on_device_estimate_state_models_with_hulls_single_config():
    """One configuration of a model WITH hulls takes the persistent loop (track_kernel) without group records, and a
    tolerance ends it early: host loop (one flash_cost_grad per iteration), device loop and the oracle's loop agree,
    for the IRB140 and for the IRB140 + squishable + table scene (hulls and an RBF skin in one loop)."""
    import ctypes as C
    from flash import _lib
    from flash._desc import SolverParams, c_double_p, c_int32_p
    rng = np.random.default_rng(13)
    m = flash.models.irb140()
    sensor = flash.depthsensors.Kinect(60, 80)
    cam = flash.depthsensors.affine((0, 1.5, 0.5), (1, 0, 0), np.pi / 2)
    pts = flash.depthsensors.raycast(flash.ManipulatorState(m), sensor, cam)
    x0 = 0.15 * rng.standard_normal(6)
    solver = flash.tracking.NaiveSolver(6, rate=20.0, max_step=0.5, iteration_limit=25)
    xh = flash.tracking.estimate_state(m, pts, x0, solver=solver)
    xd = flash.tracking.estimate_state(m, pts, x0, solver=solver, on_device=True)
    xo, co, it = ol.Oracle(m).estimate_state(x0, pts, rate=20.0, max_step=0.5, iteration_limit=25, mode=ol.GJK)
    assert np.abs(xh - xd).max() < 1e-7
    assert np.abs(xd - xo).max() < 1e-4
    # early exit on the gradient tolerance: iterations reported, cost is the cost at the returned state
    X = x0.reshape(1, 6).copy(); cost = np.zeros(1); iters = np.zeros(1, np.int32)
    sp = SolverParams(20.0, 0.5, 200, 1e-3, 10.0)
    m.check(_lib.lib().flash_estimate_state(m.ctx(), X.ctypes.data_as(c_double_p), 1, C.byref(sp), None, cost.ctypes.data_as(c_double_p), iters.ctypes.data_as(c_int32_p)))
    f = flash.gradientdescent.CostFunctor(m, pts)
    c1, g1, _ = f.batch(X, gradient=True)
    assert 0 < iters[0] < 200 and np.linalg.norm(g1[0] / len(pts)) < 1e-3
    assert abs(c1[0] - cost[0]) <= 1e-9 * abs(c1[0])
    # the combined scene: 63 states, 8 hulls + a deformable skin
    import bench
    ms = bench.scene_model(flash)
    xs = bench.scene_states(flash, ms, 1)[0]
    st = flash.ManipulatorState(ms)
    flash.gradientdescent.unflatten_(st, xs)
    cloud = flash.depthsensors.raycast(st, flash.depthsensors.Kinect(60, 80), flash.depthsensors.affine((0.3, 1.8, 1.0), (1, 0, 0), np.pi / 2 + 0.25))
    xs0 = xs.copy(); xs0[:flash.num_positions(ms)] += 0.01 * rng.standard_normal(flash.num_positions(ms))
    solver = flash.tracking.NaiveSolver(len(xs), rate=0.01, max_step=0.1, iteration_limit=15)
    xh = flash.tracking.estimate_state(ms, cloud, xs0, solver=solver)
    xd = flash.tracking.estimate_state(ms, cloud, xs0, solver=solver, on_device=True)
    assert np.abs(xh - xd).max() < 1e-6 * max(1.0, np.abs(xh).max())


def test_on_device_estimate_state_edge_cases_of_the_persistent_loop():
    """track_kernel's corners: no iterations at all (cost at the start state, state untouched); a full-resolution frame
    on a skin-only model (65 groups per block: the dynamic hand-out instantiation) against the host loop; the same
    context re-used on a smaller frame afterwards; a cloud smaller than one group."""
    import ctypes as C
    from flash import _lib
    from flash._desc import SolverParams, c_double_p, c_int32_p
    rng = np.random.default_rng(14)
    m = flash.models.squishable()
    L = _lib.lib()
    x0 = _state(m, rng, qs=0.05, ds=0.005)
    big = _points(rng, 307200, [-0.35, -0.35, -0.3], [0.35, 0.35, 0.3])
    small = big[:1000]
    div = np.r_[np.ones(7), 1e-3 * np.ones(36)]
    # (1) iteration_limit = 0
    f = flash.gradientdescent.CostFunctor(m, small)
    c_ref = f(x0)
    X = x0.reshape(1, -1).copy(); cost = np.zeros(1); iters = np.full(1, -1, np.int32)
    sp = SolverParams(0.01, 0.1, 0, 0.0, 10.0)
    m.check(L.flash_estimate_state(m.ctx(), X.ctypes.data_as(c_double_p), 1, C.byref(sp), div.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), iters.ctypes.data_as(c_int32_p)))
    assert np.array_equal(X[0], x0) and iters[0] == 0 and abs(cost[0] - c_ref) <= 1e-9 * abs(c_ref)
    # (2) a full frame, then (3) the small one again on the same context
    # (4 iterations: on a cloud that is no surface the iteration is unstable -- host and device loops, identical to 1e-15 after
    # one step, are 1e-13 apart after 4 and 1e-3 after 12, with the two-kernel device loop just the same: tools/track_diff_probe.py)
    solver = flash.tracking.NaiveSolver(43, rate=0.01, max_step=0.1, iteration_limit=4, precondition_divisors=div)
    for cloud in (big, small, big[:20]):
        xh = flash.tracking.estimate_state(m, cloud, x0, solver=solver)
        xd = flash.tracking.estimate_state(m, cloud, x0, solver=solver, on_device=True)
        assert np.isfinite(xd).all() and np.abs(xh - xd).max() < 1e-10, (len(cloud), np.abs(xh - xd).max())
        assert np.abs(xd - x0).max() > 1e-3


# ------------------------------------------------------------------ grid mode (north_star's gather design)
def _trilinear_f64(vals, org, h, y):
    g = (y - org) / h
    i = np.floor(g).astype(int); f = g - i
    def at(dx, dy, dz):
        return vals[i[:, 2] + dz, i[:, 1] + dy, i[:, 0] + dx].astype(np.float64)
    c00 = at(0, 0, 0) * (1 - f[:, 0]) + at(1, 0, 0) * f[:, 0]; c10 = at(0, 1, 0) * (1 - f[:, 0]) + at(1, 1, 0) * f[:, 0]
    c01 = at(0, 0, 1) * (1 - f[:, 0]) + at(1, 0, 1) * f[:, 0]; c11 = at(0, 1, 1) * (1 - f[:, 0]) + at(1, 1, 1) * f[:, 0]
    c0 = c00 * (1 - f[:, 1]) + c10 * f[:, 1]; c1 = c01 * (1 - f[:, 1]) + c11 * f[:, 1]
    return c0 * (1 - f[:, 2]) + c1 * f[:, 2]


def test_grid_mode_matches_same_grid_in_fp64_and_reports_deviation():
    """Grid-mode parity: the kernel's manual fp32 trilinear lookup == the same trilinear formula in
    fp64 on the same (downloaded) grid; its deviation from the exact SDF is bounded by the bake report."""
    rng = np.random.default_rng(13)
    m = flash.models.table()
    m.set_options(sdf_mode=1, grid_spacing=0.005, grid_margin=0.1)
    vals, org, h, g2b = m.grid_download(0)
    dims, sp, err = m.grid_report(0)
    assert vals.shape == (dims[2], dims[1], dims[0]) and np.isclose(sp, 0.005) and 0 < err < 0.005
    x = np.array([1.0, 0, 0, 0, 0, 0, 0])                       # identity pose: world == body == geometry + centroid
    pts = _points(rng, 20000, [-0.33, -0.33, -0.13], [0.33, 0.33, 0.13])
    st = flash.ManipulatorState(m)
    d_grid = flash.skin(st).eval(pts)
    y = pts - g2b[:, 3]
    ref = _trilinear_f64(vals, org, h, y)
    assert np.abs(d_grid - ref).max() < 2e-6
    exact = flash.models.table()
    d_exact = flash.skin(flash.ManipulatorState(exact)).eval(pts)
    assert np.abs(d_grid - d_exact).max() <= err * 1.5 + 1e-6
    far = _points(rng, 1000, [0.6, 0.6, 0.6], [0.9, 0.9, 0.9])  # outside the grid: falls back to the exact path
    assert np.abs(flash.skin(st).eval(far) - flash.skin(flash.ManipulatorState(exact)).eval(far)).max() < 1e-6


def test_grid_mode_cost_close_to_exact_on_combined_scene():
    rng = np.random.default_rng(14)
    mg, me = flash.models.irb_and_squishable(), flash.models.irb_and_squishable()
    mg.set_options(sdf_mode=1, grid_spacing=0.004)
    x0 = flash.gradientdescent.flatten(flash.ManipulatorState(me))
    x0[4:7] = [0.0, 0.0, 0.75]; x0[17:20] = [0.55, 0.45, 0.8]; x0[24:27] = [0.4, 0.0, 0.6]
    X = np.stack([x0, x0 + np.r_[0.03 * rng.standard_normal(27), np.zeros(36)]])
    pts = _points(rng, 40000, [-0.5, -0.5, 0.5], [1.0, 0.5, 2.0])
    cg, gg, ag = flash.gradientdescent.CostFunctor(mg, pts).batch(X, argmin=True)
    ce, ge, ae = flash.gradientdescent.CostFunctor(me, pts).batch(X, argmin=True)
    assert np.abs(cg - ce).max() <= 2e-3 * np.abs(ce).max()
    assert np.abs(gg - ge).max() <= 2e-2 * np.abs(ge).max()
    assert (ag != ae).mean() < 5e-3


def test_contour_mesh_lies_on_the_zero_level_set():
    """Row F4: the headless replacement of draw()'s contour_mesh.  Every vertex of the extracted mesh must
    be a zero of the ORACLE's field up to the linear-interpolation error of one cell, the mesh must be
    closed (area check on the table box) and oriented outwards."""
    m = flash.models.irb_and_squishable()
    st = flash.ManipulatorState(m)
    x = flash.gradientdescent.flatten(st)
    o = ol.Oracle(m)
    # (a) one convex surface: the table box 0.5 x 0.5 x 0.1
    table = m.surfaces[8]
    lb, ub = flash.drawing_region(st, table)
    h = 0.01
    tris = flash.contour_mesh(st, lb, ub, 0.0, h, surface=table)
    assert len(tris) > 1000
    v = tris.reshape(-1, 3).astype(np.float64)
    per = o.surfaces(x, v, mode=ol.GJK)
    assert np.abs(per[:, 8]).max() < 0.5 * h
    e1, e2 = tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0]
    nrm = np.cross(e1, e2)
    area = 0.5 * np.linalg.norm(nrm, axis=1).sum()
    assert abs(area - 2 * (0.5 * 0.5 + 2 * 0.5 * 0.1)) < 0.03 * area
    cen = tris.mean(1) - v.mean(0)
    assert ((nrm * cen).sum(1) > 0).mean() > 0.98          # outward orientation
    # (b) the deformable skin alone, and (c) skin(state) = min over everything
    sk = m.surfaces[7]
    lb, ub = flash.drawing_region(st, sk)
    t2 = flash.contour_mesh(st, lb, ub, 0.0, 0.01, surface=sk)
    per = o.surfaces(x, t2.reshape(-1, 3).astype(np.float64), mode=ol.GJK)
    assert len(t2) > 500 and np.abs(per[:, 7]).max() < 5e-3
    t3 = flash.contour_mesh(st, [-0.6, -0.6, -0.1], [0.9, 0.6, 1.0], 0.0, 0.02)
    d = o.skin(x, t3.reshape(-1, 3).astype(np.float64), mode=ol.GJK)
    assert len(t3) > 2000 and np.abs(d).max() < 1e-2
    # count-only call and a too-small buffer report the same total
    import ctypes as C
    from flash import _lib
    from flash._desc import c_double_p
    n = C.c_int64(0)
    lo = np.array([-0.6, -0.6, -0.1]); hi = np.array([0.9, 0.6, 1.0])
    m.check(_lib.lib().flash_contour_mesh(m.ctx(), x.ctypes.data_as(c_double_p), -1, lo.ctypes.data_as(c_double_p), hi.ctypes.data_as(c_double_p), 0.02, 0.0,
                                          None, 0, C.byref(n)))
    assert n.value == len(t3)


def test_two_functors_on_one_manipulator_keep_their_own_points():
    """The reference's CostFunctor captures its own sensed_points (src/gradientdescent.jl:41-47); the ctx holds
    one frame, so a displaced functor must re-upload instead of evaluating on the other functor's cloud."""
    rng = np.random.default_rng(21)
    m = flash.models.irb140()
    x = rng.uniform(-1, 1, 6)
    pa = _points(rng, 5000, [-0.6, -0.7, -0.2], [0.9, 0.7, 1.1])
    pb = _points(rng, 12000, [-0.6, -0.7, -0.2], [0.9, 0.7, 1.1])
    fa = flash.gradientdescent.CostFunctor(m, pa)
    ca = fa(x)
    fb = flash.gradientdescent.CostFunctor(m, pb)      # replaces the ctx's cloud
    cb = fb(x)
    # (fp64 atomics: repeated calls agree to rounding unless the ctx is deterministic)
    assert np.isclose(fa(x), ca, rtol=1e-12) and np.isclose(fb(x), cb, rtol=1e-12) and abs(ca - cb) > 1e-3 * ca
    _, _, arg = fa.batch(x, argmin=True)                # sized from fa's own cloud
    assert arg.shape == (1, 5000)
    assert np.isclose(flash.gradientdescent.cost(flash.ManipulatorState(m), pb), fb(np.zeros(6)), rtol=1e-12)
    assert np.isclose(fa(x), ca, rtol=1e-12)


def test_frames_batch_equals_frame_by_frame_and_async_swap_pipelines():
    """flash_frames_set + flash_cost_grad_frames (F frames x C candidates per frame, one launch) == one
    flash_points_set + flash_cost_grad per frame; flash_points_set_async / flash_points_swap hand over the frames in
    order while the previous one is still being evaluated."""
    import ctypes as C
    from flash import _lib
    from flash._desc import c_double_p
    rng = np.random.default_rng(31)
    m = flash.models.irb_and_squishable()
    nx = flash.num_states(m)
    x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    x0[4:7] = [0.0, 0.0, 0.75]; x0[17:20] = [0.55, 0.45, 0.8]; x0[24:27] = [0.4, 0.0, 0.6]
    frames = [_points(rng, n, [-0.5, -0.5, 0.5], [1.0, 0.5, 2.0]) for n in (9000, 20011, 33, 15000)]
    X = x0 + 0.03 * rng.standard_normal((4, 3, nx))
    ff = flash.gradientdescent.FramesCostFunctor(m, frames)
    cost, grad = ff.batch(X)
    o = ol.Oracle(m)
    for f, pts in enumerate(frames):
        c1, g1, _ = flash.gradientdescent.CostFunctor(m, pts).batch(X[f])
        # the batch sorts all frames in one pass (coarser keys): other groups of 32, so other points take the certified-
        # feature path instead of GJK -- same distances to fp32 rounding (1e-7 m), not bit for bit
        assert np.allclose(cost[f], c1, rtol=1e-7)
        assert np.abs(grad[f] - g1).max() <= 1e-6 * np.abs(g1).max()
        co, go = o.cost_grad(X[f, 1], pts, method="adjoint", mode=ol.GJK)
        assert abs(cost[f, 1] - co) <= RTOL * abs(co) and np.abs(grad[f, 1] - go).max() <= RTOL * np.abs(go).max()
    # async pipeline: upload frame k+1 while frame k is evaluated
    L = _lib.lib(); ctx = m.ctx()
    m.check(L.flash_points_set_async(ctx, frames[0].ctypes.data_as(c_double_p), len(frames[0])))
    got = []
    for k in range(4):
        m.check(L.flash_points_swap(ctx))
        if k + 1 < 4:
            m.check(L.flash_points_set_async(ctx, frames[k + 1].ctypes.data_as(c_double_p), len(frames[k + 1])))
        c = np.empty(3); g = np.empty((3, nx))
        Xk = np.ascontiguousarray(X[k])
        m.check(L.flash_cost_grad(ctx, Xk.ctypes.data_as(c_double_p), 3, 10.0, c.ctypes.data_as(c_double_p), g.ctypes.data_as(c_double_p), None))
        assert L.flash_num_points(ctx) == len(frames[k])
        got.append(c)
    m._points_owner = None
    assert np.allclose(np.array(got), cost, rtol=1e-7)
    assert L.flash_points_swap(ctx) != 0          # nothing pending: FLASH_ERR_STATE, reported not thrown


def test_deterministic_mode_is_bit_reproducible():
    """flash_options.deterministic = 1: static group -> warp schedule and per-tile partial sums combined in tile order
    instead of fp64 atomics.  Two flash_cost_grad calls, and two whole on-device estimate_state loops with the
    notebook's (chaotic) squishable solver settings on the real Kinect cloud, are bit-identical; the default mode
    agrees to rounding."""
    rng = np.random.default_rng(41)
    m = flash.models.irb_and_squishable()
    nx = flash.num_states(m)
    x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    x0[4:7] = [0.0, 0.0, 0.75]; x0[17:20] = [0.55, 0.45, 0.8]; x0[24:27] = [0.4, 0.0, 0.6]
    X = x0 + 0.03 * rng.standard_normal((3, nx))
    pts = _points(rng, 400_000, [-0.5, -0.5, 0.5], [1.0, 0.5, 2.0])
    c_def, g_def, _ = flash.gradientdescent.CostFunctor(m, pts).batch(X)
    m.set_options(deterministic=1)
    f = flash.gradientdescent.CostFunctor(m, pts)
    runs = [f.batch(X) for _ in range(3)]
    for c, g, _ in runs[1:]:
        assert np.array_equal(c, runs[0][0]) and np.array_equal(g, runs[0][1])
    assert np.allclose(runs[0][0], c_def, rtol=1e-12) and np.abs(runs[0][1] - g_def).max() <= 1e-10 * np.abs(g_def).max()
    m.set_options(deterministic=0)
    # the tracking loop: 50 iterations of the squishable notebook's solver on the real cloud (chaotic --
    # the default mode's fp64 atomics make the final cost differ from run to run)
    real = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "squishable_unsquished_xyz.npy")
    cloud = np.load(real).astype(np.float64)
    ms = flash.models.squishable()
    ms.set_options(deterministic=1)
    solver = flash.tracking.NaiveSolver(flash.num_states(ms), rate=0.01, max_step=0.1, iteration_limit=50,
                                        precondition_divisors=np.r_[np.ones(7), 1e-3 * np.ones(36)])
    xs = flash.gradientdescent.flatten(flash.ManipulatorState(ms))
    xs[4:7] = cloud.mean(0)
    ends = [flash.tracking.estimate_state(ms, cloud, xs, solver=solver, on_device=True) for _ in range(3)]
    assert np.array_equal(ends[0], ends[1]) and np.array_equal(ends[0], ends[2])


def test_model_from_urdf_and_points_from_file_through_the_c_abi(tmp_path):
    """flash_model_load_urdf (C++ URDF / STL ingestion, row F2) and flash_points_set_file (row F3: the text reader
    feeding the pinned staging ring) give the same cost and gradient as the Python-built model with points handed
    over in memory -- and the oracle's."""
    import ctypes as C
    from flash import _lib
    from flash._desc import Options, c_double_p
    L = _lib.lib()
    rng = np.random.default_rng(51)
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    urdf = os.path.join(gold, "IRB140", "urdf", "irb_140_convhull.urdf")
    opt = Options(); L.flash_options_default(C.byref(opt))
    ctx = C.c_void_p()
    arr = (C.c_char_p * 1)(gold.encode())
    assert L.flash_model_load_urdf(urdf.encode(), arr, 1, C.byref(opt), C.byref(ctx)) == 0, L.flash_last_error(None)
    assert L.flash_num_states(ctx) == 6 and L.flash_num_surfaces(ctx) == 7
    pts = _points(rng, 30011, [-0.6, -0.6, -0.2], [0.8, 0.6, 1.0])
    path = tmp_path / "cloud.txt"
    with open(path, "w") as f:
        f.write("0.0, 1.5, 0.5\n")
        for p in pts:
            f.write("%r, %r, %r, 0.5, 0.5, 0.5\n" % tuple(float(v) for v in p))
    origin = np.zeros(3)
    assert L.flash_points_set_file(ctx, str(path).encode(), origin.ctypes.data_as(c_double_p)) == 0, L.flash_last_error(ctx)
    assert np.allclose(origin, [0.0, 1.5, 0.5]) and L.flash_num_points(ctx) == len(pts)
    X = rng.uniform(-1.0, 1.0, (2, 6))
    cost = np.empty(2); grad = np.empty((2, 6))
    assert L.flash_cost_grad(ctx, X.ctypes.data_as(c_double_p), 2, 10.0, cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p), None) == 0
    m = flash.models.irb140()
    c_ref, g_ref, _ = flash.gradientdescent.CostFunctor(m, pts).batch(X)
    assert np.allclose(cost, c_ref, rtol=1e-12) and np.abs(grad - g_ref).max() <= 1e-9 * np.abs(g_ref).max()
    o = ol.Oracle(m)
    co, go = o.cost_grad(X[1], pts, method="dual", mode=ol.GJK)
    assert abs(cost[1] - co) <= RTOL * abs(co) and np.abs(grad[1] - go).max() <= RTOL * np.abs(go).max()
    assert L.flash_points_set_file(ctx, b"/nonexistent/cloud.txt", None) == -7 and b"cannot open" in L.flash_last_error(ctx)
    L.flash_destroy(ctx)


def _multi_devices():
    """two sub-contexts: on two GPUs when the box has them, else twice on device 0 (the sharding, the peer reduction and the
    gathers run all the same)"""
    import torch
    return (0, 1) if torch.cuda.device_count() >= 2 else (0, 0)


def test_single_process_multi_device_context_matches_one_device():
    """flash_options.n_devices: ONE context that shards points (contiguous ranges), frames (round-robin) and rays
    (contiguous ranges) over its devices and reduces the accumulators over peer memory -- the form the reference's
    single Julia process needs (src/tracking.jl:8-27).  Cost, gradient, argmin, frame batches, the on-device solver
    loop and ray casting against the same calls on one device, and the cost against the oracle."""
    rng = np.random.default_rng(61)
    devs = _multi_devices()
    m1 = flash.models.irb_and_squishable()
    m2 = flash.models.irb_and_squishable()
    m2.set_options(devices=devs)
    from flash import _lib
    assert _lib.lib().flash_num_devices(m2.ctx()) == 2 and _lib.lib().flash_num_devices(m1.ctx()) == 1
    nx = flash.num_states(m1)
    x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m1))
    x0[4:7] = [0.0, 0.0, 0.75]; x0[17:20] = [0.55, 0.45, 0.8]; x0[24:27] = [0.4, 0.0, 0.6]
    X = x0 + 0.03 * rng.standard_normal((3, nx))
    pts = _points(rng, 200_003, [-0.5, -0.5, 0.5], [1.0, 0.5, 2.0])
    c1, g1, a1 = flash.gradientdescent.CostFunctor(m1, pts).batch(X, argmin=True)
    c2, g2, a2 = flash.gradientdescent.CostFunctor(m2, pts).batch(X, argmin=True)
    # (the shards are sorted and grouped on their own: other groups of 32, so the fp32 warp sums of the hull wrenches round differently)
    assert np.allclose(c2, c1, rtol=1e-9) and np.abs(g2 - g1).max() <= 1e-7 * np.abs(g1).max()
    assert np.array_equal(a1, a2)                      # per-point results do not depend on the shard a point lands in
    o = ol.Oracle(m1)
    co, go = o.cost_grad(X[2], pts, method="adjoint", mode=ol.GJK)
    assert abs(c2[2] - co) <= RTOL * abs(co) and np.abs(g2[2] - go).max() <= RTOL * np.abs(go).max()
    # frames: 5 frames over 2 devices (3 + 2), 2 candidates each
    frames = [_points(rng, n, [-0.5, -0.5, 0.5], [1.0, 0.5, 2.0]) for n in (9000, 20011, 33, 15000, 4096)]
    XF = x0 + 0.03 * rng.standard_normal((5, 2, nx))
    cf1, gf1 = flash.gradientdescent.FramesCostFunctor(m1, frames).batch(XF)
    cf2, gf2 = flash.gradientdescent.FramesCostFunctor(m2, frames).batch(XF)
    assert np.allclose(cf2, cf1, rtol=1e-9) and np.abs(gf2 - gf1).max() <= 1e-7 * np.abs(gf1).max()
    # the solver loop: every device takes the same steps on the same sums
    ms1, ms2 = flash.models.squishable(), flash.models.squishable()
    ms2.set_options(devices=devs)
    cloud = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "squishable_unsquished_xyz.npy")).astype(np.float64)
    xs = flash.gradientdescent.flatten(flash.ManipulatorState(ms1)); xs[4:7] = cloud.mean(0)
    solver = flash.tracking.NaiveSolver(flash.num_states(ms1), rate=0.1, max_step=0.05, iteration_limit=8)
    e1 = flash.tracking.estimate_state(ms1, cloud, xs, solver=solver, on_device=True)
    e2 = flash.tracking.estimate_state(ms2, cloud, xs, solver=solver, on_device=True)
    assert np.abs(e2 - e1).max() < 1e-9
    # ray casting: rays in contiguous ranges, depths in place, hits concatenated in ray order
    sensor = flash.depthsensors.Kinect(60, 80)
    cam = flash.depthsensors.affine((0, 1.5, 0.5), (1, 0, 0), np.pi / 2)
    mi1, mi2 = flash.models.irb140(), flash.models.irb140()
    mi2.set_options(devices=devs)
    d1 = flash.depthsensors.raycast_depths(flash.skin(flash.ManipulatorState(mi1)), sensor, cam)
    d2 = flash.depthsensors.raycast_depths(flash.skin(flash.ManipulatorState(mi2)), sensor, cam)
    assert np.array_equal(np.isnan(d1), np.isnan(d2)) and np.array_equal(d1[~np.isnan(d1)], d2[~np.isnan(d2)])
    h1 = flash.depthsensors.raycast(flash.ManipulatorState(mi1), sensor, cam)
    h2 = flash.depthsensors.raycast(flash.ManipulatorState(mi2), sensor, cam)
    assert h1.shape == h2.shape and len(h1) > 100 and np.array_equal(h1, h2)
