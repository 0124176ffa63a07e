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
