"""The independent numpy / scipy restatement (tests/indep.py) against the C++ oracle on all golden scenes, against the
product's model builders, and (GPU) against the CUDA path -- two implementations by different means that agree, with
the models read straight from the reference's sources (src/models.jl, the URDF text and STL bytes of IRB140)."""
import os

import numpy as np
import pytest

import flash
import indep
import oracle_lib as ol

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_v1.npz"))
SCENES = ["two_link_arm", "beanbag", "squishable", "irb140", "irb_and_squishable"]


def _state_and_points(name, n=600, seed=3):
    """the golden scene's state, and a cloud hugging the model (golden points) plus a box of random points around it"""
    rng = np.random.default_rng(seed)
    x = np.array(GOLD[name + "/x"], float)
    g = np.array(GOLD[name + "/pts"], float)
    lo, hi = g.min(0) - 0.3, g.max(0) + 0.3
    pts = np.concatenate([g[rng.choice(len(g), min(len(g), n // 2), replace=False)], rng.uniform(lo, hi, (n // 2, 3))])
    return x, pts.astype(np.float32).astype(np.float64)


@pytest.mark.parametrize("name", SCENES)
def test_independent_models_equal_the_products_builders(name):
    """same bodies (parents, joint types, axes, joint frames), same surfaces (kinds, frames, points in the same order):
    the product's builders and URDF/STL loader against the restatement from the reference sources"""
    mi = indep.MODELS[name]()
    mf = ol.model_dict({"two_link_arm": flash.models.two_link_arm, "beanbag": flash.models.beanbag, "squishable": flash.models.squishable,
                        "irb140": flash.models.irb140, "irb_and_squishable": flash.models.irb_and_squishable}[name]())
    assert len(mi["bodies"]) == len(mf["bodies"]) and len(mi["surfaces"]) == len(mf["surfaces"])
    for a, b in zip(mi["bodies"], mf["bodies"]):
        assert (a["parent"], a["jtype"]) == (b["parent"], b["jtype"])
        assert np.allclose(a["j2p"], b["j2p"], rtol=0, atol=1e-15)
        if a["jtype"] == indep.REVOLUTE:
            assert np.allclose(a["axis"], b["axis"])
    for a, b in zip(mi["surfaces"], mf["surfaces"]):
        assert a["kind"] == b["kind"]
        if a["kind"] == "convex":
            assert a["body"] == b["body"] and np.allclose(a["frame"], b["frame"], rtol=0, atol=1e-15) and np.array_equal(a["verts"], b["verts"])
        else:
            assert [q[0] for q in a["surface"]] == [q[0] for q in b["surface"]] and [q[0] for q in a["skeleton"]] == [q[0] for q in b["skeleton"]]
            assert np.allclose([q[1] for q in a["surface"]], [q[1] for q in b["surface"]], rtol=0, atol=1e-15)
            assert np.allclose([q[1] for q in a["skeleton"]], [q[1] for q in b["skeleton"]], rtol=0, atol=1e-15)
    assert indep.nx_of(mi) == flash.num_states({"two_link_arm": flash.models.two_link_arm, "beanbag": flash.models.beanbag, "squishable": flash.models.squishable,
                                                "irb140": flash.models.irb140, "irb_and_squishable": flash.models.irb_and_squishable}[name]())


@pytest.mark.parametrize("name", SCENES)
def test_oracle_agrees_with_the_independent_restatement(name):
    """every surface's value at every point, the min / argmin and the cost: liboracle.so (C++: its own FK, hull builder,
    GJK and brute-force distances, RBF LU) against numpy FK + scipy ConvexHull + numpy.linalg.solve"""
    m = indep.MODELS[name]()
    x, pts = _state_and_points(name)
    o = ol.Oracle(m)                       # the oracle's model comes from the independent description too
    assert o.nx == indep.nx_of(m) and o.nq == indep.nq_of(m)
    vi = indep.surfaces(m, x, pts)
    for mode in (ol.BRUTE, ol.GJK):
        vo = o.surfaces(x, pts, mode=mode)
        assert vo.shape == vi.shape
        assert np.abs(vo - vi).max() <= 1e-9 * (1 + np.abs(vi).max()), (name, mode, np.abs(vo - vi).max())
    ci, ai = indep.cost(m, x, pts)
    co, ao = o.cost(x, pts, mode=ol.GJK, argmin=True)
    assert abs(co - ci) <= 1e-9 * abs(ci)
    flips = np.nonzero(ao != ai)[0]
    srt = np.sort(vi[flips], axis=1)
    assert len(flips) == 0 or (srt[:, 1] - srt[:, 0] < 1e-9).all()            # argmin differs only on exact ties


def test_oracle_gradient_against_central_differences_of_the_independent_cost():
    """d cost / d x of the oracle (forward-mode duals, the reference's method) against central differences of the numpy /
    scipy cost on the combined scene: kinematics, quaternion normalisation inside the cost, RBF weights' dependence on the
    deformations -- all of it restated twice"""
    name = "irb_and_squishable"
    m = indep.MODELS[name]()
    x, pts = _state_and_points(name, n=240, seed=5)
    o = ol.Oracle(m)
    _, g = o.cost_grad(x, pts, method="dual", mode=ol.GJK)
    rng = np.random.default_rng(1)
    for k in list(rng.choice(len(x), 10, replace=False)) + [0, 4, 13, 27]:
        e = np.zeros(len(x)); e[k] = 1e-6
        fd = (indep.cost(m, x + e, pts)[0] - indep.cost(m, x - e, pts)[0]) / 2e-6
        assert abs(fd - g[k]) <= 2e-5 * max(1.0, np.abs(g).max()), (k, fd, g[k])


def test_irb140_aabbs_of_survey_table_from_the_stl_bytes():
    """SURVEY 9's per-mesh AABB table and 52-vertex counts, from the STL bytes through the independent reader"""
    m = indep.load_urdf()
    aabb = [((-0.3132, -0.1525, -0.2019), (0.1784, 0.0005, 0.2018)), ((-0.1556, -0.4543, -0.1365), (0.1943, -0.0889, 0.2044)),
            ((-0.0293, -0.8027, -0.1755), (0.1540, -0.2831, 0.0989)), ((-0.0790, -0.8091, -0.0832), (0.3088, -0.6090, 0.0850)),
            ((0.2885, -0.7681, -0.0645), (0.5078, -0.6559, 0.0705)), ((0.4079, -0.7543, -0.0155), (0.4922, -0.6697, 0.0166)),
            ((0.4696, -0.7370, -0.0250), (0.5150, -0.6867, 0.0250))]
    assert len(m["bodies"]) == 7 and len(m["surfaces"]) == 7
    for s_, (lo, hi) in zip(m["surfaces"], aabb):
        assert len(s_["verts"]) == 52 and np.allclose(s_["verts"].min(0), lo, atol=6e-5) and np.allclose(s_["verts"].max(0), hi, atol=6e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENES)
def test_cuda_path_agrees_with_the_independent_restatement(name):
    """the CUDA path (product builders -> libflashsdf) against the numpy / scipy restatement (independent models): per-point
    distances through flash_skin_eval, cost through flash_cost_grad, argmin up to ties"""
    mi = indep.MODELS[name]()
    mf = {"two_link_arm": flash.models.two_link_arm, "beanbag": flash.models.beanbag, "squishable": flash.models.squishable,
          "irb140": flash.models.irb140, "irb_and_squishable": flash.models.irb_and_squishable}[name]()
    x, pts = _state_and_points(name, n=2000, seed=9)
    vi = indep.surfaces(mi, x, pts)
    ci = float((vi.min(1) ** 2).sum() + 10.0 * (x[indep.nq_of(mi):] ** 2).sum())
    cost, _, arg = flash.gradientdescent.CostFunctor(mf, pts).batch(x, gradient=False, argmin=True)
    assert abs(cost[0] - ci) <= 1e-5 * abs(ci)
    flips = np.nonzero(arg[0] != vi.argmin(1))[0]
    srt = np.sort(vi[flips], axis=1)
    assert len(flips) == 0 or (len(flips) <= 4 and (srt[:, 1] - srt[:, 0] < 2e-6).all())
