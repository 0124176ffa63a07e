"""CPU tests of the PRODUCT's __host__ __device__ math (csrc/flash_dev.h, csrc/flash_host.h) compiled
with g++ and replayed serially by tests/hostcheck -- the same formulas the CUDA kernels execute --
against the oracle.  (The CUDA path itself is tested in test_gpu_parity.py.)"""
import os

import numpy as np
import pytest

import flash
import hostcheck_lib as hl
import oracle_lib as ol

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_v1.npz"))
BUILDERS = {"two_link_arm": flash.models.two_link_arm, "beanbag": flash.models.beanbag, "squishable": flash.models.squishable,
            "irb140": flash.models.irb140, "irb_and_squishable": flash.models.irb_and_squishable}


@pytest.mark.parametrize("name", sorted(BUILDERS))
def test_kernel_math_matches_golden(name):
    h = hl.HostCheck(BUILDERS[name]())
    x, pts = GOLD[name + "/x"], GOLD[name + "/pts"]
    c, g, dist, arg = h.cost_grad(x, pts)
    assert abs(c - float(GOLD[name + "/cost"])) <= 1e-5 * abs(float(GOLD[name + "/cost"]))
    gg = GOLD[name + "/grad"]
    assert np.abs(g - gg).max() <= 1e-5 * np.abs(gg).max()
    per = GOLD[name + "/per_surface"]
    assert np.abs(dist - per.min(1)).max() < 2e-6
    bad = np.nonzero(arg != GOLD[name + "/argmin"])[0]
    for i in bad:  # only exact-tie flips within fp32 rounding are tolerated
        s = np.sort(per[i]); assert s[1] - s[0] < 2e-6


def test_kernel_math_large_random_scene():
    rng = np.random.default_rng(3)
    m = flash.models.irb_and_squishable()
    h, o = hl.HostCheck(m), ol.Oracle(m)
    x = GOLD["irb_and_squishable/x"]
    pts = rng.uniform([-0.8, -0.8, -0.3], [1.0, 0.8, 1.2], (30000, 3)).astype(np.float32).astype(np.float64)
    c, g, dist, arg = h.cost_grad(x, pts)
    co, go = o.cost_grad(x, pts, method="adjoint", mode=ol.GJK)
    do, ao = o.skin(x, pts, mode=ol.GJK, argmin=True)
    assert abs(c - co) <= 1e-6 * co and np.abs(g - go).max() <= 1e-6 * np.abs(go).max()
    assert np.abs(dist - do).max() < 2e-6 and (arg != ao).sum() <= 3


def test_hull_builder_agrees_with_oracle_hull():
    m = flash.models.irb_and_squishable()
    h, o = hl.HostCheck(m), ol.Oracle(m)
    conv = [s for s in range(9) if s != 7]
    for j, s in enumerate(conv):
        nv, npl, lo, hi, depth = h.convex_info(j)
        onv, onp, _ = o.hull_info(s)
        assert (nv, npl) == (onv, onp)
        assert (hi > lo).all() and depth > 0


def test_hull_sd_exact_on_box_and_lower_bound_property():
    m = flash.models.table()
    h = hl.HostCheck(m)
    rng = np.random.default_rng(5)
    y = rng.uniform(-0.6, 0.6, (5000, 3)).astype(np.float32)
    d, n = h.hull_sd(0, y)
    hw = np.array([0.25, 0.25, 0.05])
    q = np.abs(y.astype(np.float64)) - hw
    ref = np.linalg.norm(np.maximum(q, 0), axis=1) + np.minimum(q.max(1), 0)
    assert np.abs(d - ref).max() < 2e-7
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-5)
    # finite-difference check of the returned gradient away from the medial axis / edges
    eps = 1e-3
    for k in range(3):
        e = np.zeros(3, np.float32); e[k] = eps
        dp, _ = h.hull_sd(0, y + e); dm, _ = h.hull_sd(0, y - e)
        fd = (dp.astype(np.float64) - dm) / (2 * eps)
        ok = np.abs(fd - n[:, k]) < 5e-2
        assert ok.mean() > 0.97


def test_linear_in_points_and_shardable():
    """acc(A u B) = acc(A) + acc(B): the property the multi-GPU sharding relies on."""
    m = flash.models.irb_and_squishable()
    h = hl.HostCheck(m)
    x, pts = GOLD["irb_and_squishable/x"], GOLD["irb_and_squishable/pts"]
    tab, aux = h.prepare(x)
    a, _, _ = h.accumulate(tab, pts)
    a1, _, _ = h.accumulate(tab, pts[:100]); a2, _, _ = h.accumulate(tab, pts[100:])
    assert np.allclose(a, a1 + a2, rtol=1e-12, atol=1e-12)


def test_model_validation_errors():
    import ctypes as C
    from flash._desc import PackedModel
    m = flash.models.table()
    m.surfaces[0].vertices = m.surfaces[0].vertices[:3]       # not a volume
    with pytest.raises(RuntimeError, match="at least 4"):
        hl.HostCheck(m)
    m2 = flash.models.two_link_arm()
    m2.mechanism.bodies[0].parent = 1                         # not parents-first
    with pytest.raises(RuntimeError, match="parents-first"):
        hl.HostCheck(m2)
    m3 = flash.models.beanbag()
    p = PackedModel(m3); p.desc.abi_version = 99
    assert not hl.lib().hc_model_create(C.byref(p.desc), None)
    assert b"abi_version" in hl.lib().hc_last_error()


def _morton_sorted(pts, bits=5):
    """Order points as libflashsdf does at upload (Morton code of a 2^bits^3 cell grid over the AABB)."""
    lo, hi = pts.min(0), pts.max(0)
    cell = np.clip(((pts - lo) / np.maximum(hi - lo, 1e-30) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    key = np.zeros(len(pts), np.int64)
    for b in range(bits):
        for a in range(3):
            key |= ((cell[:, a] >> b) & 1) << (3 * b + a)
    return pts[np.argsort(key, kind="stable")]


def test_warp_group_culling_changes_nothing_and_culls():
    """The per-32-point bounding-ball culling of cost_kernel (group_candidates) must leave every
    distance, argmin and accumulator bit-identical, and must actually remove hulls on a dense cloud."""
    rng = np.random.default_rng(11)
    m = flash.models.irb_and_squishable()
    h = hl.HostCheck(m)
    x = GOLD["irb_and_squishable/x"]
    tab, _ = h.prepare(x)
    # a dense slab (K5 density: ~6,700 points per litre) so that 32 Morton neighbours span ~2 cm
    pts = rng.uniform([-0.4, -0.3, 0.0], [0.6, 0.3, 0.3], (1_200_000, 3)).astype(np.float32)
    pts = _morton_sorted(pts, bits=6)[: 32 * 3000]
    try:
        hl.set_exact(False)                        # culling alone (no start nodes): bit-identical
        hl.set_group_cull(False)
        a0, d0, g0 = h.accumulate(tab, pts)
        hl.set_group_cull(True); hl.group_stats()
        a1, d1, g1 = h.accumulate(tab, pts)
        st = hl.group_stats()
    finally:
        hl.set_exact(True); hl.set_group_cull(True)
    assert np.array_equal(d0, d1) and np.array_equal(g0, g1) and np.array_equal(a0, a1)
    assert st["groups"] == 3000 and st["candidates"] / st["groups"] < 5.0   # 8 hulls in the model


def test_group_exact_prepass_and_bsp_start_nodes_change_nothing():
    """group_prepass with exact centre distances: tighter candidate sets than the AABB bounds, and every point
    finishes its BSP descent from the group's start node (bsp_descend_ball) -- same distances as a descent from the
    root to fp32 rounding (a point within rounding distance of a cell boundary may take the neighbouring cell), same
    argmin as the fp64 oracle, far fewer levels per point."""
    rng = np.random.default_rng(12)
    m = flash.models.irb_and_squishable()
    h = hl.HostCheck(m)
    x = GOLD["irb_and_squishable/x"]
    tab, _ = h.prepare(x)
    # tight clusters (1 cm) at all distances from the scene, plus clusters ON the hull surfaces (depth-frame regime)
    ctr = rng.uniform([-0.8, -0.8, -0.3], [1.0, 0.8, 1.2], (3000, 1, 3))
    pts = (ctr + rng.uniform(-0.01, 0.01, (3000, 32, 3))).reshape(-1, 3).astype(np.float32)
    o = ol.Oracle(m)
    d_or, a_or = o.skin(x, pts.astype(np.float64), mode=ol.GJK, argmin=True)
    near = np.abs(d_or) < 0.01
    pts = np.concatenate([pts, pts[near][: 32 * (near.sum() // 32)]])
    d_or, a_or = o.skin(x, pts.astype(np.float64), mode=ol.GJK, argmin=True)
    try:
        hl.set_exact(False); hl.counters(); hl.group_stats()
        a0, d0, g0 = h.accumulate(tab, pts)
        c0 = hl.counters(); s0 = hl.group_stats()
        hl.set_exact(True)
        a1, d1, g1 = h.accumulate(tab, pts)
        c1 = hl.counters(); s1 = hl.group_stats()
    finally:
        hl.set_exact(True)
    assert np.abs(d1 - d0).max() < 3e-7 and np.abs(d1 - d_or).max() < 3e-7
    flips = np.nonzero(g1 != a_or)[0]
    per = np.sort(o.surfaces(x, pts[flips].astype(np.float64), mode=ol.GJK), axis=1) if len(flips) else np.zeros((0, 2))
    assert len(flips) <= 4 and (per[:, 1] - per[:, 0] < 2e-6).all()
    assert np.abs(a1 - a0).max() <= 1e-6 * np.abs(a0).max()
    assert s1["candidates"] < 0.8 * s0["candidates"]                  # exact bounds cull more hulls than the AABB bounds
    assert c1["narrow"] <= c0["narrow"]
    assert c1["bsp_levels"] / c1["narrow"] < 0.5 * c0["bsp_levels"] / c0["narrow"]   # ... and the descents start deep in the tree
    print("candidates/group %.2f -> %.2f, narrow/point %.2f -> %.2f, BSP levels/narrow %.1f -> %.1f" % (
        s0["candidates"] / s0["groups"], s1["candidates"] / s1["groups"], c0["narrow"] / c0["points"], c1["narrow"] / c1["points"],
        c0["bsp_levels"] / c0["narrow"], c1["bsp_levels"] / c1["narrow"]))


@pytest.mark.parametrize("name", ["irb_and_squishable", "squishable", "two_link_arm", "beanbag"])
def test_fp32_skin_interval_contains_fp64_value(name):
    """skin_bound_f32 is a CERTIFIED interval: the pre-filter may only skip the fp64 evaluation when the
    skin provably cannot be the minimum."""
    rng = np.random.default_rng(2)
    m = BUILDERS[name]()
    h = hl.HostCheck(m)
    x = GOLD[name + "/x"]
    tab, _ = h.prepare(x)
    centre = GOLD[name + "/pts"].mean(0)
    for scale in (0.05, 0.3, 1.0, 3.0):
        pts = (centre + scale * rng.standard_normal((50000, 3))).astype(np.float32)
        lo, hi, s = h.skin_bound(0, tab, pts)
        assert (lo <= s).all() and (s <= hi).all()
    # and it is tight enough to be useful where it matters (points within ~1 m of the skin)
    pts = (centre + 0.3 * rng.standard_normal((50000, 3))).astype(np.float32)
    lo, hi, s = h.skin_bound(0, tab, pts)
    assert np.median((hi - lo) / np.maximum(np.abs(s), 1e-9)) < 0.25


@pytest.mark.parametrize("name", ["irb_and_squishable", "squishable", "two_link_arm", "beanbag"])
def test_skin_ball_lower_bound_holds_at_every_point_of_the_ball(name):
    """skin_ball_lower_f32 (group-level skin cull) is a CERTIFIED lower bound of the skin value over the whole
    ball: checked against the fp64 value at 400 points of each ball (surface and interior), for balls of the
    sizes the Morton groups have (millimetres to decimetres) near and far from the skin."""
    rng = np.random.default_rng(11)
    m = BUILDERS[name]()
    h = hl.HostCheck(m)
    tab, _ = h.prepare(GOLD[name + "/x"])
    centre = GOLD[name + "/pts"].mean(0)
    certified = 0
    for scale in (0.1, 0.4, 1.5):
        for rho in (0.002, 0.01, 0.05, 0.2):
            for c in (centre + scale * rng.standard_normal((40, 3))).astype(np.float32):
                lb = h.skin_ball_lower(0, tab, c, rho)
                if lb < -1e37:
                    continue
                certified += 1
                d = rng.standard_normal((400, 3)); d /= np.linalg.norm(d, axis=1)[:, None]
                d *= rho * np.concatenate([np.ones(200), rng.uniform(0, 1, 200) ** (1 / 3)])[:, None]
                pts = (c.astype(np.float64) + d).astype(np.float32)
                pts = pts[np.linalg.norm(pts.astype(np.float64) - c, axis=1) <= rho]
                _, _, s = h.skin_bound(0, tab, pts)
                assert lb <= s.min(), (name, scale, rho, lb, s.min())
    assert certified > 100   # and it does certify something


def test_group_skin_cull_never_changes_the_result():
    """Replay with and without the group-level skin cull: distances, argmins and accumulators are identical,
    and the cull fires on most groups of a scene where the hulls dominate."""
    m = flash.models.irb_and_squishable()
    h = hl.HostCheck(m)
    tab, _ = h.prepare(GOLD["irb_and_squishable/x"])
    rng = np.random.default_rng(6)
    # groups of 32 points as the Morton order makes them: 2000 clusters of 1.5 cm, plus the golden frame's points
    ctr = rng.uniform([-0.8, -0.8, -0.3], [1.0, 0.8, 1.2], (2000, 1, 3))
    pts = np.concatenate([(ctr + rng.uniform(-0.015, 0.015, (2000, 32, 3))).reshape(-1, 3), GOLD["irb_and_squishable/pts"]]).astype(np.float32)
    try:
        hl.set_skin_cull(True); hl.skin_groups_skipped()
        a1, d1, g1 = h.accumulate(tab, pts)
        skipped = hl.skin_groups_skipped()
        hl.set_skin_cull(False)
        a0, d0, g0 = h.accumulate(tab, pts)
    finally:
        hl.set_skin_cull(True)
    assert np.array_equal(d0, d1) and np.array_equal(g0, g1) and np.array_equal(a0, a1)
    assert skipped > 0.2 * len(pts) / 32


def test_fp32_prefilter_defers_most_fp64_skin_evaluations():
    rng = np.random.default_rng(4)
    m = flash.models.irb_and_squishable()
    h = hl.HostCheck(m)
    x = GOLD["irb_and_squishable/x"]
    tab, _ = h.prepare(x)
    pts = rng.uniform([-0.8, -0.8, -0.3], [1.0, 0.8, 1.2], (20000, 3)).astype(np.float32)
    hl.counters()
    _, _, arg = h.accumulate(tab, pts)
    c = hl.counters()
    skin_surface = [i for i, s_ in enumerate(m.surfaces) if not isinstance(s_, flash.ConvexGeometry)][0]
    winners = int((arg == skin_surface).sum())
    assert winners <= c["skin64"] <= winners + 0.1 * len(pts)


def test_bsp_distance_equals_brute_force_everywhere():
    """The closest-feature BSP descent (flash_dev.h hull_sd, trees of flash_bsp.h) must equal the oracle's
    brute-force point-polytope distance everywhere: hugging the surfaces (points walked onto the zero level set and
    pushed out / in by 1 mm and 1 cm), mid range, and far outside the scene."""
    rng = np.random.default_rng(12)
    m = flash.models.irb_and_squishable()
    h, o = hl.HostCheck(m), ol.Oracle(m)
    x = GOLD["irb_and_squishable/x"]
    base = rng.uniform([-0.8, -0.8, -0.3], [1.0, 0.8, 1.2], (6000, 3))
    d0, _ = o.skin(x, base, mode=ol.BRUTE, argmin=True)
    # walk the points onto the zero level set along a finite-difference gradient, then offset them
    eps = 1e-4
    g = np.stack([(o.skin(x, base + eps * e, mode=ol.BRUTE) - o.skin(x, base - eps * e, mode=ol.BRUTE)) / (2 * eps) for e in np.eye(3)], 1)
    g /= np.maximum(np.linalg.norm(g, axis=1, keepdims=True), 1e-9)
    near = np.concatenate([base - (d0 - off)[:, None] * g for off in (1e-3, -1e-3, 1e-2, -1e-2)])
    far = rng.uniform(-1, 1, (6000, 3)) * np.array([8.0, 8.0, 8.0])
    pts = np.concatenate([near, base, far]).astype(np.float32).astype(np.float64)
    tab, _ = h.prepare(x)
    hl.counters()
    _, dist, arg = h.accumulate(tab, pts)
    cn = hl.counters()
    do, ao = o.skin(x, pts, mode=ol.BRUTE, argmin=True)
    assert np.abs(dist - do).max() < 2e-6 * (1 + np.abs(do).max())
    per = o.surfaces(x, pts, mode=ol.BRUTE)
    for i in np.nonzero(arg != ao)[0]:
        s = np.sort(per[i]); assert s[1] - s[0] < 4e-6
    # one descent costs ~12 plane tests on the 100-face IRB140 hulls
    assert cn["bsp_levels"] < 16 * cn["narrow"]


def _one_hull_model(verts):
    """a convex body fixed to the world (body -1): its geometry frame is the world frame shifted to the centroid"""
    from flash.core import Mechanism
    return flash.Manipulator(Mechanism([]), [flash.ConvexGeometry(np.asarray(verts, float), -1)])


def test_bsp_trees_on_awkward_hulls():
    """flash_bsp.h on hulls that stress the builder: coplanar many-vertex faces (prism caps), collinear and interior
    input points, duplicates, a tetrahedron, a thin slab, a sliver-faced random hull.  Signed distance and gradient
    from the tree against the oracle's brute-force distance; the trees stay small and shallow."""
    rng = np.random.default_rng(21)
    th = np.linspace(0, 2 * np.pi, 24, endpoint=False)
    ring = np.stack([0.2 * np.cos(th), 0.2 * np.sin(th)], 1)
    sph = rng.standard_normal((60, 3)); sph /= np.linalg.norm(sph, axis=1, keepdims=True)
    box = np.array([[x, y, z] for x in (-0.3, 0.3) for y in (-0.2, 0.2) for z in (-0.1, 0.1)])
    hulls = {
        "prism24": np.concatenate([np.c_[ring, np.full(24, -0.15)], np.c_[ring, np.full(24, 0.15)]]),
        "box+interior+collinear": np.concatenate([box, 0.5 * box, [[0.0, 0.2, 0.1], [0.3, 0.0, 0.1], [0.0, 0.0, 0.1]], box[:3]]),
        "tetrahedron": np.array([[0, 0, 0], [0.4, 0, 0], [0, 0.3, 0], [0, 0, 0.2]]),
        "slab": np.array([[x, y, z] for x in (-0.5, 0.5) for y in (-0.4, 0.4) for z in (-0.002, 0.002)]),
        "sphere60": sph * np.array([0.3, 0.2, 0.1]),
        "cone": np.concatenate([np.c_[ring, np.zeros(24)], [[0, 0, 0.5]]]),
    }
    for name, v in hulls.items():
        m = _one_hull_model(v)
        h, o = hl.HostCheck(m), ol.Oracle(m)
        st = h.bsp_stats(0)
        assert 0 < st["nodes"] < 20000 and st["max_depth"] <= 40, (name, st)
        ext = np.abs(v).max()
        pts = np.concatenate([rng.uniform(-1.5, 1.5, (6000, 3)) * ext, rng.uniform(-3, 3, (2000, 3)), v + 1e-3 * rng.standard_normal(v.shape),
                              (v[rng.integers(0, len(v), 4000)] + v[rng.integers(0, len(v), 4000)]) / 2 + 2e-3 * rng.standard_normal((4000, 3))]).astype(np.float32)
        x = np.zeros(0)
        do = o.skin(x, pts.astype(np.float64), mode=ol.BRUTE)
        nv, npl, lo, hi, db = h.convex_info(0)
        cen = np.unique(np.round(v, 12), axis=0).mean(0)             # geometry frame = input frame shifted to the vertex centroid
        d, n, leaf = h.hull_sd_bsp(0, (pts.astype(np.float64) - cen).astype(np.float32))
        assert np.abs(d - do).max() < 1e-6 * (1 + np.abs(do).max()), (name, np.abs(d - do).max())
        # gradient: unit length off the surface, and consistent with central differences of the oracle away from the
        # feature boundaries
        far = np.abs(do) > 1e-3
        assert np.allclose(np.linalg.norm(n[far], axis=1), 1.0, atol=1e-4), name
        eps = 1e-4
        fd = np.stack([(o.skin(x, pts + eps * e, mode=ol.BRUTE) - o.skin(x, pts - eps * e, mode=ol.BRUTE)) / (2 * eps) for e in np.eye(3)], 1)
        ok = np.abs(fd - n).max(1) < 2e-2
        assert ok[far].mean() > 0.95, (name, ok[far].mean())
