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
    hl.set_group_cull(False)
    a0, d0, g0 = h.accumulate(tab, pts)
    hl.set_group_cull(True); hl.group_stats()
    a1, d1, g1 = h.accumulate(tab, pts)
    st = hl.group_stats()
    assert np.array_equal(d0, d1) and np.array_equal(g0, g1) and np.array_equal(a0, a1)
    assert st["groups"] == 3000 and st["candidates"] / st["groups"] < 5.0   # 8 hulls in the model


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


def test_support_direction_map_is_exact():
    """The cube-map candidate lists must contain the true support vertex for EVERY direction (also on cell
    borders and cube edges): the map query has to return the same support value as the 56-vertex scan."""
    rng = np.random.default_rng(8)
    m = flash.models.irb_and_squishable()
    h = hl.HostCheck(m)
    q = rng.standard_normal((200000, 3)).astype(np.float32)
    # directions exactly on cell borders / cube edges / axes
    g = np.linspace(-1, 1, 17)
    U, V = np.meshgrid(g, g)
    border = np.concatenate([np.stack([np.ones_like(U), U, V], -1).reshape(-1, 3)[:, p] * s for p in ([0, 1, 2], [1, 2, 0], [2, 0, 1]) for s in (1, -1)])
    q = np.concatenate([q, border.astype(np.float32), np.eye(3, dtype=np.float32), -np.eye(3, dtype=np.float32)])
    for j in range(h.n_convex):
        im, isc, vm, vs = h.support(j, q)
        assert np.array_equal(vm, vs), (j, np.abs(vm - vs).max())
        assert (im == isc).mean() > 0.999      # differ only on exact ties


def test_plane_hint_map_never_changes_the_distance():
    """hull_plane_hint (flash_dev.h) only picks GJK's start face and rejects by a valid lower bound; the
    distance must equal the oracle's brute-force point-polytope distance everywhere: hugging the surfaces
    (ray-cast hits pushed out / in by 1 mm and 1 cm), mid range, and far outside the scene."""
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
    # and it does its job: far fewer plane tests than one full 100-plane scan per narrow-phase evaluation
    assert cn["plane_tests"] < 45 * cn["narrow"]
