"""GPU parity at the sizes that are BENCHMARKED (BASELINE.json configs[1], [3], [4]) -- the CUDA path through
the C ABI against the fp64 oracle on the same seeded inputs:

  K5  10,000,000 points x 64 configs (the headline): configs 0 and 37 against the oracle's single-pass adjoint on
      ALL points, the reference-faithful chunked-dual gradient (7 passes of Dual{9}, src/gradientdescent.jl:59-65 +
      ForwardDiff) on a 50,000-point subsample, argmin of config 0 on all points (flips only inside the tie band);
  K2  one 640 x 480 depth frame (307,200 points) of the IRB140, dual gradient on the whole frame;
  K4  8 frames x 307,200 points x 16 configs of the combined scene, two frames checked;
  the > 1,024-groups-per-tile fallback of cost_kernel (FLASHSDF_TILE), flash_points_set_device at scale, and the
  NCCL path (2 ranks, when the box has 2 GPUs).

Tolerance (BASELINE.json north_star): cost and gradient 1e-5 relative (gradient relative to its max-norm)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import flash
import oracle_lib as ol
import bench

pytestmark = pytest.mark.gpu
RTOL = 1e-5
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _check_cost_grad(o, x, pts, cost, grad, method, tag):
    co, go = o.cost_grad(x, pts, method=method, mode=ol.GJK)
    rc = abs(cost - co) / abs(co)
    rg = np.abs(grad - go).max() / np.abs(go).max()
    print("%s [%s]: cost rel %.2e grad rel %.2e" % (tag, method, rc, rg))
    assert rc <= RTOL, (tag, method, cost, co)
    assert rg <= RTOL, (tag, method, rg)
    return rc, rg


def _check_argmin(o, x, pts, arg, tag, tie_band=2e-6):
    _, ao = o.skin(x, pts, mode=ol.GJK, argmin=True)
    bad = np.nonzero(arg != ao)[0]
    print("%s: argmin flips %d of %d" % (tag, len(bad), len(pts)))
    if len(bad):
        per = o.surfaces(x, pts[bad], mode=ol.GJK)
        srt = np.sort(per, axis=1)
        gap = srt[:, 1] - srt[:, 0]
        assert (gap <= tie_band * np.maximum(1.0, np.abs(srt[:, 0]))).all(), (tag, gap.max())
        # a flip must still name one of the two tied surfaces
        two = np.argsort(per, axis=1)[:, :2]
        assert ((arg[bad][:, None] == two).any(1)).all()
    assert len(bad) <= max(2, len(pts) // 2000)
    return len(bad)


def test_k5_headline_size_against_oracle():
    m = bench.scene_model(flash)
    X = bench.scene_states(flash, m, 64)
    pts = bench.scene_points(10_000_000)
    f = flash.gradientdescent.CostFunctor(m, pts)
    cost, grad, _ = f.batch(X)
    o = ol.Oracle(m)
    for c in (0, 37):
        _check_cost_grad(o, X[c], pts, cost[c], grad[c], "adjoint", "K5 config %d, 10M points" % c)
    # the reference's own gradient algorithm on a subsample (a full pass is 7 x 10M dual evaluations)
    sub = pts[:: len(pts) // 50_000][:50_000]
    cs, gs, _ = flash.gradientdescent.CostFunctor(m, sub).batch(X[[0, 37]])
    for k, c in enumerate((0, 37)):
        _check_cost_grad(o, X[c], sub, cs[k], gs[k], "dual", "K5 config %d, 50k subsample" % c)
    # argmin on all 10M points, one config (the [C][N] buffer of all 64 would be 2.5 GB)
    f = flash.gradientdescent.CostFunctor(m, pts)
    _, _, arg = f.batch(X[:1], gradient=False, argmin=True)
    _check_argmin(o, X[0], pts, arg[0], "K5 config 0")


def _k2_frame():
    m = flash.models.irb140()
    cam = flash.depthsensors.affine((0, 1.5, 0.5), (1, 0, 0), np.pi / 2)
    pts, hits = bench.frame_cloud(flash, flash.ManipulatorState(m), cam)
    assert len(pts) == 307200 and hits > 10000
    return m, pts


def test_k2_full_frame_dual_gradient():
    m, pts = _k2_frame()
    rng = np.random.default_rng(2)
    X = rng.uniform(-np.pi / 2, np.pi / 2, (2, 6))
    cost, grad, arg = flash.gradientdescent.CostFunctor(m, pts).batch(X, argmin=True)
    o = ol.Oracle(m)
    for c in range(2):
        _check_cost_grad(o, X[c], pts, cost[c], grad[c], "dual", "K2 frame config %d" % c)
        _check_cost_grad(o, X[c], pts, cost[c], grad[c], "adjoint", "K2 frame config %d" % c)
    _check_argmin(o, X[0], pts, arg[0], "K2 frame")


def test_k4_frames_batch_against_oracle():
    rng = np.random.default_rng(2)
    m = bench.scene_model(flash)
    X = bench.scene_states(flash, m, 16)
    st = flash.ManipulatorState(m)
    nq = flash.num_positions(m)
    o = ol.Oracle(m)
    for k in (0, 5):
        x = X[0].copy()
        x[:nq] += 0.02 * rng.standard_normal(nq)
        flash.gradientdescent.unflatten_(st, x)
        cam = flash.depthsensors.affine((0.3 + 0.05 * k, 1.8, 1.0), (1, 0, 0), np.pi / 2 + 0.25)
        pts, hits = bench.frame_cloud(flash, st, cam)
        assert len(pts) == 307200 and hits > 5000
        cost, grad, arg = flash.gradientdescent.CostFunctor(m, pts).batch(X, argmin=True)
        for c in (0, 11):
            _check_cost_grad(o, X[c], pts, cost[c], grad[c], "adjoint", "K4 frame %d config %d" % (k, c))
        sub = pts[::6]
        cs, gs, _ = flash.gradientdescent.CostFunctor(m, sub).batch(X[[11]])
        _check_cost_grad(o, X[11], sub, cs[0], gs[0], "dual", "K4 frame %d config 11, 51k subsample" % k)
        _check_argmin(o, X[0], pts, arg[0], "K4 frame %d" % k)


_SNIPPET = r"""
import json, sys, numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, %(pkg)r); sys.path.insert(0, %(tests)r)
import flash, bench
m = bench.scene_model(flash)
X = bench.scene_states(flash, m, 2)
pts = bench.scene_points(2_000_000, seed=77)
c, g, a = flash.gradientdescent.CostFunctor(m, pts).batch(X, argmin=True)
print(json.dumps({"cost": c.tolist(), "grad": g.tolist(), "arg_sum": int(a.astype(np.int64).sum())}))
"""


def test_tile_fallback_more_than_1024_groups_per_tile():
    """FLASHSDF_TILE is read once per process: a child process with 65,536-point tiles (2,048 groups) takes the
    per-warp hull cull without the group pre-pass (flash_kernels.cuh) and must still match the oracle."""
    env = dict(os.environ, FLASHSDF_TILE="65536")
    code = _SNIPPET % {"root": ROOT, "pkg": os.path.join(ROOT, "point-cloud-signed-distance_b200"), "tests": HERE}
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    got = json.loads(out.stdout.strip().splitlines()[-1])
    m = bench.scene_model(flash)
    X = bench.scene_states(flash, m, 2)
    pts = bench.scene_points(2_000_000, seed=77)
    o = ol.Oracle(m)
    for c in range(2):
        _check_cost_grad(o, X[c], pts, got["cost"][c], np.array(got["grad"][c]), "adjoint", "tile fallback config %d" % c)
    # and the default tiling gives the same argmin sum on the same cloud
    _, _, a = flash.gradientdescent.CostFunctor(m, pts).batch(X, gradient=False, argmin=True)
    assert abs(int(a.astype(np.int64).sum()) - got["arg_sum"]) <= 4


def test_points_set_device_at_scale_unsorted():
    import ctypes as C
    import torch
    from flash import _lib
    m = bench.scene_model(flash)
    X = bench.scene_states(flash, m, 2)
    pts = bench.scene_points(2_000_000, seed=78)
    rng = np.random.default_rng(1)
    pts = pts[rng.permutation(len(pts))]          # caller's order: no Morton neighbours, wide group balls
    d4 = torch.zeros((len(pts), 4), dtype=torch.float32, device="cuda")
    d4[:, :3] = torch.from_numpy(pts.astype(np.float32)).cuda()
    torch.cuda.synchronize()
    f = flash.gradientdescent.CostFunctor(m, pts[:8])
    L = _lib.lib()
    m.check(L.flash_points_set_device(m.ctx(), C.c_void_p(d4.data_ptr()), len(pts)))
    f.sensed_points = pts
    m._points_owner = f              # the ctx now holds the device buffer on this functor's behalf
    cost, grad, arg = f.batch(X, argmin=True)
    o = ol.Oracle(m)
    for c in range(2):
        _check_cost_grad(o, X[c], pts, cost[c], grad[c], "adjoint", "points_set_device config %d" % c)
    _check_argmin(o, X[0], pts, arg[0], "points_set_device")


# ------------------------------------------------------------------ NCCL path (2 ranks)
def _nccl_worker(rank, world, port, q):
    sys.path.insert(0, HERE); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
    import ctypes as C
    import torch
    import torch.distributed as dist
    import flash as fl
    import bench as bn
    from flash import _lib
    from flash._desc import c_double_p
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    L = _lib.lib()
    m = bn.scene_model(fl)
    m.set_options(device=rank)
    ctx = m.ctx()
    raw = (C.c_ubyte * 128)()
    if rank == 0:
        m.check(L.flash_comm_unique_id(raw))
    t = torch.tensor(list(raw), dtype=torch.uint8)
    dist.broadcast(t, 0)
    raw = (C.c_ubyte * 128)(*t.tolist())
    m.check(L.flash_comm_init(ctx, rank, world, raw))
    X = np.ascontiguousarray(bn.scene_states(fl, m, 3))
    pts = np.ascontiguousarray(bn.scene_points(1_000_000, rank, world, seed=79))
    m.check(L.flash_points_set(ctx, pts.ctypes.data_as(c_double_p), len(pts)))
    cost = np.empty(3); grad = np.empty((3, 63))
    m.check(L.flash_cost_grad(ctx, X.ctypes.data_as(c_double_p), 3, 10.0, cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p), None))
    q.put((rank, cost, grad))
    dist.barrier()
    dist.destroy_process_group()


def test_nccl_two_ranks_equal_single_rank_and_oracle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + os.getpid() % 200
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=600) for _ in range(2)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    m = bench.scene_model(flash)
    X = bench.scene_states(flash, m, 3)
    pts = bench.scene_points(1_000_000, seed=79)
    c1, g1, _ = flash.gradientdescent.CostFunctor(m, pts).batch(X)
    for rank, c, g in res:
        # every rank returns the GLOBAL result; fp64 sums in a different order
        assert np.allclose(c, c1, rtol=1e-12), (rank, c, c1)
        assert np.abs(g - g1).max() <= 1e-6 * np.abs(g1).max()
    o = ol.Oracle(m)
    _check_cost_grad(o, X[1], pts, res[0][1][1], res[0][2][1], "adjoint", "NCCL 2 ranks config 1")
