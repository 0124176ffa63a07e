"""N > 1 path on CPU (gloo, world_size 2): points sharded across ranks, per-rank accumulators summed
with one all-reduce, epilogue on the sum == single-rank result.  The per-rank arithmetic is the
product's kernel math replayed on the host (tests/hostcheck); on GPUs the same sum is ncclAllReduce
inside flash_cost_grad."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _worker(rank, world, port, q):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "point-cloud-signed-distance_b200"))
    import torch
    import torch.distributed as dist
    import flash
    import hostcheck_lib as hl
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    gold = np.load(os.path.join(HERE, "golden", "golden_v1.npz"))
    x, pts = gold["irb_and_squishable/x"], gold["irb_and_squishable/pts"]
    h = hl.HostCheck(flash.models.irb_and_squishable())
    # bench.py's shard rule: the cloud is cut into x-slabs and rank r owns the slabs s with s % world == r
    order = np.argsort(pts[:, 0], kind="stable")
    slabs = np.array_split(order, 40)
    mine = np.concatenate([slabs[s] for s in range(40) if s % world == rank])
    tab, aux = h.prepare(x)
    acc, _, _ = h.accumulate(tab, pts[mine])
    t = torch.from_numpy(acc)
    dist.all_reduce(t)                                           # the one collective of the path
    cost, grad = h.epilogue(x, tab, aux, t.numpy())
    q.put((rank, cost, grad, len(mine)))
    dist.destroy_process_group()


def test_sharded_points_allreduce_matches_single_rank():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    sys.path.insert(0, HERE)
    import flash
    import hostcheck_lib as hl
    gold = np.load(os.path.join(HERE, "golden", "golden_v1.npz"))
    x, pts = gold["irb_and_squishable/x"], gold["irb_and_squishable/pts"]
    c1, g1, _, _ = hl.HostCheck(flash.models.irb_and_squishable()).cost_grad(x, pts)
    assert res[0][3] + res[1][3] == len(pts)
    for _, c, g, _ in res:                       # every rank holds the global result
        assert np.isclose(c, c1, rtol=1e-12) and np.allclose(g, g1, rtol=1e-10, atol=1e-12)


def test_shard_rule_covers_all_points_once():
    """bench.scene_points: under every sharding rule the union of the ranks' shards is the N = 1 cloud (same points,
    every one once), the shards are balanced and each spans the whole scene; blobs are spatially compact; the cloud is
    uniform in the scene box."""
    sys.path.insert(0, os.path.dirname(HERE))
    import bench
    key = lambda a: np.sort(a.view([("x", "f8"), ("y", "f8"), ("z", "f8")]).ravel())
    for n in (0, 1, 7, 100_003, 600_011):
        whole = bench.scene_points(n)
        assert whole.shape == (n, 3)
        for rule in ("blobs", "slabs", "index"):
            for world in (1, 2, 4, 8):
                parts = [np.ascontiguousarray(bench.scene_points(n, r, world, shard=rule)) for r in range(world)]
                assert sum(len(p) for p in parts) == n
                if n:
                    assert np.array_equal(key(np.concatenate(parts)), key(whole))
                if n > 1000:
                    sizes = [len(p) for p in parts]
                    assert max(sizes) - min(sizes) <= max(bench.SLABS // world + 1, bench.BLOB)
                    for p in parts:      # every rank sees (nearly) the full x-extent of the scene (blobs: once it holds a few of them)
                        if rule == "blobs" and len(p) < 8 * bench.BLOB:
                            continue
                        assert p[:, 0].max() - p[:, 0].min() > 0.75 * (bench.SCENE_HI[0] - bench.SCENE_LO[0])
    big = bench.scene_points(100_003)
    assert (big >= bench.SCENE_LO).all() and (big <= bench.SCENE_HI).all()
    assert np.array_equal(big, big.astype(np.float32).astype(np.float64))          # fp32-representable
    h = np.histogram(big[:, 0], bins=10, range=(bench.SCENE_LO[0], bench.SCENE_HI[0]))[0]
    assert h.max() - h.min() < 0.05 * h.mean()
    # a blob is a compact lump: its extent is far below the scene's
    part = bench.scene_points(2_000_000, 3, 8)
    blob = part[: bench.BLOB]
    assert (blob.max(0) - blob.min(0)).max() < 0.6     # a piece of the Z-curve: ~20 cm of the 1.5 m scene at this density
