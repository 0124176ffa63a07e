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
    n = len(pts)
    lo, hi = (n * rank) // world, (n * (rank + 1)) // world     # bench.py's shard rule
    tab, aux = h.prepare(x)
    acc, _, _ = h.accumulate(tab, pts[lo:hi])
    t = torch.from_numpy(acc)
    dist.all_reduce(t)                                           # the one collective of the path
    cost, grad = h.epilogue(x, tab, aux, t.numpy())
    q.put((rank, cost, grad, hi - lo))
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
    for n in (0, 1, 7, 307200, 10_000_000):
        for world in (1, 2, 4, 8):
            spans = [((n * r) // world, (n * (r + 1)) // world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1
