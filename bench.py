#!/usr/bin/env python
"""bench.py -- point x configuration SDF cost+gradient evaluations per second (BASELINE.json metric).

Workload (BASELINE.json configs[4], "K5"): the IRB140(floating base)+squishable+table scene of
examples/irb_and_squishable.ipynb (63 states, 9 surfaces), 10,000,000 synthetic points uniform in the
scene AABB [-0.5,1]x[-0.5,0.5]x[0.5,2], 64 candidate configurations (warm start + N(0,0.05)).
One "step" = one flash_cost_grad call: cost[64] + grad[64][63] over all points.
With N GPUs the POINTS of the same global cloud are sharded across ranks (total work fixed -> "strong" scaling;
scene_points: Morton-ordered blobs dealt round-robin by default); each rank's
partial accumulators are summed by one ncclAllReduce inside the library before the epilogue.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode exact|grid]

--impl reference times the CPU oracle (the fp64 C++ restatement of the Julia reference; Julia 0.5 and
its six un-vendored packages cannot run on this box) on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "point-cloud-signed-distance_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "point_config_cost_grad_evals_per_sec"
UNIT = "evals/s"
SCENE_LO, SCENE_HI = np.array([-0.5, -0.5, 0.5]), np.array([1.0, 0.5, 2.0])


def scene_model(flash):
    return flash.models.irb_and_squishable()


def scene_states(flash, m, n_configs, seed=5):
    """Warm start of examples/irb_and_squishable.ipynb cell 6 + N(0, 0.05) perturbations."""
    rng = np.random.default_rng(seed)
    nx, nq = flash.num_states(m), flash.num_positions(m)
    x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m))
    x0[4:7] = [0.0, 0.0, 0.75]            # irb floating base
    x0[13 + 4:13 + 7] = [0.55, 0.45, 0.8]  # squishable
    x0[20 + 4:20 + 7] = [0.4, 0.0, 0.6]    # table
    X = np.repeat(x0[None], n_configs, 0)
    X[1:, :nq] += 0.05 * rng.standard_normal((n_configs - 1, nq))
    X[1:, nq:] += 0.005 * rng.standard_normal((n_configs - 1, nx - nq))
    return X


SLABS = 40   # divisible by 1, 2, 4, 8 (and 5, 10, 20)


def slab_counts(n_total):
    base, rem = divmod(n_total, SLABS)
    return [base + (1 if s < rem else 0) for s in range(SLABS)]


def _global_cloud(n_total, seed):
    """The global seeded cloud: fp32-representable points uniform in the scene AABB, generated as SLABS equal-count
    slabs along x (stratified uniform sampling: slab s is uniform in its x-range, drawn from its own generator
    (seed, s)) -- the same cloud at every rank count."""
    counts = slab_counts(n_total)
    pts = np.empty((n_total, 3), np.float64)
    at = 0
    w = (SCENE_HI[0] - SCENE_LO[0]) / SLABS
    for s_ in range(SLABS):
        rng = np.random.default_rng([seed, s_])
        u = rng.random((counts[s_], 3), dtype=np.float32)
        lo = np.array([SCENE_LO[0] + s_ * w, SCENE_LO[1], SCENE_LO[2]])
        ext = np.array([w, SCENE_HI[1] - SCENE_LO[1], SCENE_HI[2] - SCENE_LO[2]])
        pts[at:at + counts[s_]] = (lo + ext * u).astype(np.float32)
        at += counts[s_]
    return pts


BLOB = 8192   # points per blob of the "blobs" sharding rule (a multiple of the kernels' 32-point groups)


def _morton_order(pts):
    """argsort by a 30-bit Morton key over the scene AABB (host side, numpy): spatial neighbours become index neighbours"""
    q = np.clip(((pts - SCENE_LO) / (SCENE_HI - SCENE_LO) * 1024).astype(np.int64), 0, 1023)
    key = np.zeros(len(pts), np.int64)
    for b_ in range(10):
        for a_ in range(3):
            key |= ((q[:, a_] >> b_) & 1) << (3 * b_ + a_)
    return np.argsort(key, kind="stable")


def scene_points(n_total, rank=0, world=1, seed=1000, shard="blobs"):
    """The points of `rank` out of `world` of the global seeded cloud (_global_cloud), by one of three sharding rules:

      blobs   (default) the cloud in Morton order is cut into blobs of BLOB consecutive points -- compact lumps a few
              centimetres across -- dealt to the ranks round-robin: every rank holds points as DENSE as the full
              cloud (the kernels' 32-point groups keep their small bounding balls) from ALL OVER the scene (the ranks'
              work is balanced).  What a caller who splits a frame into image tiles gets.
      slabs   rank r owns the x-slabs s with s % world == r: dense, but the ranks whose slabs cut through the robot work
              17 % longer than the others at world = 8 (round-2 measurement, per_rank in the bench line).
      index   contiguous index ranges of the cloud in random order: balanced, but every shard is a thin sample of the
              whole scene -- 1/world of the density, bounding balls world^(1/3) times wider: +28 % kernel time per
              evaluation at world = 8 (tools/density_probe.py).
    world = 1 returns the whole cloud in generation order under every rule."""
    pts = _global_cloud(n_total, seed)
    if world == 1:
        return pts
    if shard == "slabs":
        counts = slab_counts(n_total)
        offs = np.concatenate([[0], np.cumsum(counts)])
        return np.concatenate([pts[offs[s_]:offs[s_ + 1]] for s_ in range(SLABS) if s_ % world == rank]) if n_total else pts
    if shard == "index":
        perm = np.random.default_rng([seed, 12345]).permutation(n_total)
        return pts[perm[(n_total * rank) // world:(n_total * (rank + 1)) // world]]
    order = _morton_order(pts)
    nblob = (n_total + BLOB - 1) // BLOB
    mine = [order[b_ * BLOB:(b_ + 1) * BLOB] for b_ in range(rank, nblob, world)]
    return pts[np.concatenate(mine)] if mine else pts[:0]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile(prefix="clocks_", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None
            return
        # nvidia-smi needs a few hundred ms to print its first line: wait for it so that even a 50 ms timed
        # region (N=8) is bracketed by samples
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < 5.0 and os.path.getsize(self.path) == 0:
            time.sleep(0.02)

    def mark(self):
        """number of samples taken so far (to tell the samples of the timed region from the warm-up's)"""
        try:
            return sum(1 for _ in open(self.path))
        except Exception:
            return 0

    def stop(self, first=0):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for li, line in enumerate(open(self.path)):
            if li < max(first - 1, 0):
                continue
            t = [v.strip() for v in line.split(",")]
            if len(t) < 9:
                continue
            try:
                sm.append(float(t[1])); mx.append(float(t[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), t[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


# -------------------------------------------------------------------------------------------------
def cpu_leg(n_points, n_configs, steps, warmup, threads=None, method="dual", seed=11):
    """Times the CPU oracle (reference-faithful chunked forward-mode gradient by default) on a bounded
    sample of the bench workload.  Returns evals/s and a description."""
    import flash
    import oracle_lib as ol
    m = scene_model(flash)
    o = ol.Oracle(m, threads=threads)
    X = scene_states(flash, m, max(n_configs, 2))[:n_configs]
    pts = scene_points(n_points, seed=seed)
    for _ in range(warmup):
        o.cost_grad(X[0], pts[: max(1000, n_points // 20)], mode=ol.GJK, method=method)
    t0 = time.perf_counter()
    for _ in range(steps):
        for c in range(n_configs):
            o.cost_grad(X[c], pts, mode=ol.GJK, method=method)
    dt = time.perf_counter() - t0
    return {"value": n_points * n_configs * steps / dt, "unit": UNIT, "cores": o.threads, "kind": "port",
            "sample": "%d points x %d configs x %d steps of the same scene; oracle/ C++ fp64 port, %s gradient (%s), %d threads"
                      % (n_points, n_configs, steps, "ForwardDiff-style chunked dual" if method == "dual" else "single-pass adjoint",
                         "7 passes of Dual{9} for 63 states" if method == "dual" else "not the reference's algorithm", o.threads),
            "seconds": dt}


def run_reference(args, rank, world):
    if rank != 0:
        return
    base = cpu_leg(args.ref_points, args.ref_configs, args.steps, min(args.warmup, 1))
    line = {"metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * base["seconds"] / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": "K5: IRB140(floating)+squishable+table, 63 states, 9 surfaces; bounded sample of 10M points x 64 configs: ALL %d configurations, %d of the points (uniform in the same scene box; the rate is per evaluation)" % (args.ref_configs, args.ref_points),
                       "points": args.ref_points, "configs": args.ref_configs},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "Julia 0.5 reference cannot run here (no julia, un-vendored deps); this is the oracle/ C++ port on the host cores"}
    print(json.dumps(line), flush=True)


# -------------------------------------------------------------------------------------------------
def run_ours(args, rank, local_rank, world):
    import torch
    import flash
    from flash import _lib
    from flash._desc import Timing, c_double_p

    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = _lib.lib()
    single_proc = None
    if world > 1 and args.extras:
        # The same box through ONE process and ONE multi-device context (flash_options.n_devices -- what a single Julia
        # process uses), measured BEFORE the ranks touch their GPUs: rank 0 runs it in a child process (bounded by a
        # timeout) while the other ranks wait on a CPU (gloo) barrier, so no NCCL kernel spins on the devices meanwhile.
        gl = dist.new_group(backend="gloo")
        if rank == 0:
            try:
                r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "multi_bench.py"), str(world), str(args.points), str(args.configs)],
                                   capture_output=True, text=True, timeout=300)
                single_proc = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else {"error": r.stderr[-400:]}
            except Exception as e:   # reported, never fatal for the headline
                single_proc = {"error": repr(e)}
        dist.barrier(group=gl)
    m = scene_model(flash)
    m.set_options(device=local_rank, sdf_mode=1 if args.mode == "grid" else 0, grid_spacing=args.grid_spacing)
    ctx = m.ctx()
    nx = flash.num_states(m)
    acc_len = 1 + 6 * sum(isinstance(s_, flash.ConvexGeometry) for s_ in m.surfaces) + sum(
        4 * (len(s_.surface_points) + len(s_.skeleton_points)) + 4 for s_ in m.surfaces if not isinstance(s_, flash.ConvexGeometry))
    Cn = args.configs
    X = np.ascontiguousarray(scene_states(flash, m, Cn))

    # shard the points: rank r owns the x-slabs s with s % world == r of the global cloud (scene_points)
    N = args.points
    shard = scene_points(N, rank, world, shard=args.shard)
    n_local = len(shard)
    pin = torch.empty((n_local, 3), dtype=torch.float64, pin_memory=True)
    pts = pin.numpy()
    pts[:] = shard
    del shard

    if world > 1:
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            raw = (C.c_ubyte * 128)()
            m.check(L.flash_comm_unique_id(raw))
            idbuf.copy_(torch.tensor(list(raw), dtype=torch.uint8))
        dist.broadcast(idbuf, 0)
        raw = (C.c_ubyte * 128)(*idbuf.cpu().tolist())
        m.check(L.flash_comm_init(ctx, rank, world, raw))

    cost = np.empty(Cn); grad = np.empty((Cn, nx))
    pX, pc, pg, pp = X.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p), pts.ctypes.data_as(c_double_p)
    m.check(L.flash_points_set(ctx, pp, n_local))
    m.check(L.flash_profiling(ctx, 1))
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))

    def e2e_step():
        # the live loop of examples/irb_and_squishable.ipynb:495-503: frame k is evaluated while frame k+1 uploads.
        # Every step hands the library one frame of HOST points (24 B x N, pinned) and reads cost/grad back on the host;
        # the upload of the next step's frame is issued first and overlaps this step's kernels (copy stream + DMA).
        m.check(L.flash_points_swap(ctx))                          # the frame uploaded during the previous step
        m.check(L.flash_points_set_async(ctx, pp, n_local))        # the next one
        m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))

    def e2e_step_serial():
        m.check(L.flash_points_set(ctx, pp, n_local))
        m.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))

    import gc
    gc.collect(); gc.disable()   # as timeit does: a collection inside one of the few timed steps of one rank would be charged to all of them
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()      # sampling starts before the warm-up and ends after the end-to-end loop
    for _ in range(max(args.warmup, 3)):
        step()
    launches0 = L.flash_launch_count(ctx)
    first_sample = sampler.mark() if rank == 0 else 0
    t_steps, k_main, k_total, k_group = [], [], [], []
    tm = Timing()
    for _ in range(args.steps):
        flush.zero_()
        barrier()
        t0 = time.perf_counter()
        step()
        torch.cuda.synchronize()
        t_steps.append(time.perf_counter() - t0)
        L.flash_last_timing(ctx, C.byref(tm))
        k_main.append(tm.ms_main); k_total.append(tm.ms_total); k_group.append(tm.ms_group)
    barrier()
    launches = L.flash_launch_count(ctx) - launches0
    checksum = float(cost.sum())

    # end-to-end: host (pinned) points -> device every step, results back on the host
    t_ser = []
    for k in range(1 + min(args.steps, 3)):
        flush.zero_()
        barrier()
        t0 = time.perf_counter()
        e2e_step_serial()
        torch.cuda.synchronize()
        if k:
            t_ser.append(time.perf_counter() - t0)
    m.check(L.flash_points_set_async(ctx, pp, n_local))
    e2e_step()
    t_e2e = []
    for _ in range(args.steps):
        flush.zero_()
        barrier()
        t0 = time.perf_counter()
        e2e_step()
        torch.cuda.synchronize()
        t_e2e.append(time.perf_counter() - t0)
    m.check(L.flash_points_swap(ctx))   # retire the last upload
    barrier()
    clocks = sampler.stop(first_sample) if rank == 0 else None
    gc.enable()

    arg0 = None
    if args.parity:
        # argmin of config 0 over this rank's shard: a COLLECTIVE call when a communicator is attached (every rank makes it);
        # rank 0 checks its shard against the oracle in parity_leg
        from flash._desc import c_int32_p
        arg0 = np.empty(n_local, np.int32)
        c1 = np.empty(1)
        m.check(L.flash_cost_grad(ctx, X[:1].ctypes.data_as(c_double_p), 1, 10.0, c1.ctypes.data_as(c_double_p), None, arg0.ctypes.data_as(c_int32_p)))
    tt = torch.tensor([sum(t_steps), sum(t_e2e), sum(k_main) / 1e3, sum(k_total) / 1e3, sum(t_ser) / len(t_ser) * args.steps, sum(k_group) / 1e3], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    T, Te, Tk, Tkt, Tser, Tg = [float(v) for v in tt.cpu()]
    per_rank = None
    if dist is not None:   # where the time goes on every rank: cost kernel, group kernel, whole device step (ms, mean over the timed steps)
        mine = torch.tensor([(sum(k_main) - sum(k_group)) / args.steps, sum(k_group) / args.steps, sum(k_total) / args.steps, 1e3 * sum(t_steps) / args.steps],
                            dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        a = np.array([t.cpu().numpy() for t in allr])
        per_rank = {"cost_kernel_ms": [round(float(v), 4) for v in a[:, 0]], "group_kernel_ms": [round(float(v), 4) for v in a[:, 1]],
                    "device_step_ms": [round(float(v), 4) for v in a[:, 2]], "host_step_ms": [round(float(v), 4) for v in a[:, 3]],
                    "cost_kernel_min_median_max": [float(np.min(a[:, 0])), float(np.median(a[:, 0])), float(np.max(a[:, 0]))]}
    units = float(N) * Cn * args.steps
    if rank == 0:
        peaks, peak_src = measured_peaks()
        sm = C.c_int32(); khz = C.c_int32(); l2 = C.c_int64()
        L.flash_device_info(ctx, C.byref(sm), C.byref(khz), C.byref(l2))
        # algorithmic HBM bytes of the main kernel per launch: every point (float4, 16 B) is read once per
        # config pass it is not L2-resident for; the algorithmic minimum is one read per launch per rank
        # plus the per-config tables -- the kernel is ALU-bound, see DESIGN.md.
        n_rank = (N + world - 1) // world
        alg_bytes = 16.5 * n_rank + 8.0 * Cn * 400 + 2 * 16.0 * Cn * n_rank / 32   # float4 point + 16 B bounding ball per 32 points; group records
        kernel_s = (Tk - Tg) / args.steps     # cost_kernel alone (CUDA events on the library's stream); group_kernel = Tg
        group_s = Tg / args.steps
        achieved = alg_bytes / kernel_s / 1e9
        # instruction and DRAM counts per evaluation come from the `ncu --set full` capture of THIS kernel build
        # (profiles/r02_ncu_constants.json, written by tools/refresh_profiles.sh from the capture of the same bench
        # command; "kernel_sha" = hash of csrc/ at capture time), scaled to this launch and divided by this run's
        # CUDA-event kernel time
        ncu = {}
        try:
            ncu = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_constants.json")))
        except Exception:
            pass
        traffic = None
        roof = {"bound": "issue", "achieved": None, "peak": None, "unit": "warp-instr/s", "frac": None, "traffic": None, "kernel": "cost_kernel<GRAD>",
                "kernel_ms": 1e3 * kernel_s, "group_kernel_ms": 1e3 * group_s, "note": "profiles/r02_ncu_constants.json missing"}
        sm_clock = (clocks.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)) * 1e6
        fp32_peak = sm.value * 128 * 2 * sm_clock / 1e12
        if ncu:
            import hashlib
            sha = hashlib.sha1()
            for fn in ("flash_dev.h", "flash_kernels.cuh"):
                sha.update(open(os.path.join(ROOT, "point-cloud-signed-distance_b200", "csrc", fn), "rb").read())
            sha_ok = sha.hexdigest()[:12] == ncu.get("kernel_sha")
            scale = (n_rank * Cn) / (ncu["points"] * ncu["configs"])
            traffic = (ncu["dram_bytes_read"] + ncu["dram_bytes_write"]) * scale
            winst = ncu["warp_instructions"] * scale
            peak_issue = sm.value * 4 * sm_clock
            roof = {"bound": "issue", "achieved": winst / kernel_s, "peak": peak_issue, "unit": "warp-instr/s", "frac": winst / kernel_s / peak_issue,
                    "traffic": traffic, "kernel": "cost_kernel<GRAD>", "kernel_ms": 1e3 * kernel_s, "group_kernel_ms": 1e3 * group_s,
                    "kernel_share_of_step": kernel_s / (Tkt / args.steps),
                    "warp_instr_per_eval": ncu["warp_instructions"] / (ncu["points"] * ncu["configs"]),
                    "lanes_active": ncu.get("lanes_active"), "warps_active_pct": ncu.get("warps_active_pct"), "registers": ncu.get("registers"),
                    "spill_bytes": ncu.get("spill_bytes"), "issue_active_pct_under_ncu": ncu.get("issue_active_pct"),
                    "fp32_flops_frac": (ncu["fp32_flops"] * scale / kernel_s / 1e12 / fp32_peak) if "fp32_flops" in ncu else None,
                    "fp64_flops_tflops": (ncu["fp64_flops"] * scale / kernel_s / 1e12) if "fp64_flops" in ncu else None,
                    "peak_source": "4 warp-instructions/clk/SM x %d SMs x SM clock sampled during the run" % sm.value,
                    "source": "profiles/r02_ncu_constants.json (ncu --set full of this kernel build, %s) x this run's evaluations / this run's kernel time" % ("device code hash %s %s the sources of this run" % (ncu.get("kernel_sha", "?"), "MATCHES" if sha_ok else "DOES NOT MATCH")),
                    "profile_matches_this_build": sha_ok,
                    "note": "exact mode is instruction-issue bound (BSP descents, certified culls, fp64 RBF adjoints), not memory bound: HBM figures in hbm_roofline"}
        line = {
            "metric": METRIC, "value": units / T, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": 1e3 * T / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 distances / f64 RBF + accumulation", "data": "synthetic",
            "config": {"workload": "K5: IRB140(floating)+squishable+table (63 states, 9 surfaces), %d points x %d configs, points sharded over %d GPU(s)" % (N, Cn, world),
                       "points": N, "configs": Cn, "sdf_mode": args.mode, "l2": "L2 flushed (512 MiB memset) between timed steps; inputs %s L2" % ("exceed" if 16 * n_rank > l2.value else "fit"),
                       "parallelism": "points sharded x%d (rule '%s': %s), one ncclAllReduce of [%d][%d] fp64 per step" % (
                           world, args.shard, {"blobs": "Morton-ordered blobs of %d points dealt round-robin" % BLOB, "slabs": "interleaved x-slabs", "index": "index ranges of the shuffled cloud"}[args.shard], Cn, acc_len)},
            "device_ms_per_step": 1e3 * Tkt / args.steps,
            "roofline": roof,
            "hbm_roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                             "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes": alg_bytes,
                             "note": "16.5 B per point per launch (float4 + its share of the group ball) + 16 B group record per (group, config) written and read once; the path is not HBM-bound"},
            "alu": {"evals_per_sm_clock": (units / world / args.steps) / kernel_s / (sm.value * sm_clock), "fp32_peak_tflops_nominal": fp32_peak},
            "clocks": clocks,
            "e2e": {"value": units / Te, "unit": UNIT, "h2d_bytes_per_step": int(n_local * 24 + X.nbytes), "d2h_bytes_per_step": int(cost.nbytes + grad.nbytes),
                    "ms_per_step": 1e3 * Te / args.steps,
                    "how": "flash_points_swap + flash_points_set_async(next frame) + flash_cost_grad per step: the upload, Hilbert sort and grouping of frame k+1 overlap the kernels of frame k",
                    "serial_value": units / Tser, "serial_ms_per_step": 1e3 * Tser / args.steps,
                    "serial_how": "flash_points_set (blocking) + flash_cost_grad per step, nothing overlapped"},
            "gpu_launches": int(launches),
            "checksum_cost_sum": checksum,
        }
        if world == 1 and args.extras:
            if args.grid_leg:
                line["grid_mode"] = grid_leg(args, flash, L, X, pts, n_local, cost.copy(), flush)
            if args.tracking_leg:
                line["tracking"] = tracking_leg(flash, L)
            if args.configs_leg:
                line["other_configs"] = configs_leg(flash, L, args.steps)
        if per_rank is not None:
            line["per_rank"] = per_rank
        if single_proc is not None:
            line["single_process_ctx"] = single_proc
        if args.parity:
            line["parity"] = parity_leg(flash, m, L, ctx, X, cost, grad, pts, n_local, N, rank, world, arg0)
        if args.cpu_baseline and world == 1:
            base = cpu_leg(args.ref_points, args.ref_configs, 1, 1)
            fast = cpu_leg(args.ref_points, args.ref_configs, 1, 0, method="adjoint")
            base["adjoint_value"] = fast["value"]
            base["adjoint_note"] = "same oracle, single-pass analytic gradient (not the reference's algorithm): best CPU algorithm on %d threads" % fast["cores"]
            line["cpu_baseline"] = base
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def parity_leg(flash, m, L, ctx, X, cost, grad, pts_local, n_local, N, rank, world, arg):
    """Config 0 of the TIMED workload against the fp64 oracle (tests/oracle_lib, the checker -- never the thing
    measured): cost and gradient over the whole global cloud, argmin over this rank's shard."""
    import oracle_lib as ol
    from flash._desc import c_double_p, c_int32_p
    t0 = time.perf_counter()
    o = ol.Oracle(m)
    full = pts_local[:n_local] if world == 1 else scene_points(N)
    co, go = o.cost_grad(X[0], full, method="adjoint", mode=ol.GJK)
    _, ao = o.skin(X[0], pts_local[:n_local], mode=ol.GJK, argmin=True)
    bad = np.nonzero(arg != ao)[0]
    outside_band = 0
    if len(bad):
        per = np.sort(o.surfaces(X[0], pts_local[bad], mode=ol.GJK), axis=1)
        outside_band = int(((per[:, 1] - per[:, 0]) > 2e-6 * np.maximum(1.0, np.abs(per[:, 0]))).sum())
    return {"config": 0, "points": int(len(full)), "cost_rel": float(abs(cost[0] - co) / abs(co)), "grad_rel": float(np.abs(grad[0] - go).max() / np.abs(go).max()),
            "argmin_flips": int(len(bad)), "argmin_flips_outside_2e-6_tie_band": outside_band, "argmin_points": int(n_local),
            "oracle": "oracle/ fp64 C++ restatement, single-pass adjoint gradient, GJK distances, %d threads" % o.threads,
            "seconds": time.perf_counter() - t0}


def grid_leg(args, flash, L, X, pts, n_local, cost_exact, flush):
    """Same workload with FLASH_SDF_GRID (hull distances from baked body-frame grids, manual fp32
    trilinear): reported beside the exact-mode headline together with its deviation from it."""
    import torch
    from flash._desc import Timing, c_double_p
    mg = scene_model(flash)
    mg.set_options(sdf_mode=1, grid_spacing=args.grid_spacing)
    ctx = mg.ctx()
    Cn, nx = X.shape
    cost = np.empty(Cn); grad = np.empty((Cn, nx))
    pX, pc, pg = X.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p)
    mg.check(L.flash_points_set(ctx, pts.ctypes.data_as(c_double_p), n_local))
    mg.check(L.flash_profiling(ctx, 1))
    for _ in range(3):
        mg.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))
    ts, km = [], []
    tm = Timing()
    for _ in range(args.steps):
        flush.zero_(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        mg.check(L.flash_cost_grad(ctx, pX, Cn, 10.0, pc, pg, None))
        ts.append(time.perf_counter() - t0)
        L.flash_last_timing(ctx, C.byref(tm)); km.append(tm.ms_main)
    errs = [mg.grid_report(s)[2] for s in range(len(mg.surfaces)) if isinstance(mg.surfaces[s], flash.ConvexGeometry)]
    dims = [mg.grid_report(s)[0] for s in range(len(mg.surfaces)) if isinstance(mg.surfaces[s], flash.ConvexGeometry)]
    nodes = int(sum(d[0] * d[1] * d[2] for d in dims))
    out = {"value": float(n_local) * Cn * args.steps / sum(ts), "unit": UNIT, "ms_per_step": 1e3 * sum(ts) / args.steps, "kernel_ms": float(np.mean(km)),
           "grid_spacing_m": args.grid_spacing, "grid_nodes": nodes, "grid_bytes": 4 * nodes,
           "max_abs_err_vs_exact_sdf_m": float(max(errs)), "cost_rel_dev_vs_exact_mode": float(np.abs(cost - cost_exact).max() / np.abs(cost_exact).max()),
           "note": "approximate mode (trilinear SDF grids for the convex bodies; RBF skin stays analytic fp64); parity is defined on the same grid, see DESIGN.md"}
    # the north_star's gather roofline: measured on this GPU model by tools/microbench/l2_gather.cu (profiles/r01_l2_gather.json)
    try:
        gm = json.load(open(os.path.join(ROOT, "profiles", "r01_l2_gather.json")))["l2_48MiB"]
        samples_per_eval = 1.2   # <= narrow-phase evaluations per point*config after culling (host replay counters: 1.11), each at most one 8-corner sample
        ach = float(n_local) * Cn * samples_per_eval * 32.0 / (float(np.mean(km)) * 1e-3) / 1e9
        out["gather_roofline"] = {"bound": "l2 gather (8 manual fp32 corner loads per trilinear sample, grids L2-resident)", "achieved": ach,
                                  "peak": gm["trilerp8_warp_coherent"]["alg_gbs"], "unit": "GB/s", "frac": ach / gm["trilerp8_warp_coherent"]["alg_gbs"],
                                  "peak_random_access": gm["trilerp8_random"]["alg_gbs"],
                                  "note": "upper bound of the achieved gather rate (<= %.1f samples per evaluation x 32 B); peak = same access pattern alone on a 48 MiB grid, "
                                          "Morton-coherent warps; the kernel is bound by culling + the RBF skin pre-filter, not by the gather (DESIGN.md 5)" % samples_per_eval}
    except Exception:
        pass
    mg.invalidate()
    return out


def frame_cloud(flash, state, cam, rows=480, cols=640, ground_z=0.0, far=3.0):
    """A full Kinect frame (rows*cols points exactly): ray-cast hits on the model, misses replaced
    by the ray's hit on the ground plane z = ground_z (or the point at `far` metres when it never lands)."""
    sensor = flash.depthsensors.Kinect(rows, cols)
    depths = flash.depthsensors.raycast_depths(flash.skin(state), sensor, cam).reshape(-1)
    rays = sensor.rays.reshape(-1, 3) @ cam[:, :3].T
    rays /= np.linalg.norm(rays, axis=1, keepdims=True)
    org = cam[:, 3]
    with np.errstate(divide="ignore", invalid="ignore"):
        t_ground = (ground_z - org[2]) / rays[:, 2]
    t_ground = np.where((t_ground > 0) & np.isfinite(t_ground) & (t_ground < 10.0), t_ground, far)
    d = np.where(np.isnan(depths), t_ground, depths)
    pts = org + d[:, None] * rays
    return pts.astype(np.float32).astype(np.float64), int((~np.isnan(depths)).sum())


def configs_leg(flash, L, steps):
    """BASELINE.json configs[1] (K2) and configs[3] (K4) as SURVEY 8(d) defines them: depth frames of the
    models themselves (points ON the surfaces -- the tracking regime, unlike K5's uniform cloud)."""
    out = {}
    rng = np.random.default_rng(2)
    # K2: IRB140, one 640x480 frame from (0, 1.5, 0.5) looking along -y, one candidate configuration
    m = flash.models.irb140()
    st = flash.ManipulatorState(m)
    cam = flash.depthsensors.affine((0, 1.5, 0.5), (1, 0, 0), np.pi / 2)
    pts, hits = frame_cloud(flash, st, cam)
    X = rng.uniform(-np.pi / 2, np.pi / 2, (1, 1, 6))
    out["k2_irb140_one_frame_one_config"] = _time_frames(flash, L, m, [pts], X, steps, hits)
    # K4: IRB140(floating)+squishable+table, 8 frames x 16 configs per frame, ONE flash_cost_grad_frames launch per batch
    m4 = scene_model(flash)
    X4 = scene_states(flash, m4, 16)
    st4 = flash.ManipulatorState(m4)
    frames = []
    hits4 = 0
    for k in range(8):
        x = X4[0].copy()
        x[:flash.num_positions(m4)] += 0.02 * rng.standard_normal(flash.num_positions(m4))
        flash.gradientdescent.unflatten_(st4, x)
        cam4 = flash.depthsensors.affine((0.3 + 0.05 * k, 1.8, 1.0), (1, 0, 0), np.pi / 2 + 0.25)
        p, h = frame_cloud(flash, st4, cam4)
        frames.append(p); hits4 += h
    out["k4_combined_8_frames_x_16_configs"] = _time_frames(flash, L, m4, frames, np.repeat(X4[None], 8, 0), steps, hits4)
    return out


def _time_frames(flash, L, m, clouds, X, steps, hits):
    """A batch of F frames x C candidate states per frame through flash_frames_set(_async) + flash_cost_grad_frames
    (the batch entry points of include/flashsdf.h).  device = CUDA-event time of one flash_cost_grad_frames call with
    the frames resident; e2e = per batch, the F frames from pinned HOST memory (upload + Hilbert sort + grouping on the
    copy stream, overlapping the previous batch's kernels) + the call + cost/grad back on the host."""
    import torch
    from flash._desc import Timing, c_double_p
    ctx = m.ctx()
    F, Cn, nx = X.shape
    X = np.ascontiguousarray(X)
    cost = np.empty((F, Cn)); grad = np.empty((F, Cn, nx))
    pX, pc, pg = X.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p)
    m.check(L.flash_profiling(ctx, 1))
    tm = Timing()
    keep, ptr_list = [], []
    for p in clouds:   # frames live in pinned host memory (bench contract: H2D from pinned memory)
        t = torch.empty(p.shape, dtype=torch.float64, pin_memory=True)
        a = t.numpy(); a[:] = p
        keep.append(t); ptr_list.append(a.ctypes.data_as(c_double_p))
    ptrs = (c_double_p * F)(*ptr_list)
    counts = (C.c_int64 * F)(*[len(p) for p in clouds])
    n = sum(len(p) for p in clouds)
    m.check(L.flash_frames_set(ctx, ptrs, counts, F))
    k_ms = []
    for k in range(2 + steps):
        m.check(L.flash_cost_grad_frames(ctx, pX, Cn, 10.0, pc, pg))
        L.flash_last_timing(ctx, C.byref(tm))
        if k >= 2:
            k_ms.append((tm.ms_total, tm.ms_main, tm.ms_group))
    torch.cuda.synchronize()
    # serial end to end
    t0 = time.perf_counter()
    for _ in range(steps):
        m.check(L.flash_frames_set(ctx, ptrs, counts, F))
        m.check(L.flash_cost_grad_frames(ctx, pX, Cn, 10.0, pc, pg))
    torch.cuda.synchronize()
    t_serial = (time.perf_counter() - t0) / steps
    # pipelined end to end: batch k+1 uploads while batch k is evaluated
    m.check(L.flash_frames_set_async(ctx, ptrs, counts, F))
    def step():
        m.check(L.flash_points_swap(ctx))
        m.check(L.flash_frames_set_async(ctx, ptrs, counts, F))
        m.check(L.flash_cost_grad_frames(ctx, pX, Cn, 10.0, pc, pg))
    step()
    torch.cuda.synchronize()
    reps = max(steps, 10)
    t0 = time.perf_counter()
    for _ in range(reps):
        step()
    torch.cuda.synchronize()
    t_pipe = (time.perf_counter() - t0) / reps
    m.check(L.flash_points_swap(ctx))
    m._points_owner = None
    dev = float(np.mean([k[0] for k in k_ms]))
    return {"points": n, "ray_hits_on_model": hits, "configs": Cn, "frames": F,
            "device_ms_per_batch": dev, "device_evals_per_sec": n * Cn / (dev / 1e3),
            "cost_kernel_ms": float(np.mean([k[1] - k[2] for k in k_ms])), "group_kernel_ms": float(np.mean([k[2] for k in k_ms])),
            "e2e_ms_per_batch": 1e3 * t_pipe, "e2e_evals_per_sec": n * Cn / t_pipe,
            "e2e_serial_ms_per_batch": 1e3 * t_serial, "e2e_serial_evals_per_sec": n * Cn / t_serial,
            "h2d_bytes_per_batch": int(24 * n + X.nbytes), "d2h_bytes_per_batch": int(cost.nbytes + grad.nbytes),
            "note": "flash_frames_set(_async) + flash_cost_grad_frames, one launch per batch; device = CUDA-event time of prepare + group + cost + epilogue with "
                    "the frames resident; e2e = frames from pinned host memory every batch, pipelined (batch k+1 uploads while batch k runs); e2e_serial = nothing overlapped"}


def single_process_leg(flash, L, devices, steps, n_points=10_000_000, n_configs=64):
    """The same workloads through ONE multi-device context (flash_options.n_devices: the form a single Julia process
    uses): K5 with the cloud handed over whole -- in slab order (contiguous index ranges are compact x-slabs) and
    shuffled (index ranges thin the cloud: the case a caller who shards naively hits) -- and K4's 8 frames x 16
    candidates with the frames round-robin over the devices."""
    import torch
    from flash._desc import Timing, c_double_p
    out = {"devices": list(devices)}
    m = scene_model(flash)
    m.set_options(devices=tuple(devices))
    ctx = m.ctx()
    nx = flash.num_states(m)
    X = np.ascontiguousarray(scene_states(flash, m, n_configs))
    cost = np.empty(n_configs); grad = np.empty((n_configs, nx))
    pX, pc, pg = X.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), grad.ctypes.data_as(c_double_p)
    m.check(L.flash_profiling(ctx, 1))
    tm = Timing()
    pin = torch.empty((n_points, 3), dtype=torch.float64, pin_memory=True)
    pts = pin.numpy()
    for order in ("slab_order", "shuffled"):
        pts[:] = scene_points(n_points)
        if order == "shuffled":
            pts[:] = pts[np.random.default_rng(9).permutation(n_points)]
        pp = pts.ctypes.data_as(c_double_p)
        m.check(L.flash_points_set(ctx, pp, n_points))
        for _ in range(3):
            m.check(L.flash_cost_grad(ctx, pX, n_configs, 10.0, pc, pg, None))
        ts, km = [], []
        for _ in range(steps):
            for d in set(devices):
                torch.cuda.synchronize(d)
            t0 = time.perf_counter()
            m.check(L.flash_cost_grad(ctx, pX, n_configs, 10.0, pc, pg, None))
            ts.append(time.perf_counter() - t0)
            L.flash_last_timing(ctx, C.byref(tm)); km.append(tm.ms_main)
        # end to end, pipelined as the N=1 leg: every device uploads and sorts its shard of the next frame while it evaluates the current one
        m.check(L.flash_points_set_async(ctx, pp, n_points))
        te = []
        for k in range(steps + 1):
            t0 = time.perf_counter()
            m.check(L.flash_points_swap(ctx))
            m.check(L.flash_points_set_async(ctx, pp, n_points))
            m.check(L.flash_cost_grad(ctx, pX, n_configs, 10.0, pc, pg, None))
            if k:
                te.append(time.perf_counter() - t0)
        m.check(L.flash_points_swap(ctx))
        out["k5_" + order] = {"value": n_points * n_configs / np.mean(ts), "unit": UNIT, "ms_per_step": 1e3 * float(np.mean(ts)), "kernel_ms_max_over_devices": float(np.mean(km)),
                              "e2e_value": n_points * n_configs / np.mean(te), "e2e_ms_per_step": 1e3 * float(np.mean(te)), "checksum_cost_sum": float(cost.sum())}
    m._points_owner = None
    m.invalidate()
    del pin
    # K4
    rng = np.random.default_rng(2)
    m4 = scene_model(flash)
    X4 = scene_states(flash, m4, 16)
    st4 = flash.ManipulatorState(m4)
    frames = []
    for k in range(8):
        x = X4[0].copy()
        x[:flash.num_positions(m4)] += 0.02 * rng.standard_normal(flash.num_positions(m4))
        flash.gradientdescent.unflatten_(st4, x)
        cam4 = flash.depthsensors.affine((0.3 + 0.05 * k, 1.8, 1.0), (1, 0, 0), np.pi / 2 + 0.25)
        p, _ = frame_cloud(flash, st4, cam4)
        frames.append(p)
    m4.set_options(devices=tuple(devices))
    r = _time_frames(flash, L, m4, frames, np.repeat(X4[None], 8, 0), steps, 0)
    r.pop("ray_hits_on_model", None)
    out["k4_combined_8_frames_x_16_configs"] = r
    m4.invalidate()
    return out


def tracking_leg(flash, L):
    """K3 (BASELINE.json configs[2]): squishable tracking, whole NaiveSolver loop on the device
    (rate 0.01, max_step 0.1, 50 iterations, divisors [1x7, 1e-3x36]: examples/squishable.ipynb cell 9)."""
    from flash._desc import SolverParams, c_double_p
    out = {}
    m = flash.models.squishable()
    rng = np.random.default_rng(3)
    true = flash.ManipulatorState(m)
    true.mechanism_state.q[4:7] = [0.05, -0.03, 0.04]
    true.deformation_data[:] = 0.03 * (rng.random(36) - 0.5)
    sensor = flash.depthsensors.Kinect(480, 640)
    cam = flash.depthsensors.affine((0, 0, 1.2), (1, 0, 0), np.pi)
    t0 = time.perf_counter()
    pts = flash.depthsensors.raycast(true, sensor, cam)
    t_ray = time.perf_counter() - t0
    clouds = {"k3_synthetic_640x480_raycast": pts}
    real = os.path.join(ROOT, "tests", "golden", "squishable_unsquished_xyz.npy")
    if os.path.exists(real):
        clouds["k3_real_kinect_cloud"] = np.load(real).astype(np.float64)
    div = np.r_[np.ones(7), 1e-3 * np.ones(36)]
    sp = SolverParams(0.01, 0.1, 50, 0.0, 10.0)
    for name, cloud in clouds.items():
        f = flash.gradientdescent.CostFunctor(m, cloud)
        x0 = flash.gradientdescent.flatten(flash.ManipulatorState(m))
        x0[4:7] = cloud.mean(0) if name.startswith("k3_real") else 0.0
        c0 = f(x0)
        times, ends = [], []
        for _ in range(4):
            X = x0.reshape(1, -1).copy(); cost = np.zeros(1)
            t0 = time.perf_counter()
            m.check(L.flash_estimate_state(m.ctx(), X.ctypes.data_as(c_double_p), 1, C.byref(sp), div.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), None))
            times.append(time.perf_counter() - t0)
            ends.append(float(cost[0]))
        t = min(times[1:])
        # end to end per frame: the frame's points from host memory (upload + Morton binning) + the whole solver loop
        # (the live loop: frame k+1 uploads on the copy stream while frame k is being solved)
        cl = np.ascontiguousarray(cloud)
        te = []
        m.check(L.flash_points_set_async(m.ctx(), cl.ctypes.data_as(c_double_p), len(cl)))
        for _ in range(6):
            X = x0.reshape(1, -1).copy()
            t0 = time.perf_counter()
            m.check(L.flash_points_swap(m.ctx()))
            m.check(L.flash_points_set_async(m.ctx(), cl.ctypes.data_as(c_double_p), len(cl)))
            m.check(L.flash_estimate_state(m.ctx(), X.ctypes.data_as(c_double_p), 1, C.byref(sp), div.ctypes.data_as(c_double_p), cost.ctypes.data_as(c_double_p), None))
            te.append(time.perf_counter() - t0)
        m.check(L.flash_points_swap(m.ctx()))
        m._points_owner = None
        out[name] = {"points": int(len(cloud)), "iterations": 50, "ms_per_frame": 1e3 * t, "frames_per_sec": 1.0 / t,
                     "frames_per_sec_with_upload": 1.0 / min(te[1:]),
                     "evals_per_sec": len(cloud) * 51 / t, "cost_start": float(c0), "cost_end": float(ends[-1]), "cost_end_of_each_repeat": ends}
    out["raycast_640x480"] = {"rays": 480 * 640, "hits": int(len(pts)), "ms": 1e3 * t_ray, "note": "first call, includes allocation and H2D/D2H of rays, depths and hits"}
    out["reference_published"] = "8.083 s per estimate_state on the real cloud (examples/squishable.ipynb:328; Julia 0.5, 1 thread, with drawing callback)"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="exact", choices=["exact", "grid"])
    ap.add_argument("--points", type=int, default=10_000_000)
    ap.add_argument("--configs", type=int, default=64)
    ap.add_argument("--grid-spacing", type=float, default=0.004)
    ap.add_argument("--shard", default="blobs", choices=["blobs", "slabs", "index"], help="how the global cloud is dealt to the ranks (scene_points)")
    ap.add_argument("--ref-points", type=int, default=100_000)
    ap.add_argument("--ref-configs", type=int, default=64)
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    ap.add_argument("--no-parity", dest="parity", action="store_false", help="skip the oracle check of config 0 after the timed loops")
    ap.add_argument("--no-extras", dest="extras", action="store_false", help="skip the grid-mode and K3 tracking legs (N=1 only)")
    ap.add_argument("--no-grid", dest="grid_leg", action="store_false", help="skip the grid-mode leg")
    ap.add_argument("--no-tracking", dest="tracking_leg", action="store_false", help="skip the K3 tracking leg")
    ap.add_argument("--no-configs", dest="configs_leg", action="store_false", help="skip the K2 / K4 legs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it so that there is one process per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus), "--master-addr", "127.0.0.1",
               "--master-port", str(29500 + os.getpid() % 1000)] + sys.argv
        sys.exit(subprocess.call(cmd))
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
