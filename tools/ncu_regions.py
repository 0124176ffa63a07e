"""Roll up an .ncu-rep source page by source REGION (function line ranges) for the first kernel."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
# regions: (file suffix, lo, hi, name)
REG = [("flash_dev.h", 110, 136, "algebra helpers (inlined: F3/D3 ops)"), ("flash_dev.h", 288, 296, "seg_closest0"), ("flash_dev.h", 297, 324, "tri_closest0"),
       ("flash_dev.h", 326, 339, "simplex bookkeeping"), ("flash_dev.h", 347, 372, "hull_sd: plane scan"), ("flash_dev.h", 373, 386, "hull_sd: start triangle"),
       ("flash_dev.h", 387, 411, "hull_sd: GJK support scan"), ("flash_dev.h", 412, 443, "hull_sd: GJK simplex update + tail"), ("flash_dev.h", 446, 453, "aabb_lower_bound"),
       ("flash_dev.h", 456, 467, "to_geometry (fp64/fp32)"), ("flash_dev.h", 470, 491, "grid_sd"), ("flash_dev.h", 497, 507, "dsqrt_fast"), ("flash_dev.h", 508, 527, "skin_eval"),
       ("flash_dev.h", 532, 553, "skin_adjoint_term"), ("flash_dev.h", 571, 600, "eval_point: skins + culling loop"), ("flash_dev.h", 601, 616, "eval_point: narrow-phase loop"),
       ("flash_kernels.cuh", 69, 78, "warp_sum"), ("flash_kernels.cuh", 109, 132, "stage_block"), ("flash_kernels.cuh", 140, 167, "cost_kernel: prologue"),
       ("flash_kernels.cuh", 168, 190, "cost_kernel: point load + cost"), ("flash_kernels.cuh", 191, 210, "cost_kernel: convex wrench reduce"),
       ("flash_kernels.cuh", 212, 246, "cost_kernel: skin adjoint queue"), ("flash_kernels.cuh", 247, 260, "cost_kernel: block reduce + atomics")]
rows = list(csv.reader(out.splitlines()))
cur = None; hdr = None
agg = collections.defaultdict(lambda: [0, 0, 0]); seen = set()
for r in rows:
    if len(r) >= 2 and r[0] in ("File Path", "File Name"): cur = r[1].split("/")[-1]; hdr = None; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or not r or not r[0].isdigit(): continue
    key = (cur, int(r[0]))
    if key in seen: continue
    seen.add(key)
    try:
        s = int(r[hdr.index("# Samples")] or 0); n = int(r[hdr.index("Instructions Executed")] or 0); t = int(r[hdr.index("Thread Instructions Executed")] or 0)
    except (ValueError, IndexError):
        continue
    name = "%s (other)" % cur
    for f, lo, hi, nm in REG:
        if cur == f and lo <= key[1] <= hi: name = nm; break
    a = agg[name]; a[0] += s; a[1] += n; a[2] += t
ts = sum(a[0] for a in agg.values()) or 1; tn = sum(a[1] for a in agg.values()) or 1
print("region roll-up: total samples %d, warp instructions %d" % (ts, tn))
for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if a[1] == 0 and a[0] == 0: continue
    print("%-42s inst %5.1f%%  samples %5.1f%%  thr/inst %4.1f" % (name, 100.0 * a[1] / tn, 100.0 * a[0] / ts, a[2] / max(a[1], 1)))
