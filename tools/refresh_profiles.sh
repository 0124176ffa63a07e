#!/bin/bash
# Runs HERE after `gpurun -- tools/profile_round2.sh TAG`: turns gpurun_out/TAG_* into the tracked profiles/r02_* files
# (summaries of the ncu captures of the committed kernels; bench.py reads profiles/r02_ncu_constants.json).
set -e
tag=${1:?tag}
R=${2:-r02}
cd "$(dirname "$0")/.."
for leg in cost:k5 cost_k4:k4 cost_grid:grid; do
  f=${leg%%:*}; name=${leg##*:}
  [ -f gpurun_out/${tag}_$f.ncu-rep ] || continue
  python tools/ncu_summary.py gpurun_out/${tag}_$f.ncu-rep 40 > profiles/${R}_${name}_cost_kernel_ncu_full.txt
  python tools/ncu_regions.py gpurun_out/${tag}_$f.ncu-rep > profiles/${R}_${name}_cost_kernel_regions.txt
  python tools/ncu_opcodes.py gpurun_out/${tag}_$f.ncu-rep > profiles/${R}_${name}_cost_kernel_opcodes.txt
done
cp gpurun_out/${tag}_launches.csv profiles/${R}_launches.csv
tail -1 gpurun_out/${tag}_plain.json > profiles/${R}_bench_n1_profiled_cmd.json
[ -f gpurun_out/${tag}_bench.json ] && tail -1 gpurun_out/${tag}_bench.json > profiles/${R}_bench_n1.json
python - "$tag" "$R" <<'P'
import csv, collections, hashlib, json, os, re, subprocess, sys
tag, R = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(l for l in open("gpurun_out/%s_launches.csv" % tag) if not l.startswith("=="))]
hdr = rows[0]; ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi: continue
    k = tot.setdefault(r[ki], [0, 0.0]); k[0] += 1; k[1] += float(r[vi].replace(",", ""))
s = sum(v[1] for v in tot.values())
with open("profiles/%s_launches_summary.txt" % R, "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none : python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-parity (B200)\n# kernel | launches | total ns | share of GPU time\n")
    for k, (n, t) in tot.items():
        f.write("%-78s %8d %14.0f %7.2f%%\n" % (k[:78], n, t, 100 * t / s))
def consts(full, opc):
    txt = open(full).read()
    def g(name):
        m = re.search(re.escape(name) + r"\s+([0-9.,]+)\s*(\S*)", txt)
        v = float(m.group(1).replace(",", "")); u = m.group(2)
        return v * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}.get(u, 1)
    out = {"dram_bytes_read": int(g("dram__bytes_read.sum")), "dram_bytes_write": int(g("dram__bytes_write.sum")),
           "warp_instructions": int(g("smsp__inst_executed.sum")), "issue_active_pct": round(g("smsp__issue_active.avg.pct_of_peak_sustained_active"), 2),
           "kernel_ms_under_ncu": round(g("gpu__time_duration.sum"), 3), "lanes_active": round(g("smsp__thread_inst_executed_per_inst_executed.ratio"), 2),
           "warps_active_pct": round(g("sm__warps_active.avg.pct_of_peak_sustained_active"), 1), "registers": int(g("launch__registers_per_thread"))}
    try:                                                  # L2 traffic: sectors of 32 B (lts__t_bytes is not in every ncu section set)
        out["lts_bytes"] = int(g("lts__t_sectors.sum")) * 32
        out["l2_hit_pct"] = round(g("lts__t_sector_hit_rate.pct"), 1)
        out["l1_hit_pct"] = round(g("l1tex__t_sector_hit_rate.pct"), 1)
    except Exception:
        pass
    # executed floating-point work from the opcode roll-up (thread-level: warp instructions x active lanes)
    ops = {}
    total = 0
    for line in open(opc):
        m = re.match(r"total warp instructions (\d+)", line)
        if m: total = int(m.group(1)); continue
        m = re.match(r"(\S+)\s+inst\s+([0-9.]+)%\s+samples\s+[0-9.]+%\s+thr/inst\s+([0-9.]+)", line)
        if m: ops[m.group(1)] = float(m.group(2)) / 100 * total * float(m.group(3))
    scale = out["warp_instructions"] / max(total, 1)     # the source page counts predicated-off slots too: normalise to the raw counter
    out["fp32_flops"] = int(scale * (2 * ops.get("FFMA", 0) + 4 * ops.get("FFMA2", 0) + ops.get("FADD", 0) + ops.get("FMUL", 0) + 2 * ops.get("FADD2", 0) + 2 * ops.get("FMUL2", 0)))
    out["fp64_flops"] = int(scale * (2 * ops.get("DFMA", 0) + ops.get("DADD", 0) + ops.get("DMUL", 0)))
    fp_share = sum(ops.get(k, 0) for k in ("FFMA", "FFMA2", "FADD", "FMUL", "FADD2", "FMUL2", "DFMA", "DADD", "DMUL", "MUFU")) / max(sum(ops.values()), 1)
    out["fp_instruction_share"] = round(fp_share, 3)
    return out
sha = hashlib.sha1()
for fn in ("flash_dev.h", "flash_kernels.cuh"):      # the device code of cost_kernel (bench.py recomputes this hash and reports whether it still matches)
    sha.update(open(os.path.join("point-cloud-signed-distance_b200/csrc", fn), "rb").read())
spill = 0
try:
    log = open("point-cloud-signed-distance_b200/lib/ptxas.log").read()
    m = re.search(r"cost_kernelILb1ELb0ELb0ELb0E.*?\n.*?\n\s*(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", log)
    if m: spill = int(m.group(2)) + int(m.group(3))
except Exception:
    pass
plain = json.loads(open("profiles/%s_bench_n1_profiled_cmd.json" % R).read())
k5 = consts("profiles/%s_k5_cost_kernel_ncu_full.txt" % R, "profiles/%s_k5_cost_kernel_opcodes.txt" % R)
k5.update({"points": plain["config"]["points"], "configs": plain["config"]["configs"], "kernel_sha": sha.hexdigest()[:12], "spill_bytes": spill,
           "kernel": "cost_kernel<GRAD> (K5, exact mode)", "captured_with": "ncu --set full --clock-control none, tools/profile_round2.sh " + tag})
for name in ("k4", "grid"):
    f = "profiles/%s_%s_cost_kernel_ncu_full.txt" % (R, name)
    if os.path.exists(f):
        k5[name] = consts(f, "profiles/%s_%s_cost_kernel_opcodes.txt" % (R, name))
json.dump(k5, open("profiles/%s_ncu_constants.json" % R, "w"), indent=1)
print(json.dumps(k5, indent=1))
P
