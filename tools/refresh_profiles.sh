#!/bin/bash
# Runs HERE after `gpurun -- tools/profile_round.sh TAG`: turns gpurun_out/TAG_* into the tracked profiles/r01_* files.
set -e
tag=${1:?tag}
cd "$(dirname "$0")/.."
python tools/ncu_summary.py gpurun_out/${tag}_cost.ncu-rep 40 > profiles/r01_cost_kernel_ncu_full.txt
python tools/ncu_regions.py gpurun_out/${tag}_cost.ncu-rep > profiles/r01_cost_kernel_regions.txt
python tools/ncu_opcodes.py gpurun_out/${tag}_cost.ncu-rep > profiles/r01_cost_kernel_opcodes.txt
cp gpurun_out/${tag}_launches.csv profiles/r01_launches.csv
python - "$tag" <<'P'
import csv, collections, json, re, sys
tag = sys.argv[1]
rows = [r for r in csv.reader(l for l in open("gpurun_out/%s_launches.csv" % tag) if not l.startswith("=="))]
hdr = rows[0]; ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
tot = collections.OrderedDict()
for r in rows[1:]:
    if len(r) <= vi: continue
    k = tot.setdefault(r[ki], [0, 0.0]); k[0] += 1; k[1] += float(r[vi].replace(",", ""))
s = sum(v[1] for v in tot.values())
with open("profiles/r01_launches_summary.txt", "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none : python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline (round 1, B200)\n# kernel | launches | total ns | share of GPU time\n")
    for k, (n, t) in tot.items():
        f.write("%-78s %8d %14.0f %7.2f%%\n" % (k[:78], n, t, 100 * t / s))
txt = open("profiles/r01_cost_kernel_ncu_full.txt").read()
def g(name):
    m = re.search(re.escape(name) + r"\s+([0-9.,]+)\s*(\S*)", txt)
    v = float(m.group(1).replace(",", "")); u = m.group(2)
    return v * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}.get(u, 1)
old = json.load(open("profiles/r01_ncu_constants.json"))
old.update({"dram_bytes_read": int(g("dram__bytes_read.sum")), "dram_bytes_write": int(g("dram__bytes_write.sum")),
            "warp_instructions": int(g("smsp__inst_executed.sum")), "issue_active_pct": round(g("smsp__issue_active.avg.pct_of_peak_sustained_active"), 2),
            "kernel_ms_under_ncu": round(g("gpu__time_duration.sum"), 2)})
try:
    old["lts_bytes"] = int(g("lts__t_bytes.sum"))
except Exception:
    pass
json.dump(old, open("profiles/r01_ncu_constants.json", "w"), indent=1)
print(old)
P
tail -1 gpurun_out/${tag}_plain.json > profiles/r01_bench_n1_profiled_cmd.json
