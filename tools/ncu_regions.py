"""Roll up an .ncu-rep source page by source REGION for the first kernel.

Regions are derived from the sources themselves: a region starts at every function definition in
csrc/*.h / *.cuh and at every comment of the form `// [region: name]`, and runs to the next start.
usage: python tools/ncu_regions.py report.ncu-rep
"""
import collections
import csv
import os
import re
import subprocess
import sys

rep = sys.argv[1]
CSRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "point-cloud-signed-distance_b200", "csrc")
FUNC = re.compile(r"^(?:template\s*<[^>]*>\s*)?(?:FLASH_HDN?|__device__|__global__|__host__|inline|static)\b[^;=]*?\b([A-Za-z_]\w*)\s*\([^;]*$")
MARK = re.compile(r"//\s*\[region:\s*([^\]]+)\]")


def regions_of(path):
    starts = []  # (line, name)
    fn = "?"
    for i, line in enumerate(open(path), 1):
        m = MARK.search(line)
        if m:
            starts.append((i, "%s: %s" % (fn, m.group(1).strip())))
            continue
        m = FUNC.match(line)
        if m and not line.strip().endswith(";"):
            fn = m.group(1)
            starts.append((i, fn))
    return starts


REG = {}
for f in os.listdir(CSRC):
    if f.endswith((".h", ".cuh", ".cu")):
        REG[f] = regions_of(os.path.join(CSRC, f))


def region(fname, line):
    name = "%s (top)" % fname
    for lo, nm in REG.get(fname, []):
        if lo <= line:
            name = nm
        else:
            break
    return name


out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None
hdr = None
agg = collections.defaultdict(lambda: [0, 0, 0])
seen = set()
for r in rows:
    if len(r) >= 2 and r[0] in ("File Path", "File Name"):
        cur = r[1].split("/")[-1]
        hdr = None
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or not r or not r[0].isdigit():
        continue
    key = (cur, int(r[0]))
    if key in seen:
        continue
    seen.add(key)
    try:
        s = int(r[hdr.index("# Samples")] or 0)
        n = int(r[hdr.index("Instructions Executed")] or 0)
        t = int(r[hdr.index("Thread Instructions Executed")] or 0)
    except (ValueError, IndexError):
        continue
    a = agg[region(cur, key[1]) if cur in REG else "%s (other)" % cur]
    a[0] += s
    a[1] += n
    a[2] += t
ts = sum(a[0] for a in agg.values()) or 1
tn = sum(a[1] for a in agg.values()) or 1
print("region roll-up: total samples %d, warp instructions %d" % (ts, tn))
for name, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    if a[1] == 0 and a[0] == 0:
        continue
    print("%-46s samples %5.1f%%  inst %5.1f%%  thr/inst %4.1f" % (name, 100.0 * a[0] / ts, 100.0 * a[1] / tn, a[2] / max(a[1], 1)))
