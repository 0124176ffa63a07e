"""Opcode histogram of the first kernel in an .ncu-rep (warp instructions executed, stall samples)."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None; nk = 0
ops = collections.Counter(); samp = collections.Counter(); thr = collections.Counter()
for r in rows:
    if r and r[0] == "Kernel Name":
        nk += 1
        if nk > 1: break
        continue
    if r and r[0] == "Address": hdr = r; continue
    if hdr is None or len(r) < 8 or not r[0].startswith("0x"): continue
    toks = r[1].split()
    if toks and toks[0].startswith("@"): toks = toks[1:]
    op = toks[0].split(".")[0] if toks else "?"
    try:
        n = int(r[hdr.index("Instructions Executed")]); s = int(r[hdr.index("# Samples")]); t = int(r[hdr.index("Thread Instructions Executed")])
    except ValueError:
        continue
    ops[op] += n; samp[op] += s; thr[op] += t
tot = sum(ops.values()); ts = sum(samp.values())
print("total warp instructions %d, samples %d" % (tot, ts))
for op, n in ops.most_common(40):
    print("%-10s inst %5.1f%%  samples %5.1f%%  thr/inst %4.1f" % (op, 100.0 * n / tot, 100.0 * samp[op] / ts, thr[op] / max(n, 1)))
