"""Per-source-line instruction shares of one file in an .ncu-rep, in line order: python tools/ncu_lines.py rep file.cuh [min_pct]"""
import csv, subprocess, sys
rep, fname = sys.argv[1], sys.argv[2]
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.15
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, stats = None, []
for r in csv.reader(src.splitlines()):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 8 and r[0].isdigit():
        try:
            stats.append((cur, int(r[0]), r[1].strip()[:110], int(r[6] or 0), int(r[7] or 0), int(r[8] or 0)))
        except ValueError:
            pass
toti = sum(s[4] for s in stats) or 1
tots = sum(s[3] for s in stats) or 1
for s in stats:
    if s[0] == fname and 100 * s[4] / toti >= minp:
        print("%4d inst %5.2f%% samp %5.2f%% thr %4.1f | %s" % (s[1], 100 * s[4] / toti, 100 * s[3] / tots, s[5] / max(s[4], 1), s[2]))
