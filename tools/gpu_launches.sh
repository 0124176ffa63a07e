#!/bin/bash
# Runs ON THE GPU BOX: per-kernel time list (ncu, serialised) for the K5 bench and the K2/K4 legs
tag=${1:-l}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-parity"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${tag}_launches_cfg.csv python tools/configs_bench.py > gpurun_out/${tag}_ncu2.log 2>&1
python - <<P
import csv, collections
for name in ("launches", "launches_cfg"):
    rows=[r for r in csv.reader(l for l in open("gpurun_out/${tag}_%s.csv"%name) if not l.startswith("=="))]
    hdr=rows[0]; ki,vi=hdr.index("Kernel Name"),hdr.index("Metric Value")
    tot=collections.OrderedDict()
    for r in rows[1:]:
        if len(r)<=vi: continue
        k=tot.setdefault(r[ki][:70],[0,0.0,[]]); k[0]+=1; v=float(r[vi].replace(",","")); k[1]+=v; k[2].append(v)
    print("==",name)
    for k,(n,t,l) in tot.items(): print("%-72s n=%4d total %10.1f us  last %9.1f us"%(k,n,t/1e3,l[-1]/1e3))
P
