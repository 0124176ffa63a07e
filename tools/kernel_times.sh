#!/bin/bash
# Runs ON THE GPU BOX: per-kernel times (ncu, serialised) of the K5 bench for each variant library given
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for so in "$@"; do
  if [ "$so" = "lib" ]; then libp=$PWD/point-cloud-signed-distance_b200/lib/libflashsdf.so; else libp=$PWD/point-cloud-signed-distance_b200/lib/variants/$so.so; fi
  FLASHSDF_LIB=$libp ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'cost_kernel|group_kernel' -c 16 --csv --log-file gpurun_out/kt_$so.csv python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-parity > /dev/null 2>&1
  python - $so <<Q
import csv,sys,collections
rows=[r for r in csv.reader(l for l in open("gpurun_out/kt_%s.csv"%sys.argv[1]) if not l.startswith("=="))]
h=rows[0]; ki,vi=h.index("Kernel Name"),h.index("Metric Value")
t=collections.OrderedDict()
for r in rows[1:]:
    if len(r)>vi: t.setdefault(r[ki][:40],[]).append(float(r[vi].replace(",",""))/1e6)
print(sys.argv[1], {k:"%.3f ms (n=%d)"%(sum(v[2:])/max(len(v[2:]),1),len(v)) for k,v in t.items()})
Q
done
