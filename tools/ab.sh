#!/bin/bash
# Runs ON THE GPU BOX: A/B of kernel variants.  usage: tools/ab.sh TAG spec...   spec = variant[,ENV=VAL...]
# (variant = a library under lib/variants/, or "lib" for lib/libflashsdf.so); prints K5 value / kernel ms / checksum and
# the K2 / K4 device times of each.
cd "$(dirname "$0")/.."
tag=$1; shift
mkdir -p gpurun_out
for spec in "$@"; do
  IFS=',' read -r -a parts <<< "$spec"
  so=${parts[0]}
  envs=("${parts[@]:1}")
  if [ "$so" = "lib" ]; then libp=$PWD/point-cloud-signed-distance_b200/lib/libflashsdf.so; else libp=$PWD/point-cloud-signed-distance_b200/lib/variants/$so.so; fi
  out=gpurun_out/${tag}_$(echo "$spec" | tr ',=/' '___').json
  env FLASHSDF_LIB=$libp "${envs[@]}" python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-parity ${AB_ARGS:---no-tracking --no-grid} > $out 2> $out.err || { echo "$spec FAILED"; tail -5 $out.err; continue; }
  python - "$spec" $out <<P
import json,sys
d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
s="%-40s K5 %.4g ev/s main %.2f ms e2e %.4g chk %.13g" % (sys.argv[1], d["value"], d["roofline"]["kernel_ms"], d["e2e"]["value"], d["checksum_cost_sum"])
for k,v in (d.get("other_configs") or {}).items(): s += " | %s dev %.3f ms e2e %.3f ms" % (k[:2], v["device_ms_per_batch"], v["e2e_ms_per_batch"])
for k,v in (d.get("tracking") or {}).items():
    if isinstance(v,dict) and "frames_per_sec" in v: s += " | %s %.0f fps (%.0f up)" % (k[:7], v["frames_per_sec"], v["frames_per_sec_with_upload"])
print(s)
P
done
