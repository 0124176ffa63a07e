#!/bin/bash
# Runs ON THE GPU BOX: GPU tests + one bench line with the K2/K4/K3 legs (gpurun -- tools/gpu_quick.sh TAG [pytest-args])
tag=${1:-q}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s "$@" > gpurun_out/${tag}_pytest.log 2>&1; tail -15 gpurun_out/${tag}_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || tail -20 gpurun_out/${tag}_bench.err
python - <<P
import json
d=json.loads(open("gpurun_out/${tag}_bench.json").read().strip().splitlines()[-1])
print("K5 value %.4g ms %.3f kernel %.3f e2e %.4g parity %s"%(d["value"],d["ms_per_step"],d["roofline"]["kernel_ms"],d["e2e"]["value"],d.get("parity")))
for k,v in (d.get("other_configs") or {}).items(): print(k, "device %.4g e2e %.4g dev_ms %.3f"%(v["device_evals_per_sec"],v["e2e_evals_per_sec"],v["device_ms_per_batch"]))
for k,v in (d.get("tracking") or {}).items():
    if isinstance(v,dict) and "frames_per_sec" in v: print(k, "fps %.1f with upload %.1f"%(v["frames_per_sec"],v["frames_per_sec_with_upload"]), v["cost_end_of_each_repeat"])
print("grid", (d.get("grid_mode") or {}).get("kernel_ms"))
P
