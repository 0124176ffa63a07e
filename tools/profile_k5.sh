#!/bin/bash
# Runs ON THE GPU BOX (gpurun -- tools/profile_k5.sh TAG): plain K5 bench line, then one `ncu --set full` capture of cost_kernel (K5)
tag=${1:-p}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-parity"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { tail -5 gpurun_out/${tag}_plain.err; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 1 -c 1 -o gpurun_out/${tag}_cost $CMD > gpurun_out/${tag}_ncu.log 2>&1
tail -2 gpurun_out/${tag}_ncu.log
