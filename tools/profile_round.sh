#!/bin/bash
# Runs ON THE GPU BOX (gpurun -- tools/profile_round.sh TAG): plain bench, ncu launch list, one full capture of cost_kernel.
set -x
tag=${1:-r01}
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 1 -c 1 -o gpurun_out/${tag}_cost $CMD > gpurun_out/${tag}_ncu2.log 2>&1
tail -2 gpurun_out/${tag}_ncu2.log
