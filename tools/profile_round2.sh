#!/bin/bash
# Runs ON THE GPU BOX (gpurun -- tools/profile_round2.sh TAG [pytest]): GPU tests, plain bench, ncu launch list, and one
# `ncu --set full` capture of cost_kernel for K5 (exact), K4 (depth frames) and K5 in grid mode.
set -x
tag=${1:-r02}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${tag}_smi.txt
if [ "$2" = "pytest" ]; then python -m pytest tests -m gpu -x -q -s > gpurun_out/${tag}_pytest.log 2>&1; tail -5 gpurun_out/${tag}_pytest.log; fi
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || { tail -20 gpurun_out/${tag}_bench.err; exit 1; }
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-parity"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 1 -c 1 -o gpurun_out/${tag}_cost $CMD > gpurun_out/${tag}_ncu2.log 2>&1
tail -2 gpurun_out/${tag}_ncu2.log
# K4: tools/configs_bench.py launches cost_kernel 19 times for K2 (5 resident + 3 serial + 11 pipelined), then the K4 batches
python tools/configs_bench.py > gpurun_out/${tag}_configs.json 2> gpurun_out/${tag}_configs.err
ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 21 -c 1 -o gpurun_out/${tag}_cost_k4 python tools/configs_bench.py > gpurun_out/${tag}_ncu3.log 2>&1
tail -2 gpurun_out/${tag}_ncu3.log
ncu --set full --clock-control none --import-source on -k regex:cost_kernel -s 1 -c 1 -o gpurun_out/${tag}_cost_grid $CMD --mode grid > gpurun_out/${tag}_ncu4.log 2>&1
tail -2 gpurun_out/${tag}_ncu4.log
ls -la gpurun_out | tail -20
