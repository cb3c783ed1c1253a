#!/bin/bash
# Run under gpurun (one GPU).  Plain run first; ncu only if it exits 0 (B200_PROFILING.md recipe).
# Outputs land in gpurun_out/; copy the summaries you want judged into profiles/.
set -u
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:csmain_accumulate -s 3 -c 2 \
    -f -o gpurun_out/prof_accumulate $CMD > gpurun_out/ncu_full.log 2>&1
tail -q -n 2 gpurun_out/ncu_launches.log gpurun_out/ncu_full.log
