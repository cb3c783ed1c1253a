#!/bin/bash
# Scene-path profiles (run under gpurun, one GPU).  Plain run first; ncu only if it exits 0.
set -u
WL=${1:-c3}; SPP=${2:-4}
CMD="python bench.py --workload $WL --steps 2 --warmup 3 --spp $SPP --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_$WL.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/launches_$WL.csv $CMD > gpurun_out/ncu_launches_$WL.log 2>&1
$CMD > gpurun_out/plain2_$WL.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bounce_kernel -s 21 -c 7 \
    -f -o gpurun_out/prof_bounce_$WL $CMD > gpurun_out/ncu_full_$WL.log 2>&1
tail -q -n 2 gpurun_out/ncu_launches_$WL.log gpurun_out/ncu_full_$WL.log
