#!/bin/bash
# Round-2 captures.  Each step is ONE gpurun call on one GPU (step "peer": two GPUs): the plain command first, ncu only after it
# exited 0 (B200_PROFILING.md).  Outputs land in gpurun_out/; profiles/summarize_r02.py turns them into the tracked summaries.
#   profiles/run_ncu_r02.sh launches|accumulate|literal|c3|c4|peer
set -u
mkdir -p gpurun_out
BENCH="python bench.py --steps 5 --warmup 3 --no-extra --no-cpu-baseline"
case "$1" in
launches)
    $BENCH > gpurun_out/r02_plain.log 2>&1 &&
    ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_c2.csv \
        $BENCH > gpurun_out/r02_ncu_launches.log 2>&1 ;;
accumulate)
    $BENCH > gpurun_out/r02_plain.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:csmain_accumulate -s 3 -c 2 -f -o gpurun_out/r02_c2_accumulate \
        $BENCH > gpurun_out/r02_ncu_accumulate.log 2>&1 ;;
literal)
    python experiments/prof_c2.py > gpurun_out/r02_plain.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -c 40 -f -o gpurun_out/r02_c2_literal \
        python experiments/prof_c2.py > gpurun_out/r02_ncu_literal.log 2>&1 ;;
c3|c4)
    PW=4; [ "$1" = c4 ] && PW=1
    python experiments/prof_scene.py $1 $PW > gpurun_out/r02_plain_$1.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:bounce_kernel -c 40 -f -o gpurun_out/r02_$1_bounce \
        python experiments/prof_scene.py $1 $PW > gpurun_out/r02_ncu_$1.log 2>&1 ;;
peer)
    python experiments/prof_peer.py > gpurun_out/r02_plain_peer.log 2>&1 && PROF_PEER_DIRECT=1 python experiments/prof_peer.py >> gpurun_out/r02_plain_peer.log 2>&1 &&
    PROF_PEER_DIRECT=1 ncu --set full --metrics nvlrx__bytes.sum,nvltx__bytes.sum --clock-control none -k regex:peer_merge -c 16 -f -o gpurun_out/r02_peer_merge_g$(nvidia-smi -L | wc -l) \
        python experiments/prof_peer.py > gpurun_out/r02_ncu_peer.log 2>&1 ;;
esac
echo "step $1 rc=$?"; tail -q -n 3 gpurun_out/r02_plain*.log gpurun_out/r02_ncu_*.log 2>/dev/null | tail -12
