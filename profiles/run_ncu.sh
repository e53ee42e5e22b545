#!/bin/bash
# Profiling recipe (run under gpurun, one GPU).  Writes into gpurun_out/; summaries are made
# by profiles/summarize.py here afterwards.  Usage: bash profiles/run_ncu.sh <round tag>
TAG=${1:-r01}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 24000 -c 800 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemv_kernel -s 20000 -c 4 \
    -o gpurun_out/prof_gemv_$TAG -f $CMD > gpurun_out/ncu_gemv_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lmhead_argmax -s 400 -c 1 \
    -o gpurun_out/prof_lmhead_$TAG -f $CMD > gpurun_out/ncu_lmhead_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_decode -s 5000 -c 1 \
    -o gpurun_out/prof_attn_$TAG -f $CMD > gpurun_out/ncu_attn_$TAG.log 2>&1
tail -2 gpurun_out/plain_$TAG.log
