#!/bin/bash
# Profiling recipe (run under gpurun, one GPU).  Writes into gpurun_out/; summaries are made
# by profiles/summarize.py here afterwards.  Usage: bash profiles/run_ncu.sh <round tag>
TAG=${1:-r01}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
mkdir -p gpurun_out
# a request = 1 prefill (~150 launches) + ONE persistent decode launch; 3 warm-up + 2 measured requests
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 450 -c 310 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:decode_mega -s 3 -c 1 \
    -o gpurun_out/prof_mega_$TAG -f $CMD > gpurun_out/ncu_mega_$TAG.log 2>&1
# the tcgen05 GEMM at a prefill-sized shape (16384 x 3072 x 768 and 8192^3), through the C ABI
GCMD="python profiles/gemm_probe.py"
$GCMD > gpurun_out/plain_gemm_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 6 -c 2 \
    -o gpurun_out/prof_gemmtc_$TAG -f $GCMD > gpurun_out/ncu_gemmtc_$TAG.log 2>&1
tail -2 gpurun_out/plain_$TAG.log; cat gpurun_out/plain_gemm_$TAG.log
