#!/bin/bash
# Profiling recipe (run under gpurun, ONE GPU).  Writes into gpurun_out/; profiles/summarize.py turns that into the
# tracked summaries under profiles/ afterwards.  Usage: bash profiles/run_ncu.sh <round tag>
# Every command first runs WITHOUT ncu (exit code checked), then under it; no number printed under ncu is a bench value.
TAG=${1:-r02}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-ids-check"
mkdir -p gpurun_out
# a request = 1 prefill (~150 launches, replayed from a CUDA graph) + ONE persistent decode launch
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 450 -c 310 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:decode_mega2 -s 3 -c 1 \
    -o gpurun_out/prof_mega_$TAG -f $CMD > gpurun_out/ncu_mega_$TAG.log 2>&1
# every stand-alone kernel at a realistic size, through the C ABI
PCMD="python profiles/kernel_probe.py"
$PCMD > gpurun_out/plain_probe_$TAG.log 2>&1 &&
for K in lmhead_argmax_kernel gemv_kernel layernorm_warp_kernel attn_decode64_kernel attn_prefill64_mma_kernel sample_topk_kernel gemm_tc_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 \
      -o gpurun_out/prof_${K}_$TAG -f $PCMD > gpurun_out/ncu_${K}_$TAG.log 2>&1
done
# the batch-32 decode step (tcgen05 GEMMs + split-K + attention kernels): launch list of the probe's last op
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_probe_$TAG.csv \
    $PCMD > gpurun_out/ncu_launches_probe_$TAG.log 2>&1
tail -2 gpurun_out/plain_$TAG.log | cut -c1-300; cat gpurun_out/plain_probe_$TAG.log
