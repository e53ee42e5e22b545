#!/bin/bash
# usage: cfg5_multi.sh <S> "<B_per_stream:P> ..."   -> appends JSON lines to gpurun_out/r02_cfg5_s$S.jsonl
S=$1; shift
OUT=gpurun_out/r02_cfg5_s$S.jsonl; : > $OUT
port=29800
for spec in $@; do
  IFS=: read B P <<< "$spec"
  port=$((port+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $S --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $S --workload custom:gpt2:$B:$P:64 --steps 2 --warmup 3 --no-cpu-baseline 2>gpurun_out/cfg5_s${S}_$B_$P.err | tail -1 >> $OUT
done
python - <<PY
import json
for l in open("$OUT"):
    l=l.strip()
    if not l.startswith("{"): continue
    d=json.loads(l); c=d["config"]; print(c["stages"], c["streams"]*c["batch_per_stream"], c["prompt"], d["value"], d["p50_token_ms"], d["ids_check"]["mismatch"] if d.get("ids_check") else None)
PY
