"""One pod, N GPUs: the coordinator API of the reference (`POST /generate`, server.py:154-210) in front of an N-stage
device-resident pipeline.  Launched with one process per GPU:

    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 -m llm_sharding_demo_b200.pod

Rank 0 serves HTTP on SHARD_PORT (default 5000, server.py:25) and owns stage 0; the other ranks own the later stages and
wait for work.  A request is broadcast to every rank (token ids and max_new_tokens, a few bytes over torch.distributed),
each rank prefills its stage and then runs ONE persistent kernel for the whole decode; hidden states and tokens travel
GPU to GPU over NVLink (gp_pipeline_decode).  This replaces the reference's three Deployments + two Services
(k8s/*.yaml: coordinator, shard A, shard B talking HTTP/JSON) by one Deployment whose pod holds every stage, so the
SPLIT_AT of the stages cannot disagree (the reference ships SPLIT_AT=2 on shard A and 1 on shard B).

Env (same names as the reference where they exist): MODEL_ID, SHARD_PORT; GP_WEIGHTS / GP_RANDOM_INIT / GP_SEED / GP_DTYPE
as in `server.py`; GP_MAX_CTX (KV positions per request, default n_positions); GP_SAMPLING=reference draws every token
with the reference's temperature 0.6 / top-k 40 sampler on the last stage (seed GP_SAMPLE_SEED + request number).
`GET /healthz` answers 200 once every stage has its weights (readiness probe)."""
from __future__ import annotations

import os
import threading

import torch
import torch.distributed as dist

from . import weights as W
from ._lib import GP_BF16, GP_F32
from .pipeline import PipelineRunner
from .server import (GenerateReq, ByteTokenizer, REF_TEMPERATURE, REF_TOP_K, load_pretrained, load_tokenizer, read_env,  # noqa: F401
                     _warn)


def _load(cfg):
    dims, weights = None, None
    ckpt = cfg["GP_WEIGHTS"] or (cfg["MODEL_ID"] if os.path.isdir(cfg["MODEL_ID"]) else "")
    if ckpt:
        dims, weights = W.load_checkpoint(ckpt)
    elif not cfg["GP_RANDOM_INIT"]:
        dims, weights = load_pretrained(cfg["MODEL_ID"])
    dims = dims or W.PRESETS.get(cfg["MODEL_ID"])
    if dims is None:
        raise ValueError(f"unknown MODEL_ID {cfg['MODEL_ID']!r}")
    if weights is None:
        _warn(f"GP_RANDOM_INIT=1: serving a RANDOM-INIT {cfg['MODEL_ID']} (seed {cfg['GP_SEED']})")
    return dims, weights


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    cfg = read_env()
    dims, weights = _load(cfg)
    dtype = GP_BF16 if cfg["GP_DTYPE"] == "bf16" else GP_F32
    max_ctx = int(os.environ.get("GP_MAX_CTX", dims.n_positions))
    runner = PipelineRunner(dims, seed=cfg["GP_SEED"], dtype=dtype, batch_per_stream=1, max_ctx=max_ctx, rank=rank,
                            world=world, device=local, weights=weights, n_streams=1)
    ctl = dist.new_group(backend="gloo")      # request broadcast (host objects) stays off the NCCL stream

    sampled = cfg["GP_SAMPLING"] == "reference"
    if sampled and dtype != GP_BF16:
        raise ValueError("GP_SAMPLING=reference in the pod server needs GP_DTYPE=bf16 (the fp32 pipeline kernel selects greedily)")

    def step(ids, n_new, seed=None):
        """Runs on every rank with the same arguments."""
        if seed is not None:      # the sampler lives on the stage with the lm_head; a no-op elsewhere
            runner.set_sampler(REF_TEMPERATURE, min(REF_TOP_K, dims.vocab_size), seed)
        prompts = torch.tensor([ids], dtype=torch.int32)
        return runner.generate(prompts, n_new)

    if rank != 0:
        while True:
            box = [None]
            dist.broadcast_object_list(box, src=0, group=ctl)
            if box[0] is None:
                break
            step(*box[0])
        dist.destroy_process_group()
        return

    import uvicorn
    from fastapi import FastAPI, HTTPException
    tokenizer = load_tokenizer(cfg["MODEL_ID"])
    lock = threading.Lock()
    served = [0]
    app = FastAPI()

    @app.get("/healthz")
    def healthz():
        return {"status": "ok", "stages": world, "model": cfg["MODEL_ID"], "stage_split": [list(s) for s in runner.stages]}

    @app.post("/generate")
    def generate(req: GenerateReq):
        ids = tokenizer.encode(req.prompt)
        n_new = max(0, int(req.max_new_tokens))
        if n_new == 0:
            return {"generated": tokenizer.decode(ids, skip_special_tokens=True)}
        if not ids or min(ids) < 0 or max(ids) >= dims.vocab_size:
            raise HTTPException(status_code=400, detail="empty prompt or token id outside the vocabulary")
        if len(ids) + n_new - 1 > max_ctx:
            raise HTTPException(status_code=400, detail=f"prompt of {len(ids)} tokens + {n_new} new tokens exceeds {max_ctx} positions")
        with lock:
            served[0] += 1
            seed = cfg["GP_SAMPLE_SEED"] + served[0] - 1 if sampled else None
            dist.broadcast_object_list([(ids, n_new, seed)], src=0, group=ctl)
            try:
                out = step(ids, n_new, seed)
            except Exception as ex:
                raise HTTPException(status_code=500, detail=str(ex))
        return {"generated": tokenizer.decode(out[0].tolist(), skip_special_tokens=True)}

    try:
        uvicorn.run(app, host="0.0.0.0", port=cfg["SHARD_PORT"], log_level="warning")
    finally:
        dist.broadcast_object_list([None], src=0, group=ctl)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
