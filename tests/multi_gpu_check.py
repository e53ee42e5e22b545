"""Multi-GPU pipeline check (needs >= 2 GPUs; launched by tests/test_gpu_multi.py or by hand):
    torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_gpu_check.py
Checks, on identical seeded weights: the device-side stage boundary ("p2p": persistent kernels
writing into each other's inbox over NVLink) and the NCCL transport produce the same token ids,
and both match the CPU oracle's cached greedy decode wherever the top-2 margin allows."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import llm_sharding_demo_b200 as L  # noqa: E402
from llm_sharding_demo_b200.pipeline import PipelineRunner  # noqa: E402
from oracle import gpt2_oracle as O  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    odims = O.PRESETS["mini"]
    dims = L.GPT2Dims(**odims.to_dict())
    w = O.init_weights(odims, 31, "dense")
    B, P, n_new = 2, 6, 12
    g = torch.Generator().manual_seed(4)
    prompts = torch.randint(0, dims.vocab_size, (B * world, P), generator=g, dtype=torch.int32)
    outs = {}
    for transport in ("nccl", "p2p"):
        runner = PipelineRunner(dims, dtype=L.GP_F32, batch_per_stream=B, max_ctx=P + n_new, rank=rank, world=world,
                                device=local, weights=w, transport=transport)
        for rep in range(2):   # twice: the second request reuses inboxes, epochs and KV slots
            out, tm = runner.generate(prompts, n_new, timings=True)
        outs[transport] = out
        if rank == 0:
            print(transport, "decode_ms", round(tm["decode_ms"], 3), "launches", tm["launches"], flush=True)
        dist.barrier()
    if rank == 0:
        assert torch.equal(outs["nccl"], outs["p2p"]), "p2p and nccl transports disagree"
        ref_ids, ref_logits = O.generate_cached(w, odims, prompts.long(), n_new, n_stages=world, return_logits=True)
        margins = O.top2_margin(ref_logits)
        checked = 0
        for b in range(B * world):
            for t in range(n_new):
                if float(margins[b, t]) <= 6e-4:   # fp32 storage: ids are bit-exact above 2x the 3e-4 logit tolerance
                    break
                assert int(outs["p2p"][b, P + t]) == int(ref_ids[b, P + t]), (b, t)
                checked += 1
        assert checked >= B * world * 3, checked
        print(f"MULTI_GPU_CHECK_OK world={world} tokens_checked={checked}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
