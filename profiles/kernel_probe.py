#!/usr/bin/env python
"""Runs every stand-alone hot-path kernel once at a realistic (GPT-2 124M / XL) size through the C ABI, so that
`ncu -k regex:<kernel>` can capture each of them (profiles/run_ncu.sh).  Prints the CUDA-event time of each op."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_sharding_demo_b200 as L  # noqa: E402
from llm_sharding_demo_b200 import ops  # noqa: E402

dev = "cuda:0"
g = torch.Generator().manual_seed(0)
bf = torch.bfloat16


def timed(name, fn, bytes_moved=None, flops=None, reps=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    extra = ""
    if bytes_moved:
        extra += f" {bytes_moved / ms / 1e6:8.1f} GB/s"
    if flops:
        extra += f" {flops / ms / 1e9:8.1f} TFLOP/s"
    print(f"{name:34s} {ms * 1e3:9.1f} us{extra}", flush=True)


H, V, nh, d, T = 768, 50257, 12, 64, 1024
x1 = torch.randn(1, H, generator=g).to(dev)
x8 = torch.randn(8, H, generator=g).to(dev)
gm, bt = (1 + 0.1 * torch.randn(H, generator=g)).to(dev), (0.1 * torch.randn(H, generator=g)).to(dev)
wte = (torch.randn(V, H, generator=g) * 0.02).to(bf).to(dev)
wfc = (torch.randn(4 * H, H, generator=g) * 0.02).to(bf).to(dev)
bfc = torch.zeros(4 * H, device=dev)
scratch = ops.lmhead_scratch(dev)
timed("lmhead_argmax B=1 (77 MB)", lambda: ops.lmhead_argmax(x1, gm, bt, wte, scratch=scratch), V * H * 2)
timed("lmhead_argmax B=8", lambda: ops.lmhead_argmax(x8, gm, bt, wte, scratch=scratch), V * H * 2)
timed("gemv c_fc B=1 (4.7 MB)", lambda: ops.gemv(x1, wfc, bfc, ln=(gm, bt), gelu=True), 4 * H * H * 2)
timed("gemv c_fc B=8", lambda: ops.gemv(x8, wfc, bfc, ln=(gm, bt), gelu=True), 4 * H * H * 2)
xm = torch.randn(16384, H, generator=g).to(dev)
timed("layernorm 16384 x 768 -> bf16", lambda: ops.layernorm(xm, gm, bt, out_bf16=True), 16384 * H * 6)
B = 32
q = torch.randn(B, H, generator=g).to(dev)
kc = (torch.randn(B, nh, T, d, generator=g) * 0.5).to(bf).to(dev)
vc = (torch.randn(B, nh, T, d, generator=g) * 0.5).to(bf).to(dev)
lens = torch.full((B,), T, dtype=torch.int32, device=dev)
timed("attn_decode B=32 T=1024 (101 MB)", lambda: ops.attn_decode(q, kc, vc, lens), 2 * B * nh * T * d * 2)
qkv = (torch.randn(8, 1024, 3 * H, generator=g) * 0.5).to(bf).to(dev)
timed("attn_prefill B=8 T=1024", lambda: ops.attn_prefill(qkv, nh), flops=4 * 8 * nh * 1024 * 1024 / 2 * d)
logits = torch.randn(8, V, generator=g).to(dev)
timed("sample_topk B=8 k=40", lambda: ops.sample_topk(logits, 0.6, 40, 1), 8 * V * 4 * 5)
a = (torch.randn(16384, H, generator=g) * 0.5).to(bf).to(dev)
timed("gemm tcgen05 16384x3072x768", lambda: ops.gemm(a, wfc, bfc, gelu=True, out_bf16=True), flops=2 * 16384 * 3072 * H)
# one batch-32 decode step of the 124M engine (tcgen05 GEMMs, split-K, attention kernels): the path B > 8 takes
dims = L.PRESETS["gpt2"]
eng = L.StageEngine(dims, dtype=L.GP_BF16, max_batch=32, max_ctx=160, max_tokens=32 * 64)
eng.load_weights(L.init_weights(dims, 1234, "hf"))
ids = torch.randint(0, V, (32, 32), generator=g, dtype=torch.int32).to(dev)
out = eng.prefill(ids, use_cache=True, want_ids=True)
tok = out["next_ids"]
timed("decode step 124M B=32 (per-op)", lambda: eng.decode(tok), reps=3)
