#!/usr/bin/env python
"""One generation of a given (model, B, P, n_new) on one GPU; with GP_MEGA_TRACE=<file> the persistent kernel dumps its
per-CTA phase trace of the last decode step (read it with profiles/mega2_trace.py <file> <B * n_head>).
GP_SAMPLE=1 switches the reference's sampler on (temperature 0.6, top-k 40)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import llm_sharding_demo_b200 as L  # noqa: E402
from bench import synthetic_prompts  # noqa: E402

model, B, P, n_new = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
dims = L.PRESETS[model]
eng = L.StageEngine(dims, dtype=L.GP_BF16, max_batch=B, max_ctx=P + n_new, max_tokens=B * P)
eng.load_weights(L.init_weights(dims, 1234, "hf"))
if os.environ.get("GP_SAMPLE"):
    eng.set_sampler(0.6, 40, seed=1)
prompts = synthetic_prompts(B, P, dims.vocab_size).pin_memory()
for _ in range(2):
    out, tm = eng.generate(prompts, n_new, timings=True)
print(model, B, P, "decode ms/step", tm["decode_ms"] / (n_new - 1), "prefill ms", tm["prefill_ms"])
