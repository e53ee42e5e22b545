import sys, os, statistics; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import llm_sharding_demo_b200 as L
from bench import synthetic_prompts
nl = int(sys.argv[1])
dims = L.GPT2Dims(n_layer=nl, n_embd=768, n_head=12)
eng = L.StageEngine(dims, dtype=L.GP_BF16, max_batch=1, max_ctx=160)
eng.load_weights(L.init_weights(dims, 1234, "hf"))
prompts = synthetic_prompts(1, 32, dims.vocab_size).pin_memory()
for _ in range(3):
    out, tm = eng.generate(prompts, 128, timings=True)
print("layers", nl, "p50 step us %.1f  decode ms %.3f" % (1000 * statistics.median(tm["step_ms"]), tm["decode_ms"]))
