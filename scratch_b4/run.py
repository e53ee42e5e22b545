import sys, os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import llm_sharding_demo_b200 as L
from bench import synthetic_prompts
model, B = sys.argv[1], int(sys.argv[2])
dims = L.PRESETS[model]
eng = L.StageEngine(dims, dtype=L.GP_BF16, max_batch=B, max_ctx=256, max_tokens=B * 64)
eng.load_weights(L.init_weights(dims, 1234, "hf"))
prompts = synthetic_prompts(B, 32, dims.vocab_size).pin_memory()
for _ in range(3):
    out, tm = eng.generate(prompts, 64, timings=True)
import statistics
print(model, B, "p50 step ms", statistics.median(tm["step_ms"]))
