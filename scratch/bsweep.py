import sys, statistics, torch
sys.path.insert(0, '.')
import llm_sharding_demo_b200 as L
from bench import synthetic_prompts
for model, Bs in (("gpt2", (1, 2, 4)), ("gpt2-xl", (1, 4))):
    dims = L.PRESETS[model]
    eng = L.StageEngine(dims, dtype=L.GP_BF16, max_batch=4, max_ctx=128, max_tokens=4*64)
    eng.load_weights(L.init_weights(dims, 1234, "hf"))
    for B in Bs:
        p = synthetic_prompts(B, 32, dims.vocab_size)
        eng.generate(p, 8)
        tms = [eng.generate(p, 32, timings=True)[1] for _ in range(3)]
        tm = min(tms, key=lambda t: t["decode_ms"])
        print(model, B, "step_ms", round(statistics.median(tm["step_ms"]), 4), flush=True)
    eng.close()
