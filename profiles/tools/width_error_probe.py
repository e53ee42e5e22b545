import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import llm_sharding_demo_b200 as L
from oracle import gpt2_oracle as O
def rounded(w, dt):
    return {k: (v.to(dt).float() if v.dim() == 2 else v) for k, v in w.items()}
for preset in sys.argv[1].split(","):
  for B in [int(x) for x in sys.argv[2].split(",")]:
    dims = O.PRESETS[preset]
    w = O.init_weights(dims, 91, "dense")
    P, n_new = 5, 4
    eng = L.StageEngine(L.GPT2Dims(**dims.to_dict()), dtype=L.GP_BF16, max_batch=B)
    eng.load_weights(w)
    g = torch.Generator().manual_seed(17)
    prompt = torch.randint(0, dims.vocab_size, (B, P), generator=g)
    ref_ids, ref_logits = O.generate_cached(w, dims, prompt, n_new, return_logits=True)
    _, ref_w = O.generate_cached(rounded(w, torch.bfloat16), dims, prompt, n_new, teacher=ref_ids, return_logits=True)
    out = eng.prefill(prompt.to("cuda", torch.int32), use_cache=True, want_logits="last", want_ids=True)
    for t in range(n_new):
        if t > 0:
            out = eng.decode(ref_ids[:, P + t - 1].to("cuda", torch.int32), want_logits=True)
        lg = out["logits"].cpu()
        e = (lg - ref_w[:, t]).abs().max(dim=-1).values
        print(preset, "B", B, "t", t, "err per seq", [round(float(x), 4) for x in e], flush=True)
    eng.close()
