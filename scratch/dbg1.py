import sys, torch
sys.path.insert(0, '.')
import llm_sharding_demo_b200 as L
from oracle import gpt2_oracle as O
dims = O.PRESETS["mini"]
pd = L.GPT2Dims(**dims.to_dict())
hf = O.init_weights(dims, 11, "hf"); dn = O.init_weights(dims, 11, "dense")
ids = torch.tensor([[5, 1040, 17, 300, 2]])
def run(w, tag, lb=0, le=1):
    eng = L.StageEngine(pd, lb, le, has_embed=True, has_head=False, dtype=L.GP_F32)
    eng.load_weights(w)
    got = eng.prefill(ids.to("cuda", torch.int32), use_cache=False)["hidden"].cpu()
    ref = O.stage_forward(w, dims, ids, lb, le, True, False)
    print(tag, "max err", float((got-ref).abs().max()))
run(hf, "hf")
run(dn, "dense")
for cat in ["ln_1.weight","ln_1.bias","ln_2.weight","ln_2.bias","c_attn.bias","attn.c_proj.bias","c_fc.bias","mlp.c_proj.bias"]:
    w = dict(hf)
    for k in w:
        if k.endswith(cat): w[k] = dn[k]
    run(w, cat)
