"""Generates `tests/golden/*.npz` by running the UNMODIFIED reference `server.py`
(through `oracle/ref_harness.py`) on seeded weights.  TEST INFRASTRUCTURE ONLY.

Run in the build container (the only place `/root/reference` exists):
    python -m oracle.make_golden
The fixtures are committed; the GPU box only ever reads them.
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import gpt2_oracle as O  # noqa: E402
from oracle.ref_harness import ReferencePipeline, restore_block_forward  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def _meta(dims, seed, recipe, split_at, prompt):
    return json.dumps({"dims": dims.to_dict(), "seed": seed, "recipe": recipe,
                       "split_at": split_at, "prompt": prompt,
                       "torch": torch.__version__,
                       "transformers": __import__("transformers").__version__})


def golden_small(name, preset, seed, recipe, split_at, prompt_ids, n_new, prompt_text=None):
    dims = O.PRESETS[preset]
    w = O.init_weights(dims, seed, recipe)
    pipe = ReferencePipeline(dims, w, split_at, tag=name)
    t0 = time.time()
    ids, last_logits = pipe.greedy_via_routes(prompt_ids, n_new)
    hidden0 = torch.tensor(pipe.forward_a(prompt_ids)["hidden_states"], dtype=torch.float32)
    logits0 = torch.tensor(pipe.forward_b(hidden0.tolist())["logits"], dtype=torch.float32)
    ids_mod, last_logits_mod = pipe.greedy_via_modules(prompt_ids, n_new)
    assert ids_mod == ids, "reference JSON path and module path disagree"
    extra = {}
    if prompt_text is not None:
        extra["generate0"] = pipe.generate(prompt_text, 0)["generated"]
        extra["generate_default_len"] = len(pipe.mod_c.tokenizer.encode(
            pipe.generate(prompt_text)["generated"]))
        # role gates (`server.py:134-135,145-146,156-157`): HTTP 200 + error JSON
        extra["gate_a"] = pipe.client_b.post("/forward", json={"input_ids": [1]}).json()
        extra["gate_b"] = pipe.client_a.post("/forward_b", json={"hidden_states": [[[0.0] * dims.n_embd]]}).json()
        extra["gate_c"] = pipe.client_a.post("/generate", json={"prompt": "x"}).json()
    np.savez_compressed(
        os.path.join(GOLD, name + ".npz"),
        meta=_meta(dims, seed, recipe, split_at, prompt_text),
        extra=json.dumps(extra),
        prompt_ids=np.asarray(prompt_ids, np.int64),
        greedy_ids=np.asarray(ids, np.int64),
        hidden_step0=hidden0.numpy(),            # [1,P,H]   /forward output
        logits_step0=(logits0[0].numpy() if dims.vocab_size < 4096 else logits0[0, -1:].numpy()),
        last_logits=(last_logits.numpy() if dims.vocab_size < 4096
                     else last_logits[[0, n_new - 1]].numpy()),
        margins=O.top2_margin(last_logits).numpy(),
    )
    print(f"{name}: {n_new} tokens via reference routes in {time.time() - t0:.1f}s; "
          f"min margin {float(O.top2_margin(last_logits).min()):.3e}")
    restore_block_forward()


def golden_124m(n_new=16):
    name, seed, recipe, split_at = "gpt2_124m", 1234, "hf", 6
    dims = O.PRESETS["gpt2"]
    w = O.init_weights(dims, seed, recipe)
    g = torch.Generator().manual_seed(7)
    prompt = torch.randint(0, dims.vocab_size, (32,), generator=g).tolist()
    pipe = ReferencePipeline(dims, w, split_at, tag=name)
    t0 = time.time()
    ids, last_logits = pipe.greedy_via_modules(prompt, n_new)
    # one step through the real HTTP/JSON routes to tie the module path to the wire path
    ids_r, ll_r = pipe.greedy_via_routes(prompt, 1)
    assert ids_r[-1] == ids[len(prompt)]
    assert float((ll_r[0] - last_logits[0]).abs().max()) < 1e-5
    hidden0 = pipe.mod_a.shard_a(torch.tensor([prompt]))
    top = torch.topk(last_logits, 8, dim=-1)
    np.savez_compressed(
        os.path.join(GOLD, name + ".npz"),
        meta=_meta(dims, seed, recipe, split_at, None), extra=json.dumps({}),
        prompt_ids=np.asarray(prompt, np.int64), greedy_ids=np.asarray(ids, np.int64),
        hidden_step0_last=hidden0[0, -1].numpy(),
        last_logits=last_logits[:1].numpy(),
        top8_idx=top.indices.numpy(), top8_val=top.values.numpy(),
        margins=O.top2_margin(last_logits).numpy(),
    )
    print(f"{name}: {n_new} tokens in {time.time() - t0:.1f}s; "
          f"min margin {float(O.top2_margin(last_logits).min()):.3e}")
    restore_block_forward()


def golden_sampled(name, preset, seed, recipe, split_at, prompt_text, n_new, torch_seeds):
    """The reference's /generate with its temperature / top-k / multinomial selection (server.py:187-205),
    repeatable because torch's global generator is seeded before each request."""
    dims = O.PRESETS[preset]
    w = O.init_weights(dims, seed, recipe)
    pipe = ReferencePipeline(dims, w, split_at, tag=name)
    ids = np.array([pipe.generate_sampled_ids(prompt_text, n_new, ts) for ts in torch_seeds], dtype=np.int64)
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), meta=_meta(dims, seed, recipe, split_at, prompt_text),
                        torch_seeds=np.array(torch_seeds, dtype=np.int64), ids=ids, n_new=n_new)
    restore_block_forward()
    print(name, ids.shape, ids[0].tolist())


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(8)
    if "--sampled-only" in sys.argv:
        golden_sampled("sampled_mini", "mini", 11, "dense", 2, "Hi, ", 12, [2024, 7, 99])
        golden_sampled("sampled_tiny_gpt2", "tiny-gpt2", 0, "hf", 1, "Hi, ", 20, [5])
        return
    # config 1 of BASELINE.json: tiny-gpt2, prompt "Hi, " (README.md:115-120), 20 tokens.
    golden_small("tiny_gpt2", "tiny-gpt2", 0, "hf", 1, list("Hi, ".encode()), 20, "Hi, ")
    golden_small("tiny_gpt2_dense", "tiny-gpt2", 3, "dense", 1, list("Hi, ".encode()), 6, "Hi, ")
    golden_small("mini", "mini", 11, "dense", 2, [5, 1040, 17, 300, 2], 8)
    golden_small("mini32", "mini32", 12, "dense", 1, [1, 2, 3, 500, 522, 0, 7], 6)
    golden_124m()
    golden_sampled("sampled_mini", "mini", 11, "dense", 2, "Hi, ", 12, [2024, 7, 99])
    golden_sampled("sampled_tiny_gpt2", "tiny-gpt2", 0, "hf", 1, "Hi, ", 20, [5])


if __name__ == "__main__":
    main()
