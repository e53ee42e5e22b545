"""Runs the UNMODIFIED reference `server.py` in-process.  TEST INFRASTRUCTURE ONLY.

The file is taken from `/root/reference/server.py` (the build container) or from `oracle/_ref/server.py`, the
git-ignored copy `oracle/build_ref.py` makes of it so that the CPU baseline can run the reference ITSELF on the GPU
box's host cores (`bench.py --impl reference`, `cpu_baseline.kind = "reference"`).  `oracle/make_golden.py` uses this
harness to produce `tests/golden/`.  Nothing under `llm_sharding_demo_b200/` imports it.

`server.py` cannot be imported as is offline (`server.py:40-41` fetch from the HF hub) and
its `block(h)[0]` (`server.py:85,100`) assumes transformers-4.x tuple returns.  Three stubs
are installed BEFORE import and removed after:
  1. `AutoTokenizer.from_pretrained`  -> byte-level stub tokenizer (UTF-8 bytes as ids);
  2. `AutoModelForCausalLM.from_pretrained` -> `GPT2LMHeadModel(GPT2Config(...))` loaded with
     the oracle's seeded weights (default sdpa attention kept: it is what makes the
     mask-less call causal, `[hf] integrations/sdpa_attention.py:65,77`);
  3. `GPT2Block.forward` wrapped to return a 1-tuple so `[0]` keeps the batch dim.
The module is loaded once per role (a / b / coordinator) and driven through FastAPI's
TestClient; the coordinator's `requests.post` is routed to the in-process shard apps.
"""
from __future__ import annotations

import importlib.util
import os
import sys
from contextlib import contextmanager

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_CANDIDATES = ("/root/reference/server.py", os.path.join(_HERE, "_ref", "server.py"))
REFERENCE_SERVER = next((c for c in _CANDIDATES if os.path.exists(c)), _CANDIDATES[0])


class ByteTokenizer:
    """Offline stand-in for the GPT-2 BPE tokenizer: ids are UTF-8 bytes."""

    def encode(self, text):
        return list(text.encode("utf-8"))

    def decode(self, ids, skip_special_tokens=True):
        return bytes(int(i) & 0xFF for i in ids).decode("utf-8", errors="replace")


def reference_available() -> bool:
    return os.path.exists(REFERENCE_SERVER)


@contextmanager
def _stubs(dims, weights):
    import transformers
    from transformers import GPT2Config, GPT2LMHeadModel
    from transformers.models.gpt2 import modeling_gpt2
    from oracle.gpt2_oracle import hf_state_dict

    def make_model(_model_id):
        cfg = GPT2Config(vocab_size=dims.vocab_size, n_positions=dims.n_positions,
                         n_embd=dims.n_embd, n_layer=dims.n_layer, n_head=dims.n_head,
                         layer_norm_epsilon=dims.layer_norm_epsilon)
        m = GPT2LMHeadModel(cfg)
        missing, unexpected = m.load_state_dict(hf_state_dict(weights), strict=False)
        assert not unexpected, unexpected
        assert all("attn.bias" in k or "masked_bias" in k for k in missing), missing
        return m

    orig_tok = transformers.AutoTokenizer.from_pretrained
    orig_mod = transformers.AutoModelForCausalLM.from_pretrained
    orig_fwd = modeling_gpt2.GPT2Block.forward

    def fwd_tuple(self, *a, **k):
        out = orig_fwd(self, *a, **k)
        return out if isinstance(out, tuple) else (out,)

    transformers.AutoTokenizer.from_pretrained = staticmethod(lambda _id: ByteTokenizer())
    transformers.AutoModelForCausalLM.from_pretrained = staticmethod(make_model)
    modeling_gpt2.GPT2Block.forward = fwd_tuple
    try:
        yield
    finally:
        transformers.AutoTokenizer.from_pretrained = orig_tok
        transformers.AutoModelForCausalLM.from_pretrained = orig_mod
        # GPT2Block.forward stays wrapped while loaded server modules are alive: their
        # ShardA/ShardB index [0] at call time.  `restore_block_forward` undoes it.
        _stubs.orig_fwd = orig_fwd


def restore_block_forward():
    from transformers.models.gpt2 import modeling_gpt2
    if getattr(_stubs, "orig_fwd", None) is not None:
        modeling_gpt2.GPT2Block.forward = _stubs.orig_fwd
        _stubs.orig_fwd = None


def load_reference_server(role: str, split_at: int, dims, weights, tag: str = ""):
    """Import `/root/reference/server.py` as a fresh module instance for one SHARD_ROLE."""
    assert reference_available(), "reference tree not present on this machine"
    env = {"SHARD_ROLE": role, "SPLIT_AT": str(split_at), "MODEL_ID": "oracle-stub"}
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        with _stubs(dims, weights):
            name = f"_ref_server_{role}{tag}"
            spec = importlib.util.spec_from_file_location(name, REFERENCE_SERVER)
            mod = importlib.util.module_from_spec(spec)
            sys.modules[name] = mod
            # server.py:27 prints its role at import; keep stdout clean for callers that emit JSON lines
            import contextlib
            with contextlib.redirect_stdout(sys.stderr):
                spec.loader.exec_module(mod)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return mod


class ReferencePipeline:
    """The three reference roles wired together in-process (`server.py:154-210` call graph)."""

    def __init__(self, dims, weights, split_at: int, tag: str = ""):
        from fastapi.testclient import TestClient
        self.dims, self.split_at = dims, split_at
        self.mod_a = load_reference_server("a", split_at, dims, weights, tag)
        self.mod_b = load_reference_server("b", split_at, dims, weights, tag)
        self.mod_c = load_reference_server("coordinator", split_at, dims, weights, tag)
        self.client_a = TestClient(self.mod_a.app)
        self.client_b = TestClient(self.mod_b.app)
        self.client_c = TestClient(self.mod_c.app)
        pipe = self

        class _Resp:
            def __init__(self, r):
                self._r = r

            def raise_for_status(self):
                self._r.raise_for_status()

            def json(self):
                return self._r.json()

        class _Requests:
            @staticmethod
            def post(url, json=None, timeout=None):
                if url.endswith("/forward"):
                    return _Resp(pipe.client_a.post("/forward", json=json))
                if url.endswith("/forward_b"):
                    return _Resp(pipe.client_b.post("/forward_b", json=json))
                raise AssertionError(url)

        self.mod_c.requests = _Requests  # `server.py:173,179` call requests.post

    # --- the two shard routes exactly as an HTTP client sees them
    def forward_a(self, ids):
        r = self.client_a.post("/forward", json={"input_ids": [int(i) for i in ids]})
        r.raise_for_status()
        return r.json()

    def forward_b(self, hidden):
        r = self.client_b.post("/forward_b", json={"hidden_states": hidden})
        r.raise_for_status()
        return r.json()

    def generate(self, prompt: str, max_new_tokens=None):
        body = {"prompt": prompt}
        if max_new_tokens is not None:
            body["max_new_tokens"] = max_new_tokens
        r = self.client_c.post("/generate", json=body)
        r.raise_for_status()
        return r.json()

    def generate_sampled_ids(self, prompt: str, max_new_tokens: int, torch_seed: int):
        """The reference's own `/generate` (sampler of server.py:187-205 included), made repeatable:
        its multinomial draws come from torch's global generator, which is seeded here right before
        the request.  Returns the token ids the reference handed to `tokenizer.decode` (the stub
        tokenizer's text is lossy for ids >= 256, so the ids are captured there)."""
        tok = self.mod_c.tokenizer
        seen = {}
        orig = tok.decode

        def spy(ids, skip_special_tokens=True):
            seen["ids"] = [int(i) for i in ids]
            return orig(ids, skip_special_tokens=skip_special_tokens)

        tok.decode = spy
        try:
            torch.manual_seed(torch_seed)
            self.generate(prompt, max_new_tokens)
        finally:
            del tok.decode
        return seen["ids"]

    # --- greedy decode through the HTTP/JSON routes (full path, what /generate would do
    #     with argmax in place of `server.py:187-205`'s sampler)
    def greedy_via_routes(self, ids, n_new):
        ids = [int(i) for i in ids]
        last_logits = []
        for _ in range(n_new):
            hidden = self.forward_a(ids)["hidden_states"]
            logits = self.forward_b(hidden)["logits"]
            row = torch.tensor(logits[0][-1], dtype=torch.float32)
            last_logits.append(row)
            ids.append(int(torch.argmax(row)))
        return ids, torch.stack(last_logits)

    # --- same, calling the reference's ShardA/ShardB modules directly (no JSON)
    @torch.no_grad()
    def greedy_via_modules(self, ids, n_new):
        ids = torch.tensor([list(map(int, ids))], dtype=torch.long)
        last_logits = []
        for _ in range(n_new):
            hidden = self.mod_a.shard_a(ids)
            logits = self.mod_b.shard_b(hidden)
            row = logits[0, -1].float()
            last_logits.append(row.clone())
            ids = torch.cat([ids, torch.argmax(row).view(1, 1)], dim=1)
        return ids[0].tolist(), torch.stack(last_logits)
