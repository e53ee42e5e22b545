"""CPU oracle for the pipeline-sharded GPT-2 hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, in plain fp32 torch on the CPU, the arithmetic that the
reference (`/root/reference/server.py`) performs for its Shard A / Shard B
forward calls and its coordinator loop.  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s CPU-baseline / `--impl reference` legs may import it; the
product path (`llm_sharding_demo_b200`) never does.

The arithmetic itself is not in the reference tree: it lives in the
third-party `transformers` package (unpinned in `requirements.txt:3`; the image
has 5.5.0) and `torch` (unpinned, `requirements.txt:4`; 2.11.0 here).  Each
function below cites the reference call site and the `[hf]` source it restates
(`[hf]` = site-packages/transformers).

Parity pinning: the reference ships no tests or golden vectors,
so this oracle is pinned against outputs of the reference itself, produced in
the build container by `oracle/make_golden.py` (which imports the unmodified
`server.py` with three offline stubs) and committed under `tests/golden/`.
`tests/test_oracle.py` checks this file against those fixtures.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, asdict
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

Weights = Dict[str, torch.Tensor]


@dataclass(frozen=True)
class GPT2Dims:
    """GPT-2 shape parameters ([hf] models/gpt2/configuration_gpt2.py defaults)."""
    n_layer: int
    n_embd: int
    n_head: int
    vocab_size: int = 50257
    n_positions: int = 1024
    layer_norm_epsilon: float = 1e-5

    @property
    def head_dim(self) -> int:
        return self.n_embd // self.n_head

    def to_dict(self):
        return asdict(self)


# Named configurations of BASELINE.json `configs`.
PRESETS: Dict[str, GPT2Dims] = {
    # sshleifer/tiny-gpt2: n_embd=2, n_layer=2, n_head=2 (server.py:20 default MODEL_ID)
    "tiny-gpt2": GPT2Dims(n_layer=2, n_embd=2, n_head=2),
    "sshleifer/tiny-gpt2": GPT2Dims(n_layer=2, n_embd=2, n_head=2),
    "gpt2": GPT2Dims(n_layer=12, n_embd=768, n_head=12),
    "gpt2-medium": GPT2Dims(n_layer=24, n_embd=1024, n_head=16),
    "gpt2-large": GPT2Dims(n_layer=36, n_embd=1280, n_head=20),
    "gpt2-xl": GPT2Dims(n_layer=48, n_embd=1600, n_head=25),
    # small shapes used only by tests (ragged vocab, d=64 and d=32 heads)
    "mini": GPT2Dims(n_layer=4, n_embd=128, n_head=2, vocab_size=1041, n_positions=64),
    "mini32": GPT2Dims(n_layer=2, n_embd=96, n_head=3, vocab_size=523, n_positions=48),
    # 8 blocks of the `mini` shape (8-stage pipelines), and the WIDTHS of gpt2-medium / gpt2-xl with two
    # blocks and a small vocabulary: row ownership, chunking and staging of the persistent kernel depend on
    # H, not on L, so these keep the oracle fast while covering configs[2] / configs[3] shapes
    "mini8": GPT2Dims(n_layer=8, n_embd=128, n_head=2, vocab_size=1041, n_positions=64),
    "wide1024": GPT2Dims(n_layer=2, n_embd=1024, n_head=16, vocab_size=1041, n_positions=64),
    "wide1600": GPT2Dims(n_layer=2, n_embd=1600, n_head=25, vocab_size=1041, n_positions=64),
}


def weight_names(dims: GPT2Dims) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(name, shape, kind) in HF `GPT2Model.state_dict()` naming without the
    `transformer.` prefix.  Conv1D weights are [in, out] ([hf] pytorch_utils.py:97-123)."""
    H, F = dims.n_embd, 4 * dims.n_embd
    out = [("wte.weight", (dims.vocab_size, H), "emb"),
           ("wpe.weight", (dims.n_positions, H), "emb")]
    for i in range(dims.n_layer):
        p = f"h.{i}."
        out += [
            (p + "ln_1.weight", (H,), "ln_w"), (p + "ln_1.bias", (H,), "ln_b"),
            (p + "attn.c_attn.weight", (H, 3 * H), "w"), (p + "attn.c_attn.bias", (3 * H,), "b"),
            (p + "attn.c_proj.weight", (H, H), "w_proj"), (p + "attn.c_proj.bias", (H,), "b"),
            (p + "ln_2.weight", (H,), "ln_w"), (p + "ln_2.bias", (H,), "ln_b"),
            (p + "mlp.c_fc.weight", (H, F), "w"), (p + "mlp.c_fc.bias", (F,), "b"),
            (p + "mlp.c_proj.weight", (F, H), "w_proj"), (p + "mlp.c_proj.bias", (H,), "b"),
        ]
    out += [("ln_f.weight", (H,), "ln_w"), ("ln_f.bias", (H,), "ln_b")]
    return out


def init_weights(dims: GPT2Dims, seed: int, recipe: str = "hf",
                 only: Optional[List[str]] = None) -> Weights:
    """Seeded random-init weights.

    recipe="hf" follows `[hf] models/gpt2/modeling_gpt2.py:432-458` (`_init_weights`):
    N(0, 0.02) for Linear/Conv1D/Embedding weights, zero biases, LayerNorm w=1 b=0,
    and `c_proj.weight` ~ N(0, 0.02/sqrt(2*n_layer)).
    recipe="dense" additionally randomises biases and the LayerNorm affine so that
    every term of every kernel is exercised by the parity tests.

    Each tensor is drawn from its own generator seeded by (seed, position in
    `weight_names`) so a pipeline stage can materialise only its own layers and
    still get exactly the tensors every other stage / the oracle sees.
    """
    assert recipe in ("hf", "dense")
    w: Weights = {}
    for idx, (name, shape, kind) in enumerate(weight_names(dims)):
        if only is not None and name not in only:
            continue
        g = torch.Generator().manual_seed(seed * 100003 + idx)
        if kind in ("emb", "w"):
            t = torch.randn(shape, generator=g) * 0.02
        elif kind == "w_proj":
            t = torch.randn(shape, generator=g) * (0.02 / math.sqrt(2 * dims.n_layer))
        elif kind == "b":
            t = torch.randn(shape, generator=g) * 0.05 if recipe == "dense" else torch.zeros(shape)
        elif kind == "ln_w":
            t = 1.0 + 0.1 * torch.randn(shape, generator=g) if recipe == "dense" else torch.ones(shape)
        elif kind == "ln_b":
            t = 0.1 * torch.randn(shape, generator=g) if recipe == "dense" else torch.zeros(shape)
        else:  # pragma: no cover
            raise AssertionError(kind)
        w[name] = t.contiguous()
    return w


def stage_weight_names(dims: GPT2Dims, layer_begin: int, layer_end: int,
                       has_embed: bool, has_head: bool) -> List[str]:
    """Names a pipeline stage needs.  The last stage needs `wte.weight` as the tied
    lm_head ([hf] modeling_gpt2.py:646,651)."""
    names = []
    for name, _, _ in weight_names(dims):
        if name.startswith("h."):
            li = int(name.split(".")[1])
            if layer_begin <= li < layer_end:
                names.append(name)
        elif name == "wte.weight":
            if has_embed or has_head:
                names.append(name)
        elif name == "wpe.weight":
            if has_embed:
                names.append(name)
        elif name.startswith("ln_f."):
            if has_head:
                names.append(name)
    return names


# --------------------------------------------------------------------------- ops

def layer_norm(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, eps: float) -> torch.Tensor:
    """nn.LayerNorm over the last dim, biased variance ([hf] modeling_gpt2.py:252,254,505)."""
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    return (x - mu) * torch.rsqrt(var + eps) * w + b


def conv1d(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """`Conv1D.forward`: addmm(bias, x.view(-1,nx), W[nx,nf]) ([hf] pytorch_utils.py:119-123)."""
    return torch.addmm(b, x.reshape(-1, x.shape[-1]), w).reshape(*x.shape[:-1], w.shape[1])


def gelu_new(x: torch.Tensor) -> torch.Tensor:
    """NewGELUActivation, tanh form ([hf] activations.py:59-66)."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x ** 3)))


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, q_pos0: int) -> torch.Tensor:
    """softmax(q k^T / sqrt(d) + causal) v on [B, nh, Tq, d] x [B, nh, Tk, d].

    Restates `F.scaled_dot_product_attention(..., is_causal = Tq>1 and mask is None)` as
    reached from `[hf] integrations/sdpa_attention.py:65,77,92-101` with
    `scaling = head_dim**-0.5` ([hf] modeling_gpt2.py:96-100).  `q_pos0` is the absolute
    position of the first query (0 for the reference's full recompute; cache length for
    the cached decode of `[hf] modeling_gpt2.py:155-199`).  Query i sees keys j <= q_pos0+i.
    """
    d = q.shape[-1]
    s = torch.matmul(q, k.transpose(-1, -2)) * (d ** -0.5)
    Tq, Tk = q.shape[-2], k.shape[-2]
    qi = torch.arange(Tq).unsqueeze(1) + q_pos0
    kj = torch.arange(Tk).unsqueeze(0)
    s = s.masked_fill(kj > qi, float("-inf"))
    return torch.matmul(torch.softmax(s, dim=-1), v)


def block_forward(w: Weights, i: int, h: torch.Tensor, dims: GPT2Dims,
                  kv: Optional[Tuple[torch.Tensor, torch.Tensor]] = None
                  ) -> Tuple[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
    """One pre-LN GPT2Block ([hf] modeling_gpt2.py:262-309; attention :144-226; MLP :229-243),
    as called by `server.py:85` (Shard A) and `server.py:100` (Shard B).
    Returns (hidden, (K, V)) with K, V covering past+new positions."""
    p = f"h.{i}."
    B, T, H = h.shape
    nh, d = dims.n_head, dims.head_dim
    x = layer_norm(h, w[p + "ln_1.weight"], w[p + "ln_1.bias"], dims.layer_norm_epsilon)
    qkv = conv1d(x, w[p + "attn.c_attn.weight"], w[p + "attn.c_attn.bias"])
    q, k, v = qkv.split(H, dim=2)                                 # modeling_gpt2.py:185
    q = q.view(B, T, nh, d).transpose(1, 2)
    k = k.view(B, T, nh, d).transpose(1, 2)
    v = v.view(B, T, nh, d).transpose(1, 2)
    past = 0
    if kv is not None:
        past = kv[0].shape[2]
        k = torch.cat([kv[0], k], dim=2)
        v = torch.cat([kv[1], v], dim=2)
    a = attention(q, k, v, past)
    a = a.transpose(1, 2).reshape(B, T, H)                       # :222
    a = conv1d(a, w[p + "attn.c_proj.weight"], w[p + "attn.c_proj.bias"])
    h = h + a                                                     # :282
    x = layer_norm(h, w[p + "ln_2.weight"], w[p + "ln_2.bias"], dims.layer_norm_epsilon)
    x = conv1d(x, w[p + "mlp.c_fc.weight"], w[p + "mlp.c_fc.bias"])
    x = gelu_new(x)
    x = conv1d(x, w[p + "mlp.c_proj.weight"], w[p + "mlp.c_proj.bias"])
    return h + x, (k, v)                                          # :307


def embed(w: Weights, ids: torch.Tensor, pos0: int = 0) -> torch.Tensor:
    """wte[ids] + wpe[pos0 + arange(T)]; dropout is identity in eval (`server.py:78-83`)."""
    T = ids.shape[-1]
    pos = torch.arange(pos0, pos0 + T, dtype=torch.long)
    return w["wte.weight"][ids] + w["wpe.weight"][pos].unsqueeze(0)


def lm_head(w: Weights, h: torch.Tensor, dims: GPT2Dims) -> torch.Tensor:
    """ln_f then tied lm_head, no bias (`server.py:101-102`; [hf] modeling_gpt2.py:505,646,651)."""
    x = layer_norm(h, w["ln_f.weight"], w["ln_f.bias"], dims.layer_norm_epsilon)
    return torch.matmul(x, w["wte.weight"].t())


# ------------------------------------------------------------------ stage / shards

@torch.no_grad()
def stage_forward(w: Weights, dims: GPT2Dims, x: torch.Tensor, layer_begin: int, layer_end: int,
                  has_embed: bool, has_head: bool, kvs: Optional[list] = None, pos0: int = 0):
    """A pipeline stage = a contiguous slice of the reference's split (`server.py:63-64`),
    generalised from 2 to S stages.  `x` is ids [B,T] (int64) when has_embed else hidden
    [B,T,H].  Returns hidden [B,T,H], or logits [B,T,V] when has_head.  If `kvs` (a list, one
    entry per local layer, None or (K,V)) is given it is used and updated in place."""
    h = embed(w, x, pos0) if has_embed else x
    for li, i in enumerate(range(layer_begin, layer_end)):
        past = kvs[li] if kvs is not None else None
        h, kv = block_forward(w, i, h, dims, past)
        if kvs is not None:
            kvs[li] = kv
    return lm_head(w, h, dims) if has_head else h


def shard_a_forward(w: Weights, dims: GPT2Dims, input_ids: torch.Tensor, split_at: int) -> torch.Tensor:
    """`ShardA.forward` (`server.py:76-86`): ids [B,T] -> hidden [B,T,H]."""
    return stage_forward(w, dims, input_ids, 0, split_at, True, False)


def shard_b_forward(w: Weights, dims: GPT2Dims, hidden: torch.Tensor, split_at: int) -> torch.Tensor:
    """`ShardB.forward` (`server.py:97-103`): hidden [B,T,H] -> logits [B,T,V] (all T rows)."""
    return stage_forward(w, dims, hidden, split_at, dims.n_layer, False, True)


def split_layers(n_layer: int, n_stages: int) -> List[Tuple[int, int]]:
    """Contiguous layer ranges; earlier stages take the remainder (S=2 with the default
    ranges reproduces `split_gpt2_model(split_at=n_layer//2)` of `server.py:51-105`)."""
    base, rem = divmod(n_layer, n_stages)
    out, b = [], 0
    for s in range(n_stages):
        e = b + base + (1 if s < rem else 0)
        out.append((b, e))
        b = e
    return out


def greedy_select(last_logits: torch.Tensor) -> torch.Tensor:
    """Greedy restatement of `server.py:187-205`: `topk_indices[0]` of a sorted top-k is the
    argmax and the positive temperature (0.6) does not move it.  Ties -> lowest index
    (torch.argmax on CPU returns the first maximum)."""
    return torch.argmax(last_logits, dim=-1)


# --------------------------------------------------------------------------- sampled selection
_PHILOX_M0, _PHILOX_M1 = 0xD2511F53, 0xCD9E8D57
_PHILOX_W0, _PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(counter, key):
    """Philox4x32 with 10 rounds (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as
    1, 2, 3", SC'11).  counter: 4 words, key: 2 words -> 4 words.  Pinned by the Random123
    known-answer vectors in tests/test_oracle.py."""
    c = [int(x) & 0xFFFFFFFF for x in counter]
    k = [int(x) & 0xFFFFFFFF for x in key]
    for _ in range(10):
        p0, p1 = _PHILOX_M0 * c[0], _PHILOX_M1 * c[2]
        c = [(p1 >> 32) ^ c[1] ^ k[0], p1 & 0xFFFFFFFF, (p0 >> 32) ^ c[3] ^ k[1], p0 & 0xFFFFFFFF]
        k = [(k[0] + _PHILOX_W0) & 0xFFFFFFFF, (k[1] + _PHILOX_W1) & 0xFFFFFFFF]
    return c


def topk_probs(last_logits: torch.Tensor, temperature: float, top_k: int):
    """`server.py:187-199` for one logits row [V]: `logits / temperature`, `torch.topk` (sorted),
    `softmax` over the k values.  Returns (token ids [k], probabilities [k]) in the reference's
    order; ties between equal logits resolve to the lower token id (torch leaves that unspecified)."""
    x = last_logits.to(torch.float32) / temperature
    kth = torch.topk(x, top_k).values[-1]
    above = torch.nonzero(x > kth).flatten()
    equal = torch.nonzero(x == kth).flatten()[: top_k - above.numel()]   # ascending ids: the lower ones
    idx = torch.cat([above, equal]).sort().values
    idx = idx[torch.sort(x[idx], descending=True, stable=True).indices]   # value desc, id asc among equals
    probs = F.softmax(x[idx], dim=-1)
    return idx, probs


def reference_sample(last_logits: torch.Tensor, temperature: float = 0.6, top_k: int = 40) -> int:
    """`server.py:187-205` line for line for one logits row, drawing from torch's GLOBAL generator
    like the reference does.  Pinned against the reference's own /generate under torch.manual_seed
    (tests/golden/sampled_*.npz)."""
    logits_tensor = last_logits.to(torch.float32) / temperature
    topk_values, topk_indices = torch.topk(logits_tensor, top_k)
    probs = F.softmax(topk_values, dim=-1)
    return int(topk_indices[torch.multinomial(probs, 1)])


@torch.no_grad()
def generate_reference_sampled(w: Weights, dims: GPT2Dims, ids, n_new: int, split_at: int, torch_seed: int,
                               temperature: float = 0.6, top_k: int = 40):
    """The coordinator loop `server.py:169-206` with its own sampler: full recompute of Shard A and
    Shard B per token, `reference_sample` on logits[0, -1].  ids: python list (batch of one)."""
    ids = [int(i) for i in ids]
    torch.manual_seed(torch_seed)
    for _ in range(n_new):
        x = torch.tensor([ids], dtype=torch.long)
        logits = shard_b_forward(w, dims, shard_a_forward(w, dims, x, split_at), split_at)
        ids.append(reference_sample(logits[0, -1], temperature, top_k))
    return ids


def sample_select(last_logits: torch.Tensor, temperature: float, top_k: int, seed: int, row: int, position: int):
    """Seeded restatement of `server.py:187-205` for one row: the candidate set and its probabilities
    are the reference's (`topk_probs`); the reference's `torch.multinomial` draw comes from torch's
    unseeded global generator and cannot be restated, so the draw here is inverse-CDF sampling with
    u = (philox4x32_10((position, row, 0, 0), seed)[0] >> 8) * 2^-24 over the UNNORMALISED weights
    exp(x_j - x_0) accumulated in fp32 in candidate order - the arithmetic of csrc/k_sample.cu.
    Returns (token id, margin) with margin = distance of u * sum to the nearest CDF step (a draw
    closer than ~1e-5 may legitimately differ between exp implementations)."""
    import numpy as np
    idx, _ = topk_probs(last_logits, temperature, top_k)
    v = last_logits.to(torch.float32)[idx].numpy().astype(np.float32)
    inv_t = np.float32(1.0) / np.float32(temperature)
    m = np.float32(v[0] * inv_t)
    w = np.exp((v * inv_t - m).astype(np.float32)).astype(np.float32)
    total = np.float32(0.0)
    for x in w:
        total = np.float32(total + x)
    word = philox4x32_10((position, row, 0, 0), (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))[0]
    u = np.float32(word >> 8) * np.float32(1.0 / 16777216.0)
    target = np.float32(u * total)
    acc = np.float32(0.0)
    pick, margin = len(w) - 1, float("inf")
    found = False
    for j, x in enumerate(w):
        acc = np.float32(acc + x)
        margin = min(margin, abs(float(target) - float(acc)))
        if not found and target < acc:
            pick, found = j, True
    return int(idx[pick]), margin / max(float(total), 1e-30)


@torch.no_grad()
def generate_recompute(w: Weights, dims: GPT2Dims, ids: torch.Tensor, n_new: int, split_at: int,
                       return_logits: bool = False):
    """The coordinator loop of `server.py:169-206` with greedy selection: every step re-runs
    Shard A then Shard B on the WHOLE sequence and uses logits[:, -1].  ids: [B,P] int64."""
    ids = ids.clone()
    logs = []
    for _ in range(n_new):
        hidden = shard_a_forward(w, dims, ids, split_at)
        logits = shard_b_forward(w, dims, hidden, split_at)
        last = logits[:, -1]
        if return_logits:
            logs.append(last.clone())
        ids = torch.cat([ids, greedy_select(last).unsqueeze(1)], dim=1)
    return (ids, torch.stack(logs, 1)) if return_logits else ids


@torch.no_grad()
def generate_cached(w: Weights, dims: GPT2Dims, ids: torch.Tensor, n_new: int,
                    n_stages: int = 1, teacher: Optional[torch.Tensor] = None,
                    return_logits: bool = False):
    """KV-cached greedy decode across `n_stages` stages (second oracle; the cache semantics
    are those of `[hf] modeling_gpt2.py:155-199,569-582`: the new token's position is the
    cache length and it attends to every cached key).  If `teacher` ([B,P+n_new]) is given
    the fed-back token is taken from it (teacher forcing) while the model's own argmax is
    still what is returned."""
    ranges = split_layers(dims.n_layer, n_stages)
    kvs = [[None] * (e - b) for b, e in ranges]
    B, P = ids.shape
    out = [ids.clone()]
    logs = []
    x = ids
    pos0 = 0
    for step in range(n_new):
        h = x
        for s, (b, e) in enumerate(ranges):
            h = stage_forward(w, dims, h, b, e, s == 0, s == n_stages - 1, kvs[s], pos0)
        last = h[:, -1]
        nxt = greedy_select(last)
        out.append(nxt.unsqueeze(1))
        if return_logits:
            logs.append(last.clone())
        pos0 += x.shape[1]
        x = (teacher[:, P + step] if teacher is not None else nxt).unsqueeze(1)
    ids_out = torch.cat(out, dim=1)
    return (ids_out, torch.stack(logs, 1)) if return_logits else ids_out


def top2_margin(logits: torch.Tensor) -> torch.Tensor:
    """Gap between the best and second-best logit (the margin rule of SURVEY.md §8(c))."""
    t = torch.topk(logits, 2, dim=-1).values
    return t[..., 0] - t[..., 1]


def hf_state_dict(w: Weights) -> Weights:
    """Map to `GPT2LMHeadModel.state_dict()` names (adds `transformer.` and the tied head)."""
    sd = {"transformer." + k: v for k, v in w.items()}
    sd["lm_head.weight"] = w["wte.weight"]
    return sd
