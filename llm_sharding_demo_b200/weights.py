"""Model shapes, the stage split and seeded random-init weights (no checkpoints are
available offline; real `GPT2LMHeadModel.state_dict()` tensors load through the same
`StageEngine.load_weights`)."""
from __future__ import annotations

import math
from dataclasses import dataclass, asdict
from typing import Dict, List, Optional, Tuple

import torch


@dataclass(frozen=True)
class GPT2Dims:
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

    def n_params(self) -> int:
        H = self.n_embd
        return self.n_layer * (12 * H * H + 13 * H) + self.vocab_size * H + self.n_positions * H + 2 * H


# MODEL_ID values (server.py:20) -> shapes; "gpt2*" are the BASELINE.json configs
PRESETS: Dict[str, GPT2Dims] = {
    "sshleifer/tiny-gpt2": GPT2Dims(n_layer=2, n_embd=2, n_head=2),
    "tiny-gpt2": GPT2Dims(n_layer=2, n_embd=2, n_head=2),
    "gpt2": GPT2Dims(n_layer=12, n_embd=768, n_head=12),
    "gpt2-medium": GPT2Dims(n_layer=24, n_embd=1024, n_head=16),
    "gpt2-large": GPT2Dims(n_layer=36, n_embd=1280, n_head=20),
    "gpt2-xl": GPT2Dims(n_layer=48, n_embd=1600, n_head=25),
}


def weight_names(dims: GPT2Dims) -> List[Tuple[str, Tuple[int, ...], str]]:
    """GPT2Model.state_dict() names (no `transformer.` prefix), shapes and init class.
    Conv1D weights are [in, out]."""
    H, F = dims.n_embd, 4 * dims.n_embd
    out = [("wte.weight", (dims.vocab_size, H), "emb"), ("wpe.weight", (dims.n_positions, H), "emb")]
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


def split_layers(n_layer: int, n_stages: int) -> List[Tuple[int, int]]:
    """Contiguous block ranges per stage; S=2 reproduces split_gpt2_model(split_at=n_layer//2)
    (server.py:51-105)."""
    if not 1 <= n_stages <= n_layer:
        raise ValueError(f"cannot split {n_layer} layers into {n_stages} stages")
    base, rem = divmod(n_layer, n_stages)
    out, b = [], 0
    for s in range(n_stages):
        e = b + base + (1 if s < rem else 0)
        out.append((b, e))
        b = e
    return out


def stage_costs_us(dims: GPT2Dims, esz: int = 2) -> Tuple[float, float]:
    """(one block, ln_f + lm_head) in microseconds of a batch-1 decode step on one B200: a fixed part for the
    dependent phases (5 exchanges per block, the argmax exchange of the head) plus the weight stream at ~5.5-6 TB/s.
    Calibrated on the persistent kernel (profiles/r02_results.md: 124M 12.8 / 17 us per token of 170.7); only the
    RATIO matters for the split."""
    H, V = dims.n_embd, dims.vocab_size
    return 10.2 + esz * (12 * H * H + 13 * H) / 5.5e6, 4.0 + esz * V * H / 6.0e6


def balanced_split(dims: GPT2Dims, n_stages: int, esz: int = 2) -> List[Tuple[int, int]]:
    """Contiguous block ranges per stage that minimise the slowest stage of a decode pipeline.  The last stage also
    carries ln_f and the tied lm_head (1.5 blocks' worth of time at 124M), so it gets fewer blocks - none at all when
    that balances best (a head-only stage: (n_layer, n_layer)).  The reference's two-shard split (server.py:51-105,
    SPLIT_AT = n_layer // 2) is `split_layers(n_layer, 2)`; this is its time-balanced generalisation."""
    L = dims.n_layer
    if n_stages == 1:
        return [(0, L)]
    if not 2 <= n_stages <= L + 1:
        raise ValueError(f"cannot split {L} layers into {n_stages} stages")
    t_layer, t_head = stage_costs_us(dims, esz)
    best = None
    for n_last in range(0, L - (n_stages - 1) + 1):
        front = L - n_last
        per = -(-front // (n_stages - 1))           # ceil: the largest front stage
        cost = max(per * t_layer, n_last * t_layer + t_head)
        if best is None or cost < best[0] - 1e-9:
            best = (cost, n_last)
    n_last = best[1]
    out, b = [], 0
    base, rem = divmod(L - n_last, n_stages - 1)
    for s in range(n_stages - 1):
        # the heavier front stages come last: stage 0 also gathers the embeddings
        e = b + base + (1 if s >= n_stages - 1 - rem else 0)
        out.append((b, e))
        b = e
    out.append((b, L))
    return out


_ATTN_HALF = ("ln_1.", "attn.")     # tensors of a block's attention half; the rest (ln_2, mlp) is its MLP half


def half_costs_us(dims: GPT2Dims, esz: int = 2) -> Tuple[float, float]:
    """(attention half, MLP half) of one block, same calibration as stage_costs_us (they add up to its block cost):
    three dependent phases + the attention itself against two phases, 4 H^2 against 8 H^2 weights."""
    H = dims.n_embd
    return 6.1 + esz * (4 * H * H + 6 * H) / 5.5e6, 4.1 + esz * (8 * H * H + 7 * H) / 5.5e6


def stage_flags(begin: float, end: float) -> Tuple[int, int, bool, bool]:
    """(layer_begin, layer_end, begin_half, end_half) of a stage given in blocks, x.5 = a cut between the attention and
    the MLP half of block x (balanced_split_halves)."""
    return int(math.floor(begin)), int(math.ceil(end)), begin != math.floor(begin), end != math.floor(end)


def balanced_split_halves(dims: GPT2Dims, n_stages: int, esz: int = 2) -> List[Tuple[float, float]]:
    """Like balanced_split, with stage boundaries allowed between the two halves of a block (x += attn(ln_1 x) |
    x += mlp(ln_2 x): the residual stream is a complete hand-off there too).  Boundaries are in blocks, 3.5 = after the
    attention half of block 3.  Minimises the slowest stage; every stage but the last gets at least one half, the last
    one carries ln_f + lm_head and may get none.  124M on 4 stages: 3.5 blocks instead of 4 on the slowest stage."""
    L = dims.n_layer
    if n_stages == 1:
        return [(0.0, float(L))]
    if not 2 <= n_stages <= 2 * L + 1:
        raise ValueError(f"cannot split {L} layers into {n_stages} stages")
    t_a, t_m = half_costs_us(dims, esz)
    _, t_head = stage_costs_us(dims, esz)
    n = 2 * L
    pre = [0.0]
    for i in range(n):
        pre.append(pre[-1] + (t_a if i % 2 == 0 else t_m))
    INF = float("inf")
    # best[s][i]: smallest possible slowest-stage cost when the first s stages cover halves [0, i)
    best = [[INF] * (n + 1) for _ in range(n_stages + 1)]
    cut = [[0] * (n + 1) for _ in range(n_stages + 1)]
    best[0][0] = 0.0
    for s in range(1, n_stages + 1):
        last = s == n_stages
        for i in range(n + 1):
            for j in range(i + 1 if not last else i + 0, -1, -1):
                if j > i or (not last and j == i) or best[s - 1][j] == INF:
                    continue
                c = max(best[s - 1][j], pre[i] - pre[j] + (t_head if last else 0.0))
                if c < best[s][i] - 1e-9:
                    best[s][i], cut[s][i] = c, j
    bounds, i = [n], n
    for s in range(n_stages, 0, -1):
        i = cut[s][i]
        bounds.append(i)
    bounds.reverse()
    return [(bounds[k] / 2.0, bounds[k + 1] / 2.0) for k in range(n_stages)]


def stage_weight_names(dims: GPT2Dims, layer_begin: int, layer_end: int, has_embed: bool, has_head: bool,
                       begin_half: bool = False, end_half: bool = False) -> List[str]:
    """Tensors a stage holds.  begin_half: block `layer_begin` contributes only its MLP half (its attention half ran
    on the previous stage); end_half: block `layer_end - 1` contributes only its attention half (GP_FLAG_*_HALF)."""
    names = []
    for name, _, _ in weight_names(dims):
        if name.startswith("h."):
            li = int(name.split(".")[1])
            if layer_begin <= li < layer_end:
                attn = name.split(".", 2)[2].startswith(_ATTN_HALF)
                if (begin_half and li == layer_begin and attn) or (end_half and li == layer_end - 1 and not attn):
                    continue
                names.append(name)
        elif name == "wte.weight":
            if has_embed or has_head:  # tied lm_head: the last stage keeps its own copy
                names.append(name)
        elif name == "wpe.weight":
            if has_embed:
                names.append(name)
        elif has_head:
            names.append(name)
    return names


def init_weights(dims: GPT2Dims, seed: int, recipe: str = "hf", only: Optional[List[str]] = None) -> Dict[str, torch.Tensor]:
    """Seeded random init following GPT2PreTrainedModel._init_weights (N(0,0.02); c_proj
    scaled by 1/sqrt(2 n_layer); LayerNorm 1/0; zero biases).  recipe="dense" also randomises
    biases and LayerNorm affine.  One generator per tensor, so a stage can build only its own
    tensors and still agree bit-for-bit with every other holder of the same seed."""
    if recipe not in ("hf", "dense"):
        raise ValueError(recipe)
    w = {}
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
        else:
            t = 0.1 * torch.randn(shape, generator=g) if recipe == "dense" else torch.zeros(shape)
        w[name] = t.contiguous()
    return w


# --------------------------------------------------------------------------- checkpoints
def dims_from_config(cfg: dict) -> GPT2Dims:
    """A Hugging Face GPT-2 `config.json` -> shapes (`[hf] models/gpt2/configuration_gpt2.py`)."""
    act = cfg.get("activation_function", "gelu_new")
    if act != "gelu_new":
        raise ValueError(f"activation_function={act!r}: only GPT-2's gelu_new is implemented")
    H = int(cfg["n_embd"])
    if cfg.get("n_inner") not in (None, 4 * H):
        raise ValueError(f"n_inner={cfg['n_inner']}: the kernels implement the GPT-2 MLP width 4*n_embd={4 * H} only")
    for flag in ("scale_attn_by_inverse_layer_idx", "reorder_and_upcast_attn"):
        if cfg.get(flag):
            raise ValueError(f"{flag}=True changes the attention arithmetic ([hf] modeling_gpt2.py) and is not implemented")
    if cfg.get("scale_attn_weights", True) is False:
        raise ValueError("scale_attn_weights=False is not implemented (scores are always scaled by head_dim**-0.5)")
    return GPT2Dims(n_layer=int(cfg["n_layer"]), n_embd=int(cfg["n_embd"]), n_head=int(cfg["n_head"]),
                    vocab_size=int(cfg.get("vocab_size", 50257)),
                    n_positions=int(cfg.get("n_positions", cfg.get("n_ctx", 1024))),
                    layer_norm_epsilon=float(cfg.get("layer_norm_epsilon", 1e-5)))


def _read_safetensors(path: str) -> Dict[str, torch.Tensor]:
    """The safetensors container: u64 header length, JSON header {name: {dtype, shape, data_offsets}}, raw data.
    Read directly so that a checkpoint loads without the `safetensors` package."""
    import json
    import numpy as np
    kinds = {"F32": (np.float32, torch.float32), "F16": (np.float16, torch.float16), "BF16": (np.uint16, torch.bfloat16),
             "F64": (np.float64, torch.float64), "I64": (np.int64, torch.int64), "I32": (np.int32, torch.int32)}
    with open(path, "rb") as f:
        n = int.from_bytes(f.read(8), "little")
        header = json.loads(f.read(n))
        base = 8 + n
        out = {}
        for name, meta in header.items():
            if name == "__metadata__":
                continue
            if meta["dtype"] not in kinds:
                continue   # e.g. the boolean causal-mask buffers `attn.bias` of old GPT-2 exports: not weights
            np_dt, t_dt = kinds[meta["dtype"]]
            b, e = meta["data_offsets"]
            f.seek(base + b)
            arr = np.frombuffer(f.read(e - b), dtype=np_dt).copy()
            t = torch.from_numpy(arr)
            if t_dt == torch.bfloat16:
                t = t.view(torch.bfloat16)
            out[name] = t.reshape(meta["shape"])
    return out


def load_checkpoint(path: str):
    """GPT-2 weights from disk -> (dims or None, state_dict).  `path` is a Hugging Face model directory
    (`config.json` + `model.safetensors` or `pytorch_model.bin`), a `.safetensors` file, or a torch `.bin/.pt`
    state_dict.  Names may carry the `transformer.` prefix; Conv1D matrices stay `[in, out]`
    (`[hf] pytorch_utils.py:119-123`) - the engine transposes them at load; `lm_head.weight` is the
    tied `wte.weight` (`[hf] modeling_gpt2.py:646`) and is ignored."""
    import json
    import os
    dims = None
    file = path
    if os.path.isdir(path):
        cfg = os.path.join(path, "config.json")
        if os.path.exists(cfg):
            with open(cfg) as f:
                dims = dims_from_config(json.load(f))
        for cand in ("model.safetensors", "pytorch_model.bin"):
            if os.path.exists(os.path.join(path, cand)):
                file = os.path.join(path, cand)
                break
        else:
            raise FileNotFoundError(f"{path}: neither model.safetensors nor pytorch_model.bin")
    if file.endswith(".safetensors"):
        sd = _read_safetensors(file)
    else:
        sd = torch.load(file, map_location="cpu", weights_only=True)
    sd = {k: v.to(torch.float32) for k, v in sd.items() if torch.is_tensor(v) and v.is_floating_point()}
    return dims, sd
