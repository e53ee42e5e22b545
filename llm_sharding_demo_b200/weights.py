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


def stage_weight_names(dims: GPT2Dims, layer_begin: int, layer_end: int, has_embed: bool, has_head: bool) -> List[str]:
    names = []
    for name, _, _ in weight_names(dims):
        if name.startswith("h."):
            if layer_begin <= int(name.split(".")[1]) < layer_end:
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
