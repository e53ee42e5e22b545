"""Host-side wrapper of one pipeline stage (`gp_engine`).  PyTorch is used only for device
buffers and streams; all arithmetic happens in libgpt2pipe.so."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from . import _lib
from ._lib import GPConfig, GPError, GP_BF16, GP_F32, GP_FLAG_BEGIN_HALF, GP_FLAG_END_HALF, check, load_library
from .weights import GPT2Dims, stage_weight_names


def _require_cuda():
    if not torch.cuda.is_available():
        raise GPError(-2, "no CUDA device: llm_sharding_demo_b200 has no CPU path")


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


class StageEngine:
    """Blocks [layer_begin, layer_end) of a GPT-2, plus embeddings (stage 0) and ln_f + tied
    lm_head (last stage).  S=2 stages are the reference's ShardA / ShardB (server.py:63-103)."""

    def __init__(self, dims: GPT2Dims, layer_begin: int = 0, layer_end: Optional[int] = None,
                 has_embed: bool = True, has_head: bool = True, dtype: int = GP_BF16, max_batch: int = 1,
                 max_ctx: Optional[int] = None, max_tokens: int = 0, device: int = 0, flags: int = 0,
                 begin_half: bool = False, end_half: bool = False):
        """begin_half / end_half: the stage starts at the MLP half of block `layer_begin` / stops after the attention
        half of block `layer_end - 1` (GP_FLAG_BEGIN_HALF / GP_FLAG_END_HALF, weights.balanced_split_halves)."""
        _require_cuda()
        self.lib = load_library()
        self.dims = dims
        self.layer_begin = layer_begin
        self.layer_end = dims.n_layer if layer_end is None else layer_end
        self.has_embed, self.has_head = bool(has_embed), bool(has_head)
        self.dtype = dtype
        self.max_batch = max_batch
        self.begin_half, self.end_half = bool(begin_half), bool(end_half)
        flags |= (GP_FLAG_BEGIN_HALF if begin_half else 0) | (GP_FLAG_END_HALF if end_half else 0)
        self.max_ctx = dims.n_positions if max_ctx is None else max_ctx
        self.device = torch.device("cuda", device)
        cfg = GPConfig(dims.n_layer, dims.n_embd, dims.n_head, dims.vocab_size, dims.n_positions,
                       dims.layer_norm_epsilon, self.layer_begin, self.layer_end, int(has_embed), int(has_head),
                       dtype, max_batch, self.max_ctx, max_tokens, device, flags)
        h = C.c_void_p()
        check(self.lib.gp_engine_create(C.byref(cfg), C.byref(h)))
        self._h = h

    # ---- lifecycle
    def close(self):
        if getattr(self, "_h", None):
            self.lib.gp_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def weight_names(self):
        return stage_weight_names(self.dims, self.layer_begin, self.layer_end, self.has_embed, self.has_head,
                                  self.begin_half, self.end_half)

    def load_weights(self, weights: Dict[str, torch.Tensor]):
        """`weights`: GPT2Model.state_dict()-style fp32 tensors (`transformer.` prefix optional)."""
        for name in self.weight_names():
            t = weights.get(name)
            if t is None:
                t = weights.get("transformer." + name)
            if t is None:
                raise KeyError(f"missing weight {name}")
            t = t.detach().to("cpu", torch.float32).contiguous()
            shape = (C.c_int64 * t.dim())(*t.shape)
            check(self.lib.gp_load_weight(self._h, name.encode(), C.c_void_p(t.data_ptr()), shape, t.dim()), self._h)
        check(self.lib.gp_weights_ready(self._h), self._h)

    # ---- cache
    def kv_reset(self, slot: int = -1):
        check(self.lib.gp_kv_reset(self._h, slot, _stream_ptr(self.device)), self._h)

    def set_sampler(self, temperature: float = 0.0, top_k: int = 0, seed: int = 0):
        """Token selection of this stage: greedy (default, also temperature <= 0 or top_k <= 0) or the
        reference's temperature / top-k / multinomial (server.py:187-205 uses 0.6 / 40).  The draw
        depends on (seed, KV slot, position of the new token) only."""
        check(self.lib.gp_set_sampler(self._h, float(temperature), int(top_k), int(seed) & (2**64 - 1)), self._h)

    def kv_lengths(self):
        buf = (C.c_int32 * self.max_batch)()
        check(self.lib.gp_kv_lengths(self._h, buf, self.max_batch, _stream_ptr(self.device)), self._h)
        return list(buf)

    # ---- forward
    def prefill(self, x: torch.Tensor, slot0: int = 0, use_cache: bool = True, want_logits: str = "none",
                want_ids: bool = False):
        """x: int32 ids [B,T] on stage 0, else fp32 hidden [B,T,H].  Returns a dict with
        `hidden` (non-last stage), `logits` ("all" -> [B,T,V], "last" -> [B,V]) and `next_ids`."""
        H, V = self.dims.n_embd, self.dims.vocab_size
        if self.has_embed:
            assert x.dtype == torch.int32 and x.dim() == 2 and x.is_cuda
            B, T = x.shape
            ids, hid = x.contiguous(), None
        else:
            assert x.dtype == torch.float32 and x.dim() == 3 and x.shape[2] == H and x.is_cuda
            B, T = x.shape[:2]
            ids, hid = None, x.contiguous()
        out = {}
        hidden_out = logits = next_ids = None
        if not self.has_head:
            hidden_out = torch.empty((B, T, H), dtype=torch.float32, device=self.device)
        else:
            if want_logits == "all":
                logits = torch.empty((B, T, V), dtype=torch.float32, device=self.device)
            elif want_logits == "last":
                logits = torch.empty((B, V), dtype=torch.float32, device=self.device)
            if want_ids:
                next_ids = torch.empty((B,), dtype=torch.int32, device=self.device)
        check(self.lib.gp_stage_prefill(self._h, _ptr(ids), _ptr(hid), B, T, slot0, int(use_cache), _ptr(hidden_out),
                                        _ptr(logits), int(want_logits == "all"), _ptr(next_ids),
                                        _stream_ptr(self.device)), self._h)
        if hidden_out is not None:
            out["hidden"] = hidden_out
        if logits is not None:
            out["logits"] = logits
        if next_ids is not None:
            out["next_ids"] = next_ids
        return out

    def decode(self, x: torch.Tensor, slot0: int = 0, want_logits: bool = False, hidden_out=None, next_ids=None):
        """One cached step. x: int32 ids [B] on stage 0, else fp32 hidden [B,H]."""
        H, V = self.dims.n_embd, self.dims.vocab_size
        B = x.shape[0]
        ids, hid = (x, None) if self.has_embed else (None, x)
        logits = None
        if not self.has_head:
            if hidden_out is None:
                hidden_out = torch.empty((B, H), dtype=torch.float32, device=self.device)
        else:
            if want_logits:
                logits = torch.empty((B, V), dtype=torch.float32, device=self.device)
            if next_ids is None:
                next_ids = torch.empty((B,), dtype=torch.int32, device=self.device)
        check(self.lib.gp_stage_decode(self._h, _ptr(ids), _ptr(hid), B, slot0, _ptr(hidden_out), _ptr(logits),
                                       _ptr(next_ids), _stream_ptr(self.device)), self._h)
        out = {}
        if hidden_out is not None:
            out["hidden"] = hidden_out
        if logits is not None:
            out["logits"] = logits
        if next_ids is not None and self.has_head:
            out["next_ids"] = next_ids
        return out

    def decode_slots(self, x: torch.Tensor, slots, want_logits: bool = False):
        """One cached step for the sequences living in KV slots `slots` (any distinct slot numbers; continuous
        batching).  x: int32 ids [B] on stage 0, else fp32 hidden [B,H]; row b belongs to slots[b]."""
        H, V = self.dims.n_embd, self.dims.vocab_size
        B = x.shape[0]
        assert len(slots) == B
        ids, hid = (x, None) if self.has_embed else (None, x)
        hidden_out = logits = next_ids = None
        if not self.has_head:
            hidden_out = torch.empty((B, H), dtype=torch.float32, device=self.device)
        else:
            if want_logits:
                logits = torch.empty((B, V), dtype=torch.float32, device=self.device)
            next_ids = torch.empty((B,), dtype=torch.int32, device=self.device)
        sl = (C.c_int32 * B)(*[int(v) for v in slots])
        check(self.lib.gp_stage_decode_slots(self._h, _ptr(ids), _ptr(hid), B, sl, _ptr(hidden_out), _ptr(logits),
                                             _ptr(next_ids), _stream_ptr(self.device)), self._h)
        out = {}
        if hidden_out is not None:
            out["hidden"] = hidden_out
        if logits is not None:
            out["logits"] = logits
        if next_ids is not None:
            out["next_ids"] = next_ids
        return out

    def generate(self, prompt_ids: torch.Tensor, n_new: int, timings: bool = False):
        """Decode on a single-stage engine (greedy unless set_sampler switched it). prompt_ids: CPU int32 [B,P] (pinned memory
        avoids a staging copy).  Returns CPU int32 [B,P+n_new] (+ timings dict)."""
        assert prompt_ids.device.type == "cpu" and prompt_ids.dtype == torch.int32 and prompt_ids.dim() == 2
        prompt_ids = prompt_ids.contiguous()
        B, P = prompt_ids.shape
        out = torch.empty((B, P + n_new), dtype=torch.int32)
        tm = (C.c_float * (2 + max(n_new, 0)))() if timings else None
        check(self.lib.gp_generate(self._h, C.c_void_p(prompt_ids.data_ptr()), B, P, n_new,
                                   C.c_void_p(out.data_ptr()), tm), self._h)
        if timings:
            return out, {"prefill_ms": tm[0], "decode_ms": tm[1], "step_ms": list(tm)[2:2 + max(n_new - 1, 0)]}
        return out

    # ---- device-side stage boundary (multi-GPU pipelines)
    def comm_export(self) -> bytes:
        buf = (C.c_ubyte * 64)()
        check(self.lib.gp_comm_export(self._h, buf), self._h)
        return bytes(buf)

    def comm_connect(self, peer_handle: bytes):
        buf = (C.c_ubyte * 64).from_buffer_copy(peer_handle)
        check(self.lib.gp_comm_connect(self._h, buf), self._h)

    def comm_connect_local(self, peer: "StageEngine"):
        """Same-process variant of comm_connect: `peer` is the engine this stage feeds."""
        check(self.lib.gp_comm_connect_local(self._h, peer._h), self._h)
        self._peer_keepalive = peer

    def pipeline_decode(self, first_ids, n_rounds: int, n_streams: int, B: int, comm_epoch0: int,
                        timeout_ms: int = 20000):
        """One persistent launch: n_rounds decode steps of n_streams micro-batches on this stage.
        Stage 0 passes `first_ids` (int32 [n_streams*B], device) and gets int32 [n_streams*B, n_rounds]."""
        tokens = None
        if self.has_embed:
            tokens = torch.empty((n_streams * B, n_rounds), dtype=torch.int32, device=self.device)
        ms = C.c_float(0.0)
        check(self.lib.gp_pipeline_decode(self._h, _ptr(first_ids), n_rounds, n_streams, B, comm_epoch0, _ptr(tokens),
                                          timeout_ms, C.byref(ms), _stream_ptr(self.device)), self._h)
        return tokens, float(ms.value)

    def last_launch_count(self) -> int:
        return int(self.lib.gp_last_launch_count(self._h))
