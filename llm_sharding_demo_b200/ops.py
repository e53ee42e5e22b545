"""Op-level entry points of the C ABI on torch CUDA tensors (used by the parity tests and for
ncu).  Weight-like operands must already be in the storage dtype (fp32 or bf16) and laid out
[N, K] (K contiguous)."""
from __future__ import annotations

import ctypes as C

import torch

from ._lib import GP_BF16, GP_F32, GP_EPI_GELU, GP_GEMM_SIMT, GP_OUT_BF16, check, load_library


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _dt(t):
    if t.dtype == torch.float32:
        return GP_F32
    if t.dtype == torch.bfloat16:
        return GP_BF16
    raise TypeError(t.dtype)


def _s(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def embed(ids, pos, wte, wpe):
    lib = load_library()
    M, H = ids.numel(), wte.shape[1]
    out = torch.empty((M, H), dtype=torch.float32, device=wte.device)
    check(lib.gp_op_embed(_p(ids), _p(pos), _p(wte), _p(wpe), _p(out), M, H, _dt(wte), _s(wte)))
    return out


def layernorm(x, gamma, beta, eps=1e-5, out_bf16=False):
    lib = load_library()
    M, H = x.shape
    out = torch.empty((M, H), dtype=torch.bfloat16 if out_bf16 else torch.float32, device=x.device)
    check(lib.gp_op_layernorm(_p(x), _p(gamma), _p(beta), _p(out), M, H, eps, GP_OUT_BF16 if out_bf16 else 0, _s(x)))
    return out


def gemv(x, Wt, bias=None, residual=None, ln=None, eps=1e-5, gelu=False):
    lib = load_library()
    B, K = x.shape
    N = Wt.shape[0]
    out = torch.empty((B, N), dtype=torch.float32, device=x.device)
    g, b = (ln if ln is not None else (None, None))
    check(lib.gp_op_gemv(_p(x), _p(Wt), _p(bias), _p(residual), _p(out), B, N, K, _p(g), _p(b), eps,
                         GP_EPI_GELU if gelu else 0, _dt(Wt), _s(x)))
    return out


def gemm(A, Wt, bias=None, residual=None, gelu=False, out_bf16=False, simt=False):
    lib = load_library()
    M, K = A.shape
    N = Wt.shape[0]
    assert A.dtype == Wt.dtype
    out = torch.empty((M, N), dtype=torch.bfloat16 if out_bf16 else torch.float32, device=A.device)
    flags = (GP_EPI_GELU if gelu else 0) | (GP_OUT_BF16 if out_bf16 else 0) | (GP_GEMM_SIMT if simt else 0)
    check(lib.gp_op_gemm(_p(A), _p(Wt), _p(bias), _p(residual), _p(out), M, N, K, flags, _dt(Wt), _s(A)))
    return out


def attn_decode(q, kcache, vcache, ctx_len):
    lib = load_library()
    B, nh, max_ctx, d = kcache.shape
    out = torch.empty((B, nh * d), dtype=torch.float32, device=q.device)
    check(lib.gp_op_attn_decode(_p(q), _p(kcache), _p(vcache), _p(ctx_len), _p(out), B, nh, d, max_ctx, _dt(kcache), _s(q)))
    return out


def attn_prefill(qkv, nh):
    lib = load_library()
    B, T, H3 = qkv.shape
    H = H3 // 3
    out = torch.empty((B, T, H), dtype=qkv.dtype, device=qkv.device)
    check(lib.gp_op_attn_prefill(_p(qkv), _p(out), B, T, nh, H // nh, _dt(qkv), _s(qkv)))
    return out


def lmhead_scratch(device):
    lib = load_library()
    n = lib.gp_lmhead_scratch_bytes(8, 0)
    return torch.zeros((n,), dtype=torch.uint8, device=device)  # the ticket must start at zero


def lmhead_argmax(x, gamma, beta, wte, eps=1e-5, want_logits=False, scratch=None):
    lib = load_library()
    B, H = x.shape
    V = wte.shape[0]
    if scratch is None:
        scratch = lmhead_scratch(x.device)
    ids = torch.empty((B,), dtype=torch.int32, device=x.device)
    logits = torch.empty((B, V), dtype=torch.float32, device=x.device) if want_logits else None
    check(lib.gp_op_lmhead_argmax(_p(x), _p(gamma), _p(beta), _p(wte), _p(ids), _p(logits), _p(scratch), B, V, H, eps,
                                  _dt(wte), _s(x)))
    return (ids, logits) if want_logits else ids


def sample_topk(logits, temperature, top_k, seed, row0=0, position=0):
    """The reference coordinator's token selection (server.py:187-205) on fp32 logits [B,V]:
    logits / temperature -> top-k -> softmax -> one draw per row; row b draws with the Philox
    counter (position, row0 + b) under key `seed`.  Returns int32 ids [B]."""
    lib = load_library()
    assert logits.dtype == torch.float32 and logits.dim() == 2 and logits.is_contiguous()
    B, V = logits.shape
    ids = torch.empty((B,), dtype=torch.int32, device=logits.device)
    check(lib.gp_op_sample_topk(_p(logits), B, V, float(temperature), int(top_k), int(seed) & (2**64 - 1), int(row0),
                                int(position), _p(ids), _s(logits)))
    return ids


def lmhead_argmax_static(x, gamma, beta, wte, eps=1e-5, want_logits=False):
    """ln_f + lm_head + argmax for callers that cannot allocate device memory: like
    `lmhead_argmax` but the kernel's cross-CTA scratch is a fixed array inside the library, so
    nothing is allocated by the call itself (the outputs here are ordinary caller tensors).
    B <= 8; calls that overlap on the GPU race on the shared array, so keep them on one stream."""
    lib = load_library()
    B, H = x.shape
    V = wte.shape[0]
    ids = torch.empty((B,), dtype=torch.int32, device=x.device)
    logits = torch.empty((B, V), dtype=torch.float32, device=x.device) if want_logits else None
    check(lib.gp_op_lmhead_argmax_static(_p(x), _p(gamma), _p(beta), _p(wte), _p(ids), _p(logits), B, V, H, eps,
                                         _dt(wte), _s(x)))
    return (ids, logits) if want_logits else ids
