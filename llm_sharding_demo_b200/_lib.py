"""ctypes binding of libgpt2pipe.so (the C ABI declared in include/gpt2pipe.h)."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgpt2pipe.so")

GP_OK = 0
GP_F32, GP_BF16 = 0, 1
GP_FLAG_NO_TCGEN05, GP_FLAG_NO_GRAPH, GP_FLAG_NO_MEGA = 1, 2, 4
GP_FLAG_BEGIN_HALF, GP_FLAG_END_HALF = 16, 32
GP_EPI_GELU, GP_OUT_BF16, GP_GEMM_SIMT = 1, 2, 4


class LibraryMissing(RuntimeError):
    """libgpt2pipe.so is not built: there is deliberately no CPU or PyTorch fallback."""


class GPError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gpt2pipe error {code}: {msg}")
        self.code = code


class GPConfig(C.Structure):
    _fields_ = [("n_layer", C.c_int32), ("n_embd", C.c_int32), ("n_head", C.c_int32), ("vocab_size", C.c_int32),
                ("n_positions", C.c_int32), ("ln_eps", C.c_float), ("layer_begin", C.c_int32),
                ("layer_end", C.c_int32), ("has_embed", C.c_int32), ("has_head", C.c_int32), ("dtype", C.c_int32),
                ("max_batch", C.c_int32), ("max_ctx", C.c_int32), ("max_tokens", C.c_int32), ("device", C.c_int32),
                ("flags", C.c_uint32)]


_P, _I, _U, _F = C.c_void_p, C.c_int32, C.c_uint32, C.c_float
# name -> (restype, argtypes); every symbol include/gpt2pipe.h declares
SIGNATURES = {
    "gp_version": (C.c_int, []),
    "gp_last_error": (C.c_char_p, [_P]),
    "gp_engine_create": (C.c_int, [C.POINTER(GPConfig), C.POINTER(_P)]),
    "gp_engine_destroy": (C.c_int, [_P]),
    "gp_load_weight": (C.c_int, [_P, C.c_char_p, _P, C.POINTER(C.c_int64), _I]),
    "gp_weights_ready": (C.c_int, [_P]),
    "gp_kv_reset": (C.c_int, [_P, _I, _P]),
    "gp_kv_lengths": (C.c_int, [_P, _P, _I, _P]),
    "gp_stage_prefill": (C.c_int, [_P, _P, _P, _I, _I, _I, _I, _P, _P, _I, _P, _P]),
    "gp_stage_decode": (C.c_int, [_P, _P, _P, _I, _I, _P, _P, _P, _P]),
    "gp_generate": (C.c_int, [_P, _P, _I, _I, _I, _P, _P]),
    "gp_set_sampler": (C.c_int, [_P, _F, _I, C.c_uint64]),
    "gp_last_launch_count": (C.c_int64, [_P]),
    "gp_comm_export": (C.c_int, [_P, _P]),
    "gp_comm_connect": (C.c_int, [_P, _P]),
    "gp_comm_connect_local": (C.c_int, [_P, _P]),
    "gp_stage_decode_slots": (C.c_int, [_P, _P, _P, _I, _P, _P, _P, _P, _P]),
    "gp_pipeline_decode": (C.c_int, [_P, _P, _I, _I, _I, _U, _P, _I, C.POINTER(C.c_float), _P]),
    "gp_op_embed": (C.c_int, [_P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "gp_op_layernorm": (C.c_int, [_P, _P, _P, _P, _I, _I, _F, _U, _P]),
    "gp_op_gemv": (C.c_int, [_P, _P, _P, _P, _P, _I, _I, _I, _P, _P, _F, _U, _I, _P]),
    "gp_op_gemm": (C.c_int, [_P, _P, _P, _P, _P, _I, _I, _I, _U, _I, _P]),
    "gp_op_attn_decode": (C.c_int, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "gp_op_attn_prefill": (C.c_int, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "gp_lmhead_scratch_bytes": (C.c_int64, [_I, _I]),
    "gp_op_sample_topk": (C.c_int, [_P, _I, _I, _F, _I, C.c_uint64, _I, _I, _P, _P]),
    "gp_op_lmhead_argmax": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _I, _P]),
    "gp_op_lmhead_argmax_static": (C.c_int, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _I, _P]),
}

_lib = None


def load_library(path: str = LIB_PATH):
    """dlopen the in-tree library and bind every exported symbol; raises LibraryMissing if it
    has not been built (`python -m llm_sharding_demo_b200.build`)."""
    global _lib
    if _lib is not None and path == LIB_PATH:
        return _lib
    if not os.path.exists(path):
        raise LibraryMissing(f"{path} not found: build it with `python -m llm_sharding_demo_b200.build` "
                             "(nvcc, sm_100a). This package has no CPU fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype, fn.argtypes = res, args
    if path == LIB_PATH:
        _lib = lib
    return lib


def check(rc: int, engine=None):
    if rc != GP_OK:
        msg = load_library().gp_last_error(engine)
        raise GPError(rc, (msg or b"").decode("utf-8", "replace"))
