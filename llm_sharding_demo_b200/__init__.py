"""B200-native pipeline-sharded GPT-2 inference: the hot path of
kanchan-rihan/llm-sharding-demo (`server.py`) rebuilt on hand-written sm_100a CUDA kernels
behind a C ABI (`include/gpt2pipe.h`, `libgpt2pipe.so`).  There is no CPU fallback: every
compute entry point raises if the CUDA library or a CUDA device is missing."""
from .weights import (GPT2Dims, PRESETS, init_weights, split_layers, balanced_split, balanced_split_halves,  # noqa: F401
                      stage_flags, stage_weight_names, load_checkpoint, dims_from_config)
from ._lib import load_library, LibraryMissing, GPError  # noqa: F401
from .engine import StageEngine, GP_F32, GP_BF16  # noqa: F401

__all__ = ["GPT2Dims", "PRESETS", "init_weights", "split_layers", "balanced_split", "balanced_split_halves", "stage_flags", "stage_weight_names", "load_checkpoint",
           "dims_from_config", "load_library",
           "LibraryMissing", "GPError", "StageEngine", "GP_F32", "GP_BF16"]
