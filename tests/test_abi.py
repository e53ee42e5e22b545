"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/gpt2pipe.h declares; the product package never routes through the oracle; no compute
is attempted without a GPU."""
import ctypes
import os
import re

import pytest

import llm_sharding_demo_b200 as L
from llm_sharding_demo_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "gpt2pipe.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = header_symbols()
    assert "gp_engine_create" in names and "gp_generate" in names and len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in gpt2pipe.h but not exported"
    # and the ctypes table binds exactly the declared set
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_error_channel():
    lib = L.load_library()
    assert lib.gp_version() == 200
    cfg = _lib.GPConfig(n_layer=0)
    h = ctypes.c_void_p()
    rc = lib.gp_engine_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc == -1 and b"invalid configuration" in lib.gp_last_error(None)
    assert lib.gp_engine_create(None, None) == -1


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(L.LibraryMissing):
        L.load_library(str(tmp_path / "libgpt2pipe.so"))


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(L.GPError):
        L.StageEngine(L.PRESETS["tiny-gpt2"])


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "llm_sharding_demo_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "/root/reference" not in text, f


def test_weights_match_oracle_generator():
    import torch
    from oracle import gpt2_oracle as O
    for preset, seed, recipe in (("tiny-gpt2", 0, "hf"), ("gpt2", 1234, "hf")):
        dims = L.PRESETS[preset]
        names = ["wpe.weight", "h.1.attn.c_proj.weight", "h.0.mlp.c_fc.bias", "ln_f.weight"]
        a = L.init_weights(dims, seed, recipe, only=names)
        b = O.init_weights(O.PRESETS[preset], seed, recipe, only=names)
        for n in names:
            assert torch.equal(a[n], b[n])
    assert L.split_layers(12, 2) == O.split_layers(12, 2) == [(0, 6), (6, 12)]
    assert L.split_layers(48, 8) == O.split_layers(48, 8)
