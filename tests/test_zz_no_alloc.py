"""gp_op_lmhead_argmax_static: the lm_head + argmax entry point for hosts that cannot allocate
device memory (their HBM is owned by another allocator, or their decode loop must not hit the
device-wide sync of cudaMalloc).  Its scratch is a fixed array inside the library instead of a
caller buffer, so it must (a) give exactly what gp_op_lmhead_argmax gives, (b) keep doing so
on back-to-back calls although the array still holds the previous call's contents, and
(c) refuse sizes the fixed array cannot serve.  Collected last (test_zz_*), and written to pass
both in a fresh process and after every other test has run."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(B, V, H, seed, dt):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, H, generator=g)
    wte = (torch.randn(V, H, generator=g) * 0.02).to(dt)
    gm = 1 + 0.1 * torch.randn(H, generator=g)
    bt = 0.1 * torch.randn(H, generator=g)
    return x.to(DEV), gm.to(DEV), bt.to(DEV), wte.to(DEV)


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float32])
def test_static_scratch_entry_matches_allocating_entry_twice_in_a_row(dt):
    from llm_sharding_demo_b200 import ops
    # two different problems back to back: the second call starts from the first one's leftovers
    # (more CTAs and more batch rows first, then fewer, so stale partials would be picked up)
    for (B, V, H, seed) in [(8, 50257, 768, 1), (3, 1041, 128, 2)]:
        x, gm, bt, wte = _inputs(B, V, H, seed, dt)
        ref_ids, ref_logits = ops.lmhead_argmax(x, gm, bt, wte, want_logits=True)
        got_ids, got_logits = ops.lmhead_argmax_static(x, gm, bt, wte, want_logits=True)
        torch.cuda.synchronize()
        # same kernel, same launch shape: the tolerance of tests/test_gpu_ops.py (2e-4 fp32,
        # 2e-3 bf16 storage) is the bar; in practice the results are bit-identical
        tol = 2e-4 if dt == torch.float32 else 2e-3
        assert float((got_logits - ref_logits).abs().max()) <= tol
        assert torch.equal(got_ids, ref_ids)
        # and against plain fp32 torch on the storage-rounded weights (ln_f -> h wte^T -> argmax, server.py:101-102)
        xn = torch.nn.functional.layer_norm(x, (H,), gm, bt, 1e-5)
        want = xn @ wte.float().t()
        assert float((got_logits - want).abs().max()) <= 5e-3
        top2 = torch.topk(want, 2, dim=-1).values
        firm = (top2[:, 0] - top2[:, 1]) > 1e-2
        assert torch.equal(got_ids[firm].long(), want.argmax(-1)[firm])
        again = ops.lmhead_argmax_static(x, gm, bt, wte)
        torch.cuda.synchronize()
        assert torch.equal(again, ref_ids)


def test_static_scratch_entry_rejects_oversize_and_allocates_nothing():
    from llm_sharding_demo_b200 import ops, load_library, GPError
    lib = load_library()
    x, gm, bt, wte = _inputs(2, 523, 96, 3, torch.bfloat16)
    ops.lmhead_argmax_static(x, gm, bt, wte)          # first call on the device does the set-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    ids = torch.empty((2,), dtype=torch.int32, device=DEV)
    p = lambda t: C.c_void_p(t.data_ptr())
    for _ in range(3):
        rc = lib.gp_op_lmhead_argmax_static(p(x), p(gm), p(bt), p(wte), p(ids), None, 2, 523, 96, 1e-5, 1, None)
        assert rc == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0                             # no device memory was taken by the calls
    # 9 rows is one more than the fixed array has partial slots for
    rc = lib.gp_op_lmhead_argmax_static(p(x), p(gm), p(bt), p(wte), p(ids), None, 9, 523, 96, 1e-5, 1, None)
    assert rc == -1
    msg = lib.gp_last_error(None).decode()
    assert "static scratch" in msg and "B <= 8" in msg
    # a hidden size whose rows do not fit shared memory is refused too
    rc = lib.gp_op_lmhead_argmax_static(p(x), p(gm), p(bt), p(wte), p(ids), None, 8, 523, 8192, 1e-5, 1, None)
    assert rc == -1
    with pytest.raises(GPError):
        ops.lmhead_argmax_static(torch.zeros((9, 96), device=DEV), gm, bt, wte)
