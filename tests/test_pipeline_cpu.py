"""Host-side pipeline logic on CPU: the systolic tick schedule and, with world_size 2 and 3
over gloo, the ring exchange (hidden forward, token ids back to stage 0) driven by the same
`run_ticks` the NCCL path uses, with a toy stage in place of the CUDA engine."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from llm_sharding_demo_b200.pipeline import run_ticks, tick_schedule


def test_tick_schedule_properties():
    for S in (1, 2, 4, 8):
        M, R = S, 5
        sched = tick_schedule(S, M, R)
        assert len(sched) == R * M + S - 1
        seen = {}
        for t, row in enumerate(sched):
            for s, item in enumerate(row):
                if item is not None:
                    seen.setdefault(item, []).append((t, s))
        assert len(seen) == R * M
        for (r, m), visits in seen.items():
            # each item visits stages 0..S-1 on consecutive ticks
            assert [s for _, s in visits] == list(range(S))
            assert [t for t, _ in visits] == list(range(visits[0][0], visits[0][0] + S))
            # ring closure: round r of stream m leaves the last stage before round r+1 enters stage 0
            if (r + 1, m) in seen:
                assert seen[(r + 1, m)][0][0] >= visits[-1][0] + 1
    with pytest.raises(ValueError):
        tick_schedule(4, 2, 3)


def toy_stage(rank, x):
    return x * 3 + (rank + 1)


def serial_reference(S, M, R, prompts):
    """what the ring must produce: per stream, ids_{r} = f(stages(ids_{r-1}))"""
    out = {m: [] for m in range(M)}
    for m in range(M):
        cur = prompts[m].clone()
        for r in range(R):
            h = cur.float()
            for s in range(S):
                h = toy_stage(s, h)
            cur = (h.long() % 1000).to(torch.int32)
            out[m].append(cur.clone())
    return out


def _worker(rank, world, port, R, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    M, B = world, 2
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    prompts = [torch.arange(B, dtype=torch.int32) + 10 * m for m in range(M)]
    collected = {}

    def exchange(send_t, recv_kind):
        reqs = []
        if send_t is not None:
            reqs.append(dist.isend(send_t, nxt))
        recv_t = None
        if recv_kind is not None:
            recv_t = torch.empty((B,), dtype=torch.int32 if recv_kind == "ids" else torch.float32)
            reqs.append(dist.irecv(recv_t, prv))
        for q in reqs:
            q.wait()
        return recv_t

    def compute(r, m, got):
        if rank == 0:
            x = (prompts[m] if r == 0 else got).float()
        else:
            x = got
        h = toy_stage(rank, x)
        return (h.long() % 1000).to(torch.int32) if rank == world - 1 else h

    def on_ids(r, m, ids):
        collected[(r, m)] = ids.clone()

    run_ticks(rank, world, M, R, exchange, compute, on_ids)
    if rank == 0:
        ret.put({k: v.tolist() for k, v in collected.items()})
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_ring_pipeline_over_gloo(world):
    R = 4
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = 29600 + world + (os.getpid() % 200)
    procs = [ctx.Process(target=_worker, args=(r, world, port, R, ret)) for r in range(world)]
    for p in procs:
        p.start()
    got = ret.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    M, B = world, 2
    prompts = [torch.arange(B, dtype=torch.int32) + 10 * m for m in range(M)]
    ref = serial_reference(world, M, R, prompts)
    assert len(got) == R * M
    for (r, m), ids in got.items():
        assert ids == ref[m][r].tolist(), (r, m)


def test_half_block_splits_cover_the_model_and_partition_its_weights():
    """weights.balanced_split_halves: contiguous stages in half-block units that cover [0, n_layer]; every stage but
    the last has work; the stage flags and the per-stage tensor lists partition the model's tensors exactly (the tied
    wte is held by the first and the last stage); a cut never makes the predicted slowest stage slower than the
    block-granular split."""
    import llm_sharding_demo_b200 as L
    from llm_sharding_demo_b200.pipeline import _slowest
    from llm_sharding_demo_b200.weights import weight_names
    for name in ("gpt2", "gpt2-medium", "gpt2-xl", "mini"):
        dims = L.PRESETS[name] if name in L.PRESETS else None
        if dims is None:
            continue
        for S in (1, 2, 4, 8):
            if S > 2 * dims.n_layer + 1:
                continue
            st = L.balanced_split_halves(dims, S)
            assert len(st) == S and st[0][0] == 0.0 and st[-1][1] == float(dims.n_layer)
            assert all(st[k][1] == st[k + 1][0] for k in range(S - 1))
            assert all(e > b for b, e in st[:-1]) and st[-1][1] >= st[-1][0]
            assert all((2 * x) == int(2 * x) for be in st for x in be)
            if 2 <= S <= dims.n_layer + 1:
                assert _slowest(dims, st) <= _slowest(dims, L.balanced_split(dims, S)) + 1e-6
            held = []
            for k, (b, e) in enumerate(st):
                lb, le, bh, eh = L.stage_flags(b, e)
                held += L.stage_weight_names(dims, lb, le, k == 0, k == S - 1, bh, eh)
            every = [n for n, _, _ in weight_names(dims)]
            extra_wte = 1 if S > 1 else 0
            assert sorted(held) == sorted(every + ["wte.weight"] * extra_wte)
    assert L.balanced_split_halves(L.PRESETS["gpt2"], 4) == [(0.0, 3.5), (3.5, 7.0), (7.0, 10.5), (10.5, 12.0)]
    assert L.stage_flags(3.5, 7.0) == (3, 7, True, False) and L.stage_flags(0.0, 3.5) == (0, 4, False, True)
    assert L.stage_flags(12.0, 12.0) == (12, 12, False, False)
