"""Pipeline of stages over GPUs: one process per GPU, stage s = blocks split_layers()[s].

world == 1 : a single-stage engine; the whole greedy loop runs on the device (`gp_generate`).
world  > 1 : the reference's A -> B handoff (server.py:172-181, HTTP/JSON through the
             coordinator) becomes an on-device exchange between neighbouring ranks, and the
             logits -> coordinator hop becomes a ring of int32 token ids from the last stage
             back to stage 0.  `world` micro-batches ("streams") are kept in flight so every
             stage works every tick.  Transport: torch.distributed (NCCL) grouped send/recv.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ._lib import GP_BF16
from .weights import (GPT2Dims, balanced_split, balanced_split_halves, half_costs_us, init_weights, split_layers, stage_costs_us,
                      stage_flags, stage_weight_names)


def tick_schedule(n_stages: int, n_streams: int, n_rounds: int) -> List[List[Optional[Tuple[int, int]]]]:
    """schedule[t][s] = (round, stream) that stage s computes at tick t, or None.
    Item (r, m) enters stage 0 at tick r*n_streams + m and stage s at that + s; with
    n_streams >= n_stages the ids a round produces are back at stage 0 before the next round
    needs them.  Pure function (unit-tested on CPU)."""
    if n_streams < n_stages and n_rounds > 1:
        raise ValueError("need at least one stream per stage to close the token ring in time")
    n_ticks = n_rounds * n_streams + n_stages - 1
    sched = []
    for t in range(n_ticks):
        row = []
        for s in range(n_stages):
            i = t - s
            row.append((i // n_streams, i % n_streams) if 0 <= i < n_rounds * n_streams else None)
        sched.append(row)
    return sched


def run_ticks(rank: int, n_stages: int, n_streams: int, n_rounds: int, exchange, compute, on_ids):
    """Drives one rank through the systolic schedule.  Transport- and engine-agnostic so the
    same code runs under NCCL on GPUs and under gloo with a toy stage on CPUs (tests).

      exchange(send_obj, recv_kind) -> received object or None
          send_obj : what this rank computed last tick (None = nothing to send)
          recv_kind: None | "ids" | "hidden_prefill" | "hidden_decode" (what the previous
                     rank in the ring computed last tick)
      compute(round, stream, received) -> object to pass on next tick
      on_ids(round, stream, ids)       -> stage 0 only: ids produced by the last stage
    """
    if n_rounds <= 0:
        return
    first = rank == 0
    prv = (rank - 1) % n_stages
    sched = tick_schedule(n_stages, n_streams, n_rounds)
    sched.append([None] * n_stages)          # drain tick: the last ids travel to stage 0
    pending = None
    for t, row in enumerate(sched):
        prev_item = sched[t - 1][prv] if t > 0 else None
        recv_kind = None
        if prev_item is not None:
            recv_kind = "ids" if first else ("hidden_prefill" if prev_item[0] == 0 else "hidden_decode")
        got = exchange(pending, recv_kind)
        pending = None
        if first and prev_item is not None:
            on_ids(prev_item[0], prev_item[1], got)
        item = row[rank]
        if item is not None:
            pending = compute(item[0], item[1], got)


def _slowest(dims: GPT2Dims, stages) -> float:
    """Predicted time of the slowest stage of a split (weights.stage_costs_us / half_costs_us), boundaries in blocks."""
    t_a, t_m = half_costs_us(dims)
    t_head = stage_costs_us(dims)[1]
    worst = 0.0
    for k, (b, e) in enumerate(stages):
        c = sum(t_a if i % 2 == 0 else t_m for i in range(int(round(2 * b)), int(round(2 * e))))
        worst = max(worst, c + (t_head if k == len(stages) - 1 else 0.0))
    return worst


class PipelineRunner:
    def __init__(self, dims: GPT2Dims, seed: int = 0, dtype: int = GP_BF16, batch_per_stream: int = 1,
                 max_ctx: Optional[int] = None, rank: int = 0, world: int = 1, device: int = 0, recipe: str = "hf",
                 weights=None, flags: int = 0, transport: str = "p2p", split="auto", n_streams: Optional[int] = None):
        """split: "even" = equal block counts (the reference's SPLIT_AT rule generalised), "balanced" = block counts
        chosen so that the stage that also carries ln_f + lm_head is not the slowest (it may get no block at all),
        "halves" = the same with boundaries also allowed between the attention and the MLP half of a block
        (self.stages then holds x.5 values), "auto" = where the bf16 persistent kernel runs (it is what executes a
        head-only stage and a cut block at full speed) halves when that makes the slowest stage at least 2 % lighter,
        else balanced; even everywhere else.  A list of (begin, end) pairs in blocks sets the boundaries explicitly."""
        from .engine import StageEngine
        self.dims, self.rank, self.world, self.B = dims, rank, world, batch_per_stream
        self.n_streams = n_streams or world
        if self.n_streams < world and transport != "p2p":
            raise ValueError("the host-driven (NCCL) transport needs at least one stream per stage to close the token ring in time")
        self.device = torch.device("cuda", device)
        if split == "auto":
            ok = dtype == GP_BF16 and dims.head_dim == 64 and dims.n_embd % 64 == 0 and batch_per_stream <= 8 and transport == "p2p"
            split = "even"
            if ok and world > 1:
                split = "halves" if _slowest(dims, balanced_split_halves(dims, world)) < 0.98 * _slowest(dims, balanced_split(dims, world)) \
                    else "balanced"
        if isinstance(split, (list, tuple)):       # explicit stage boundaries in blocks (x.5 = inside block x)
            self.stages = [(float(b), float(e)) for b, e in split]
            if len(self.stages) != world or self.stages[0][0] != 0 or self.stages[-1][1] != dims.n_layer or \
                    any(self.stages[k][1] != self.stages[k + 1][0] for k in range(world - 1)):
                raise ValueError(f"stage boundaries {split} do not cover blocks [0, {dims.n_layer}] contiguously in {world} stages")
        elif split == "halves":
            self.stages = balanced_split_halves(dims, world)
        else:
            self.stages = balanced_split(dims, world) if split == "balanced" else split_layers(dims.n_layer, world)
        b, e, begin_half, end_half = stage_flags(*self.stages[rank])
        if any(x != int(x) for be in self.stages for x in be) and dtype != GP_BF16 and transport == "p2p" and world > 1:
            raise ValueError("stage boundaries inside a block need bf16 storage with the device-side transport (the fp32 "
                             "persistent kernel runs whole blocks only); use transport='nccl' or a block-granular split")
        self.first, self.last = rank == 0, rank == world - 1
        self.engine = StageEngine(dims, b, e, has_embed=self.first, has_head=self.last, dtype=dtype,
                                  max_batch=batch_per_stream * self.n_streams, max_ctx=max_ctx,
                                  max_tokens=batch_per_stream * (max_ctx or dims.n_positions), device=device,
                                  flags=flags, begin_half=begin_half, end_half=end_half)
        names = self.engine.weight_names()
        w = weights if weights is not None else init_weights(dims, seed, recipe, only=names)
        self.engine.load_weights(w)
        if world > 1:
            import torch.distributed as dist
            if not dist.is_initialized():
                dist.init_process_group("nccl", device_id=self.device)
            self.dist = dist
            self.transport = transport
            self.comm_epoch = 1
            if transport == "p2p":
                handles = [None] * world
                dist.all_gather_object(handles, self.engine.comm_export())
                self.engine.comm_connect(handles[(rank + 1) % world])
                dist.barrier()
        else:
            self.dist = None
            self.transport = "none"

    def set_sampler(self, temperature: float = 0.0, top_k: int = 0, seed: int = 0):
        """The reference coordinator's temperature / top-k / multinomial selection instead of greedy (engine.set_sampler);
        it lives on the stage that carries the lm_head, the other stages have nothing to switch."""
        if self.last:
            self.engine.set_sampler(temperature, top_k, seed)

    # ---- small collectives for the bench harness
    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def max_over_ranks(self, v: float) -> float:
        if self.dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    # ---- generation
    def generate(self, prompts: torch.Tensor, n_new: int, timings: bool = False):
        """prompts: CPU int32 [B*n_streams, P].  Returns CPU int32 [B*n_streams, P+n_new] on
        rank 0 (other ranks return None) and, with timings=True, a dict of device times."""
        if self.world == 1:
            if timings:
                out, tm = self.engine.generate(prompts, n_new, timings=True)
                tm["launches"] = self.engine.last_launch_count()
                return out, tm
            return self.engine.generate(prompts, n_new)
        return self._generate_pipelined(prompts, n_new, timings)

    def _generate_pipelined(self, prompts, n_new, timings):
        dist, eng, B, S, M = self.dist, self.engine, self.B, self.world, self.n_streams
        H = self.dims.n_embd
        P = prompts.shape[1]
        if P + max(n_new, 1) - 1 > eng.max_ctx:
            raise ValueError(f"prompt of {P} tokens + {n_new} new tokens exceeds the KV cache of {eng.max_ctx} positions")
        dev = self.device
        nxt, prv = (self.rank + 1) % S, (self.rank - 1) % S
        eng.kv_reset(-1)
        ev = lambda: torch.cuda.Event(enable_timing=True)
        e_start, e_prefill, e_end = ev(), ev(), ev()
        e_start.record()
        seq = None
        if self.first:
            seq = torch.empty((B * M, P + n_new), dtype=torch.int32, device=dev)
            seq[:, :P].copy_(prompts, non_blocking=True)
        bufs = {"ids": torch.empty((B,), dtype=torch.int32, device=dev),
                "hidden_prefill": torch.empty((B, P, H), dtype=torch.float32, device=dev),
                "hidden_decode": torch.empty((B, H), dtype=torch.float32, device=dev)}
        state = {"launches": 0, "step_events": []}

        def exchange(send_t, recv_kind):
            ops = []
            if send_t is not None:
                ops.append(dist.P2POp(dist.isend, send_t, nxt))
            recv_t = bufs[recv_kind] if recv_kind is not None else None
            if recv_t is not None:
                ops.append(dist.P2POp(dist.irecv, recv_t, prv))
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            return recv_t

        def on_ids(r, m, ids):
            seq[m * B:(m + 1) * B, P + r] = ids

        def compute(r, m, got):
            slot0 = m * B
            if r == 0:
                x = seq[slot0:slot0 + B, :P].contiguous() if self.first else got
                out = eng.prefill(x, slot0=slot0, use_cache=True, want_ids=self.last)
            else:
                out = eng.decode(got, slot0=slot0)
            state["launches"] += eng.last_launch_count()
            if self.last and timings and m == M - 1:
                e = ev()
                e.record()
                state["step_events"].append(e)
            if r == 0 and m == M - 1:
                e_prefill.record()
            return out["next_ids"] if self.last else out["hidden"]

        p2p = self.transport == "p2p" and n_new > 1
        run_ticks(self.rank, S, M, min(n_new, 1) if p2p else n_new, exchange, compute, on_ids)
        if n_new <= 0:
            e_prefill.record()
        kernel_ms = None
        if p2p:
            # decode: one persistent launch per stage; the stage boundary is peer-memory writes
            rounds = n_new - 1
            first = seq[:, P].contiguous() if self.first else None
            tokens, kernel_ms = eng.pipeline_decode(first, rounds, M, B, self.comm_epoch)
            self.comm_epoch += rounds
            state["launches"] += eng.last_launch_count()
            if self.first:
                seq[:, P + 1:] = tokens
        e_end.record()
        torch.cuda.synchronize(dev)
        result = seq.cpu() if self.first else None
        if not timings:
            return result
        tm = {"prefill_ms": e_start.elapsed_time(e_prefill), "decode_ms": e_prefill.elapsed_time(e_end),
              "step_ms": [], "launches": state["launches"]}
        if kernel_ms is not None:
            # every round advances all streams by one token; the persistent kernel has no per-round events
            tm["step_ms"] = [kernel_ms / (n_new - 1)] * (n_new - 1)
        se = state["step_events"]
        if kernel_ms is None and self.last and len(se) > 1:
            # one entry per round: time for ALL streams to advance one token
            tm["step_ms"] = [se[i - 1].elapsed_time(se[i]) for i in range(1, len(se))]
        gathered = [None] * S
        dist.all_gather_object(gathered, tm["step_ms"])
        tm["step_ms"] = gathered[S - 1]
        return result, tm
