#!/usr/bin/env python
"""bench.py — decode tokens/s of the pipeline-sharded GPT-2 hot path on N B200s.

Contract: `python bench.py --gpus N --steps K --warmup W [--impl reference]` prints ONE JSON
line (rank 0).  A "step" is one greedy generation request of the workload: B prompts of P
tokens -> n_new tokens each (BASELINE.json configs[1]: GPT-2 124M, batch 1, 128 new tokens,
random-init seeded weights, synthetic prompt).  N GPUs = N pipeline stages; with N > 1 the
job carries N sequences (one micro-batch per stage in flight), i.e. weak scaling.

 value : generated tokens/s with the prompt already resident in HBM (device time, CUDA events)
 e2e   : the same through the public API with HOST buffers (pinned prompt H2D + ids D2H
         inside the timed region)
 roofline / cpu_baseline : see DESIGN.md §Measurement.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

class _Workloads(dict):
    """`custom:<preset>:<B per stream>:<P>:<n_new>` names a workload on the fly (configs[4] sweeps)."""
    def __missing__(self, key):
        if isinstance(key, str) and key.startswith("custom:"):
            _, preset, B, P, n = key.split(":")
            return (preset, int(B), int(P), int(n), 1234)
        raise KeyError(key)

    def __contains__(self, key):
        return dict.__contains__(self, key) or (isinstance(key, str) and key.startswith("custom:"))
METRIC = "decode tokens/s & p50 per-token latency, GPT-2 pipeline at 1/2/4/8 B200"


WORKLOADS = _Workloads({
    # name: (preset, B per stream, P, n_new, weight seed).  N GPUs = N stages x N streams of B sequences.
    "tiny-gpt2-hi-n20": ("tiny-gpt2", 1, 4, 20, 0),               # BASELINE.json configs[0]: prompt "Hi, " (byte ids), 20 tokens
    "gpt2-124m-b1-p32-n128": ("gpt2", 1, 32, 128, 1234),          # BASELINE.json configs[1] (the metric's config)
    "gpt2-medium-b8-p512-n128": ("gpt2-medium", 2, 512, 128, 1234),  # configs[2]: 4 stages x 4 streams x 2 = batch 8
    "gpt2-xl-b32-p896-n128": ("gpt2-xl", 4, 896, 128, 1234),      # configs[3]: 8 stages x 8 streams x 4 = batch 32, ctx 1024
})
DEFAULT_WORKLOAD = "gpt2-124m-b1-p32-n128"


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the persistent decode kernel, from the newest
    `profiles/rNN_traffic.json` (written by profiles/summarize.py from a `ncu --set full` capture of the default
    workload).  Returns (bytes or None, source)."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")))
    if not files:
        return None, None
    try:
        d = json.load(open(files[-1]))
        return int(d["dram_bytes_per_launch"]), os.path.relpath(files[-1], ROOT) + " (" + d.get("kernel", "?") + ")"
    except Exception:
        return None, None


def synthetic_prompts(B, P, vocab, seed=7):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab, (B, P), generator=g, dtype=torch.int32)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def algorithmic_bytes_per_step(dims, B, t_ctx, n_stages, esz=2):
    """weights once + KV read + KV write + stage hops, per decode step."""
    H, L, V = dims.n_embd, dims.n_layer, dims.vocab_size
    w = esz * (L * (12 * H * H + 13 * H) + V * H + 2 * H)
    kv = esz * 2 * L * H * B * t_ctx + esz * 2 * L * H * B
    hops = 4 * B * H * (n_stages - 1)
    return w + kv + hops


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------
_REF_WEIGHTS = {}
_REF_MODULES = {}


def workload_prompts(workload, n_seq, P, vocab):
    if workload == "tiny-gpt2-hi-n20":           # the reference README's example prompt under the byte tokenizer
        return torch.tensor([list("Hi, ".encode())] * n_seq, dtype=torch.int32)
    return synthetic_prompts(n_seq, P, vocab)


def reference_modules(preset, seed):
    """The reference's own ShardA / ShardB (server.py:68-103) built from oracle/_ref/server.py (or /root/reference)
    on the oracle's seeded weights, or None when the file is not there (then the oracle port is timed)."""
    from oracle import gpt2_oracle as O
    from oracle import ref_harness as RH
    key = (preset, seed)
    if key in _REF_MODULES:
        return _REF_MODULES[key]
    mods = None
    if RH.reference_available():
        dims = O.PRESETS[preset]
        if key not in _REF_WEIGHTS:
            _REF_WEIGHTS[key] = O.init_weights(dims, seed, "hf")
        split = O.split_layers(dims.n_layer, 2)[0][1]
        a = RH.load_reference_server("a", split, dims, _REF_WEIGHTS[key], tag=f"_{preset}")
        b = RH.load_reference_server("b", split, dims, _REF_WEIGHTS[key], tag=f"_{preset}")
        mods = (a, b)
    _REF_MODULES[key] = mods
    return mods


def cpu_reference_tokens_per_s(preset, seed, B, P, n_new, n_samples, full_path_probe=False, workload=None):
    """The reference's path on the host cores: full recompute of Shard A then Shard B per token (server.py:169-206),
    compute only (no HTTP/JSON): the reference's own ShardA / ShardB modules when oracle/_ref/server.py is present
    (kind "reference"), else the oracle port.  The reference's cost per token grows with the sequence, so `n_samples`
    steps evenly spread over the n_new-token run are timed (each one token at its true sequence length) and
    averaged.  Returns (tokens/s, p50 s/token, info)."""
    from oracle import gpt2_oracle as O
    dims = O.PRESETS[preset]
    key = (preset, seed)
    mods = reference_modules(preset, seed)
    if key not in _REF_WEIGHTS:
        _REF_WEIGHTS[key] = O.init_weights(dims, seed, "hf")
    w = _REF_WEIGHTS[key]
    split = O.split_layers(dims.n_layer, 2)[0][1]
    lat = []
    steps = sorted(set(int(round(i * (n_new - 1) / max(1, n_samples - 1))) for i in range(n_samples)))
    # bound the CPU work: at most 2048 token positions per timed forward (the per-sequence cost is what is
    # measured; the reference has no batching on /forward anyway, server.py:137)
    B_full, B = B, max(1, min(B, 2048 // (P + n_new)))
    with torch.no_grad():
        for t in steps:
            ids = synthetic_prompts(B, P + t, dims.vocab_size, seed=7 + t).long()
            t0 = time.perf_counter()
            if mods is not None:
                hidden = mods[0].shard_a(ids)
                logits = mods[1].shard_b(hidden)
            else:
                hidden = O.shard_a_forward(w, dims, ids, split)
                logits = O.shard_b_forward(w, dims, hidden, split)
            O.greedy_select(logits[:, -1])
            lat.append(time.perf_counter() - t0)
    info = {"kind": "reference" if mods is not None else "port", "sampled_steps": steps, "sequences_timed": B,
            "sequences_per_stream": B_full}
    if full_path_probe and hidden.shape[1] <= 200:
        # one token of the reference's real wire format: hidden and ALL-position logits as JSON (one sequence;
        # skipped for long sequences, where it alone would take minutes)
        t0 = time.perf_counter()
        hidden, logits = hidden[:1], logits[:1]
        h_json = json.dumps({"hidden_states": hidden.tolist()})
        hidden2 = torch.tensor(json.loads(h_json)["hidden_states"], dtype=torch.float32)
        l_json = json.dumps({"logits": logits.tolist()})
        _ = json.loads(l_json)["logits"][0][-1]
        info["json_overhead_s_per_token"] = round(time.perf_counter() - t0, 3)
        info["json_T"] = int(hidden2.shape[1])
    return B * len(lat) / sum(lat), statistics.median(lat), info


def reference_full_path(preset, seed, prompt_text, n_new):
    """the reference's REAL route path - coordinator /generate driving /forward and /forward_b
    through FastAPI, JSON bodies and all (in-process TestClient instead of three pods), its own sampler included -
    timed end to end.  Needs oracle/_ref/server.py.  Returns (tokens/s, seconds) or None."""
    from oracle import gpt2_oracle as O
    from oracle import ref_harness as RH
    if not RH.reference_available():
        return None
    key = ("full", preset, seed)
    if key not in _REF_MODULES:
        dims = O.PRESETS[preset]
        w = O.init_weights(dims, seed, "hf")
        _REF_MODULES[key] = RH.ReferencePipeline(dims, w, O.split_layers(dims.n_layer, 2)[0][1], tag=f"_full_{preset}")
        _REF_MODULES[key].generate(prompt_text, 2)   # warm-up
    pipe = _REF_MODULES[key]
    t0 = time.perf_counter()
    pipe.generate(prompt_text, n_new)
    dt = time.perf_counter() - t0
    return n_new / dt, dt


def ids_check(workload, dims, seed, dtype, out, P, n_new, device):
    """Parity of the ids the timed runs produced (rank 0, outside every timed region):
      (1) default workload: sequence 0 is the prompt of tests/golden/gpt2_124m.npz, whose continuation came from
          the reference's own /forward + /forward_b routes (oracle/make_golden.py): while the two sequences share
          their prefix, tokens must be equal at every position whose reference top-2 margin exceeds twice the
          storage tolerance (a flip below that margin is allowed and ends the comparison);
      (2) every sequence: a STATELESS full-sequence pass over the finished sequence on a separate single-stage
          engine (prefill kernels: tcgen05 GEMMs + causal attention, no KV cache, no persistent kernel, no stage
          boundary) must have the generated token as its argmax wherever its top-2 margin exceeds twice the tolerance.
    Returns {"checked": n, "mismatch": m, ...}; the caller fails the run when m > 0."""
    import numpy as np
    import llm_sharding_demo_b200 as L
    tol = 0.06 if dtype == L.GP_BF16 else 3e-4
    res = {"checked": 0, "mismatch": 0, "golden_checked": 0, "golden_mismatch": 0, "margin_rule": 2 * tol}
    gpath = os.path.join(ROOT, "tests", "golden", "gpt2_124m.npz")
    if workload == DEFAULT_WORKLOAD and os.path.exists(gpath):
        z = np.load(gpath)
        if z["prompt_ids"].tolist() == out[0, :P].tolist():
            ref = z["greedy_ids"][P:]
            for t in range(min(len(ref), n_new)):
                firm = float(z["margins"][t]) > 2 * tol
                same = int(out[0, P + t]) == int(ref[t])
                if firm:
                    res["golden_checked"] += 1
                    res["golden_mismatch"] += int(not same)
                if not same:     # a flip (allowed only below the margin rule) ends the common prefix
                    break
    if n_new > 0:
        eng = L.StageEngine(dims, dtype=dtype, max_batch=1, max_ctx=P + n_new, max_tokens=P + n_new, device=device)
        eng.load_weights(L.init_weights(dims, seed, "hf"))
        dev = torch.device("cuda", device)
        for b in range(out.shape[0]):
            full = eng.prefill(out[b:b + 1, :-1].to(dev), use_cache=False, want_logits="all")["logits"][0, P - 1:]
            top2 = torch.topk(full, 2, dim=-1)
            firm = ((top2.values[:, 0] - top2.values[:, 1]) > 2 * tol).cpu()
            pred = top2.indices[:, 0].cpu().to(torch.int32)
            res["checked"] += int(firm.sum())
            res["mismatch"] += int((pred[firm] != out[b, P:][firm]).sum())
        eng.close()
    res["sequences"] = int(out.shape[0])
    res["checked"] += res["golden_checked"]
    res["mismatch"] += res["golden_mismatch"]
    return res


def workload_config(workload, n_gpus, n_streams=None):
    """The `config` object of the JSON line - the same for both arms: N stages, `streams` micro-batches of
    `batch_per_stream` sequences in flight (default: one per stage; the reference arm works through the same sequences
    one request at a time)."""
    import llm_sharding_demo_b200.weights as W
    preset, B, P, n_new, _ = WORKLOADS[workload]
    dims = W.PRESETS[preset]
    bytes_step = algorithmic_bytes_per_step(dims, B, P + n_new // 2, 1, 2)
    return {"workload": workload, "model": preset, "batch_per_stream": B, "streams": n_streams or n_gpus, "prompt": P,
            "new_tokens": n_new, "stages": n_gpus,
            "l2": "inputs larger than L2: %.0f MB of weights and KV cache are streamed every decode step" % (bytes_step / 1e6)}


def run_reference(args):
    preset, B, P, n_new, seed = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    cores = torch.get_num_threads()
    n_streams = args.streams or args.gpus
    if args.workload == "tiny-gpt2-hi-n20":
        # configs[0] is small enough for the reference's FULL path: /generate -> /forward -> /forward_b over JSON
        vals = []
        for _ in range(args.warmup):
            reference_full_path(preset, seed, "Hi, ", 2)
        t_all = time.perf_counter()
        for _ in range(args.steps):
            r = reference_full_path(preset, seed, "Hi, ", n_new)
            if r is None:
                break
            vals.append(r)
        wall = time.perf_counter() - t_all
        if vals:
            value = n_new * len(vals) / sum(v[1] for v in vals)
            line = {"impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": "tokens/s", "n_gpus": args.gpus,
                    "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1000 * wall / args.steps, 3),
                    "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                    "config": workload_config(args.workload, args.gpus, n_streams),
                    "p50_token_ms": round(1000 * statistics.median(v[1] for v in vals) / n_new, 3),
                    "cpu_baseline": {"value": round(value, 3), "unit": "tokens/s", "cores": cores, "kind": "reference",
                                     "sample": f"the reference's own /generate (server.py:154-210, its sampler included) driving "
                                               f"/forward and /forward_b through FastAPI + JSON in one process, {n_new} tokens per step"},
                    "e2e": {"value": round(value, 3), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    "gpu_launches": 0}
            print(json.dumps(line))
            return
    sample = min(n_new, 8 if (P + n_new) <= 256 else 2)
    for _ in range(args.warmup):
        cpu_reference_tokens_per_s(preset, seed, B, P, n_new, 2)
    vals, lats = [], []
    t_all = time.perf_counter()
    info = {}
    for i in range(args.steps):
        v, p50, inf = cpu_reference_tokens_per_s(preset, seed, B, P, n_new, sample, full_path_probe=(i == 0))
        info.update(inf)
        vals.append(v)
        lats.append(p50)
    wall = time.perf_counter() - t_all
    value = len(vals) / sum(1.0 / v for v in vals)
    info["sequences_in_workload"] = B * n_streams
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1000 * wall / args.steps, 3),
        "ms_per_step_is": f"wall time of one bench step = {sample} sampled tokens (not a {n_new}-token request)",
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, args.gpus, n_streams),
        "p50_token_ms": round(1000 * statistics.median(lats), 3),
        "cpu_baseline": {"value": round(value, 3), "unit": "tokens/s", "cores": cores, "kind": info.get("kind", "port"),
                         "sample": f"{sample} of the {n_new} decode steps per bench step, evenly spread over the run, "
                                   f"each one token at its true sequence length; full recompute of both shards "
                                   f"per token (server.py:169-181), compute only (no HTTP/JSON); the CPU serves the "
                                   f"workload's {B * n_streams} sequences one after the other, so its aggregate rate is "
                                   f"this per-sequence rate; {info}"},
        "e2e": {"value": round(value, 3), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------
def run_b200(args):
    import llm_sharding_demo_b200 as L
    from llm_sharding_demo_b200.pipeline import PipelineRunner

    preset, B, P, n_new, seed = WORKLOADS[args.workload]
    B = args.batch_per_stream or B
    dims = L.PRESETS[preset]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torchrun for N>1)"
    torch.cuda.set_device(local)
    dtype = L.GP_BF16 if args.dtype == "bf16" else L.GP_F32
    runner = PipelineRunner(dims, seed=seed, dtype=dtype, batch_per_stream=B, max_ctx=P + n_new,
                            rank=rank, world=world, device=local, n_streams=args.streams or None)
    n_streams = runner.n_streams            # default = world: one micro-batch per stage in flight
    prompts = workload_prompts(args.workload, B * n_streams, P, dims.vocab_size).pin_memory()

    for _ in range(args.warmup):
        runner.generate(prompts, n_new)
    runner.barrier()
    torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # --- device-timed steps (value)
    dev_ms, step_tok_ms, launches, decode_ms = [], [], 0, []
    for _ in range(args.steps):
        out, tm = runner.generate(prompts, n_new, timings=True)
        dev_ms.append(tm["prefill_ms"] + tm["decode_ms"])
        decode_ms.append(tm["decode_ms"])
        step_tok_ms += tm["step_ms"]
        launches += tm["launches"]
    runner.barrier()
    torch.cuda.synchronize()
    # --- end-to-end steps through the public API, host buffers, wall clock
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = runner.generate(prompts, n_new)
    torch.cuda.synchronize()
    runner.barrier()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None

    dev_total_ms = runner.max_over_ranks(sum(dev_ms))
    e2e_s = runner.max_over_ranks(e2e_s)
    check = None
    if rank == 0 and not args.no_ids_check:
        check = ids_check(args.workload, dims, seed, dtype, out, P, n_new, local)
    runner.barrier()
    if rank != 0:
        return
    tokens_per_step = B * n_streams * n_new
    value = tokens_per_step * args.steps / (dev_total_ms / 1000.0)
    e2e = tokens_per_step * args.steps / e2e_s
    p50_tok_ms = statistics.median(step_tok_ms) if step_tok_ms else None

    # roofline of the dominant kernel.  world == 1: decode_mega_kernel, ONE launch per request that runs
    # all n_new-1 decode steps; its duration is the CUDA-event time between the end of the prefill and the
    # end of the loop on the engine's stream (gp_generate's own events), its algorithmic bytes are
    # (n_new-1) x the per-step figure of SURVEY.md 8(d) at the mean context of the run.
    # world > 1: the same kernel runs one step per launch on every stage; per-stage bytes / per-stage time.
    peak, peak_kind = measured_peaks()
    traffic, traffic_src = ncu_traffic()
    esz = 2 if dtype == L.GP_BF16 else 4
    t_ctx = P + n_new // 2
    bytes_step = algorithmic_bytes_per_step(dims, B, t_ctx, world, esz)
    roof = None
    decode_ms_mean = statistics.mean(decode_ms) if decode_ms else None
    if decode_ms_mean and n_new > 1:
        if world == 1:
            bytes_launch = bytes_step * (n_new - 1)
            achieved = bytes_launch / (decode_ms_mean / 1000.0) / 1e9
            mega = B <= 8 and dims.head_dim == 64 and dims.n_embd % 64 == 0
            kern = ("decode_mega2_kernel (persistent, %d decode steps per launch)" % (n_new - 1)) if mega else \
                "decode step (one kernel per op in a CUDA graph)"
        else:
            # one persistent launch per stage for the whole request: every round each stage streams its
            # share of the weights once per stream plus its share of the KV cache
            H_, L_, V_ = dims.n_embd, dims.n_layer, dims.vocab_size
            w_bytes = esz * (L_ * (12 * H_ * H_ + 13 * H_) + V_ * H_ + 2 * H_)
            kv_bytes = bytes_step - w_bytes
            bytes_round_all = n_streams * (w_bytes + kv_bytes)           # all stages together, per round
            bytes_launch = bytes_round_all * (n_new - 1) / world           # per stage launch
            achieved = bytes_launch / (decode_ms_mean / 1000.0) / 1e9
            kern = "decode_mega2_kernel (one persistent launch per stage, %d rounds x %d streams)" % (n_new - 1, n_streams)
        roof = {"bound": "hbm", "kernel": kern, "achieved": round(achieved, 1), "peak": peak, "peak_kind": peak_kind,
                "unit": "GB/s", "frac": round(achieved / peak, 4),
                "traffic": traffic if (world == 1 and args.workload == DEFAULT_WORKLOAD) else None,
                "traffic_source": traffic_src if (world == 1 and args.workload == DEFAULT_WORKLOAD) else None,
                "algorithmic_bytes_per_launch": int(bytes_launch),
                "note": "per-GPU figure; small-batch decode is latency-bound (dependent phases), see DESIGN.md"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
        sample = 16
        v, p50, info = cpu_reference_tokens_per_s(preset, seed, B, P, n_new, sample, full_path_probe=True)
        cpu = {"value": round(v, 3), "unit": "tokens/s", "cores": torch.get_num_threads(), "kind": info["kind"],
               "sample": f"{sample} of the {n_new} decode steps evenly spread over the run, each one token at its true "
                         f"sequence length, full recompute per token, compute only (no HTTP/JSON), p50 "
                         f"{1000 * p50:.1f} ms/token; the reference's JSON wire format would add "
                         f"{info.get('json_overhead_s_per_token')} s per token at T={info.get('json_T')}"}

    line = {
        "metric": METRIC, "value": round(value, 2), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(dev_total_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": dict(workload_config(args.workload, world, n_streams), batch_per_stream=B),
        "p50_token_ms": None if p50_tok_ms is None else round(p50_tok_ms, 4),
        "decode_tokens_per_s": None if not p50_tok_ms else round(B * n_streams * 1000.0 / p50_tok_ms, 1),
        "e2e": {"value": round(e2e, 2), "unit": "tokens/s", "h2d_bytes_per_step": int(prompts.numel() * 4),
                "d2h_bytes_per_step": int(B * n_streams * (P + n_new) * 4)},
        "gpu_launches": int(launches),
        "clocks": clocks, "roofline": roof, "cpu_baseline": cpu, "ids_check": check,
        "stage_split": [list(x) for x in runner.stages],
    }
    print(json.dumps(line), flush=True)
    if check is not None and check["mismatch"] > 0:
        sys.stderr.write("bench.py: ids_check FAILED: %s\n" % json.dumps(check))
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD,
                    help="one of %s, or custom:<preset>:<B per stream>:<P>:<n_new>" % ", ".join(sorted(WORKLOADS)))
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ids-check", action="store_true")
    ap.add_argument("--batch-per-stream", type=int, default=0,
                    help="override the workload's sequences per micro-batch (micro-batch sweeps of configs[2] / [3])")
    ap.add_argument("--streams", type=int, default=0,
                    help="micro-batches in flight (default: one per stage; more hide the stage-boundary latency)")
    args = ap.parse_args()
    if args.workload not in WORKLOADS:
        ap.error(f"unknown workload {args.workload!r}")
    if args.warmup < 3:
        args.warmup = 3   # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
