#!/usr/bin/env python
"""Sweep of the decode path over models / batch sizes / prompt lengths on ONE GPU (configs 3-5 of
BASELINE.json reduced to what one B200 holds): decode tokens/s, p50 step time, prefill time and
the fraction of the HBM roofline (algorithmic bytes of SURVEY.md 8(d) / measured step time).
Writes one JSON object per line; `python bench_sweep.py > profiles/rNN_sweep.jsonl`.
Not the contract benchmark (that is bench.py)."""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import llm_sharding_demo_b200 as L  # noqa: E402
from bench import algorithmic_bytes_per_step, measured_peaks, synthetic_prompts  # noqa: E402

CASES = [  # model, B, P, n_new
    ("gpt2", 1, 32, 64), ("gpt2", 1, 512, 64), ("gpt2", 2, 32, 64), ("gpt2", 4, 32, 64), ("gpt2", 8, 32, 64), ("gpt2", 32, 32, 64),
    ("gpt2", 32, 512, 32), ("gpt2-medium", 1, 32, 64), ("gpt2-medium", 2, 512, 32), ("gpt2-medium", 8, 512, 32),
    ("gpt2-xl", 1, 32, 32), ("gpt2-xl", 4, 32, 32), ("gpt2-xl", 4, 896, 32), ("gpt2-xl", 8, 32, 32), ("gpt2-xl", 32, 128, 16),
]


CFG5 = [("gpt2", B, P, 64) for P in (32, 512, 960) for B in (1, 2, 4, 8, 16, 32, 64, 128, 256)]   # BASELINE.json configs[4], one stage


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--cfg5", action="store_true", help="the 124M batch x prompt grid of BASELINE.json configs[4] on one GPU")
    ap.add_argument("--min-batch", type=int, default=1, help="only the cases with at least this many sequences")
    args = ap.parse_args()
    peak, kind = measured_peaks()
    engines = {}
    for model, B, P, n_new in (CFG5 if args.cfg5 else CASES):
        if (args.only and args.only not in model) or B < args.min_batch:
            continue
        dims = L.PRESETS[model]
        key = model
        if key not in engines:
            engines.clear()          # one model resident at a time
            torch.cuda.empty_cache()
            mb = 256 if args.cfg5 else 32
            eng = L.StageEngine(dims, dtype=L.GP_BF16, max_batch=mb, max_ctx=1024, max_tokens=mb * (960 if args.cfg5 else 512))
            eng.load_weights(L.init_weights(dims, 1234, "hf"))
            engines[key] = eng
        eng = engines[key]
        prompts = synthetic_prompts(B, P, dims.vocab_size).pin_memory()
        for _ in range(2):
            eng.generate(prompts, n_new)
        runs = [eng.generate(prompts, n_new, timings=True)[1] for _ in range(3)]
        tm = min(runs, key=lambda t: t["decode_ms"])
        step_ms = statistics.median(tm["step_ms"]) if tm["step_ms"] else tm["decode_ms"] / max(1, n_new - 1)
        t_ctx = P + n_new // 2
        bytes_step = algorithmic_bytes_per_step(dims, B, t_ctx, 1, 2)
        achieved = bytes_step / (step_ms * 1e-3) / 1e9
        cpu = None
        if args.cfg5 and B == 1:
            # the reference's compute-only CPU figure for the same point (its own ShardA / ShardB when oracle/_ref is there)
            from bench import cpu_reference_tokens_per_s
            torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
            v, _, info = cpu_reference_tokens_per_s(model, 1234, 1, P, n_new, 2)
            cpu = {"tokens_per_s": round(v, 2), "kind": info["kind"], "cores": torch.get_num_threads()}
        print(json.dumps({"model": model, "B": B, "P": P, "n_new": n_new, "stages": 1, "cpu_reference": cpu,
                          "prefill_ms": round(tm["prefill_ms"], 3),
                          "decode_tokens_per_s": round(B * 1000.0 / step_ms, 1), "p50_step_ms": round(step_ms, 4),
                          "launches_per_request": eng.last_launch_count(),
                          "algorithmic_mb_per_step": round(bytes_step / 1e6, 1), "achieved_gbs": round(achieved, 1),
                          "roofline_frac": round(achieved / peak, 4), "peak": peak, "peak_kind": kind,
                          "decode_path": "persistent kernel (mma.sync rows)" if B <= 8 else "tcgen05 GEMM + graph"}), flush=True)


if __name__ == "__main__":
    main()
