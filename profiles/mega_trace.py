#!/usr/bin/env python
"""Reads the raw per-CTA phase trace written with GP_MEGA_TRACE=<file>: int64 [n_cta][512], each stamp
= SM clock (low 56 bits) | tag << 56 with tag = phase * 8 + event, taken by thread 0 of every CTA at
the phase boundaries of the LAST decode step of a launch.  Prints, per (tag -> next tag) segment, the
median / min / max duration over CTAs and layers, in microseconds at the given SM clock.
Usage: mega_trace.py file [sm_mhz=1965] [n_attn_ctas=12]"""
import sys
from collections import defaultdict
import numpy as np

CAP = 512
raw = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(-1, CAP)
mhz = float(sys.argv[2]) if len(sys.argv) > 2 else 1965.0
n_attn = int(sys.argv[3]) if len(sys.argv) > 3 else 12
PH = {0: "qkv", 1: "attn", 2: "proj", 3: "fc", 4: "proj2", 5: "head", 7: "step"}
EV = {0: "enter", 1: "x-ok", 2: "staged", 3: "w-full", 4: "rows-done", 5: "dot", 6: "reduced", 7: "stored"}


def name(tag):
    ph, ev = tag >> 3, tag & 7
    if ph == 1:
        return "attn:" + ["entered", "q-ok", "loop-done", "done"][ev]
    if ph == 7:
        return "step:" + ["start", "end", "partials"][ev]
    return f"{PH.get(ph, ph)}:{EV.get(ev, ev)}"


def segments(rows):
    seg = defaultdict(list)
    order = []
    total = []
    for r in rows:
        r = r[r != 0]
        tags = (r >> np.uint64(56)).astype(int)
        clk = (r & np.uint64((1 << 56) - 1)).astype(np.int64)
        total.append((clk[-1] - clk[0]) / mhz)
        for i in range(1, len(r)):
            key = (tags[i - 1], tags[i])
            if key not in seg:
                order.append(key)
            seg[key].append((clk[i] - clk[i - 1]) / mhz)
    return seg, order, total


for label, rows in (("CTAs with an attention task", raw[:n_attn]), ("other CTAs", raw[n_attn:])):
    seg, order, total = segments(rows)
    print(f"== {label}: {len(rows)} CTAs, traced span median {np.median(total):.1f} us")
    print(f"{'segment':>34} {'n':>6} {'median':>8} {'p10':>8} {'p90':>8} {'sum(med)':>9}")
    acc = 0.0
    for key in order:
        d = np.array(seg[key])
        per_cta = len(d) / len(rows)
        acc += np.median(d) * per_cta
        print(f"{name(key[0]) + ' -> ' + name(key[1]):>34} {len(d):6d} {np.median(d):8.2f} {np.percentile(d, 10):8.2f} "
              f"{np.percentile(d, 90):8.2f} {acc:9.1f}")
