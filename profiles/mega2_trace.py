#!/usr/bin/env python
"""Reads the per-CTA phase trace of the second-generation persistent kernel (GP_MEGA_TRACE=<file>): int64
[n_cta][512], each stamp = %globaltimer ns (low 56 bits) | tag << 56 with tag = kind * 16 + event, taken by thread 0
of every CTA during the LAST decode step of a launch.  %globaltimer is one clock for the whole GPU, so stamps of
different CTAs are comparable.  Prints
  (1) per (tag -> next tag) segment the median / p10 / p90 duration over CTAs and layers (us), for CTAs with and
      without an attention task;
  (2) per event the SPREAD over CTAs (latest - earliest occurrence, median over layers) and which CTAs are last.
Usage: mega2_trace.py file [n_attn_ctas=12]"""
import sys
from collections import defaultdict
import numpy as np

CAP = 1024
full = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(-1, CAP)
raw = full[:, :CAP - 128]
dbg = full[:, CAP - 128:].astype(np.int64).reshape(-1, 8, 16)
n_attn = int(sys.argv[2]) if len(sys.argv) > 2 else 12
PH = {0: "qkv", 1: "attn", 2: "proj", 3: "fc", 4: "proj2", 5: "head", 7: "step"}
EV = {0: "enter", 1: "x-ok", 2: "staged", 3: "bar", 4: "rows-done", 5: "mma-done", 6: "bar2", 7: "stored",
      8: "released", 9: "reduced"}


def name(tag):
    ph, ev = tag >> 4, tag & 15
    if ph == 1:
        return "attn:" + ["entered", "q-ok", "loop-done", "done"][ev]
    if ph == 7:
        return "step:" + ["start", "end", "partials", "item-begin"][ev]
    return f"{PH.get(ph, ph)}:{EV.get(ev, ev)}"


def parse(row):
    r = row[row != 0]
    tags = (r >> np.uint64(56)).astype(int)
    t = (r & np.uint64((1 << 56) - 1)).astype(np.int64)
    return tags, t


def segments(rows):
    seg = defaultdict(list)
    order = []
    total = []
    for row in rows:
        tags, t = parse(row)
        total.append((t[-1] - t[0]) / 1e3)
        for i in range(1, len(t)):
            key = (tags[i - 1], tags[i])
            if key not in seg:
                order.append(key)
            seg[key].append((t[i] - t[i - 1]) / 1e3)
    return seg, order, total


for label, rows in (("CTAs with an attention task", raw[:n_attn]), ("other CTAs", raw[n_attn:])):
    seg, order, total = segments(rows)
    print(f"== {label}: {len(rows)} CTAs, traced span median {np.median(total):.1f} us")
    print(f"{'segment':>38} {'n':>6} {'median':>8} {'p10':>8} {'p90':>8} {'sum(med)':>9}")
    acc = 0.0
    for key in order:
        d = np.array(seg[key])
        per_cta = len(d) / len(rows)
        acc += np.median(d) * per_cta
        print(f"{name(key[0]) + ' -> ' + name(key[1]):>38} {len(d):6d} {np.median(d):8.2f} {np.percentile(d, 10):8.2f} "
              f"{np.percentile(d, 90):8.2f} {acc:9.1f}")

# ---- cross-CTA view: occurrence k of a tag in CTA c happens at time T[c][tag][k]
occ = defaultdict(lambda: defaultdict(list))   # tag -> k -> [(time, cta)]
for c, row in enumerate(raw):
    tags, t = parse(row)
    seen = defaultdict(int)
    for tg, tt in zip(tags, t):
        occ[tg][seen[tg]].append((tt, c))
        seen[tg] += 1
print("== spread over CTAs per event (us): median over occurrences of (latest - earliest), (latest - median); last CTAs")
for tg in sorted(occ):
    if tg >> 4 == 1:
        continue
    spreads, tails, last = [], [], defaultdict(int)
    for k, lst in occ[tg].items():
        if len(lst) < len(raw) // 2:
            continue
        ts = np.array([x[0] for x in lst])
        spreads.append((ts.max() - ts.min()) / 1e3)
        tails.append((ts.max() - np.median(ts)) / 1e3)
        for x in sorted(lst)[-3:]:
            last[x[1]] += 1
    if spreads:
        top = sorted(last.items(), key=lambda kv: -kv[1])[:6]
        print(f"{name(tg):>18} spread {np.median(spreads):6.2f}  tail {np.median(tails):6.2f}  last: {top}")

# ---- per-warp marks (middle layer): when each warp of a CTA finished polling the input of a phase
print("== per-warp poll-finish times in the middle layer, relative to the CTA's earliest warp (us): median over CTAs")
for which, nm in ((0, "qkv"), (2, "proj"), (3, "fc"), (4, "proj2")):
    d = dbg[:, which, :15]
    ok = (d > 0).all(axis=1)
    if not ok.any():
        continue
    rel = (d[ok] - d[ok].min(axis=1, keepdims=True)) / 1e3
    print(f"{nm:>6}: " + " ".join(f"{x:5.2f}" for x in np.median(rel, axis=0)), " | max over warps: median %.2f p90 %.2f" % (np.median(rel.max(axis=1)), np.percentile(rel.max(axis=1), 90)))
    # global: earliest and latest finish over all CTAs and warps
    print(f"        all CTAs: span of finish times {(d[ok].max() - d[ok].min()) / 1e3:.2f} us; per-warp latest-CTA offsets: " +
          " ".join(f"{x:5.2f}" for x in (d[ok].max(axis=0) - d[ok].min()) / 1e3))
