#!/usr/bin/env python
"""Reads the raw per-CTA phase trace written with GP_MEGA_TRACE=<file> (int64 [n_cta][256]
%globaltimer ns at every phase boundary of the last decode step) and prints, per boundary, when
the first / median / last CTA got there and which CTA was last.  Usage: mega_trace.py file [n]"""
import sys
import numpy as np
t = np.fromfile(sys.argv[1], dtype=np.int64).reshape(-1, 256)
n_show = int(sys.argv[2]) if len(sys.argv) > 2 else 24
valid = (t > 0).all(axis=0)
n = int(valid.sum())
t = t[:, :n].astype(np.float64)
t0 = t[:, 0].min()
t = (t - t0) / 1000.0  # us
print(f"{t.shape[0]} CTAs, {n} stamps, step length {t[:, -1].max():.1f} us")
names9 = ["lnP1", "cP1", "attn", "stP3", "cP3", "lnP4", "cP4", "stP5", "cP5"]
print(f"{'i':>3} {'phase':>6} {'first':>8} {'median':>8} {'last':>8} {'lastCTA':>7} {'dur(med)':>8}")
prev = t[:, 0]
for i in range(1, min(n, n_show + 1)):
    col = t[:, i]
    print(f"{i:3d} {names9[(i - 1) % 9]:>6} {col.min():8.2f} {np.median(col):8.2f} {col.max():8.2f} {int(col.argmax()):7d} "
          f"{np.median(col - prev):8.2f}")
    prev = col
