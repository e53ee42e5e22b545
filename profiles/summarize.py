#!/usr/bin/env python
"""Turns the raw ncu outputs under gpurun_out/ into the tracked summaries under profiles/.
Usage: python profiles/summarize.py r01"""
import collections
import csv
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TAG = sys.argv[1] if len(sys.argv) > 1 else "r01"
GO = os.path.join(ROOT, "gpurun_out")
OUT = os.path.join(ROOT, "profiles")


def launches():
    p = os.path.join(GO, f"launches_{TAG}.csv")
    if not os.path.exists(p):
        return
    lines = [l for l in open(p) if not l.startswith("==")]
    agg = collections.OrderedDict()
    n = 0
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void unnamed>::", "").replace("unnamed>::", "")
        key = (name, row["Grid Size"], row["Block Size"])
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(row["Metric Unit"], 1e-3)
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(OUT, f"{TAG}_launch_list_summary.txt"), "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none, {n} launches of\n"
                f"# `python bench.py --steps 1 --warmup 3 --no-cpu-baseline` (cold-cache, serialised: compare SHARES)\n")
        f.write(f"{'kernel':58s} {'grid':>14s} {'n':>5s} {'avg_us':>8s} {'total_us':>9s} {'share':>6s}\n")
        for (name, grid, blk), a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{name[:58]:58s} {grid:>14s} {a[0]:5d} {a[1] / a[0]:8.2f} {a[1]:9.1f} {a[1] / tot:6.3f}\n")
    print("wrote launch summary")


WANT = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__waves_per_multiprocessor", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__cycles_active.avg", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warp_latency_issue_stalled_barrier.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor.sum", "sm__cycles_elapsed.max"]


def full():
    with open(os.path.join(OUT, f"{TAG}_ncu_full_summary.txt"), "w") as f:
        for fn in sorted(os.listdir(GO)):
            if not (fn.endswith(f"_{TAG}.ncu-rep") and fn.startswith("prof_")):
                continue
            r = subprocess.run(["ncu", "-i", os.path.join(GO, fn), "--page", "raw", "--csv"], capture_output=True, text=True)
            rows = list(csv.reader(r.stdout.splitlines()))
            if len(rows) < 3:
                continue
            hdr, units = rows[0], rows[1]
            f.write(f"## {fn}  (ncu --set full --clock-control none)\n")
            for row in rows[2:]:
                for w in WANT:
                    if w in hdr:
                        i = hdr.index(w)
                        f.write(f"  {w:68s} {row[i][:90]} {units[i]}\n")
                f.write("\n")
    print("wrote full summary")


def traffic():
    """dram bytes per launch of the persistent decode kernel -> profiles/<tag>_traffic.json (bench.py reads the newest)."""
    import json
    fn = os.path.join(GO, f"prof_mega_{TAG}.ncu-rep")
    if not os.path.exists(fn):
        return
    r = subprocess.run(["ncu", "-i", fn, "--page", "raw", "--csv"], capture_output=True, text=True)
    rows = list(csv.reader(r.stdout.splitlines()))
    if len(rows) < 3:
        return
    hdr, units, row = rows[0], rows[1], rows[2]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}

    def val(name):
        i = hdr.index(name)
        return float(row[i].replace(",", "")) * scale.get(units[i], 1)
    rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
    i = hdr.index("gpu__time_duration.sum")
    out = {"kernel": row[hdr.index("Kernel Name")][:80], "dram_bytes_per_launch": int(rd + wr), "dram_bytes_read": int(rd),
           "dram_bytes_written": int(wr), "gpu_time_duration": row[i] + " " + units[i],
           "source": f"ncu --set full --clock-control none, gpurun_out/prof_mega_{TAG}.ncu-rep (python bench.py --steps 1 --warmup 3)",
           "steps_per_launch": 127}
    with open(os.path.join(OUT, f"{TAG}_traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote traffic json", out["dram_bytes_per_launch"])


def sass():
    """Opcode histogram of the built library: proves which kernels use tcgen05 / TMA / TMEM / mma.sync."""
    so = os.path.join(ROOT, "llm_sharding_demo_b200", "libgpt2pipe.so")
    if not os.path.exists(so):
        return
    r = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True)
    cur, hist = None, collections.OrderedDict()
    ops = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UBLKCP", "HMMA", "LDSM", "SYNCS", "UTCBAR"]
    for ln in r.stdout.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            continue
        if cur is None:
            continue
        for o in ops:
            if re.search(r"\b" + o + r"[\. ]", ln):
                hist.setdefault(cur, collections.Counter())[o] += 1
    with open(os.path.join(OUT, f"{TAG}_sass_histogram.txt"), "w") as f:
        f.write("# cuobjdump -sass llm_sharding_demo_b200/libgpt2pipe.so: occurrences of tensor-core / TMA / TMEM / mbarrier opcodes per kernel\n")
        f.write("# (UTCHMMA = tcgen05.mma, UTMALDG = cp.async.bulk.tensor, LDTM = tcgen05.ld, UBLKCP = cp.async.bulk, HMMA = mma.sync)\n")
        tot = collections.Counter()
        for k, c in hist.items():
            name = subprocess.run(["c++filt", k], capture_output=True, text=True).stdout.strip()[:110]
            f.write(f"{name}\n    " + "  ".join(f"{o}={n}" for o, n in sorted(c.items())) + "\n")
            tot.update(c)
        f.write("TOTAL  " + "  ".join(f"{o}={n}" for o, n in sorted(tot.items())) + "\n")
    print("wrote sass histogram")


launches()
full()
traffic()
sass()
