#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): per-launch key metrics, stall mix, hottest SASS.
usage: ncu_summary.py report.ncu-rep [n_hot]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
nhot = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.avg.per_cycle_active"]
idx = [hdr.index(w) for w in want if w in hdr]
for r in rows[2:]:
    print(" | ".join(f"{hdr[i].split('.')[0][-28:]}={r[i][:40]}" for i in idx))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
for k, st in enumerate(starts):
    end = starts[k + 1] - 1 if k + 1 < len(starts) else len(rows)
    h = rows[st]
    ix = {n: i for i, n in enumerate(h)}
    data = [r for r in rows[st + 1:end] if len(r) == len(h)]

    def f(r, n):
        try:
            return float(r[ix[n]])
        except Exception:
            return 0.0
    tot = sum(f(r, "# Samples") for r in data) or 1.0
    print(f"--- launch {k}: samples {tot:.0f}")
    stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
    agg = sorted(((sum(f(r, n) for r in data), n) for n in stalls), reverse=True)[:7]
    print("   stalls: " + ", ".join(f"{n[6:]} {v / tot:.2f}" for v, n in agg))
    if k % 2 == 0 or True:
        for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:nhot]:
            print(f"   {f(r, '# Samples'):6.0f} thr={r[ix['Avg. Threads Executed']]:>4s} exec={r[ix['Instructions Executed']]:>9s} {r[ix['Source']][:80]}")
