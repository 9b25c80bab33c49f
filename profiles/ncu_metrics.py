#!/usr/bin/env python
"""Per-launch metric table of an .ncu-rep (`ncu --set full`), read on the CPU box.
Writes the counters DESIGN.md cites (warp execution efficiency, pipe utilisation, L1/L2 hit
rates, DRAM bytes, stall mix) and, with --traffic-json, the per-launch DRAM traffic of each
kernel for bench.py's `roofline.traffic`.

usage: ncu_metrics.py report.ncu-rep [--traffic-json profiles/traffic.json --paths N --commit HASH]
(--paths: camera paths of the captured wave; --commit: the commit the capture ran — both recorded in the json)"""
import csv
import io
import json
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units, data = rows[0], rows[1], rows[2:]
M = [
    ("time_us", "gpu__time_duration.sum"),
    ("grid", "launch__grid_size"),
    ("regs", "launch__registers_per_thread"),
    ("warps_active_%", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("threads/inst (warp exec eff, of 32)", "smsp__thread_inst_executed_per_inst_executed.ratio"),
    ("issue_active_%", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("sm_throughput_%", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("pipe_alu_%", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    ("pipe_fma(fp32)_%", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    ("pipe_fp64_%", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
    ("pipe_lsu_%", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    ("pipe_xu_%", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    ("warp_inst", "smsp__inst_executed.sum"),
    ("l1_hit_%", "l1tex__t_sector_hit_rate.pct"),
    ("l2_hit_%", "lts__t_sector_hit_rate.pct"),
    ("dram_read", "dram__bytes_read.sum"),
    ("dram_write", "dram__bytes_write.sum"),
    ("dram_throughput_%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("l1tex_throughput_%", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("lts_throughput_%", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("stall_long_scoreboard", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
    ("stall_wait", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"),
    ("stall_no_instruction", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"),
    ("stall_not_selected", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"),
    ("stall_math_throttle", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"),
    ("stall_short_scoreboard", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"),
    ("stall_branch_resolving", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio"),
]
ki = h.index("Kernel Name")
names = [r[ki].split("(")[0].replace("void ", "") for r in data]
print(f"{'metric':38s} {'unit':8s} " + " ".join(f"{n[:22]:>22s}" for n in names))
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
traffic = {}
for label, key in M:
    if key not in h:
        continue
    i = h.index(key)
    print(f"{label:38s} {units[i][:8]:8s} " + " ".join(f"{r[i][:22]:>22s}" for r in data))
for r, n in zip(data, names):
    rd, wr = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
    b = float(r[rd]) * SCALE.get(units[rd], 1.0) + float(r[wr]) * SCALE.get(units[wr], 1.0)
    traffic.setdefault(n.split("<")[0], []).append(b)
print()
for k, v in traffic.items():
    print(f"dram bytes per launch, {k}: " + ", ".join(f"{x/1e6:.1f} MB" for x in v))
if "--traffic-json" in sys.argv:
    out = sys.argv[sys.argv.index("--traffic-json") + 1]
    def arg(name, default):
        return sys.argv[sys.argv.index(name) + 1] if name in sys.argv else default
    paths, commit = int(arg("--paths", "0")), arg("--commit", "unknown")
    json.dump({k: {"per_launch_bytes": v, "max": max(v), "source": rep.split("/")[-1], "commit": commit, "paths_per_wave": paths,
                   "note": f"dram__bytes_read.sum + dram__bytes_write.sum per launch; one wave of {paths} paths of the headline scene; launches in capture order (bounce 0 first)"}
               for k, v in traffic.items()}, open(out, "w"), indent=1)
