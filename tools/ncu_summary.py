#!/usr/bin/env python
"""Summarise ncu outputs into profiles/ (run here, no GPU needed).

    python tools/ncu_summary.py launches gpurun_out/launches.csv "<command>"  > profiles/rNN_launches_summary.txt
    python tools/ncu_summary.py full     gpurun_out/prof.ncu-rep              > profiles/rNN_kernels_ncu_full.txt
"""
import collections
import csv
import subprocess
import sys

FULL_METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
    "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_uniform.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
]
STALLS = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio"
STALL_NAMES = ["wait", "math_pipe_throttle", "short_scoreboard", "long_scoreboard", "not_selected", "barrier",
               "branch_resolving", "dispatch_stall", "lg_throttle", "mio_throttle", "no_instruction", "selected"]


def launches(path, cmd):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v, u = float(r[vi].replace(",", "")), r[ui]
        v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v * 1e6 if u == "s" else v
        a = agg.setdefault(r[ki], [0.0, 0])
        a[0] += v
        a[1] += 1
    tot = sum(a[0] for a in agg.values())
    print(f"ncu --metrics gpu__time_duration.sum --clock-control none, {cmd} (cold-cache, serialised: compare shares)")
    for k, (t, c) in sorted(agg.items(), key=lambda x: -x[1][0]):
        print(f"{t:12.1f} us {100 * t / tot:5.1f}% x{c:4d} avg {t / c:8.1f} us  {k[:110]}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    print(f"ncu --set full --clock-control none --import-source on  ({path}); per launch")
    for r in rows[2:]:
        print("---", r[hdr.index("Kernel Name")][:120])
        for m in FULL_METRICS + [STALLS % s for s in STALL_NAMES]:
            if m in hdr:
                i = hdr.index(m)
                print(f"  {m:78s} {r[i]:>16s} {units[i]}")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    else:
        full(sys.argv[2])
