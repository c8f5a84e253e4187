#!/usr/bin/env python
"""Summarise ncu outputs into profiles/ (run here, no GPU needed).

    python tools/ncu_summary.py launches gpurun_out/launches.csv "<command>"  > profiles/rNN_launches_summary.txt
    python tools/ncu_summary.py full     gpurun_out/prof.ncu-rep              > profiles/rNN_kernels_ncu_full.txt
    python tools/ncu_summary.py metrics  gpurun_out/prof.ncu-rep 65536 profiles/rNN_kernels_ncu_full.txt
                                         -> profiles/ncu_metrics.json + profiles/ncu_traffic.json (what bench.py quotes)
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


def metrics(path, n_envs, source):
    """Per-kernel figures of one env step from an `ncu --set full` capture, averaged over the profiled launches of each
    kernel: duration, warp-instructions, issue-slot utilisation, FP64-pipe activity and the FP64 warp-instructions that
    activity amounts to (the pipe takes one warp-instruction per two cycles per SM sub-partition), DRAM bytes."""
    import json
    import os
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    units = rows[1]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name, scale=None):
        v = float(r[col[name]].replace(",", ""))
        u = units[col[name]]
        if scale == "bytes":
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        if scale == "us":
            v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}[u]
        return v

    per = collections.OrderedDict()
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").strip()
        name = "usv::" + name.split("<")[0] if not name.startswith("usv::") else name.split("<")[0]
        cyc = val(r, "sm__cycles_active.avg")
        sms = val(r, "launch__sm_count")
        fp64_pct = val(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active")
        rec = dict(us=val(r, "gpu__time_duration.sum", "us"), warp_inst=val(r, "smsp__inst_executed.sum"),
                   issue_slot_util_pct=val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                   fp64_pipe_active_pct=fp64_pct, fp64_warp_inst=fp64_pct / 100 * cyc * sms * 2.0,
                   dram_bytes=val(r, "dram__bytes_read.sum", "bytes") + val(r, "dram__bytes_write.sum", "bytes"),
                   warps_active_pct=val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                   registers=val(r, "launch__registers_per_thread"))
        per.setdefault(name, []).append(rec)
    kern = {k: {m: sum(x[m] for x in v) / len(v) for m in v[0]} for k, v in per.items()}
    step_us = sum(k["us"] for k in kern.values())
    doc = {
        "source": f"{source} (ncu --set full --clock-control none, {n_envs} envs, average of the profiled launches per kernel)",
        "n_envs": n_envs,
        "kernels": kern,
        "fp64_warp_inst_per_env_step": sum(k["fp64_warp_inst"] for k in kern.values()) / n_envs,
        "warp_inst_per_env_step": sum(k["warp_inst"] for k in kern.values()) / n_envs,
        "issue_slot_util": {"per_kernel_pct": {n: k["issue_slot_util_pct"] for n, k in kern.items()},
                            "step_time_weighted_pct": sum(k["issue_slot_util_pct"] * k["us"] for k in kern.values()) / step_us},
        "fp64_pipe_active": {"per_kernel_pct": {n: k["fp64_pipe_active_pct"] for n, k in kern.items()},
                             "step_time_weighted_pct": sum(k["fp64_pipe_active_pct"] * k["us"] for k in kern.values()) / step_us},
    }
    prof = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles")
    with open(os.path.join(prof, "ncu_metrics.json"), "w") as f:
        json.dump(doc, f, indent=1)
    tpath = os.path.join(prof, "ncu_traffic.json")
    try:
        with open(tpath) as f:
            t = json.load(f)
    except Exception:  # noqa: BLE001
        t = {}
    t["source"] = doc["source"] + ": dram__bytes_read.sum + dram__bytes_write.sum per launch"
    for n, k in kern.items():
        t[n] = int(k["dram_bytes"])
    with open(tpath, "w") as f:
        json.dump(t, f, indent=1)
    print(json.dumps(doc, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    elif sys.argv[1] == "metrics":
        metrics(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 65536, sys.argv[4] if len(sys.argv) > 4 else sys.argv[2])
    else:
        full(sys.argv[2])
