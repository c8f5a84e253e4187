#!/usr/bin/env python
"""Per-source-line instruction and stall breakdown of one kernel from an `ncu --set full --import-source on` capture
(run here, no GPU needed):

    python tools/ncu_source.py gpurun_out/prof.ncu-rep [n_envs] [top] [kernel-substring] > profiles/rNN_<kernel>_source_hotspots.txt

Reads `ncu -i <rep> --page source --csv --print-source cuda,sass` (first profiled launch), sums the executed warp-
instructions and the stall samples per source file and per source line, and prints them per env-step.
"""
import collections
import csv
import subprocess
import sys

STALLS = ["stall_long_sb", "stall_short_sb", "stall_wait", "stall_math", "stall_not_selected", "stall_selected",
          "stall_barrier", "stall_branch_resolving", "stall_no_inst", "stall_mio", "stall_lg", "stall_dispatch",
          "stall_sleep", "stall_membar", "stall_misc", "stall_drain", "stall_tex"]


def main():
    rep = sys.argv[1]
    n_envs = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
    want = sys.argv[4] if len(sys.argv) > 4 else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    files = collections.OrderedDict()   # file -> {line: dict}
    cur, hdr, func = None, None, None
    seen_funcs = []
    sections = collections.Counter()   # (function, file) -> how often the section appears = profiled launches
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1]
            continue
        if r[0] == "Function Name":
            func = r[1]
            if (want is None or want in func) and func not in seen_funcs:
                seen_funcs.append(func)
            sections[(func, cur)] += 1
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or cur is None or not r[0].strip().isdigit():
            continue
        if not seen_funcs or func != seen_funcs[0]:
            continue  # one kernel of the capture only
        d = dict(zip(hdr, r))
        try:
            inst = float(d["Instructions Executed"] or 0)
            smp = float(d["# Samples"] or 0)
        except ValueError:
            continue
        rec = files.setdefault(cur, {}).setdefault(int(r[0]), {"inst": 0.0, "samples": 0.0, "src": r[1].strip(),
                                                               **{s: 0.0 for s in STALLS}})
        rec["inst"] += inst
        rec["samples"] += smp
        for s in STALLS:
            try:
                rec[s] += float(d.get(s, 0) or 0)
            except ValueError:
                pass
    launches = max([c for (fn, _), c in sections.items() if seen_funcs and fn == seen_funcs[0]] or [1])
    for f in files.values():   # the capture holds `launches` launches of the kernel: report one
        for v in f.values():
            for k in v:
                if k != "src":
                    v[k] /= launches
    tot_i = sum(v["inst"] for f in files.values() for v in f.values())
    tot_s = sum(v["samples"] for f in files.values() for v in f.values())
    print(f"{seen_funcs[0][:140] if seen_funcs else '?'}")
    print(f"average of {launches} profiled launch(es) at {n_envs} envs: {tot_i:.0f} warp-instructions attributed to source lines = {tot_i / n_envs:.1f} per env, "
          f"{tot_s:.0f} stall samples")
    tot_st = {s: sum(v[s] for f in files.values() for v in f.values()) for s in STALLS}
    print("stall samples by reason: " + ", ".join(f"{s[6:]} {100 * c / max(tot_s, 1):.1f}%" for s, c in
                                                    sorted(tot_st.items(), key=lambda x: -x[1]) if c > 0))
    print("per file: instr/env, % instr, % samples")
    for f, lines in sorted(files.items(), key=lambda x: -sum(v["inst"] for v in x[1].values())):
        fi = sum(v["inst"] for v in lines.values())
        fs = sum(v["samples"] for v in lines.values())
        fst = {s: sum(v[s] for v in lines.values()) for s in STALLS}
        top3 = ", ".join(f"{s[6:]} {100 * c / max(fs, 1):.0f}%" for s, c in sorted(fst.items(), key=lambda x: -x[1])[:4])
        print(f"  {fi / n_envs:8.1f} {100 * fi / tot_i:5.1f}% {100 * fs / max(tot_s, 1):5.1f}%  {f.split('/')[-1]:28s} {top3}")
    flat = [(f.split("/")[-1], ln, v) for f, lines in files.items() for ln, v in lines.items()]
    print(f"top {top} source lines by instructions: instr/env, % instr, % samples, top stall, file:line, source")
    for f, ln, v in sorted(flat, key=lambda x: -x[2]["inst"])[:top]:
        st = max(STALLS, key=lambda s: v[s])
        print(f"  {v['inst'] / n_envs:7.1f} {100 * v['inst'] / tot_i:5.1f}% {100 * v['samples'] / max(tot_s, 1):5.1f}% "
              f"{st[6:]:>12s} {f}:{ln}  {v['src'][:100]}")
    print(f"top {top} source lines by stall samples: % samples, instr/env, top stalls, file:line, source")
    for f, ln, v in sorted(flat, key=lambda x: -x[2]["samples"])[:top]:
        sts = sorted(STALLS, key=lambda s: -v[s])[:2]
        print(f"  {100 * v['samples'] / max(tot_s, 1):5.1f}% {v['inst'] / n_envs:7.1f} "
              f"{sts[0][6:]}:{v[sts[0]]:.0f} {sts[1][6:]}:{v[sts[1]]:.0f}  {f}:{ln}  {v['src'][:100]}")


if __name__ == "__main__":
    main()
