#!/usr/bin/env python
"""Summarise an ncu report: headline metrics per kernel and the source page grouped into address segments
of equal execution count (loop bodies), so the share of issue slots per loop is visible at a glance.
    python tools/ncu_segments.py gpurun_out/prof.ncu-rep [kernel-regex]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else None
WANT = ["gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__cycles_elapsed.avg.per_second",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H = rows[0]
for r in rows[2:]:
    name = r[H.index("Kernel Name")]
    if pat and pat not in name:
        continue
    print("=====", name[:100])
    for w in WANT:
        if w in H:
            print(f"  {w:86s} {r[H.index(w)]} {rows[1][H.index(w)]}")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + name.split("(")[0].split()[-1].split("<")[0]],
                         capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    hi = [k for k, x in enumerate(srows) if x and x[0] == "Address"]
    if not hi:
        continue
    SH = srows[hi[0]]
    data = [x for x in srows[hi[0] + 1:] if len(x) == len(SH)]
    ia, isrc, ie, it, iss = (SH.index(c) for c in ("Address", "Source", "Instructions Executed",
                                                    "Avg. Threads Executed", "# Samples"))
    base = int(data[0][ia], 16)
    tot = sum(int(x[ie]) for x in data)
    seg, cur = [], None
    for x in data:
        e, a = int(x[ie]), int(x[ia], 16) - base
        if cur and abs(e - cur["e"]) <= 0.02 * max(e, cur["e"], 1):
            cur["n"] += 1; cur["sum"] += e; cur["end"] = a; cur["smp"] += int(x[iss]); cur["mufu"] += "MUFU" in x[isrc]
        else:
            cur = dict(start=a, end=a, e=e, n=1, sum=e, smp=int(x[iss]), mufu=int("MUFU" in x[isrc]), thr=x[it])
            seg.append(cur)
    print(f"  total warp instructions {tot/1e6:.1f}M")
    for s in seg:
        if s["sum"] > 0.004 * tot:
            print(f"  {s['start']:5x}-{s['end']:5x} n={s['n']:3d} exec={s['e']/1e6:7.2f}M sum={s['sum']/1e6:8.1f}M "
                  f"({100*s['sum']/tot:4.1f}%) mufu={s['mufu']} thr={s['thr']} samples={s['smp']}")
