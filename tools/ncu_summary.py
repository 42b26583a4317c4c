#!/usr/bin/env python
"""Text summary of an `ncu --set full` report for profiles/: per kernel the headline metrics (duration, DRAM
traffic, pipe utilisation, occupancy limits) and, for kernels compiled with -lineinfo, the CUDA source lines
ranked by executed warp instructions and stall samples.

    python tools/ncu_summary.py gpurun_out/x.ncu-rep [--lines 25] > profiles/rNN_x_summary.txt
"""
import argparse
import csv
import io
import re
import subprocess

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "memory throughput %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots active %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU (MUFU) pipe %"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "FMA pipe %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "shared-memory wavefronts %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__occupancy_limit_registers", "blocks/SM (register limit)"),
    ("launch__occupancy_limit_shared_mem", "blocks/SM (shared limit)"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("sm__cycles_elapsed.avg.per_second", "SM clock"),
]


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--lines", type=int, default=25)
    a = ap.parse_args()
    rows = list(csv.reader(io.StringIO(run(["-i", a.report, "--page", "raw", "--csv"]))))
    H, U = rows[0], rows[1]
    seen = set()
    for r in rows[2:]:
        name = r[H.index("Kernel Name")]
        m = re.search(r"(\w+)\s*(<[^()]*>)?\s*\(", name)  # "void ggs::k<0, unsigned int, 1>(args)" -> k<0,unsigned int,1>
        short = (m.group(1) + (m.group(2) or "").replace(", ", ",")) if m else name.split("(")[0].split()[-1]
        grid = r[H.index("Grid Size")] if "Grid Size" in H else ""
        print(f"===== {short}  grid {grid}")
        for key, label in WANT:
            if key in H:
                v = r[H.index(key)]
                try:
                    v = f"{float(v):.4g}"
                except ValueError:
                    pass
                print(f"    {label:32s} {v} {U[H.index(key)]}")
        base = short.split("<")[0]
        if base in seen:
            continue
        seen.add(base)
        src = run(["-i", a.report, "--page", "source", "--csv", "--kernel-name", "regex:" + base, "--print-source",
                   "cuda,sass"])
        srows = list(csv.reader(io.StringIO(src)))
        hi = [k for k, x in enumerate(srows) if x and x[0] == "Line No"]
        if not hi:
            continue
        SH = srows[hi[0]]
        i_s, i_e = SH.index("# Samples"), SH.index("Instructions Executed")
        lines = {}
        for x in srows[hi[0] + 1:]:
            if len(x) > i_e and x[0].isdigit():
                ln = int(x[0])
                if ln in lines:
                    if x[1] == lines[ln][0]:
                        continue  # second launch of the same kernel
                try:
                    lines.setdefault(ln, (x[1], int(x[i_s]), int(x[i_e])))
                except ValueError:
                    pass
        tot = sum(v[2] for v in lines.values()) or 1
        ts = sum(v[1] for v in lines.values()) or 1
        print(f"    -- source lines by executed warp instructions (total {tot / 1e6:.1f} M, {ts} stall samples)")
        for ln, v in sorted(lines.items(), key=lambda kv: -kv[1][2])[:a.lines]:
            print(f"    {ln:5d}  inst {100 * v[2] / tot:5.1f}%  samples {100 * v[1] / ts:5.1f}%   {v[0].strip()[:100]}")
        print("    -- source lines by stall samples")
        for ln, v in sorted(lines.items(), key=lambda kv: -kv[1][1])[:10]:
            print(f"    {ln:5d}  inst {100 * v[2] / tot:5.1f}%  samples {100 * v[1] / ts:5.1f}%   {v[0].strip()[:100]}")


if __name__ == "__main__":
    main()
