"""Executed warp instructions and stall samples of one kernel of an ncu report, summed over source-line ranges:
    python tools/ncu_regions.py report.ncu-rep kernel_regex name:lo-hi [name:lo-hi ...]   (line numbers of the .cu at capture time)"""
import csv, io, subprocess, sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern,
                          "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hi = [k for k, x in enumerate(rows) if x and x[0] == "Line No"]
    SH = rows[hi[0]]
    i_s, i_e = SH.index("# Samples"), SH.index("Instructions Executed")
    end = hi[1] if len(hi) > 1 else len(rows)
    lines = {}
    for x in rows[hi[0] + 1:end]:
        if len(x) > i_e and x[0].isdigit():
            try:
                lines[int(x[0])] = (x[1], int(x[i_s]), int(x[i_e]))
            except ValueError:
                pass
    tot = sum(v[2] for v in lines.values()) or 1
    ts = sum(v[1] for v in lines.values()) or 1
    print(f"total {tot / 1e6:.1f} M warp instructions, {ts} stall samples")
    for spec in [a for a in sys.argv[3:] if ":" in a]:
        name, rng = spec.split(":")
        a, b = (int(v) for v in rng.split("-"))
        i = sum(v[2] for l, v in lines.items() if a <= l <= b)
        s = sum(v[1] for l, v in lines.items() if a <= l <= b)
        print(f"{name:24s} {a:4d}-{b:4d}  inst {i / 1e6:7.1f} M {100 * i / tot:5.1f}%   samples {100 * s / ts:5.1f}%")
    if "--all" in sys.argv:
        for l in sorted(lines):
            v = lines[l]
            if v[2] / tot > 0.002:
                print(f"{l:5d} {v[2] / 1e6:7.1f} M {100 * v[2] / tot:5.1f}% {100 * v[1] / ts:5.1f}%  {v[0].strip()[:100]}")


if __name__ == "__main__":
    main()
