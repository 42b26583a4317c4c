"""profiles/rNN_full_vM_summary.txt (tools/ncu_summary.py over one `ncu --set full` capture of bench.py) ->
profiles/rNN_traffic_vM.json: per kernel and launch the measured DRAM bytes (dram__bytes_read.sum +
dram__bytes_write.sum), the duration under ncu and the executed pipe utilisation.  bench.py reads the newest such file
for the `traffic` and `executed` fields of its roofline object.
    python tools/traffic_from_summary.py profiles/r02_full_v7_summary.txt profiles/r02_traffic_v7.json"""
import json
import re
import sys

UNITS = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TIME = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
FIELDS = {"issue slots active %": "issue_slots_active_pct", "XU (MUFU) pipe %": "xu_pipe_pct",
          "FMA pipe %": "fma_pipe_pct", "ALU pipe %": "alu_pipe_pct", "achieved occupancy %": "achieved_occupancy_pct",
          "memory throughput %": "memory_throughput_pct"}


def main(src, dst):
    kernels, cur, seen = {}, None, {}
    for line in open(src):
        m = re.match(r"===== (\S+?)(<.*>)?\s+grid \((.*)\)", line)
        if m:
            base = m.group(1)
            seen[base + m.group(3)] = seen.get(base + m.group(3), 0) + 1
            name = base
            if base == "onesweep_kernel":  # depth pre-sort passes (one wave) and tile passes (thousands of blocks)
                name = "onesweep_kernel_tile_pass" if int(m.group(3).split(",")[0]) > 1000 else "onesweep_kernel_depth_pass"
            cur = kernels.setdefault(name, None)
            if cur is None:  # the first launch of each kernel carries the full metric set
                cur = kernels[name] = {"grid": m.group(3), "dram_bytes": 0.0}
                cur["_open"] = True
            else:
                cur = {"_open": False}
            continue
        if cur is None or not cur.get("_open"):
            continue
        m = re.match(r"\s+dram (read|write)\s+([0-9.e+]+) (\w+)", line)
        if m:
            cur["dram_bytes"] += float(m.group(2)) * UNITS[m.group(3)]
            continue
        m = re.match(r"\s+duration\s+([0-9.e+]+) (\w+)", line)
        if m:
            cur["duration_ms_under_ncu"] = float(m.group(1)) * TIME[m.group(2)]
            continue
        for label, key in FIELDS.items():
            if line.strip().startswith(label):
                cur[key] = float(line.split()[-2])
    for v in kernels.values():
        v.pop("_open", None)
    out = {"source": f"ncu --set full --clock-control none, {src} (C3 workload, per launch: dram__bytes_read.sum + "
                     "dram__bytes_write.sum; pipe and issue utilisation as executed)", "kernels": kernels}
    with open(dst, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps({k: (round(v["dram_bytes"] / 1e6, 1), v.get("duration_ms_under_ncu")) for k, v in kernels.items()}))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
