"""The committed measurement artefacts bench.py reads stay consistent with the tools that produce them (no GPU):
profiles/rNN_traffic_vM.json is what tools/traffic_from_summary.py makes of the committed ncu summary, and it names
every kernel bench.py looks up for the `traffic` / `executed` fields of its roofline object."""
import glob
import json
import os
import subprocess
import sys

import helpers as H


def _latest_traffic():
    cands = sorted(glob.glob(os.path.join(H.ROOT, "profiles", "r*_traffic_*.json")))
    assert cands, "no profiles/r*_traffic_*.json committed"
    return cands[-1]


def test_traffic_file_names_every_stage_kernel():
    tj = json.load(open(_latest_traffic()))
    for kernel in ("preprocess_kernel", "scan_kernel", "duplicate_keys_kernel", "tile_ranges_kernel",
                   "render_forward_kernel", "render_backward_kernel", "gaussian_backward_kernel"):
        k = tj["kernels"][kernel]
        assert k["dram_bytes"] > 0 and k["duration_ms_under_ncu"] > 0
        assert 0 < k["issue_slots_active_pct"] <= 100
    # the tile kernels are L2-resident: DRAM traffic far below their 641 / 787 MB of algorithmic bytes
    assert tj["kernels"]["render_forward_kernel"]["dram_bytes"] < 100e6
    assert tj["kernels"]["render_backward_kernel"]["dram_bytes"] < 200e6


def test_traffic_file_regenerates_from_the_committed_summary(tmp_path):
    path = _latest_traffic()
    tag = os.path.basename(path).replace("_traffic_", "_full_").replace(".json", "_summary.txt")
    summary = os.path.join(H.ROOT, "profiles", tag)
    if not os.path.exists(summary):  # round 1's file came from an older summary layout
        return
    out = tmp_path / "traffic.json"
    subprocess.check_call([sys.executable, os.path.join(H.ROOT, "tools", "traffic_from_summary.py"),
                           os.path.relpath(summary, H.ROOT), str(out)], cwd=H.ROOT, stdout=subprocess.DEVNULL)
    assert json.load(open(out))["kernels"] == json.load(open(path))["kernels"]
