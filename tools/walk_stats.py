"""Where the forward's bisection walk spends its list positions (development build with -DGGS_WALK_STATS):

    make -C geometry-grounded-gaussian-splatting_b200/csrc variant NAME=stats DEFS=-DGGS_WALK_STATS
    GGS_RASTERIZER_LIB=$PWD/geometry-grounded-gaussian-splatting_b200/diff_gaussian_rasterization/variants/libggs_rasterizer_stats.so \
        python tools/walk_stats.py [scene ...] > gpurun_out/walk_stats.json

Per bisection iteration: mask words visited, list positions walked, positions no pixel of the warp is active on,
positions evaluated with exp / rsqrt (and how many lanes were active / not quiet there), positions handled by the
constant-factor path.
"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

import helpers as H  # noqa: E402
import scenes  # noqa: E402


def main():
    import diff_gaussian_rasterization as ours

    lib = ours._abi.lib
    fn = lib.ggs_debug_walk_stats
    fn.argtypes = [C.POINTER(C.c_uint64), C.c_int]
    fn.restype = None
    names = ["walked", "no_lane_active", "evaluated", "constant_path", "lanes_active_on_evaluated",
             "lanes_not_quiet_on_evaluated", "mask_words", "spare"]
    out = {}
    for name in (sys.argv[1:] or ["c3"]):
        sc = scenes.make_scene(name)
        buf = (C.c_uint64 * 40)()
        fn(buf, 1)
        H.run_module(ours, sc, "cuda:0", require_depth=True, backward=False)
        torch.cuda.synchronize()
        fn(buf, 1)
        out[name] = {f"iteration_{it}": {names[k]: int(buf[it * 8 + k]) for k in range(7)} for it in range(5)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
