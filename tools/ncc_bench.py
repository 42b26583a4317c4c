#!/usr/bin/env python
"""Patch-NCC kernel: our library against the compiled reference module on one GPU (run on the box).

    python tools/ncc_bench.py [--points 500000] [--height 822] [--width 1237] [--iters 200]

Times `warp_patch_ncc._run` (ours) and `ref_wpn._C.warp_patch_ncc` (oracle/_ref, when shipped) with CUDA events after
warm-up, on pixels in raster order (what `torch.argwhere(mask)` gives the trainer, utils/loss_utils.py:229-231) and in
random order, and prints one JSON object with the times, the parity statistics and the share of valid patches."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

import ncc_scenes  # noqa: E402


def timed(fn, iters):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=500000)
    ap.add_argument("--height", type=int, default=822)
    ap.add_argument("--width", type=int, default=1237)
    ap.add_argument("--iters", type=int, default=200)
    args = ap.parse_args()
    import warp_patch_ncc as ours
    from oracle import ref_loader

    dev = "cuda:0"
    sc = ncc_scenes.make_pair(height=args.height, width=args.width, n=args.points, seed=3)
    out = {"points": args.points, "image": [args.height, args.width]}
    for order in ("raster", "random"):
        if order == "raster":  # sort the pixels row-major
            key = sc["uvs"][:, 1].long() * args.width + sc["uvs"][:, 0].long()
            idx = torch.argsort(key)
            for k in ("uvs", "depths", "normals"):
                sc[k] = sc[k][idx].contiguous()
        else:
            idx = torch.randperm(args.points, generator=torch.Generator().manual_seed(1))
            for k in ("uvs", "depths", "normals"):
                sc[k] = sc[k][idx].contiguous()
        a = ncc_scenes.args_of(sc, dev)
        params = ours.WarpParams(a[3], a[4], *a[7:], False)
        res = {"ours_ms": timed(lambda: ours._run(a[0], a[1], a[2], a[5], a[6], params), args.iters)}
        got = ours._run(a[0], a[1], a[2], a[5], a[6], params)
        res["valid_share"] = float(got[3].float().mean())
        if ref_loader.ncc_available():
            ref = ref_loader.load_ncc()
            res["reference_ms"] = timed(lambda: ref._C.warp_patch_ncc(*a, False), args.iters)
            want = ref._C.warp_patch_ncc(*a, False)
            both = got[3] & want[3]
            res["valid_flips"] = int((got[3] != want[3]).sum())
            res["bit_equal"] = bool(all(torch.equal(x, y) for x, y in zip(got, want)))
            res["ncc_max_abs"] = float((got[0] - want[0]).abs()[both].max())
            for i, name in ((1, "grad_depths"), (2, "grad_normals")):
                res[name + "_max_rel"] = float((got[i] - want[i]).abs()[both].max() / want[i][both].abs().max())
            res["speedup"] = res["reference_ms"] / res["ours_ms"]
        out[order] = res
    print(json.dumps(out))


if __name__ == "__main__":
    main()
