#!/usr/bin/env python
"""Fused SSIM loss: our library against the reference's torch formulation on one GPU (run on the box).

    python tools/ssim_bench.py [--height 822] [--width 1237] [--iters 200]

Times forward + backward of `fused_ssim.fused_ssim(img1, img2, padding="valid")` and of the reference's conv2d-based
`_ssim` (utils/loss_utils.py:52-73, restated in tests/ssim_reference.py, float32) with CUDA events after warm-up, and
prints one JSON object with the times, the value / gradient differences against float64, and the achieved fraction of
the HBM roofline for our two kernels (44 algorithmic bytes per pixel and channel)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

import ssim_reference as R  # noqa: E402


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
    ap.add_argument("--height", type=int, default=822)
    ap.add_argument("--width", type=int, default=1237)
    ap.add_argument("--iters", type=int, default=200)
    args = ap.parse_args()
    import fused_ssim as ours

    dev = "cuda:0"
    img1, img2 = R.image_pair((1, 3, args.height, args.width), seed=1)
    img1, img2 = img1.to(dev), img2.to(dev)
    x = img1.clone().requires_grad_(True)

    def step_ours():
        x.grad = None
        ours.fused_ssim(x, img2, padding="valid").backward()

    def step_torch():
        x.grad = None
        R.ssim_mean(x, img2, "valid", dtype=torch.float32).backward()

    out = {"image": [3, args.height, args.width], "ours_ms": timed(step_ours, args.iters),
           "torch_conv_ms": timed(step_torch, args.iters)}
    out["speedup_vs_torch_conv"] = out["torch_conv_ms"] / out["ours_ms"]
    want, want_grad = R.value_and_grad(img1, img2, "valid")
    step_ours()
    got = ours.fused_ssim(img1, img2, padding="valid")
    out["value_rel_err"] = abs(float(got) - float(want)) / abs(float(want))
    out["grad_rel_err"] = float((x.grad.double() - want_grad).abs().max() / want_grad.abs().max())
    peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = None
    if os.path.exists(peaks):
        with open(peaks) as fh:
            pj = json.load(fh)
        hbm = next((float(v) for k, v in pj.items() if "hbm" in k.lower() and isinstance(v, (int, float))), None)
    alg_bytes = 44.0 * 3 * args.height * args.width
    out["alg_bytes"] = alg_bytes
    out["achieved_gbs"] = alg_bytes / out["ours_ms"] / 1e6
    if hbm:
        out["hbm_peak_gbs"] = hbm
        out["frac_hbm"] = out["achieved_gbs"] / hbm
    print(json.dumps(out))


if __name__ == "__main__":
    main()
