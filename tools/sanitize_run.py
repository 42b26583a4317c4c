"""One pass over every product kernel on config 1, meant to be run under compute-sanitizer on the GPU box:

    compute-sanitizer --tool memcheck  python tools/sanitize_run.py
    compute-sanitizer --tool racecheck python tools/sanitize_run.py

Image forward + backward (with and without require_depth, SH and SG colour paths), sample_depth forward + backward,
integrate, evaluate_sdf, markVisible, the fused activations, the patch-NCC and fused-SSIM kernels.  SURVEY.md section 5.
"""
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

    dev = "cuda:0"
    small = "--small" in sys.argv
    for kw in (dict(name="c1", P=3000 if small else 10_000), dict(name="tiny_sg"),
               dict(name="tiny", kernel_size=0.1, random_bg=True, seed=4242)):
        sc = scenes.make_scene(**kw)
        H.run_module(ours, sc, dev, require_depth=True)
        H.run_module(ours, sc, dev, require_depth=False)
    sc = scenes.make_scene("c1", P=3000 if small else 10_000, scale_mult=2.0)
    g = torch.Generator().manual_seed(1)
    pts = (torch.randn(20_000, 3, generator=g) * 0.7).to(dev)
    up = torch.randn(20_000, 3, generator=g).to(dev)
    x = {k: getattr(sc, k).to(dev).clone().requires_grad_(True) for k in ("means3D", "opacities", "scales", "rotations")}
    rs = H.settings_for(ours, sc, dev)
    rast = ours.GaussianRasterizer(rs._replace(bg=None))
    p = pts.clone().requires_grad_(True)
    cam, inside = ours.GaussianRasterizer(rs._replace(bg=0)).sample_depth(p, x["means3D"], x["opacities"], x["scales"],
                                                                         x["rotations"])
    (cam * up).sum().backward()
    with torch.no_grad():
        rast.integrate(pts, x["means3D"], x["opacities"], x["scales"], x["rotations"], None, None)
        rast.evaluate_sdf(pts, x["means3D"], x["opacities"], x["scales"], x["rotations"], None, None)
        rast.markVisible(x["means3D"])
    raw = [torch.randn(sc.P, 3, device=dev, requires_grad=True), torch.randn(sc.P, 1, device=dev, requires_grad=True),
           torch.randn(sc.P, 4, device=dev, requires_grad=True)]
    s, o, r = ours.activate_gaussians(raw[0], raw[1], raw[2], torch.rand(sc.P, 1, device=dev) * 0.01)
    (s.sum() + o.sum() + r.sum()).backward()
    torch.cuda.synchronize()
    print("rasterizer kernels done")

    import fused_ssim

    a = torch.rand(1, 3, 120, 150, device=dev, requires_grad=True)
    b = torch.rand(1, 3, 120, 150, device=dev)
    for pad in ("same", "valid"):
        fused_ssim.fused_ssim(a, b, padding=pad).backward()
    torch.cuda.synchronize()
    print("ssim kernels done")

    import ncc_scenes
    import warp_patch_ncc as wpn

    nc = ncc_scenes.make_pair(height=120, width=160, n=4001, seed=7)
    a = ncc_scenes.args_of(nc, dev)
    wpn._run(a[0], a[1], a[2], a[5], a[6], wpn.WarpParams(a[3], a[4], *a[7:], False))
    torch.cuda.synchronize()
    print("ncc kernel done")
    print("sanitize_run complete")


if __name__ == "__main__":
    main()
