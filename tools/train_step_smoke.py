#!/usr/bin/env python
"""A few training iterations through every kernel of this repository on one GPU (run on the box).

    python tools/train_step_smoke.py [--workload c2] [--iters 30] [--no-ncc]

The reference's train.py cannot run on the box (its sources are not there), so this is the same
iteration restated on synthetic data: fused activations (N2) -> rasterizer forward (colour, median depth, normal)
-> loss = 0.8 L1 + 0.2 (1 - fused_ssim) (train.py:189) + 0.15 patch-NCC against a second view
(utils/loss_utils.py:227-264) -> backward through all of them -> Adam.  The target images are renders of the
unperturbed scene.  Prints one JSON object: loss at the first and last iteration, ms per iteration, the share of
valid NCC patches.  The loss must go down; that is the check."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import scenes  # noqa: E402


def settings(mod, sc, dev):
    c = sc.cam
    return mod.GaussianRasterizationSettings(
        image_height=c.H, image_width=c.W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, kernel_size=sc.kernel_size,
        bg=sc.bg.to(dev), scale_modifier=1.0, viewmatrix=c.viewmatrix.to(dev), projmatrix=c.projmatrix.to(dev),
        sh_degree=sc.sh_degree, sg_degree=0, campos=c.campos.to(dev), prefiltered=False, require_depth=True,
        debug=False)


def render(mod, sc, dev, means, opac, scales, rots, shs):
    rast = mod.GaussianRasterizer(settings(mod, sc, dev))
    means2D = torch.zeros(means.shape[0], 3, device=dev, requires_grad=True)
    return rast(means3D=means, means2D=means2D, opacities=opac, shs=shs, sg_axis=sc.sg_axis.to(dev),
                sg_sharpness=sc.sg_sharpness.to(dev), sg_color=sc.sg_color.to(dev), colors_precomp=None,
                scales=scales, rotations=rots, cov3D_precomp=None)


def grey(img):
    return (0.299 * img[0] + 0.587 * img[1] + 0.114 * img[2]).contiguous()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--no-ncc", action="store_true")
    args = ap.parse_args()
    import diff_gaussian_rasterization as dgr
    import fused_ssim
    import warp_patch_ncc

    dev = "cuda:0"
    views = [scenes.make_scene(args.workload, view=v, n_views=24) for v in (0, 1)]  # neighbours 15 degrees apart
    sc = views[0]
    P = sc.P
    gt = {k: getattr(sc, k).to(dev) for k in ("means3D", "opacities", "scales", "rotations", "shs")}
    sg0 = sc.sg_color.shape[1]
    if sg0:
        raise SystemExit("pick a workload without spherical-Gaussian lobes")
    with torch.no_grad():
        targets = [render(dgr, v, dev, gt["means3D"], gt["opacities"], gt["scales"], gt["rotations"], gt["shs"])[0]
                   for v in views]
    g = torch.Generator(device=dev).manual_seed(1)
    # raw parameters as the trainer keeps them (scene/gaussian_model.py:158-212), perturbed
    raw = {
        "xyz": (gt["means3D"] + 0.004 * torch.randn(P, 3, device=dev, generator=g)),
        "scaling": torch.log(gt["scales"]) + 0.05 * torch.randn(P, 3, device=dev, generator=g),
        "opacity": torch.logit(gt["opacities"].clamp(1e-4, 1 - 1e-4)) + 0.1 * torch.randn(P, 1, device=dev, generator=g),
        "rotation": gt["rotations"] + 0.02 * torch.randn(P, 4, device=dev, generator=g),
        "shs": gt["shs"] + 0.02 * torch.randn(gt["shs"].shape, device=dev, generator=g),
    }
    params = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
    filter_3D = torch.zeros(P, 1, device=dev)
    opt = torch.optim.Adam([{"params": [params["xyz"]], "lr": 1e-4}, {"params": [params["scaling"]], "lr": 2e-3},
                            {"params": [params["opacity"]], "lr": 1e-2}, {"params": [params["rotation"]], "lr": 1e-3},
                            {"params": [params["shs"]], "lr": 2e-3}])
    c0, c1 = views[0].cam, views[1].cam
    fx, fy = c0.W / (2 * c0.tanfovx), c0.H / (2 * c0.tanfovy)
    cx, cy = (c0.W - 1) / 2.0, (c0.H - 1) / 2.0
    w0, w1 = c0.viewmatrix.to(dev), c1.viewmatrix.to(dev)  # transposed world->view, as Camera.world_view_transform
    rel_r = w1[:3, :3].transpose(-1, -2) @ w0[:3, :3]       # utils/loss_utils.py:232-233
    rel_t = -rel_r @ w0[3, :3] + w1[3, :3]
    grey_r, grey_n = grey(targets[0]), grey(targets[1])
    ys, xs = torch.meshgrid(torch.arange(4, c0.H - 4, 3, device=dev), torch.arange(4, c0.W - 4, 3, device=dev),
                            indexing="ij")
    pixels = torch.stack([xs.reshape(-1), ys.reshape(-1)], dim=1).to(torch.int32).contiguous()
    flat = (pixels[:, 1].long() * c0.W + pixels[:, 0].long())

    losses, valid_share = [], None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(args.iters):
        if it == 3:
            e0.record()
        scales, opac, rots = dgr.activate_gaussians(params["scaling"], params["opacity"], params["rotation"], filter_3D)
        color, radii, mdepth, alpha, normal = render(dgr, sc, dev, params["xyz"], opac, scales, rots, params["shs"])
        l1 = (color - targets[0]).abs().mean()
        ssim = fused_ssim.fused_ssim(color.unsqueeze(0), targets[0].unsqueeze(0), padding="valid")
        loss = 0.8 * l1 + 0.2 * (1.0 - ssim)
        if not args.no_ncc:
            depth_sel = mdepth.reshape(-1)[flat]
            normal_sel = F.normalize(normal.reshape(3, -1)[:, flat].transpose(1, 0), dim=-1)
            keep = depth_sel > 0
            cc, valid = warp_patch_ncc.warp_patch_ncc(depth_sel[keep], normal_sel[keep].contiguous(), pixels[keep],
                                                      rel_r.T, rel_t, grey_r, grey_n, fx, fy, cx, cy, fx, fy, cx, cy,
                                                      False)
            ncc = torch.clamp(1 - cc, 0.0, 2.0)
            mask = (ncc < 0.9) & valid
            valid_share = float(valid.float().mean())
            if bool(mask.any()):
                loss = loss + 0.15 * ncc[mask].mean()
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    e1.record()
    torch.cuda.synchronize()
    out = {"workload": args.workload, "gaussians": P, "image": [c0.H, c0.W], "iters": args.iters,
           "loss_first": losses[0], "loss_last": losses[-1], "loss_min": min(losses),
           "ms_per_iter": e0.elapsed_time(e1) / max(args.iters - 3, 1), "ncc_valid_share": valid_share,
           "ncc_points": int(pixels.shape[0]) if not args.no_ncc else 0}
    print(json.dumps(out))
    if not (losses[-1] < losses[0]):
        raise SystemExit("loss did not go down")


if __name__ == "__main__":
    main()
