"""Generate golden vectors from the UNMODIFIED reference CUDA rasterizer (oracle/_ref) on a GPU box.

    python tests/golden/make_golden.py            # writes tests/golden/<scene>.npz (+ gpurun_out copy)

The reference ships no fixtures for this path, so these are outputs of the reference itself,
run on a B200 through its own pybind entry points on the seeded scenes of scenes.py.  Each file holds the
forward outputs, the bit-exact integer intermediates cut out of the reference's workspaces, and the
gradients for the scene's fixed upstream gradients.  Committed so that CPU-only runs can pin the oracle.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H  # noqa: E402

import scenes  # noqa: E402
from oracle import ref_loader  # noqa: E402

SCENES = {
    "tiny": dict(name="tiny"),
    "tiny_sg": dict(name="tiny_sg"),
    "tiny_k01_bg": dict(name="tiny", kernel_size=0.1, random_bg=True, seed=4242),
    "tiny_dense": dict(name="tiny", scale_mult=4.0, seed=77),
}


def main():
    ref = ref_loader.load()
    dev = "cuda:0"
    out_dirs = [HERE]
    if os.path.isdir(os.path.join(H.ROOT, "gpurun_out")):
        out_dirs.append(os.path.join(H.ROOT, "gpurun_out", "golden"))
    for tag, kw in SCENES.items():
        sc = scenes.make_scene(**kw)
        outs, inter = H.ref_forward_raw(ref, sc, dev)
        o2, grads = H.run_module(ref, sc, dev)
        vis = (outs["radii"] > 0).cpu().numpy()
        d = {"meta_P": np.int64(sc.P), "meta_D": np.int64(inter["num_rendered"])}
        for k in ("color", "alpha", "mdepth", "normal", "radii"):
            d["out_" + k] = outs[k].cpu().numpy()
        for k in ("tiles_touched", "point_offsets", "sorted_keys", "point_list", "ranges", "max_contributor",
                  "n_contrib"):
            d["int_" + k] = inter[k].cpu().numpy()
        # per-Gaussian floats only where the Gaussian is visible (the reference leaves the rest uninitialised)
        rec = inter["records"].cpu().numpy().copy()
        rec[~vis] = 0
        d["rec"] = rec
        dep = inter["depths"].cpu().numpy().copy()
        dep[~vis] = 0
        d["depths"] = dep
        cl = inter["clamped"].cpu().numpy().copy()
        cl[~vis] = 0
        d["clamped"] = cl
        for k, v in grads.items():
            d["grad_" + k] = v.cpu().numpy()
        for od in out_dirs:
            os.makedirs(od, exist_ok=True)
            np.savez_compressed(os.path.join(od, tag + ".npz"), **d)
        print(tag, "P", sc.P, "D", inter["num_rendered"], "visible", int(vis.sum()))


POINT_SCENES = {
    "tiny_dense": dict(scene=dict(name="tiny", scale_mult=4.0, seed=77), n=4000, seed=2),
    "c1": dict(scene=dict(name="c1"), n=6000, seed=3),
}


def point_queries(n, seed):
    """Query points of the point-query fixtures: most inside the unit scene, a tenth far away (outside the image)."""
    g = torch.Generator().manual_seed(seed)
    pts = torch.randn(n, 3, generator=g) * 0.7
    far = torch.rand(n, generator=g) < 0.1
    pts[far] = (torch.rand(int(far.sum()), 3, generator=g) - 0.5) * 12.0
    up = torch.randn(n, 3, generator=g)
    return pts, up


def points_main():
    """Golden vectors of the point-query entry points (sample_depth forward + backward, evaluate_sdf, integrate)
    from the reference: tests/golden/points/<scene>.npz."""
    ref = ref_loader.load()
    dev = "cuda:0"
    out_dirs = [os.path.join(HERE, "points")]
    if os.path.isdir(os.path.join(H.ROOT, "gpurun_out")):
        out_dirs.append(os.path.join(H.ROOT, "gpurun_out", "golden", "points"))
    for tag, case in POINT_SCENES.items():
        sc = scenes.make_scene(**case["scene"])
        pts, up = point_queries(case["n"], case["seed"])
        rs = H.settings_for(ref, sc, dev)
        rast = ref.GaussianRasterizer(rs)
        x = {k: getattr(sc, k).to(dev).clone().requires_grad_(True) for k in ("means3D", "opacities", "scales", "rotations")}
        p = pts.to(dev).clone().requires_grad_(True)
        cam, inside = rast.sample_depth(p, x["means3D"], x["opacities"], x["scales"], x["rotations"])
        (cam * up.to(dev)).sum().backward()
        d = {"points": pts.numpy(), "upstream": up.numpy(), "sd_camera_points": cam.detach().cpu().numpy(),
             "sd_inside": inside.cpu().numpy(), "sd_grad_points3D": p.grad.cpu().numpy()}
        for k, v in x.items():
            d["sd_grad_" + k] = v.grad.cpu().numpy()
        with torch.no_grad():
            a = [getattr(sc, k).to(dev) for k in ("means3D", "opacities", "scales", "rotations")]
            depth, sdf, ins = rast.evaluate_sdf(pts.to(dev), *a)
            alpha, ins2 = rast.integrate(pts.to(dev), *a)
        d.update(sdf_depth=depth.cpu().numpy(), sdf_sdf=sdf.cpu().numpy(), sdf_inside=ins.cpu().numpy(),
                 int_alpha=alpha.cpu().numpy(), int_inside=ins2.cpu().numpy())
        for od in out_dirs:
            os.makedirs(od, exist_ok=True)
            np.savez_compressed(os.path.join(od, tag + ".npz"), **d)
        print("points", tag, "queries", case["n"], "with a median depth", int(inside.sum()))


if __name__ == "__main__":
    if "--points-only" not in sys.argv:
        main()
    points_main()
