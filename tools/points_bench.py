"""Point-query paths on the GPU box: ours vs the compiled reference (Oracle A), times and parity.

    python tools/points_bench.py [--what integrate,sample_depth,evaluate_sdf] [--scene c5] [--points 7500000]
                                 [--iters 5] [--out gpurun_out/points_bench.json]

integrate    : the mesh-extraction opacity query (mesh_extract_tetrahedra.py:65-87): one chunk of the 15-point
               stencil (scenes.tetra_points) against one view -- what the reference runs per (chunk, view).
sample_depth : the multi-view loss query (utils/loss_utils.py:148-166): every pixel of view A unprojected at its
               rendered median depth, sampled in neighbouring view B (15 degrees apart), forward + backward.
Times are CUDA-event medians on the current stream after warm-up.  Not run under a profiler.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import helpers as H  # noqa: E402

import scenes  # noqa: E402
from oracle import ref_loader  # noqa: E402


def timed(fn, iters, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def bench_integrate(mods, scene, npts, iters, device):
    sc = scenes.make_scene(scene)
    pts_all = scenes.tetra_points(sc, device)
    chunks = torch.chunk(pts_all, pts_all.shape[0] // 10_000_000 + 1)  # the reference's chunking (:70)
    pts = chunks[0] if npts <= 0 else pts_all[:npts].contiguous()
    x = [getattr(sc, k).to(device) for k in ("means3D", "opacities", "scales", "rotations")]
    out = {"scene": scene, "P": sc.P, "points": int(pts.shape[0]), "W": sc.cam.W, "H": sc.cam.H}
    res = {}
    for name, mod in mods.items():
        rast = mod.GaussianRasterizer(H.settings_for(mod, sc, device)._replace(bg=None))

        def run():
            return rast.integrate(pts, x[0], x[1], x[2], x[3], None, None)

        res[name] = run()
        torch.cuda.synchronize()
        out[name + "_ms"] = timed(run, iters)
        out[name + "_Mpoints_per_s"] = pts.shape[0] / out[name + "_ms"] / 1e3
    if "ref" in res:
        (a, ia), (b, ib) = res["ours"], res["ref"]
        out["inside_mismatch"] = int((ia != ib).sum())
        out["inside_count"] = int(ib.sum())
        out["alpha_max_abs_diff"] = float((a - b).abs().max())
        out["alpha_bit_equal"] = bool(torch.equal(a, b))
        out["speedup"] = out["ref_ms"] / out["ours_ms"]
    return out


def _pixel_queries(mod, scene, device):
    """World points of every pixel of view A at its rendered median depth; returns (scene of view B, points [H,W,3])."""
    n_views = 24
    sa = scenes.make_scene(scene, view=0, n_views=n_views)
    sb = scenes.make_scene(scene, view=1, n_views=n_views)
    out, _ = H.run_module(mod, sa, device, backward=False)
    md = out["mdepth"][0]  # ray-space median depth * rln == z depth along the camera axis
    Hh, W = md.shape
    c = sa.cam
    fx, fy = W / (2 * c.tanfovx), Hh / (2 * c.tanfovy)
    ix = (torch.arange(W, device=device, dtype=torch.float32) - (W - 1) / 2) / fx
    iy = (torch.arange(Hh, device=device, dtype=torch.float32) - (Hh - 1) / 2) / fy
    pts_cam = torch.stack([md * ix[None, :], md * iy[:, None], md], dim=-1)  # [H,W,3]
    view = c.viewmatrix.to(device)  # transposed world->view: p_view = p_world @ view[:3,:3] + view[3,:3]
    Rm, t = view[:3, :3], view[3, :3]
    pts_world = (pts_cam - t) @ Rm.t()
    return sb, pts_world.contiguous()


def bench_sample_depth(mods, scene, iters, device):
    sb, pts = _pixel_queries(mods["ours"], scene, device)
    g = torch.Generator().manual_seed(5)
    up = torch.randn(pts.shape, generator=g).to(device)
    out = {"scene": scene, "P": sb.P, "points": int(pts.shape[0] * pts.shape[1]), "W": sb.cam.W, "H": sb.cam.H}
    res = {}
    for name, mod in mods.items():
        rast = mod.GaussianRasterizer(H.settings_for(mod, sb, device)._replace(bg=0))
        x = {k: getattr(sb, k).to(device).clone().requires_grad_(True)
             for k in ("means3D", "opacities", "scales", "rotations")}
        p = pts.clone().requires_grad_(True)

        def fwd():
            return rast.sample_depth(p, x["means3D"], x["opacities"], x["scales"], x["rotations"])

        def fwdbwd():
            for v in x.values():
                v.grad = None
            p.grad = None
            cam, inside = fwd()
            (cam * up).sum().backward()
            return cam, inside

        cam, inside = fwdbwd()
        torch.cuda.synchronize()
        res[name] = (cam.detach(), inside.detach(), {k: v.grad.detach().clone() for k, v in x.items()},
                     p.grad.detach().clone())
        with torch.no_grad():
            out[name + "_fwd_ms"] = timed(fwd, iters)
        out[name + "_fwdbwd_ms"] = timed(fwdbwd, iters)
    if "ref" in res:
        o, r = res["ours"], res["ref"]
        out["inside_mismatch"] = int((o[1] != r[1]).sum())
        out["inside_count"] = int(r[1].sum())
        out["camera_points_max_rel"] = H.image_rel_err(o[0], r[0])[0]
        out["camera_points_bit_equal"] = bool(torch.equal(o[0], r[0]))
        for k in o[2]:
            out["grad_" + k] = H.grad_rel_err(o[2][k], r[2][k])
        out["grad_points3D"] = H.grad_rel_err(o[3], r[3])
        out["speedup_fwd"] = out["ref_fwd_ms"] / out["ours_fwd_ms"]
        out["speedup_fwdbwd"] = out["ref_fwdbwd_ms"] / out["ours_fwdbwd_ms"]
    return out


def bench_evaluate_sdf(mods, scene, npts, iters, device):
    sc = scenes.make_scene(scene)
    pts_all = scenes.tetra_points(sc, device)
    pts = pts_all[:npts if npts > 0 else 2_000_000].contiguous()
    x = [getattr(sc, k).to(device) for k in ("means3D", "opacities", "scales", "rotations")]
    out = {"scene": scene, "P": sc.P, "points": int(pts.shape[0])}
    res = {}
    for name, mod in mods.items():
        rast = mod.GaussianRasterizer(H.settings_for(mod, sc, device)._replace(bg=None))

        def run():
            return rast.evaluate_sdf(pts, x[0], x[1], x[2], x[3], None, None)

        res[name] = run()
        torch.cuda.synchronize()
        out[name + "_ms"] = timed(run, iters)
    if "ref" in res:
        (od, os_, oi), (rd, rs_, ri) = res["ours"], res["ref"]
        out["inside_mismatch"] = int((oi != ri).sum())
        out["depth_max_rel"] = H.image_rel_err(od, rd)[0]
        out["sdf_max_rel"] = H.image_rel_err(os_, rs_)[0]
        out["speedup"] = out["ref_ms"] / out["ours_ms"]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="integrate,sample_depth")
    ap.add_argument("--scene", default="c5")
    ap.add_argument("--sample-scene", default="c2")
    ap.add_argument("--points", type=int, default=0)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--out", default="gpurun_out/points_bench.json")
    args = ap.parse_args()
    device = "cuda:0"
    import diff_gaussian_rasterization as ours

    mods = {"ours": ours}
    if not args.no_ref and ref_loader.available():
        mods["ref"] = ref_loader.load()
    report = {}
    for what in args.what.split(","):
        if what == "integrate":
            report[what] = bench_integrate(mods, args.scene, args.points, args.iters, device)
        elif what == "sample_depth":
            for s in args.sample_scene.split("+"):
                report[what + "_" + s] = bench_sample_depth(mods, s, args.iters, device)
        elif what == "evaluate_sdf":
            report[what] = bench_evaluate_sdf(mods, args.scene, args.points, args.iters, device)
        print(what, json.dumps({k: v for k, v in report.items() if k.startswith(what)}), flush=True)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
