"""GPU parity + timing report: product vs the compiled reference rasterizer (Oracle A).

    python tools/parity_report.py [--scenes tiny,c1,c2] [--out gpurun_out/parity.json] [--time]

Runs on a GPU box only.  Writes one JSON object per scene with mismatch counts for the bit-exact
integer outputs, relative errors for images and gradients (same metrics as tests/helpers.py), the
reference's own run-to-run gradient noise (atomic ordering) and, with --time, fwd / fwd+bwd ms.
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import helpers as H  # noqa: E402

import scenes  # noqa: E402
from oracle import ref_loader  # noqa: E402


def time_module(mod, sc, device, iters=10, warmup=3, backward=True):
    rs = H.settings_for(mod, sc, device, True)
    x = H.leaf_inputs(sc, device)
    g = {k: v.to(device) for k, v in sc.grads.items()}
    rast = mod.GaussianRasterizer(rs)

    def step():
        for v in x.values():
            v.grad = None
        color, radii, mdepth, alpha, normal = rast(
            means3D=x["means3D"], means2D=x["means2D"], opacities=x["opacities"], shs=x["shs"], sg_axis=x["sg_axis"],
            sg_sharpness=x["sg_sharpness"], sg_color=x["sg_color"], colors_precomp=None, scales=x["scales"],
            rotations=x["rotations"], cov3D_precomp=None)
        if backward:
            torch.autograd.backward([color, mdepth, alpha, normal], [g["color"], g["mdepth"], g["alpha"], g["normal"]])

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenes", default="tiny,tiny_sg,c1")
    ap.add_argument("--out", default="gpurun_out/parity.json")
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--kernel-size", type=float, default=0.0)
    args = ap.parse_args()
    device = "cuda:0"
    import diff_gaussian_rasterization as ours

    ref = ref_loader.load()
    report = {}
    for name in args.scenes.split(","):
        t0 = time.time()
        sc = scenes.make_scene(name, kernel_size=args.kernel_size)
        r = {"P": sc.P, "W": sc.cam.W, "H": sc.cam.H}
        o_out, o_int = H.ours_forward_raw(sc, device)
        r_out, r_int = H.ref_forward_raw(ref, sc, device)
        r["num_rendered"] = [o_int["num_rendered"], r_int["num_rendered"]]
        for k in ("tiles_touched", "sorted_keys", "point_list", "ranges", "max_contributor",
                  "n_contrib", "clamped"):
            r["int_" + k] = H.int_mismatch(o_int[k], r_int[k])
        r["int_radii"] = H.int_mismatch(o_out["radii"], r_out["radii"])
        r["int_depth_bits"] = H.int_mismatch(o_int["depths"].view(torch.int32)[r_out["radii"] > 0],
                                             r_int["depths"].view(torch.int32)[r_out["radii"] > 0])
        vis = (r_out["radii"] > 0) & (o_out["radii"] > 0)
        rec_o, rec_r = o_int["records"][vis], r_int["records"][vis]
        names = ["xy"] * 2 + ["conic"] * 3 + ["opac"] + ["ray_plane"] * 4 + ["rgb"] * 3 + ["normal"] * 3
        # our record order: xy(2) conic(3) opac(1) ray_plane(4) rgb(3) normal(3)  == reference concat order
        for grp in ("xy", "conic", "opac", "ray_plane", "rgb", "normal"):
            cols = [i for i, n in enumerate(names) if n == grp]
            a, b = rec_o[:, cols].double(), rec_r[:, cols].double()
            den = torch.clamp(b.abs(), min=1e-3 * float(b.abs().max()) + 1e-30)
            r["rec_" + grp] = float(((a - b).abs() / den).max()) if a.numel() else 0.0
            r["rec_bits_" + grp] = int((rec_o[:, cols].contiguous().view(torch.int32) !=
                                        rec_r[:, cols].contiguous().view(torch.int32)).sum())
        for k in ("color", "alpha", "mdepth", "normal"):
            e, n = H.image_rel_err(o_out[k], r_out[k])
            r["img_" + k] = e
            r["img_bad_" + k] = n
        # module-level forward+backward
        og_out, og = H.run_module(ours, sc, device)
        rg_out, rg = H.run_module(ref, sc, device)
        rg2_out, rg2 = H.run_module(ref, sc, device)
        for k in H.GRAD_NAMES:
            r["grad_" + k] = H.grad_rel_err(og[k], rg[k])
            r["refnoise_" + k] = H.grad_rel_err(rg2[k], rg[k])
        # no-depth variant
        o2, g2 = H.run_module(ours, sc, device, require_depth=False)
        r2, gr2 = H.run_module(ref, sc, device, require_depth=False)
        r["nodepth_img_color"] = H.image_rel_err(o2["color"], r2["color"])[0]
        r["nodepth_grad_max"] = max(H.grad_rel_err(g2[k], gr2[k]) for k in H.GRAD_NAMES)
        if sc.P <= 20000:
            # CPU restatement (Oracle B) against the compiled reference
            from oracle.oracle import Oracle, scene_kwargs

            orc = Oracle()
            oo = orc.forward(**scene_kwargs(sc))
            oi = orc.intermediates()
            r["oracleB_num_rendered"] = oo["num_rendered"]
            r["oracleB_int_radii"] = H.int_mismatch(oo["radii"], r_out["radii"])
            r["oracleB_int_tiles_touched"] = int((oi["tiles_touched"].astype("int64") !=
                                                  r_int["tiles_touched"].cpu().numpy().astype("int64")).sum())
            if oo["num_rendered"] == r_int["num_rendered"]:
                r["oracleB_int_point_list"] = int((oi["point_list"].astype("int64") !=
                                                   r_int["point_list"].cpu().numpy().astype("int64")).sum())
                r["oracleB_int_n_contrib"] = int((oi["n_contrib"].astype("int64") !=
                                                  r_int["n_contrib"].cpu().numpy().astype("int64")).sum())
            for k in ("color", "alpha", "mdepth", "normal"):
                r["oracleB_img_" + k] = H.image_rel_err(oo[k], r_out[k])[0]
            gsc = sc.grads
            og_b = orc.backward(gsc["color"], gsc["mdepth"], gsc["alpha"], gsc["normal"])
            for k in H.GRAD_NAMES:
                r["oracleB_grad_" + k] = H.grad_rel_err(og_b[k].reshape(rg[k].shape), rg[k])
        if args.time:
            r["ms_fwd_ours"] = time_module(ours, sc, device, backward=False)
            r["ms_fwd_ref"] = time_module(ref, sc, device, backward=False)
            r["ms_fwdbwd_ours"] = time_module(ours, sc, device)
            r["ms_fwdbwd_ref"] = time_module(ref, sc, device)
        r["wall_s"] = time.time() - t0
        report[name] = r
        print(name, json.dumps(r), flush=True)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
