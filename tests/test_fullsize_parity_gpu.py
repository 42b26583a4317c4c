"""Parity against the compiled reference (Oracle A) at the sizes the numbers are quoted on.

BASELINE.json configs 2-5 at their full sizes (300 k / 1 M / 3 M / 2 M Gaussians; SH2 + 7 spherical-Gaussian lobes), a
scene whose tile lists overflow the forward's shared-memory acceptance mask (MASK_WORDS * 32 = 8192 entries per warp),
and config 2 with the 2D filter and a random background.  Same bars as test_parity_gpu.py: integer outputs bit-exact,
images <= 1e-5, gradients <= 1e-4 of the tensor's max-norm -- and every comparison is RECORDED, element counts of the
two tolerance clauses included, in gpurun_out/r02_parity.json (copied to profiles/ when run by the builder).
"""
import json
import os

import pytest
import torch

import helpers as H
import scenes

pytestmark = pytest.mark.gpu

INT_KEYS = ("tiles_touched", "sorted_keys", "point_list", "ranges", "max_contributor", "n_contrib")
MASK_ENTRIES = 8192  # ggs_common.cuh MASK_WORDS * 32

CASES = {
    "c2_300k": dict(name="c2"),
    "c2_300k_k01_bg": dict(name="c2", kernel_size=0.1, random_bg=True, seed=2101),
    "c3_1m": dict(name="c3"),
    "c3_3m": dict(name="c3_3m"),
    "c4_2m": dict(name="c4"),
    "c5_sh2_sg7_1m": dict(name="c5"),
    # 60 k nearly transparent splats 40x larger than usual on a 256x256 image: tile lists of up to 56 k entries, pixels
    # that blend 47 k of them (CPU oracle) -- far past the 8192 entries the forward's shared-memory mask holds
    "long_lists": dict(name="c1", P=60_000, scale_mult=40.0, seed=31, opacity_range=(0.0045, 0.0085)),
}

REPORT_PATH = os.path.join(H.ROOT, "gpurun_out", "r02_parity.json")
_report = {}


def _record(tag, section, data):
    _report.setdefault(tag, {})[section] = data
    os.makedirs(os.path.dirname(REPORT_PATH), exist_ok=True)
    try:
        with open(REPORT_PATH) as f:
            old = json.load(f)
    except Exception:
        old = {}
    for k, v in _report.items():
        old.setdefault(k, {}).update(v)
    with open(REPORT_PATH, "w") as f:
        json.dump(old, f, indent=1, sort_keys=True)


def _free():
    import gc

    gc.collect()
    torch.cuda.empty_cache()


@pytest.mark.parametrize("tag", list(CASES))
def test_forward_full_size(tag, device, ref_module):
    if ref_module is None:
        pytest.skip("oracle/_ref not shipped")
    sc = scenes.make_scene(**CASES[tag])
    o_out, o_int = H.ours_forward_raw(sc, device)
    r_out, r_int = H.ref_forward_raw(ref_module, sc, device)
    rec = dict(P=sc.P, W=sc.cam.W, H=sc.cam.H, sh_degree=sc.sh_degree, sg_degree=sc.sg_degree,
               kernel_size=sc.kernel_size, num_rendered=[o_int["num_rendered"], r_int["num_rendered"]])
    lens = (r_int["ranges"][:, 1] - r_int["ranges"][:, 0]).to(torch.int64)
    rec["tile_list_max"] = int(lens.max())
    rec["tile_list_mean"] = float(lens.double().mean())
    rec["n_contrib_max"] = int(r_int["n_contrib"].max())
    rec["visible"] = int((r_out["radii"] > 0).sum())
    mism = {"radii": H.int_mismatch(o_out["radii"], r_out["radii"])}
    same_D = o_int["num_rendered"] == r_int["num_rendered"]
    for k in INT_KEYS:
        mism[k] = H.int_mismatch(o_int[k], r_int[k]) if (same_D or k in ("tiles_touched",)) else -1
    vis = r_out["radii"] > 0
    mism["depth_bits"] = H.int_mismatch(o_int["depths"].view(torch.int32)[vis], r_int["depths"].view(torch.int32)[vis])
    mism["clamped"] = H.int_mismatch(o_int["clamped"][vis], r_int["clamped"][vis])
    mism["record_bits"] = H.int_mismatch(o_int["records"][vis].view(torch.int32),
                                         r_int["records"][vis].view(torch.int32))
    rec["int_mismatches"] = mism
    img = {}
    for k in ("color", "alpha", "mdepth", "normal"):
        err, bad = H.image_rel_err(o_out[k], r_out[k])
        img[k] = dict(max_rel=err, pixels_above_tol=bad, bit_equal=bool(torch.equal(o_out[k], r_out[k])))
    rec["images"] = img
    rec["mdepth_pixels"] = int((r_out["mdepth"] > 0).sum())
    _record(tag, "forward", rec)
    if tag == "long_lists":
        assert rec["tile_list_max"] > 3 * MASK_ENTRIES, rec["tile_list_max"]
        assert rec["n_contrib_max"] > 2 * MASK_ENTRIES, rec["n_contrib_max"]
    assert same_D, rec["num_rendered"]
    for k, v in mism.items():
        assert v == 0, (k, v)
    H.check_depth_order(o_int, o_out["radii"])
    for k, v in img.items():
        assert v["max_rel"] <= H.IMG_TOL, (k, v)
    del o_out, o_int, r_out, r_int
    _free()


@pytest.mark.parametrize("tag", list(CASES))
def test_backward_full_size(tag, device, ref_module):
    if ref_module is None:
        pytest.skip("oracle/_ref not shipped")
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene(**CASES[tag])
    o_out, o_g = H.run_module(ours, sc, device)
    r_out, r_g = H.run_module(ref_module, sc, device)
    _, r_g2 = H.run_module(ref_module, sc, device)
    rec = {}
    cancel = None
    failures = []
    for k in H.GRAD_NAMES:
        a = o_g[k].detach().double().cpu()
        b = r_g[k].detach().double().cpu()
        b2 = r_g2[k].detach().double().cpu()
        if b.numel() == 0:
            continue
        den = float(b.abs().max())
        if den == 0.0:
            rec[k] = dict(max_norm=0.0, max_abs_err=float(a.abs().max()))
            if float(a.abs().max()) != 0.0:
                failures.append((k, "reference gradient is zero, ours is not"))
            continue
        diff = (a - b).abs()
        err = float(diff.max()) / den
        noise = float((b2 - b).abs().max()) / den
        over = diff > H.GRAD_TOL * den  # elements that do not meet the plain 1e-4 bar
        n_over = int(over.sum())
        r = dict(max_norm=den, max_rel=err, reference_run_to_run=noise, elements=int(b.numel()),
                 elements_above_1e4=n_over, bit_equal=bool(torch.equal(o_g[k], r_g[k])))
        if n_over:
            # clause 1: within 3x the reference's own run-to-run (float-atomic ordering) noise
            within_noise = diff <= 3.0 * noise * den
            need_cond = over & ~within_noise
            r["elements_needing_noise_clause"] = int((over & within_noise).sum())
            r["elements_needing_conditioning_clause"] = int(need_cond.sum())
            r["worst_excess_over_1e4"] = float((diff[over] / den).max()) - H.GRAD_TOL
            if int(need_cond.sum()):
                if cancel is None:
                    cancel = H.grad_cancellation(ours, sc, device)
                bound = H.GRAD_TOL * den + 32.0 * 5.96e-8 * cancel[k].reshape(a.shape)
                exc = (diff - bound)[need_cond]
                r["worst_excess_over_conditioning_bound"] = float(exc.max())
                if float(exc.max()) > 0:
                    failures.append((k, f"rel err {err:.3e}: {float(exc.max()):.3e} above the conditioning bound"))
        if noise > 1e-3:
            failures.append((k, f"reference run-to-run noise {noise:.3e}"))
        rec[k] = r
    img = {k: H.image_rel_err(o_out[k], r_out[k])[0] for k in ("color", "alpha", "mdepth", "normal")}
    rec["images_max_rel"] = img
    _record(tag, "backward", rec)
    for k, v in img.items():
        assert v <= H.IMG_TOL, (k, v)
    assert not failures, failures
    del o_out, o_g, r_out, r_g, r_g2
    _free()
