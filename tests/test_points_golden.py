"""Point-query entry points (sample_depth, evaluate_sdf, integrate) against golden vectors the reference produced
on a B200 (tests/golden/points/*.npz, tests/golden/make_golden.py): the C oracle on the CPU, the CUDA path on the GPU."""
import glob
import os

import numpy as np
import pytest
import torch

import helpers as H
import scenes
from golden.make_golden import POINT_SCENES

GOLDEN = sorted(glob.glob(os.path.join(H.ROOT, "tests", "golden", "points", "*.npz")))
IDS = [os.path.basename(p)[:-4] for p in GOLDEN]


def test_point_golden_present():
    assert len(GOLDEN) >= 2, "point-query golden vectors missing: run tests/golden/make_golden.py --points-only on a GPU box"


def _kw(sc, pts):
    return dict(points3D=pts, means3D=sc.means3D, opacities=sc.opacities, scales=sc.scales, rotations=sc.rotations,
                viewmatrix=sc.cam.viewmatrix, projmatrix=sc.cam.projmatrix, campos=sc.cam.campos,
                tanfovx=sc.cam.tanfovx, tanfovy=sc.cam.tanfovy, image_height=sc.cam.H, image_width=sc.cam.W)


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_oracle_point_queries_match_reference_golden(path):
    """IEEE C restatement vs the fast-math reference: a handful of knife-edge queries may flip `inside`."""
    from oracle.oracle import Oracle

    tag = os.path.basename(path)[:-4]
    gold = np.load(path)
    sc = scenes.make_scene(**POINT_SCENES[tag]["scene"])
    pts = torch.from_numpy(gold["points"])
    orc = Oracle()
    sd = orc.sample_depth(**_kw(sc, pts))
    gin = torch.from_numpy(gold["sd_inside"])
    flips = int((sd["inside"] != gin).sum())
    assert flips <= max(2, len(pts) // 1000), f"{flips} queries differ in `inside`"
    both = sd["inside"] & gin
    assert int(both.sum()) > len(pts) // 50
    err, _ = H.image_rel_err(sd["camera_points"][both], torch.from_numpy(gold["sd_camera_points"])[both])
    assert err < 2e-3, err
    sdf = orc.evaluate_sdf(**_kw(sc, pts))
    gin2 = torch.from_numpy(gold["sdf_inside"])
    both2 = sdf["inside"] & gin2
    assert int((sdf["inside"] != gin2).sum()) <= max(2, len(pts) // 1000)
    for k in ("depth", "sdf"):
        err, _ = H.image_rel_err(sdf[k][both2], torch.from_numpy(gold["sdf_" + k])[both2])
        assert err < 2e-3, (k, err)
    it = orc.integrate(**_kw(sc, pts))
    assert torch.equal(it["inside"], torch.from_numpy(gold["int_inside"]))
    err, _ = H.image_rel_err(it["alpha"], torch.from_numpy(gold["int_alpha"]))
    assert err < 2e-3, err


@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_oracle_sample_depth_backward_matches_reference_golden(path):
    """The C restatement of the sample_depth backward against the reference's gradients for the same upstream.
    The median depth comes out of a bisection, so IEEE vs fast-math moves a few queries by ~1e-3 (measured: worst
    element 1.9e-3 of the largest gradient, median element 2e-5 relative); a wrong or missing term would show in the
    median."""
    from oracle.oracle import Oracle

    tag = os.path.basename(path)[:-4]
    gold = np.load(path)
    sc = scenes.make_scene(**POINT_SCENES[tag]["scene"])
    pts = torch.from_numpy(gold["points"])
    up = torch.from_numpy(gold["upstream"])
    g = Oracle().sample_depth_backward(up, kernel_size=sc.kernel_size, **_kw(sc, pts))
    for k, v in g.items():
        ref = torch.from_numpy(gold["sd_grad_" + k]).reshape(v.shape).double()
        assert H.grad_rel_err(v, ref) < 4e-3, k
        big = ref.abs() > 1e-3 * ref.abs().max()
        assert int(big.sum()) > 100, k
        rel = ((v.double() - ref).abs() / ref.abs())[big]
        assert float(rel.median()) < 1e-4, (k, float(rel.median()))
        # queries outside the image / Gaussians never blended get exactly no gradient in both
        quiet = ref.abs().reshape(ref.shape[0], -1).sum(1) == 0
        assert float(v.reshape(v.shape[0], -1)[quiet].abs().max()) < 1e-30, k


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=IDS)
def test_cuda_point_queries_match_reference_golden(path, device):
    import diff_gaussian_rasterization as ours

    tag = os.path.basename(path)[:-4]
    gold = np.load(path)
    sc = scenes.make_scene(**POINT_SCENES[tag]["scene"])
    pts = torch.from_numpy(gold["points"]).to(device)
    up = torch.from_numpy(gold["upstream"]).to(device)
    rast = ours.GaussianRasterizer(H.settings_for(ours, sc, device))
    x = {k: getattr(sc, k).to(device).clone().requires_grad_(True) for k in ("means3D", "opacities", "scales", "rotations")}
    p = pts.clone().requires_grad_(True)
    cam, inside = rast.sample_depth(p, x["means3D"], x["opacities"], x["scales"], x["rotations"])
    (cam * up).sum().backward()
    assert np.array_equal(inside.cpu().numpy(), gold["sd_inside"])
    err, bad = H.image_rel_err(cam, torch.from_numpy(gold["sd_camera_points"]))
    assert err <= H.IMG_TOL, (err, bad)
    grads = dict(points3D=p.grad, **{k: v.grad for k, v in x.items()})
    for k, g in grads.items():
        e = H.grad_rel_err(g, torch.from_numpy(gold["sd_grad_" + k]))
        assert e <= H.GRAD_TOL, (k, e)
    a = [getattr(sc, k).to(device) for k in ("means3D", "opacities", "scales", "rotations")]
    depth, sdf, ins = rast.evaluate_sdf(pts, *a)
    assert np.array_equal(ins.cpu().numpy(), gold["sdf_inside"])
    for v, k in ((depth, "sdf_depth"), (sdf, "sdf_sdf")):
        err, bad = H.image_rel_err(v, torch.from_numpy(gold[k]))
        assert err <= H.IMG_TOL, (k, err, bad)
    alpha, ins2 = rast.integrate(pts, *a)
    assert np.array_equal(ins2.cpu().numpy(), gold["int_inside"])
    err, bad = H.image_rel_err(alpha, torch.from_numpy(gold["int_alpha"]))
    assert err <= H.IMG_TOL, (err, bad)
