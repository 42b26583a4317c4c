"""The reference's OWN callers run unchanged on the drop-in (north_star: "so gaussian_renderer/__init__.py, train.py,
render.py and mesh_extract*.py run unchanged").

oracle/build_ref.py installs the reference's gaussian_renderer/__init__.py, scene/gaussian_model.py, scene/cameras.py and
utils/loss_utils.py verbatim (git-ignored install artefacts, like the compiled reference).  Here they are imported twice
-- once with `diff_gaussian_rasterization` / `warp_patch_ncc` bound to this repo's packages, once bound to the compiled
reference -- and fed the same GaussianModel (the reference's class, parameters from a scenes.Scene, 3D filter on) and the
same Camera objects:

    render()              gaussian_renderer/__init__.py:18-98     + autograd through the model's activations
    integrate()           :101-160      (bg=None, the model's active SH / SG degrees in the settings)
    evaluate_sdf()        :163-222
    sample_depth()        :225-277      (bg=0) + autograd
    PatchMatch.__call__   utils/loss_utils.py:126-268  (sample_depth + warp_patch_ncc: the multi-view losses)
    GaussianModel.get_tetra_points + the opacity test of mesh_extract_tetrahedra.py:65-87 through parallel.alpha_cull

The returned dictionaries must agree within the usual bars (integers exact, images 1e-5, gradients 1e-4).
"""
import math
import os
import types

import numpy as np
import pytest
import torch

import helpers as H
import scenes
from oracle import ref_loader

pytestmark = pytest.mark.gpu

SHIMS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims")
PIPE = types.SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)


def _need_app(ref_module):
    if ref_module is None or not ref_loader.app_available() or not ref_loader.ncc_available():
        pytest.skip("oracle/_ref (compiled reference + installed callers) not shipped")


def _inverse_softplus(y):
    return torch.log(torch.expm1(y.clamp_min(1e-6)))


def _model(app, sc, device):
    """The reference's GaussianModel holding the scene's Gaussians as its raw parameters."""
    gm = app.gaussian_model.GaussianModel(3, sc.sg_axis.shape[1])
    gm.active_sh_degree, gm.active_sg_degree = sc.sh_degree, sc.sg_degree
    par = lambda t: torch.nn.Parameter(t.to(device).contiguous())  # noqa: E731
    gm._xyz = par(sc.means3D)
    gm._features_dc = par(sc.shs[:, :1])
    gm._features_rest = par(sc.shs[:, 1:])
    gm._scaling = par(torch.log(sc.scales))
    gm._rotation = par(sc.rotations * 1.7)  # not unit length: get_rotation normalises
    gm._opacity = par(torch.logit(sc.opacities.clamp(1e-4, 1 - 1e-4)))
    gm._sg_axis = par(sc.sg_axis * 2.0)
    gm._sg_sharpness = par(_inverse_softplus(sc.sg_sharpness))
    gm._sg_color = par(sc.sg_color)
    g = torch.Generator().manual_seed(11)
    gm.filter_3D = (0.004 * torch.rand(sc.P, 1, generator=g)).to(device)
    return gm


def _camera(app, cam, seed, uid):
    """scene/cameras.py Camera from a scenes.Camera: R is stored transposed (camera-to-world), T = world-to-view shift."""
    Rt = cam.viewmatrix.t().double().numpy()
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(cam.H, dtype=torch.float32), torch.arange(cam.W, dtype=torch.float32),
                            indexing="ij")
    tex = 0.5 + 0.25 * torch.sin(0.21 * xx + 0.13 * yy) + 0.2 * torch.sin(0.07 * xx - 0.17 * yy)
    image = (tex[None].repeat(3, 1, 1) + 0.05 * torch.rand(3, cam.H, cam.W, generator=g)).clamp(0, 1)
    return app.cameras.Camera(colmap_id=uid, R=Rt[:3, :3].T.copy(), T=Rt[:3, 3].copy(),
                              FoVx=2 * math.atan(cam.tanfovx), FoVy=2 * math.atan(cam.tanfovy), image=image,
                              gt_alpha_mask=None, gt_mask=None, image_name=f"v{uid}", uid=uid)


def _params(gm):
    return dict(xyz=gm._xyz, f_dc=gm._features_dc, f_rest=gm._features_rest, scaling=gm._scaling, rotation=gm._rotation,
                opacity=gm._opacity, sg_axis=gm._sg_axis, sg_sharpness=gm._sg_sharpness, sg_color=gm._sg_color)


def _grads(gm):
    out = {}
    for k, p in _params(gm).items():
        out[k] = torch.zeros_like(p) if p.grad is None else p.grad.detach().clone()
        p.grad = None
    return out


def _run_everything(app, sc, device):
    """One pass through every rasterizer caller of the reference; returns plain tensors."""
    torch.manual_seed(0)
    gm = _model(app, sc, device)
    cam_a = _camera(app, scenes.config_camera(sc.name, 0, 24), 1, 0)
    cam_b = _camera(app, scenes.config_camera(sc.name, 1, 24), 2, 1)
    R = app.gaussian_renderer
    bg = torch.tensor([0.1, 0.2, 0.3], device=device)
    res = {}

    pkg = R.render(cam_a, gm, PIPE, bg, kernel_size=0.1)
    up = {k: v.to(device) for k, v in sc.grads.items()}
    loss = (pkg["render"] * up["color"]).sum() + (pkg["mask"] * up["alpha"]).sum() + \
        (pkg["median_depth"] * up["mdepth"]).sum() + (pkg["normal"] * up["normal"]).sum()
    loss.backward()
    for k in ("render", "mask", "median_depth", "normal", "radii", "visibility_filter"):
        res["render." + k] = pkg[k].detach().clone()
    res["render.viewspace_grad"] = pkg["viewspace_points"].grad.detach().clone()
    for k, v in _grads(gm).items():
        res["render.grad." + k] = v

    with torch.no_grad():
        pts, _scale = gm.get_tetra_points()
        pts = pts[torch.randperm(pts.shape[0], generator=torch.Generator().manual_seed(3))[:300_000].to(device)].contiguous()
        it = R.integrate(pts, cam_a, gm, PIPE, 0.1)
        res["integrate.alpha"], res["integrate.inside"] = it["alpha_integrated"], it["inside"]
        sd = R.evaluate_sdf(pts[:100_000].contiguous(), cam_a, gm, PIPE, 0.1)
        res["sdf.depth"], res["sdf.sdf"], res["sdf.inside"] = sd["depth"], sd["sdf"], sd["inside"]
    res["tetra_points"] = pts

    # the multi-view losses: sample_depth of view A's pixels in view B, then the patch NCC
    pkg = R.render(cam_a, gm, PIPE, bg, kernel_size=0.1)
    pm = app.loss_utils.PatchMatch(3, 1.0, 0.1, PIPE, debug=False)
    ncc_loss, geo_loss = pm(gm, pkg, cam_a, cam_b)
    (ncc_loss + geo_loss).sum().backward()
    res["patchmatch.ncc_loss"], res["patchmatch.geo_loss"] = ncc_loss.detach().reshape(-1), geo_loss.detach().reshape(-1)
    for k, v in _grads(gm).items():
        res["patchmatch.grad." + k] = v

    # sample_depth on its own, N-d query tensor
    q = (torch.randn(64, 50, 3, generator=torch.Generator().manual_seed(5)) * 0.6).to(device).requires_grad_(True)
    s = R.sample_depth(q, cam_b, gm, PIPE, 0.1)
    (s["sampled_depth"] * torch.randn(64, 50, 3, generator=torch.Generator().manual_seed(6)).to(device)).sum().backward()
    res["sample.points"], res["sample.inside"] = s["sampled_depth"].detach().clone(), s["inside"].clone()
    res["sample.grad_points"] = q.grad.detach().clone()
    for k, v in _grads(gm).items():
        res["sample.grad." + k] = v
    torch.cuda.synchronize()
    return res, gm, (cam_a, cam_b)


def _compare(ours, ref, grad_tol=H.GRAD_TOL):
    bad = []
    for k, b in ref.items():
        a = ours[k]
        assert a.shape == b.shape and a.dtype == b.dtype, (k, a.shape, b.shape, a.dtype, b.dtype)
        if b.dtype in (torch.bool, torch.int32, torch.int64):
            if H.int_mismatch(a, b):
                bad.append((k, "integer / mask mismatch", H.int_mismatch(a, b)))
        elif ".grad" in k or k.endswith("_grad") or k.endswith("loss"):
            e = H.grad_rel_err(a, b)
            if e > grad_tol:
                bad.append((k, "grad rel err", e))
        else:
            e, n = H.image_rel_err(a, b)
            if e > H.IMG_TOL:
                bad.append((k, "rel err", e, n))
    return bad


def test_reference_callers_unchanged_on_the_drop_in(device, ref_module):
    _need_app(ref_module)
    import diff_gaussian_rasterization as ours
    import fused_ssim as ours_ssim
    import warp_patch_ncc as ours_ncc

    sc = scenes.make_scene("c2", P=60_000, seed=21, sg_degree=2)
    with ref_loader.ReferenceApp(SHIMS, diff_gaussian_rasterization=ours, warp_patch_ncc=ours_ncc,
                                 fused_ssim=ours_ssim) as app:
        assert app.gaussian_renderer.GaussianRasterizer is ours.GaussianRasterizer
        got, _, _ = _run_everything(app, sc, device)
    with ref_loader.ReferenceApp(SHIMS, diff_gaussian_rasterization=ref_module, warp_patch_ncc=ref_loader.load_ncc(),
                                 fused_ssim=ours_ssim) as app:
        assert app.gaussian_renderer.GaussianRasterizer is ref_module.GaussianRasterizer
        want, _, _ = _run_everything(app, sc, device)
    assert int(want["integrate.inside"].sum()) > 10_000 and int(want["sample.inside"].sum()) > 100
    assert float(want["patchmatch.ncc_loss"].abs().sum() + want["patchmatch.geo_loss"].abs().sum()) > 0
    # the patch-NCC kernels of the two sides round differently on low-contrast patches (tests/test_patch_ncc_gpu.py):
    # the losses and what flows back through them are compared at 2e-3, everything else at the usual bars
    ncc_keys = [k for k in want if k.startswith("patchmatch.") and k != "patchmatch.geo_loss"]
    bad = _compare({k: v for k, v in got.items() if k not in ncc_keys}, {k: v for k, v in want.items() if k not in ncc_keys})
    bad += _compare({k: got[k] for k in ncc_keys}, {k: want[k] for k in ncc_keys}, grad_tol=2e-3)
    assert not bad, bad


def test_reference_tetra_points_through_alpha_cull(device, ref_module):
    """GaussianModel.get_tetra_points (the reference's own) -> the opacity test over several views: parallel.alpha_cull
    (prepared views, view-major) against the reference's chunk-major loop on the reference rasterizer."""
    _need_app(ref_module)
    import diff_gaussian_rasterization as ours
    import fused_ssim as ours_ssim
    import warp_patch_ncc as ours_ncc
    from diff_gaussian_rasterization import parallel

    sc = scenes.make_scene("c2", P=40_000, seed=23)
    V = 4

    def settings(mod, gm, cam):
        return mod.GaussianRasterizationSettings(
            image_height=int(cam.image_height), image_width=int(cam.image_width), tanfovx=math.tan(cam.FoVx * 0.5),
            tanfovy=math.tan(cam.FoVy * 0.5), kernel_size=0.1, bg=None, scale_modifier=1.0,
            viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, sh_degree=gm.active_sh_degree,
            sg_degree=gm.active_sg_degree, campos=cam.camera_center, prefiltered=False, debug=False, require_depth=True)

    with ref_loader.ReferenceApp(SHIMS, diff_gaussian_rasterization=ours, warp_patch_ncc=ours_ncc,
                                 fused_ssim=ours_ssim) as app:
        gm = _model(app, sc, device)
        cams = [_camera(app, scenes.config_camera("c2", v, V), v, v) for v in range(V)]
        with torch.no_grad():
            pts, _ = gm.get_tetra_points()
            x = (gm.get_xyz, gm.get_opacity_with_3D_filter, gm.get_scaling_with_3D_filter, gm.get_rotation)
            sdf, valid = parallel.alpha_cull(pts, [settings(ours, gm, c) for c in cams], *x, chunk_size=250_000)
            # the reference's loop (mesh_extract_tetrahedra.py:65-87) through the reference's integrate()
            w = torch.ones(pts.shape[0], device=device)
            anyv = torch.zeros(pts.shape[0], dtype=torch.bool, device=device)
            for c in cams:
                r = ref_module.GaussianRasterizer(settings(ref_module, gm, c)).integrate(pts, *x, None, None)
                anyv |= r[1]
                w = torch.where(r[1], torch.min(r[0], w), w)
            w[~anyv] = 0
    assert pts.shape[0] == 15 * sc.P and int(anyv.sum()) > pts.shape[0] // 2
    assert torch.equal(valid, anyv)
    assert float((sdf - (0.5 - w)).abs().max()) <= 1e-5
