"""Edge cases of the drop-in surface on the GPU: empty / fully culled inputs, require_depth=False,
precomputed colours, ragged image sizes, markVisible, error behaviour."""
import pytest
import torch

import helpers as H
import scenes

pytestmark = pytest.mark.gpu


def _rast(sc, device, **kw):
    import diff_gaussian_rasterization as ours

    return ours, ours.GaussianRasterizer(H.settings_for(ours, sc, device, **kw))


def test_empty_scene(device):
    sc = scenes.make_scene("tiny", P=0)
    sc.bg = torch.tensor([0.1, 0.2, 0.3])
    ours, r = _rast(sc, device)
    m = torch.zeros(0, 3, device=device, requires_grad=True)
    color, radii, mdepth, alpha, normal = r(m, torch.zeros(0, 3, device=device), torch.zeros(0, 1, device=device),
                                            shs=torch.zeros(0, 16, 3, device=device),
                                            scales=torch.zeros(0, 3, device=device),
                                            rotations=torch.zeros(0, 4, device=device))
    assert radii.numel() == 0 and color.shape == (3, sc.cam.H, sc.cam.W)
    # rasterize_points.cu:95: with P == 0 nothing runs and the zero-initialised outputs come back
    assert float(color.abs().max()) == 0 and float(alpha.abs().max()) == 0 and float(mdepth.abs().max()) == 0


def test_everything_culled(device):
    sc = scenes.make_scene("tiny", P=500)
    sc.means3D[:, 2] = -10.0
    sc.bg = torch.tensor([0.25, 0.5, 0.75])
    import diff_gaussian_rasterization as ours

    out, g = H.run_module(ours, sc, device)
    assert int(out["radii"].abs().sum()) == 0
    assert torch.allclose(out["color"][2], torch.full_like(out["color"][2], 0.75))
    assert float(out["alpha"].abs().max()) == 0 and float(out["normal"].abs().max()) == 0
    for k in ("means3D", "opacities", "scales", "rotations", "shs", "means2D"):
        assert float(g[k].abs().max()) == 0, k


def test_require_depth_false_leaves_geometry_zero(device):
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene("c1")
    a, _ = H.run_module(ours, sc, device, require_depth=False)
    b, _ = H.run_module(ours, sc, device, require_depth=True)
    assert float(a["mdepth"].abs().max()) == 0 and float(a["normal"].abs().max()) == 0
    assert torch.equal(a["color"], b["color"]) and torch.equal(a["alpha"], b["alpha"])
    assert torch.equal(a["radii"], b["radii"])


def test_colors_precomp_path(device, ref_module):
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene("tiny")
    cols = torch.rand(sc.P, 3, generator=torch.Generator().manual_seed(3))

    def run(mod):
        rs = H.settings_for(mod, sc, device)
        x = H.leaf_inputs(sc, device)
        cp = cols.to(device).requires_grad_(True)
        r = mod.GaussianRasterizer(rs)
        color, radii, mdepth, alpha, normal = r(
            means3D=x["means3D"], means2D=x["means2D"], opacities=x["opacities"], shs=None, sg_axis=x["sg_axis"],
            sg_sharpness=x["sg_sharpness"], sg_color=x["sg_color"], colors_precomp=cp, scales=x["scales"],
            rotations=x["rotations"], cov3D_precomp=None)
        g = {k: v.to(device) for k, v in sc.grads.items()}
        ((color * g["color"]).sum() + (alpha * g["alpha"]).sum() + (mdepth * g["mdepth"]).sum()).backward()
        return color.detach(), cp.grad.detach(), x["means3D"].grad.detach()

    c, gc, gm = run(ours)
    assert float(gc.abs().max()) > 0
    if ref_module is not None:
        c2, gc2, gm2 = run(ref_module)
        assert H.image_rel_err(c, c2)[0] <= H.IMG_TOL
        assert H.grad_rel_err(gc, gc2) <= H.GRAD_TOL and H.grad_rel_err(gm, gm2) <= H.GRAD_TOL


@pytest.mark.parametrize("wh", [(1, 1), (17, 5), (16, 16), (33, 47)])
def test_ragged_image_sizes(wh, device, ref_module):
    import diff_gaussian_rasterization as ours

    W, Hh = wh
    sc = scenes.make_scene("tiny", W=W, H=Hh, scale_mult=6.0, seed=W * 100 + Hh)
    o, g = H.run_module(ours, sc, device)
    assert o["color"].shape == (3, Hh, W)
    if ref_module is not None:
        r, rg = H.run_module(ref_module, sc, device)
        assert H.int_mismatch(o["radii"], r["radii"]) == 0
        for k in ("color", "alpha", "mdepth", "normal"):
            assert H.image_rel_err(o[k], r[k])[0] <= H.IMG_TOL, k
        for k in H.GRAD_NAMES:
            assert H.grad_rel_err(g[k], rg[k]) <= H.GRAD_TOL, k


def test_mark_visible(device, ref_module):
    sc = scenes.make_scene("c1")
    ours, r = _rast(sc, device)
    m = r.markVisible(sc.means3D.to(device))
    view = sc.cam.viewmatrix
    z = sc.means3D @ view[:3, 2] + view[3, 2]
    assert m.dtype == torch.bool
    # knife-edge points aside, identical to the z > 0.2 test
    exp = z > 0.2
    sure = (z - 0.2).abs() > 1e-5
    assert torch.equal(m.cpu()[sure], exp[sure])
    if ref_module is not None:
        rr = ref_module.GaussianRasterizer(H.settings_for(ref_module, sc, device))
        assert torch.equal(rr.markVisible(sc.means3D.to(device)), m)


def test_unsupported_paths_are_loud(device):
    sc = scenes.make_scene("tiny")
    ours, r = _rast(sc, device)
    x = H.leaf_inputs(sc, device)
    with pytest.raises(NotImplementedError):
        r(x["means3D"], x["means2D"], x["opacities"], shs=x["shs"], cov3D_precomp=torch.zeros(sc.P, 6, device=device))
    with pytest.raises(RuntimeError, match="sg_degree"):
        import diff_gaussian_rasterization as o2

        rs = H.settings_for(o2, sc, device)._replace(sg_degree=3)
        o2.GaussianRasterizer(rs)(x["means3D"], x["means2D"], x["opacities"], shs=x["shs"], sg_axis=x["sg_axis"],
                                  sg_sharpness=x["sg_sharpness"], sg_color=x["sg_color"], scales=x["scales"],
                                  rotations=x["rotations"])


def test_non_default_stream_and_noncontiguous_grads(device):
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene("tiny")
    base, gbase = H.run_module(ours, sc, device)
    s = torch.cuda.Stream(device=device)
    with torch.cuda.stream(s):
        out, g = H.run_module(ours, sc, device)
    s.synchronize()
    assert torch.equal(base["color"], out["color"])
    assert H.grad_rel_err(g["means3D"], gbase["means3D"]) < 1e-5


def test_point_queries_accept_the_reference_callers_bg(device):
    """gaussian_renderer/__init__.py passes bg=None to integrate (:117) and evaluate_sdf (:178) and bg=0 to sample_depth
    (:238); none of those paths reads it."""
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene("tiny", scale_mult=3.0)
    g = torch.Generator().manual_seed(3)
    pts = (torch.randn(4000, 3, generator=g) * 0.6).to(device)
    x = [getattr(sc, k).to(device) for k in ("means3D", "opacities", "scales", "rotations")]
    base = H.settings_for(ours, sc, device)
    want = ours.GaussianRasterizer(base)
    none = ours.GaussianRasterizer(base._replace(bg=None))
    zero = ours.GaussianRasterizer(base._replace(bg=0))
    a0, i0 = want.integrate(pts, *x, None, None)
    a1, i1 = none.integrate(pts, *x, None, None)
    assert torch.equal(a0, a1) and torch.equal(i0, i1) and int(i0.sum()) > 0
    d0 = want.evaluate_sdf(pts, *x, None, None)
    d1 = none.evaluate_sdf(pts, *x, None, None)
    assert all(torch.equal(u, v) for u, v in zip(d0, d1))
    c0, j0 = want.sample_depth(pts, *x)
    c1, j1 = zero.sample_depth(pts, *x)
    assert torch.equal(c0, c1) and torch.equal(j0, j1)


@pytest.mark.parametrize("scene", [dict(name="tiny", scale_mult=4.0, seed=77), dict(name="c1")],
                         ids=["tiny_dense", "c1"])
def test_capacity_hint_does_not_change_results(device, scene):
    """GgsForwardState.capacity (include/ggs_rasterizer.h): with a hint the host no longer blocks on the pair count
    before it launches duplicate / sort / ranges, the kernels take the count from device memory and the workspace is
    laid out for the hint.  Generous hint, exact hint, hint too small (binned again at the exact size), hint 1: images,
    sorted keys / list / ranges are bit-identical to the blocking call."""
    sc = scenes.make_scene(**scene)
    ref_out, ref_it = H.ours_forward_raw(sc, device)
    D = ref_it["num_rendered"]
    assert ref_it["capacity"] == D and D > 0
    for hint in (2 * D + 12345, D + 1, D, D - 1, D // 2, 1):
        out, it = H.ours_forward_raw(sc, device, capacity_hint=hint)
        assert it["num_rendered"] == D
        assert it["capacity"] == (hint if hint >= D else D), (hint, it["capacity"])
        for k in ("color", "alpha", "mdepth", "normal", "radii"):
            assert torch.equal(out[k], ref_out[k]), (hint, k)
        for k in ("sorted_keys", "point_list", "ranges", "max_contributor", "n_contrib"):
            assert torch.equal(it[k], ref_it[k]), (hint, k)


def test_hinted_forward_backward_through_the_module(device):
    """The Python surface keeps a per-(device, P, H, W) hint: the first call blocks, the next ones do not; forward
    and backward results are the same bits either way (the backward takes its offsets from the capacity)."""
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene("c1")
    ours._CAPACITY_HINTS.clear()
    out0, g0 = H.run_module(ours, sc, device)
    assert len(ours._CAPACITY_HINTS) == 1 and next(iter(ours._CAPACITY_HINTS.values())) > 0
    out1, g1 = H.run_module(ours, sc, device)
    for k in out0:
        assert torch.equal(out0[k], out1[k]), k
    for k in g0:
        assert torch.equal(g0[k], g1[k]), k
    # a stale hint that is far too small
    key = next(iter(ours._CAPACITY_HINTS))
    ours._CAPACITY_HINTS[key] = 7
    out2, g2 = H.run_module(ours, sc, device)
    for k in out0:
        assert torch.equal(out0[k], out2[k]), k
    for k in g0:
        assert torch.equal(g0[k], g2[k]), k
