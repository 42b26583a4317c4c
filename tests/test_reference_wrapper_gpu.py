"""INTEGRATION.md option 2, executed: the reference's UNMODIFIED Python wrapper
(DGR/diff_gaussian_rasterization/__init__.py, as installed next to the compiled reference by oracle/build_ref.py) running
on top of this repo's ctypes `_C` module (diff_gaussian_rasterization/_C.py) instead of its compiled extension, compared
with the same wrapper on its own extension: forward + backward, integrate, evaluate_sdf, sample_depth + backward,
markVisible."""
import importlib
import os
import shutil
import sys

import pytest
import torch

import helpers as H
import scenes
from oracle import ref_loader

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def wrapper_on_shim(tmp_path_factory):
    if not ref_loader.available():
        pytest.skip("oracle/_ref not shipped")
    import diff_gaussian_rasterization._C as shim

    root = tmp_path_factory.mktemp("wrapped")
    pkg = root / "ref_wrapper_on_ggs"
    pkg.mkdir()
    # the reference's wrapper, byte for byte (an install artefact of oracle/build_ref.py), in a package of its own ...
    shutil.copyfile(os.path.join(os.path.dirname(ref_loader.__file__), "_ref", "ref_dgr", "__init__.py"),
                    pkg / "__init__.py")
    sys.path.insert(0, str(root))
    # ... whose `from . import _C` finds this repo's shim
    sys.modules["ref_wrapper_on_ggs._C"] = shim
    try:
        mod = importlib.import_module("ref_wrapper_on_ggs")
        assert mod._C is shim
        yield mod
    finally:
        sys.path.remove(str(root))
        sys.modules.pop("ref_wrapper_on_ggs", None)
        sys.modules.pop("ref_wrapper_on_ggs._C", None)


@pytest.mark.parametrize("tag", ["tiny_sg", "c1", "c1_k01_bg"])
def test_forward_backward_through_the_reference_wrapper(tag, device, ref_module, wrapper_on_shim):
    case = dict(tiny_sg=dict(name="tiny_sg"), c1=dict(name="c1"),
                c1_k01_bg=dict(name="c1", kernel_size=0.1, random_bg=True, seed=17))[tag]
    sc = scenes.make_scene(**case)
    for require_depth in (True, False):
        o_out, o_g = H.run_module(wrapper_on_shim, sc, device, require_depth=require_depth)
        r_out, r_g = H.run_module(ref_module, sc, device, require_depth=require_depth)
        assert H.int_mismatch(o_out["radii"], r_out["radii"]) == 0
        for k in ("color", "alpha", "mdepth", "normal"):
            err, bad = H.image_rel_err(o_out[k], r_out[k])
            assert err <= H.IMG_TOL, (k, err, bad)
        for k in H.GRAD_NAMES:
            assert o_g[k].shape == r_g[k].shape, k
            err = H.grad_rel_err(o_g[k], r_g[k])
            assert err <= H.GRAD_TOL, (k, err)


def test_point_queries_through_the_reference_wrapper(device, ref_module, wrapper_on_shim):
    sc = scenes.make_scene("c1", scale_mult=2.0)
    g = torch.Generator().manual_seed(4)
    pts = (torch.randn(60_000, 3, generator=g) * 0.7).to(device)
    up = torch.randn(48, 25, 3, generator=g).to(device)
    res = {}
    for name, mod in (("shim", wrapper_on_shim), ("ref", ref_module)):
        x = {k: getattr(sc, k).to(device).clone().requires_grad_(True)
             for k in ("means3D", "opacities", "scales", "rotations")}
        rs = H.settings_for(mod, sc, device)
        rast = mod.GaussianRasterizer(rs._replace(bg=None, kernel_size=0.1))
        with torch.no_grad():
            a, inside = rast.integrate(pts, x["means3D"], x["opacities"], x["scales"], x["rotations"], None, None)
            d, s, ins2 = rast.evaluate_sdf(pts, x["means3D"], x["opacities"], x["scales"], x["rotations"], None, None)
            vis = rast.markVisible(x["means3D"])
        q = pts[:1200].reshape(48, 25, 3).clone().requires_grad_(True)
        cam, ins3 = rast.sample_depth(q, x["means3D"], x["opacities"], x["scales"], x["rotations"])
        (cam * up).sum().backward()
        torch.cuda.synchronize()
        res[name] = dict(alpha=a, inside=inside, depth=d, sdf=s, inside_sdf=ins2, visible=vis, cam=cam.detach(),
                         inside_sample=ins3, g_points=q.grad.detach().clone(),
                         **{"g_" + k: v.grad.detach().clone() for k, v in x.items()})
    a, b = res["shim"], res["ref"]
    for k in ("inside", "inside_sdf", "inside_sample", "visible"):
        assert a[k].shape == b[k].shape and a[k].dtype == b[k].dtype and H.int_mismatch(a[k], b[k]) == 0, k
    assert int(b["inside"].sum()) > 1000 and int(b["inside_sample"].sum()) > 50
    for k in ("alpha", "depth", "sdf", "cam"):
        assert a[k].shape == b[k].shape, k
        err, bad = H.image_rel_err(a[k], b[k])
        assert err <= H.IMG_TOL, (k, err, bad)
    for k in ("g_points", "g_means3D", "g_opacities", "g_scales", "g_rotations"):
        assert a[k].shape == b[k].shape, k
        assert H.grad_rel_err(a[k], b[k]) <= H.GRAD_TOL, k
