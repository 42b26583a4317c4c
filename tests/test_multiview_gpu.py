"""View-parallel training: the SH gradient summed over views from each view's 3-float colour gradient
(ggs_sh_gradient_from_views, parallel.factored_sh_gradients) equals the sum of the per-view SH gradients."""
import ctypes as C

import pytest
import torch

import helpers as H
import scenes

pytestmark = pytest.mark.gpu


def test_factored_sh_gradient_single_view(device):
    """One view: the factored path (colour gradient -> ggs_sh_gradient_from_views) against the plain backward.  Every
    other gradient is the same kernel's output and must be the same bits; the SH gradient comes from two different
    kernels built with --use_fast_math (the compiler may associate basis * dL/dRGB differently in each), so it is held
    to a few ulps of the tensor's largest element instead."""
    import diff_gaussian_rasterization as ours
    from diff_gaussian_rasterization import parallel

    sc = scenes.make_scene("c1")
    _, plain = H.run_module(ours, sc, device)
    with parallel.factored_sh_gradients():
        _, fact = H.run_module(ours, sc, device)
    for k in H.GRAD_NAMES:
        if k == "shs":
            err = float((plain[k] - fact[k]).abs().max() / plain[k].abs().max())
            assert err <= 5e-7, err
        else:
            assert torch.equal(plain[k], fact[k]), k


@pytest.mark.parametrize("sh_degree", [0, 1, 2, 3])
def test_sh_gradient_from_two_views(device, sh_degree):
    from diff_gaussian_rasterization import _abi

    views = [scenes.make_scene("c1", view=v, n_views=2, sh_degree=sh_degree) for v in range(2)]
    per_view = [H.ours_backward_raw(sc, device)[0] for sc in views]
    want = per_view[0]["shs"].double() + per_view[1]["shs"].double()
    P, SHM = views[0].P, views[0].shs.shape[1]
    rgb = torch.stack([g["rgb"] for g in per_view]).contiguous()
    cam = torch.stack([sc.cam.campos.to(device).reshape(3) for sc in views]).contiguous()
    means = views[0].means3D.to(device).contiguous()
    out = torch.full((P, SHM, 3), float("nan"), device=device)
    rc = _abi.lib.ggs_sh_gradient_from_views(P, sh_degree, SHM, 2, _abi.ptr(means), _abi.ptr(cam), _abi.ptr(rgb),
                                             _abi.ptr(out), _abi.stream_handle(torch.device(device)))
    _abi.check(rc, "ggs_sh_gradient_from_views")
    torch.cuda.synchronize()
    assert not torch.isnan(out).any()
    assert float(want.abs().max()) > 0
    err = float((out.double() - want).abs().max()) / float(want.abs().max())
    assert err <= 1e-6, err
    # rows beyond the active degree stay zero
    assert float(out[:, (sh_degree + 1) ** 2:].abs().max() if SHM > (sh_degree + 1) ** 2 else 0.0) == 0.0


def test_sh_gradient_from_views_argument_errors(device):
    from diff_gaussian_rasterization import _abi

    z = torch.zeros(4, 3, device=device)
    assert _abi.lib.ggs_sh_gradient_from_views(4, 3, 9, 1, _abi.ptr(z), _abi.ptr(z), _abi.ptr(z), _abi.ptr(z), None) < 0
    assert _abi.lib.ggs_sh_gradient_from_views(4, 1, 4, 0, _abi.ptr(z), _abi.ptr(z), _abi.ptr(z), _abi.ptr(z), None) < 0
    assert _abi.lib.ggs_sh_gradient_from_views(0, 1, 4, 1, None, None, None, None, None) == 0
