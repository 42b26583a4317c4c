"""`GaussianRasterizer.integrate` (mesh-extraction opacity queries, reference __init__.py:338-402,
rasterizer_impl.cu:594-815, sample_forward.cu:10-169) against the compiled reference and the C oracle."""
import pytest
import torch

import helpers as H
import scenes

pytestmark = pytest.mark.gpu


def _points(sc, n, seed):
    g = torch.Generator().manual_seed(seed)
    idx = torch.randint(0, sc.P, (n,), generator=g)
    jitter = torch.randn(n, 3, generator=g) * 0.05
    pts = sc.means3D[idx] + jitter
    pts[: n // 10] = (torch.rand(n // 10, 3, generator=g) - 0.5) * 8.0  # some far outside / behind
    return pts.contiguous()


@pytest.mark.parametrize("name,n", [("tiny", 5_000), ("c1", 200_000)])
def test_integrate_vs_reference(name, n, device, ref_module):
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene(name, scale_mult=3.0)
    pts = _points(sc, n, 11).to(device)

    def run(mod):
        r = mod.GaussianRasterizer(H.settings_for(mod, sc, device))
        a, inside = r.integrate(pts, sc.means3D.to(device), sc.opacities.to(device), sc.scales.to(device),
                                sc.rotations.to(device), None, None)
        torch.cuda.synchronize()
        return a, inside

    a, inside = run(ours)
    assert a.shape == (n,) and inside.dtype == torch.bool
    assert float(a.min()) >= 0 and float(a.max()) <= 1
    assert bool((a[~inside] == 1).all()) or bool((a[~inside] == 0).all())
    if ref_module is not None:
        ra, rinside = run(ref_module)
        assert torch.equal(inside, rinside)
        assert float((a - ra).abs().max()) <= 1e-5
    from oracle.oracle import Oracle

    c = sc.cam
    o = Oracle().integrate(points3D=pts.cpu(), means3D=sc.means3D, opacities=sc.opacities, scales=sc.scales,
                           rotations=sc.rotations, viewmatrix=c.viewmatrix, projmatrix=c.projmatrix, campos=c.campos,
                           tanfovx=c.tanfovx, tanfovy=c.tanfovy, kernel_size=0.0, image_height=c.H, image_width=c.W)
    same = o["inside"] == inside.cpu()
    assert float(same.float().mean()) > 0.999
    assert float((o["alpha"][same] - a.cpu()[same]).abs().max()) < 5e-3


def test_integrate_empty(device):
    import diff_gaussian_rasterization as ours

    sc = scenes.make_scene("tiny")
    r = ours.GaussianRasterizer(H.settings_for(ours, sc, device))
    a, inside = r.integrate(torch.zeros(0, 3, device=device), sc.means3D.to(device), sc.opacities.to(device),
                            sc.scales.to(device), sc.rotations.to(device), None, None)
    assert a.numel() == 0 and inside.numel() == 0
