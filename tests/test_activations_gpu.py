"""Fused caller-side activations against the reference's torch expressions
(scene/gaussian_model.py:203-212 get_scaling_n_opacity_with_3D_filter, :158-160 get_rotation), forward and autograd."""
import pytest
import torch

import helpers as H  # noqa: F401  (sets up sys.path)
import diff_gaussian_rasterization as dgr

pytestmark = pytest.mark.gpu


def _reference(raw_scaling, raw_opacity, raw_rotation, filter_3D):
    opacity = torch.sigmoid(raw_opacity)
    scales = torch.exp(raw_scaling)
    scales_square = torch.square(scales)
    det1 = scales_square.prod(dim=1)
    scales_after_square = scales_square + torch.square(filter_3D)
    det2 = scales_after_square.prod(dim=1)
    coef = det1.sqrt() * det2.rsqrt()
    scales = scales_after_square.sqrt()
    return scales, opacity * coef[..., None], torch.nn.functional.normalize(raw_rotation)


@pytest.mark.parametrize("P", [1, 33, 10_000, 1_000_003])
@pytest.mark.parametrize("filtered", [False, True])
def test_activations_match_torch(P, filtered, device):
    g = torch.Generator().manual_seed(P + int(filtered))
    raw = dict(s=torch.randn(P, 3, generator=g) * 0.6 - 4.4, o=torch.randn(P, 1, generator=g) * 2,
               q=torch.randn(P, 4, generator=g))
    f = (torch.rand(P, 1, generator=g) * 0.02 if filtered else torch.zeros(P, 1)).to(device)
    ups = [torch.randn(P, w, generator=g).to(device) for w in (3, 1, 4)]
    a = {k: v.to(device).requires_grad_(True) for k, v in raw.items()}
    b = {k: v.to(device).requires_grad_(True) for k, v in raw.items()}
    ours = dgr.activate_gaussians(a["s"], a["o"], a["q"], f)
    ref = _reference(b["s"], b["o"], b["q"], f)
    for x, y, name in zip(ours, ref, ("scales", "opacities", "rotations")):
        assert x.shape == y.shape
        err = ((x - y).abs() / y.abs().clamp_min(1e-30)).max().item()
        assert err <= 1e-6, (name, err)  # a few ulps: IEEE kernels on both sides (opacity chains six roundings)
    torch.autograd.backward(list(ours), ups)
    torch.autograd.backward(list(ref), ups)
    for k in ("s", "o", "q"):
        ga, gb = a[k].grad, b[k].grad
        err = ((ga - gb).abs().max() / gb.abs().max().clamp_min(1e-30)).item()
        assert err <= 1e-5, (k, err)


def test_activations_feed_the_rasterizer(device):
    """End to end: raw parameters -> activate_gaussians -> rasterizer, gradients w.r.t. the raw parameters agree with
    the torch activation chain in front of the same rasterizer."""
    import scenes

    sc = scenes.make_scene("tiny")
    raw_s = torch.log(sc.scales).to(device)
    raw_o = torch.logit(sc.opacities.clamp(1e-4, 1 - 1e-4)).to(device)
    raw_q = (sc.rotations * 1.7).to(device)
    f = torch.full((sc.P, 1), 0.003, device=device)

    def run(act):
        ps = [t.clone().requires_grad_(True) for t in (raw_s, raw_o, raw_q)]
        scales, opac, rot = act(ps[0], ps[1], ps[2], f)
        out = H.rasterize_with(dgr, sc, device, scales=scales, opacities=opac, rotations=rot)
        loss = sum((o * g.to(device)).sum() for o, g in zip(out, (sc.grads["color"], sc.grads["mdepth"],
                                                                  sc.grads["alpha"], sc.grads["normal"])))
        loss.backward()
        return [p.grad for p in ps]

    ga = run(dgr.activate_gaussians)
    gb = run(_reference)
    for x, y in zip(ga, gb):
        assert ((x - y).abs().max() / y.abs().max().clamp_min(1e-30)).item() <= 1e-4
