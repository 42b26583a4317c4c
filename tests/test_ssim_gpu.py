"""Fused SSIM on the GPU: `fused_ssim.fused_ssim` against the reference's torch statement of SSIM
(utils/loss_utils.py:34-73) evaluated in float64 on the same device.  Value within 1e-5 relative, gradient within
1e-4 of its largest element (measured under emulation on the CPU: 5e-8 and 1e-6)."""
import pytest
import torch

import helpers as H  # noqa: F401  (sys.path)
import ssim_reference as R

pytestmark = [pytest.mark.gpu, pytest.mark.first_hardware_run(order=1)]

CASES = [
    (1, 3, 822, 1237, "valid"),  # what train.py:189 passes at the headline resolution
    (1, 3, 545, 980, "same"),
    (2, 3, 97, 131, "valid"),
    (1, 1, 33, 32, "same"),
    (1, 3, 11, 11, "valid"),     # a 1 x 1 map
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
def test_fused_ssim_vs_torch_float64(case, device):
    import fused_ssim as ours

    B, Ch, Hh, W, padding = case
    img1, img2 = R.image_pair((B, Ch, Hh, W), seed=Hh)
    img1, img2 = img1.to(device), img2.to(device)
    want, want_grad = R.value_and_grad(img1, img2, padding, upstream=-0.2)
    x = img1.clone().requires_grad_(True)
    got = ours.fused_ssim(x, img2, padding=padding)
    assert got.shape == () and got.dtype == torch.float32
    (got * -0.2).backward()  # the trainer's lambda_dssim * (1 - ssim)
    assert abs(float(got) - float(want)) <= 1e-5 * abs(float(want)), (float(got), float(want))
    err = float((x.grad.double() - want_grad).abs().max() / want_grad.abs().max())
    assert err <= 1e-4, err
    assert x.grad.shape == img1.shape and bool(torch.isfinite(x.grad).all())


def test_fused_ssim_surface(device):
    import fused_ssim as ours

    img1, img2 = R.image_pair((1, 3, 120, 160), seed=2)
    img1, img2 = img1.to(device), img2.to(device)
    a = ours.fused_ssim(img1, img2, padding="valid")
    b = ours.fused_ssim(img1, img2, padding="valid", train=False)
    assert torch.equal(a, b)  # same kernel, the derivative maps are just not written
    assert torch.equal(a, ours.fused_ssim(img1, img2, padding="valid"))  # no atomics: bit-reproducible
    assert float(ours.fused_ssim(img1, img1.clone(), padding="same")) == pytest.approx(1.0, abs=1e-6)
    # default padding is "same" as in the package; a non-contiguous view is accepted
    full = torch.rand(1, 3, 120, 200, device=device)
    view = full[..., 20:180]
    assert not view.is_contiguous()
    assert torch.equal(ours.fused_ssim(view, img2), ours.fused_ssim(view.contiguous(), img2, padding="same"))
    # gradient only flows to img1, and scales with the incoming gradient
    x = img1.clone().requires_grad_(True)
    y = img2.clone().requires_grad_(True)
    ours.fused_ssim(x, y, padding="valid").backward()
    g1 = x.grad.clone()
    assert y.grad is None
    x.grad = None
    (3.0 * ours.fused_ssim(x, y, padding="valid")).backward()
    assert float((x.grad - 3.0 * g1).abs().max()) <= 1e-5 * float((3.0 * g1).abs().max())
    with pytest.raises(RuntimeError, match="no derivative maps"):
        x.grad = None
        ours.fused_ssim(x, y, padding="valid", train=False).backward()
    with pytest.raises(RuntimeError):
        ours.fused_ssim(img1[0], img2[0])  # [C,H,W]: the trainer unsqueezes (train.py:189)
    with pytest.raises(RuntimeError):
        ours.fused_ssim(img1[..., :10, :], img2[..., :10, :], padding="valid")  # no map left
