"""oracle/activations.py (what csrc/ggs_activations.cu computes, restated in numpy) against torch autograd on the
reference's own expressions (scene/gaussian_model.py:158-160, 203-212), on the CPU in float64."""
import numpy as np
import pytest
import torch

import helpers as H  # noqa: F401  (sys.path)
from oracle import activations as A


def _reference(raw_scaling, raw_opacity, raw_rotation, filter_3D):
    opacity = torch.sigmoid(raw_opacity)
    scales = torch.exp(raw_scaling)
    scales_square = torch.square(scales)
    det1 = scales_square.prod(dim=1)
    scales_after_square = scales_square + torch.square(filter_3D)
    det2 = scales_after_square.prod(dim=1)
    coef = det1.sqrt() * det2.rsqrt()
    return scales_after_square.sqrt(), opacity * coef[..., None], torch.nn.functional.normalize(raw_rotation)


@pytest.mark.parametrize("filtered", [False, True])
def test_activation_oracle_matches_torch_autograd(filtered):
    rng = np.random.default_rng(11 + int(filtered))
    P = 257
    rs = rng.normal(-4.4, 0.6, (P, 3))
    ro = rng.normal(0, 2, (P, 1))
    rq = rng.normal(0, 1, (P, 4))
    f = rng.uniform(0, 0.02, (P, 1)) if filtered else np.zeros((P, 1))
    ups = [rng.normal(0, 1, (P, w)) for w in (3, 1, 4)]
    t = [torch.tensor(x, dtype=torch.float64, requires_grad=True) for x in (rs, ro, rq)]
    out = _reference(t[0], t[1], t[2], torch.tensor(f))
    torch.autograd.backward(list(out), [torch.tensor(u) for u in ups])
    fwd = A.forward(rs, ro, rq, f)
    for a, b in zip(fwd, out):
        np.testing.assert_allclose(a, b.detach().numpy(), rtol=1e-12, atol=0)
    bwd = A.backward(rs, ro, rq, f, *ups)
    for a, b in zip(bwd, t):
        np.testing.assert_allclose(a, b.grad.numpy(), rtol=1e-9, atol=1e-15)
