"""CPU restatement (numpy, float64) of the caller-side activations the fused kernels of
csrc/ggs_activations.cu replace, and of their analytic backward.  TEST INFRASTRUCTURE ONLY: imported by tests/, never by
the product package.

Follows the reference's torch expressions:
  scene/gaussian_model.py:203-212  get_scaling_n_opacity_with_3D_filter
  scene/gaussian_model.py:158-160  get_rotation  (torch.nn.functional.normalize, eps = 1e-12)
  scene/gaussian_model.py:52,57,62 scaling_activation = exp, opacity_activation = sigmoid, rotation_activation = normalize
Pinned by tests/test_activations_cpu.py against torch autograd run on those expressions."""
import numpy as np


def forward(raw_scaling, raw_opacity, raw_rotation, filter_3D):
    s = np.exp(raw_scaling)
    u = s * s
    a = u + filter_3D * filter_3D
    det1 = u.prod(axis=1)
    det2 = a.prod(axis=1)
    coef = np.sqrt(det1) / np.sqrt(det2)
    o = 1.0 / (1.0 + np.exp(-raw_opacity))
    n = np.maximum(np.linalg.norm(raw_rotation, axis=1, keepdims=True), 1e-12)
    return np.sqrt(a), o * coef[:, None], raw_rotation / n


def backward(raw_scaling, raw_opacity, raw_rotation, filter_3D, g_scales, g_opacity, g_rotation):
    """Gradients w.r.t. (raw_scaling, raw_opacity, raw_rotation); filter_3D is a buffer (no gradient)."""
    s = np.exp(raw_scaling)
    u = s * s
    f2 = filter_3D * filter_3D
    a = u + f2
    coef = (np.sqrt(u.prod(axis=1)) / np.sqrt(a.prod(axis=1)))[:, None]
    o = 1.0 / (1.0 + np.exp(-raw_opacity))
    d_opacity = g_opacity * coef * o * (1.0 - o)
    # dS_k/dr_k = u_k / S_k ;  d ln(coef) / dr_k = 1 - u_k / a_k = f^2 / a_k
    d_scaling = g_scales * u / np.sqrt(a) + (g_opacity * o * coef) * f2 / a
    norm = np.linalg.norm(raw_rotation, axis=1, keepdims=True)
    n = np.maximum(norm, 1e-12)
    R = raw_rotation / n
    dot = np.where(norm > 1e-12, (R * g_rotation).sum(axis=1, keepdims=True), 0.0)
    d_rotation = (g_rotation - R * dot) / n
    return d_scaling, d_opacity, d_rotation
