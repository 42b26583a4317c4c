"""The reference's own torch statement of SSIM (utils/loss_utils.py:34-73: 11-tap Gaussian window, sigma 1.5,
C1 = 0.01^2, C2 = 0.03^2) with the padding modes of the `fused_ssim` package it calls (:48-49): "same" is the
`_ssim` of :52-73 (padding = 5), "valid" is the same formula without padding.  Test infrastructure: the checker for
the fused kernels, in float64 unless told otherwise."""
import math

import torch
import torch.nn.functional as F


def gaussian(window_size=11, sigma=1.5):
    g = torch.Tensor([math.exp(-((x - window_size // 2) ** 2) / float(2 * sigma ** 2)) for x in range(window_size)])
    return g / g.sum()


def ssim_mean(img1, img2, padding="same", dtype=torch.float64):
    """Mean of the SSIM map of [B,C,H,W] images; differentiable in img1."""
    ch = img1.shape[1]
    w1 = gaussian().to(device=img1.device, dtype=dtype).unsqueeze(1)
    window = (w1 @ w1.t()).unsqueeze(0).unsqueeze(0).expand(ch, 1, 11, 11).contiguous()
    a, b = img1.to(dtype), img2.to(dtype)
    pad = 5 if padding == "same" else 0

    def conv(t):
        return F.conv2d(t, window, padding=pad, groups=ch)

    mu1, mu2 = conv(a), conv(b)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = conv(a * a) - mu1_sq
    sigma2_sq = conv(b * b) - mu2_sq
    sigma12 = conv(a * b) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    return ssim_map.mean()


def value_and_grad(img1, img2, padding, upstream=1.0, dtype=torch.float64):
    a = img1.detach().clone().to(dtype).requires_grad_(True)
    val = ssim_mean(a, img2.detach(), padding, dtype)
    (val * upstream).backward()
    return val.detach(), a.grad


def image_pair(shape, seed, noise=0.1):
    g = torch.Generator().manual_seed(seed)
    img1 = torch.rand(*shape, generator=g)
    img2 = (img1 + noise * torch.randn(*shape, generator=g)).clamp(0, 1).contiguous()
    return img1, img2
