"""Seeded synthetic inputs for the patch-NCC kernel (the reference has no fixtures for it): two grey images of the same
analytic texture seen from two nearby pinhole cameras, pixels of the reference view with a depth and a unit normal.
Everything is generated on the CPU with a seeded generator so that every arm sees identical bits."""
import math

import torch


def _texture(x, y, g):
    """Smooth texture with enough local contrast for the variance test, plus a perfectly flat band."""
    v = (0.5 + 0.22 * torch.sin(0.31 * x + 0.17 * y) + 0.18 * torch.sin(0.11 * x - 0.23 * y + 1.0)
         + 0.08 * torch.sin(0.53 * x + 0.41 * y + 2.0))
    return v


def make_pair(height=120, width=160, n=4000, seed=7, neighbour_size=None, shift=(3.2, -1.7)):
    g = torch.Generator().manual_seed(seed)
    hn, wn = neighbour_size or (height, width)
    yr, xr = torch.meshgrid(torch.arange(height, dtype=torch.float32), torch.arange(width, dtype=torch.float32),
                            indexing="ij")
    yn, xn = torch.meshgrid(torch.arange(hn, dtype=torch.float32), torch.arange(wn, dtype=torch.float32), indexing="ij")
    image_r = _texture(xr, yr, g) + 0.02 * torch.randn(height, width, generator=g)
    image_n = _texture(xn - shift[0], yn - shift[1], g) + 0.02 * torch.randn(hn, wn, generator=g)
    # a flat band: patches inside it fail the variance test (exactly constant pixels)
    image_r[height // 2 - 6:height // 2 + 6, : width // 3] = 0.4
    image_r, image_n = image_r.clamp(0, 1).contiguous(), image_n.clamp(0, 1).contiguous()

    fx_r, fy_r, cx_r, cy_r = 1.9 * width, 1.9 * width, width / 2 - 0.5, height / 2 - 0.5
    fx_n, fy_n, cx_n, cy_n = 1.9 * wn + 3.0, 1.9 * wn - 2.0, wn / 2 + 0.25, hn / 2 - 0.75
    # small rotation about an oblique axis, small baseline
    axis = torch.tensor([0.3, -0.8, 0.52])
    axis = axis / axis.norm()
    ang = 0.006
    K = torch.tensor([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    R = torch.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * (K @ K)
    T = torch.tensor([0.035, -0.02, 0.01])

    # pixels: mostly interior, some on / near the border (invalid), some in the flat band
    u = torch.randint(0, width, (n,), generator=g)
    v = torch.randint(0, height, (n,), generator=g)
    uvs = torch.stack([u, v], dim=1).to(torch.int32).contiguous()
    depths = (2.0 + 2.5 * torch.rand(n, generator=g)).contiguous()
    nrm = torch.tensor([0.0, 0.0, -1.0]) + 0.35 * torch.randn(n, 3, generator=g)
    normals = (nrm / nrm.norm(dim=1, keepdim=True)).contiguous()
    # what the caller passes as R: the row-major tensor R^T, i.e. the columns of R in memory order
    return dict(depths=depths, normals=normals, uvs=uvs, R=R.t().contiguous(), T=T.contiguous(), image_r=image_r,
                image_n=image_n, fx_r=fx_r, fy_r=fy_r, cx_r=cx_r, cy_r=cy_r, fx_n=fx_n, fy_n=fy_n, cx_n=cx_n, cy_n=cy_n)


ARG_ORDER = ("depths", "normals", "uvs", "R", "T", "image_r", "image_n", "fx_r", "fy_r", "cx_r", "cy_r", "fx_n", "fy_n",
             "cx_n", "cy_n")


def args_of(sc, device=None):
    out = []
    for k in ARG_ORDER:
        v = sc[k]
        out.append(v.to(device) if (device is not None and torch.is_tensor(v)) else v)
    return out
