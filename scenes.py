"""Deterministic synthetic scenes for parity tests and bench.

Everything is generated on the CPU with a seeded ``torch.Generator`` so the same
tensors exist in this container, on the GPU box and in the committed golden
fixtures.  The reference ships no scenes or fixtures for the rasterizer; these
are ours.

The camera follows the reference's construction exactly
(``utils/graphics_utils.py:72-92`` getProjectionMatrix, ``scene/cameras.py:70-73``):
``viewmatrix`` is the transposed world->view matrix, ``projmatrix`` is
``viewmatrix @ P^T`` and ``campos`` is row 3 of the inverse view matrix.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch

# name -> (P, W, H, focal_px, sh_degree, sg_degree)   
CONFIGS = {
    "tiny": (2_000, 96, 80, 90.0, 3, 0),
    "tiny_sg": (1_500, 80, 96, 90.0, 2, 3),
    "c1": (10_000, 256, 256, 220.0, 3, 0),
    "c2": (300_000, 800, 600, 1446.0, 3, 0),
    "c3": (1_000_000, 1237, 822, 1040.0, 3, 0),
    "c3_3m": (3_000_000, 1237, 822, 1040.0, 3, 0),
    "c4": (2_000_000, 980, 545, 580.0, 3, 0),
    "c5": (1_000_000, 1237, 822, 1040.0, 2, 7),
}


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor  # [4,4] transposed world->view
    projmatrix: torch.Tensor  # [4,4] full projection (transposed convention)
    campos: torch.Tensor  # [3]


@dataclass
class Scene:
    name: str
    sh_degree: int
    sg_degree: int
    kernel_size: float
    cam: Camera
    bg: torch.Tensor
    means3D: torch.Tensor
    scales: torch.Tensor
    rotations: torch.Tensor
    opacities: torch.Tensor
    shs: torch.Tensor
    sg_axis: torch.Tensor
    sg_sharpness: torch.Tensor
    sg_color: torch.Tensor
    grads: Dict[str, torch.Tensor] = field(default_factory=dict)

    @property
    def P(self) -> int:
        return self.means3D.shape[0]

    def to(self, device) -> "Scene":
        def mv(t):
            return t.to(device) if isinstance(t, torch.Tensor) else t

        cam = Camera(self.cam.W, self.cam.H, self.cam.tanfovx, self.cam.tanfovy,
                     mv(self.cam.viewmatrix), mv(self.cam.projmatrix), mv(self.cam.campos))
        return Scene(self.name, self.sh_degree, self.sg_degree, self.kernel_size, cam, mv(self.bg),
                     mv(self.means3D), mv(self.scales), mv(self.rotations), mv(self.opacities),
                     mv(self.shs), mv(self.sg_axis), mv(self.sg_sharpness), mv(self.sg_color),
                     {k: mv(v) for k, v in self.grads.items()})


def projection_matrix(znear: float, zfar: float, tanx: float, tany: float) -> torch.Tensor:
    """Same entries as the reference's getProjectionMatrix (utils/graphics_utils.py:72-92)."""
    top = tany * znear
    right = tanx * znear
    P = torch.zeros(4, 4, dtype=torch.float32)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def look_at_camera(W: int, H: int, focal: float, eye, target=(0.0, 0.0, 0.0)) -> Camera:
    """World->view with +z forward, built like scene/cameras.py:70-73."""
    eye_t = torch.tensor(eye, dtype=torch.float64)
    tgt = torch.tensor(target, dtype=torch.float64)
    fwd = tgt - eye_t
    fwd = fwd / fwd.norm()
    up = torch.tensor([0.0, -1.0, 0.0], dtype=torch.float64)
    if abs(float(fwd @ up)) > 0.99:
        up = torch.tensor([1.0, 0.0, 0.0], dtype=torch.float64)
    right = torch.linalg.cross(up, fwd)
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    R = torch.stack([right, down, fwd], dim=0)  # rows = camera axes in world
    Rt = torch.eye(4, dtype=torch.float64)
    Rt[:3, :3] = R
    Rt[:3, 3] = -R @ eye_t
    view = Rt.to(torch.float32).transpose(0, 1).contiguous()
    tanx = W / (2.0 * focal)
    tany = H / (2.0 * focal)
    proj = projection_matrix(0.01, 100.0, tanx, tany).transpose(0, 1)
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return Camera(W, H, tanx, tany, view, full, campos)


def ring_camera(W: int, H: int, focal: float, view: int = 0, n_views: int = 1) -> Camera:
    """Camera `view` of `n_views` on a circle of radius 4 around the origin, looking at it (view 0 = the single-view
    camera at z = -4)."""
    if n_views > 1:
        ang = 2.0 * math.pi * view / n_views
        eye = (4.0 * math.sin(ang), 0.0, -4.0 * math.cos(ang))
    else:
        eye = (0.0, 0.0, -4.0)
    return look_at_camera(W, H, focal, eye)


def config_camera(name: str, view: int = 0, n_views: int = 1) -> Camera:
    """The camera make_scene(name, view=view, n_views=n_views) would use, without generating the Gaussians."""
    _, W, H, focal, _, _ = CONFIGS[name]
    return ring_camera(W, H, focal, view, n_views)


def make_scene(name: str = "c1", *, seed: Optional[int] = None, view: int = 0, n_views: int = 1,
               P: Optional[int] = None, W: Optional[int] = None, H: Optional[int] = None,
               kernel_size: float = 0.0, random_bg: bool = False, sh_degree: Optional[int] = None,
               sg_degree: Optional[int] = None, scale_mult: float = 1.0,
               mask_depth_grad: bool = False, opacity_range: Optional[tuple] = None) -> Scene:
    cP, cW, cH, focal, csh, csg = CONFIGS[name]
    P = cP if P is None else P
    if W is not None and W != cW:
        focal = focal * W / cW
    W = cW if W is None else W
    H = cH if H is None else H
    sh_degree = csh if sh_degree is None else sh_degree
    sg_degree = csg if sg_degree is None else sg_degree
    if seed is None:
        seed = 1000 * (sorted(CONFIGS).index(name) + 1)
    g = torch.Generator().manual_seed(seed)

    def randn(*shape):
        return torch.randn(*shape, generator=g, dtype=torch.float32)

    def rand(*shape):
        return torch.rand(*shape, generator=g, dtype=torch.float32)

    n_far = P // 10
    n_near = P - n_far
    near = randn(n_near, 3) * 0.6
    r = near.norm(dim=1, keepdim=True).clamp_min(1e-6)
    near = near * torch.clamp(2.0 / r, max=1.0)
    far = (rand(n_far, 3) - 0.5) * 8.0
    means = torch.cat([near, far], 0)
    means = means[torch.randperm(P, generator=g)].contiguous()

    scales = torch.exp(math.log(0.012 * scale_mult) + 0.6 * randn(P, 3)).contiguous()
    rot = randn(P, 4)
    rot = (rot / rot.norm(dim=1, keepdim=True).clamp_min(1e-8)).contiguous()
    opac = torch.sigmoid(2.0 * randn(P, 1)).contiguous()
    if opacity_range is not None:
        # nearly transparent splats: the blend walks thousands of list entries before T < 1e-4 (long-list scenes)
        lo, hi = opacity_range
        opac = (lo + (hi - lo) * rand(P, 1)).contiguous()
    n_coef = 16  # allocated like the reference model (max_sh_degree 3): SHM = 16 regardless of active degree
    shs = torch.cat([randn(P, 1, 3), 0.15 * randn(P, n_coef - 1, 3)], 1).contiguous()
    SGM = sg_degree
    if SGM > 0:
        ax = randn(P, SGM, 3)
        ax = (ax / ax.norm(dim=2, keepdim=True).clamp_min(1e-8)).contiguous()
        sharp = torch.nn.functional.softplus(randn(P, SGM)).contiguous()
        sgc = (0.15 * randn(P, SGM, 3)).contiguous()
    else:
        ax = torch.zeros(P, 0, 3)
        sharp = torch.zeros(P, 0)
        sgc = torch.zeros(P, 0, 3)

    cam = ring_camera(W, H, focal, view, n_views)
    bg = rand(3) if random_bg else torch.zeros(3)

    gg = torch.Generator().manual_seed(seed + 7919 + view)
    scale = 1.0 / (H * W)
    grads = {
        "color": torch.randn(3, H, W, generator=gg) * scale,
        "mdepth": torch.randn(1, H, W, generator=gg) * scale,
        "alpha": torch.randn(1, H, W, generator=gg) * scale,
        "normal": torch.randn(3, H, W, generator=gg) * scale,
    }
    sc = Scene(name, sh_degree, sg_degree, kernel_size, cam, bg, means, scales, rot, opac, shs, ax, sharp, sgc, grads)
    sc.mask_depth_grad = mask_depth_grad
    return sc


def rotation_matrices(q: torch.Tensor) -> torch.Tensor:
    """[P,4] (r,x,y,z) -> [P,3,3]; normalises like utils/general_utils.py build_rotation."""
    q = q / q.norm(dim=1, keepdim=True)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.view(-1, 3, 3)


def tetra_points(sc: Scene, device="cpu") -> torch.Tensor:
    """The query points of the tetrahedral mesh extraction: per Gaussian the 8 corners of its box at 1.5 sigma, the
    6 axis points at 3 sigma (all 14 per Gaussian first, Gaussian-major), then the centres -- the 15-point stencil of
    GaussianModel.get_tetra_points (scene/gaussian_model.py:496-519).  [15 P, 3]."""
    xyz = sc.means3D.to(device)
    scale = sc.scales.to(device)
    R = rotation_matrices(sc.rotations.to(device))
    box = torch.tensor([[sx, sy, sz] for sx in (-1.0, 1.0) for sy in (-1.0, 1.0) for sz in (-1.0, 1.0)],
                       dtype=torch.float32, device=device) * 1.5
    axes = torch.tensor([[1, 0, 0], [0, 1, 0], [0, 0, 1], [-1, 0, 0], [0, -1, 0], [0, 0, -1]],
                        dtype=torch.float32, device=device) * 3.0
    verts = torch.cat([box, axes], 0).t().unsqueeze(0)  # [1,3,14]
    v = verts * scale.unsqueeze(-1)  # [P,3,14]
    v = torch.bmm(R, v) + xyz.unsqueeze(-1)
    v = v.permute(0, 2, 1).reshape(-1, 3)
    return torch.cat([v, xyz], 0).contiguous()
