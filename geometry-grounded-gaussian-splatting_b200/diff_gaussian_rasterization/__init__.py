"""Drop-in replacement for the reference package `diff_gaussian_rasterization`
(XiShuFan/Geometry-Grounded-Gaussian-Splatting, submodules/diff-gaussian-rasterization,
DGR/diff_gaussian_rasterization/__init__.py).

Same public names, argument order, return order and exceptions:

    GaussianRasterizationSettings   NamedTuple, 15 fields            (reference :254-269)
    GaussianRasterizer(settings)    nn.Module                        (reference :272-483)
        .forward(means3D, means2D, opacities, shs, sg_axis, sg_sharpness, sg_color,
                 colors_precomp, scales, rotations, cov3D_precomp)
              -> (color [3,H,W], radii [P] int32, mdepth [1,H,W], alpha [1,H,W], normal [3,H,W])
        .integrate(points3D, means3D, opacities, scales, rotations, ...) -> (1 - T [PN], inside [PN] bool)
        .markVisible(positions) -> bool [P]

so gaussian_renderer/__init__.py, train.py, render.py and mesh_extract*.py import it unchanged.
All computation happens in hand-written sm_100a CUDA kernels behind the C ABI of
include/ggs_rasterizer.h; this file only marshals tensors.  No CPU / PyTorch fallback exists.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _abi
from ._abi import lib as _lib

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "PreparedView", "rasterize_gaussians",
           "activate_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    kernel_size: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    sg_degree: int
    campos: torch.Tensor
    prefiltered: bool
    require_depth: bool
    debug: bool


# View-parallel training (parallel.factored_sh_gradients): when set, the backward does not write the per-view SH
# gradient; it all-gathers the 3-float masked colour gradient of every view and forms the SUM over views locally
# (ggs_sh_gradient_from_views).  Process-wide on purpose: autograd runs the backward on its own thread.
# "exchange": a parallel.PeerGradientExchange -- the sums then go through peer memory in one fused launch instead of NCCL.
_SH_FACTORING = {"on": False, "group": None, "exchange": None}

# Pair-count hints for ggs_rasterize_forward (GgsForwardState.capacity): with a hint the host does not block on the
# count before it launches the rest of the forward (the reference does, rasterizer_impl.cu:384).  Per (device, P, H, W):
# 1.25 x the largest count of recent frames, decaying slowly; a frame that exceeds its hint is binned again at its exact
# size, so results never depend on it.  A cache of sizes, nothing else: losing it costs one blocking frame.
_CAPACITY_HINTS: dict = {}
_USE_HINTS = os.environ.get("GGS_CAPACITY_HINTS", "1") != "0"  # 0: every forward blocks on the count (A/B timing)


def _remember_pair_count(key, n):
    old = _CAPACITY_HINTS.get(key, 0)
    _CAPACITY_HINTS[key] = max(int(n * 1.25) + 4096, int(old * 0.97))
    if len(_CAPACITY_HINTS) > 64:
        _CAPACITY_HINTS.pop(next(iter(_CAPACITY_HINTS)))


def _dev_f32(t, device):
    """Contiguous float32 tensor on `device` (the reference calls .contiguous() on everything)."""
    if t is None:
        return None
    if t.device != device or t.dtype != torch.float32:
        t = t.to(device=device, dtype=torch.float32)
    return t.contiguous()


def _present(t):
    return t is not None and t.numel() != 0


def _frame(rs: GaussianRasterizationSettings, device, keep: list, kernel_size=None) -> _abi.GgsFrame:
    view = _dev_f32(rs.viewmatrix, device)
    proj = _dev_f32(rs.projmatrix, device)
    campos = _dev_f32(rs.campos, device)
    bg = rs.bg
    if not isinstance(bg, torch.Tensor):
        # the reference's point-query callers pass bg=None (gaussian_renderer/__init__.py:117,178) or bg=0 (:238);
        # those paths never read it
        bg = torch.full((3,), 0.0 if bg is None else float(bg), dtype=torch.float32, device=device)
    bg = _dev_f32(bg, device)
    keep += [view, proj, campos, bg]
    f = _abi.GgsFrame()
    f.image_width = int(rs.image_width)
    f.image_height = int(rs.image_height)
    f.tanfovx = float(rs.tanfovx)
    f.tanfovy = float(rs.tanfovy)
    f.kernel_size = float(rs.kernel_size if kernel_size is None else kernel_size)
    f.scale_modifier = float(rs.scale_modifier)
    f.sh_degree = int(rs.sh_degree)
    f.sg_degree = int(rs.sg_degree)
    f.prefiltered = int(bool(rs.prefiltered))
    f.require_depth = int(bool(rs.require_depth))
    f.debug = int(bool(rs.debug))
    f.viewmatrix = _abi.ptr(view)
    f.projmatrix = _abi.ptr(proj)
    f.campos = _abi.ptr(campos)
    f.background = _abi.ptr(bg)
    return f


def _gaussians(means3D, opacities, scales, rotations, sh, sg_axis, sg_sharpness, sg_color, colors_precomp,
               sg_degree) -> _abi.GgsGaussians:
    g = _abi.GgsGaussians()
    g.count = means3D.shape[0]
    g.means3D = _abi.ptr(means3D)
    g.opacities = _abi.ptr(opacities)
    g.scales = _abi.ptr(scales)
    g.rotations = _abi.ptr(rotations)
    g.shs = _abi.ptr(sh)
    g.sh_coeffs = sh.shape[1] if _present(sh) else 0
    has_sg = _present(sg_color) and _present(sg_axis) and _present(sg_sharpness)
    g.sg_axis = _abi.ptr(sg_axis) if has_sg else None
    g.sg_sharpness = _abi.ptr(sg_sharpness) if has_sg else None
    g.sg_color = _abi.ptr(sg_color) if has_sg else None
    g.sg_lobes = sg_color.shape[1] if has_sg else 0
    g.colors_precomp = _abi.ptr(colors_precomp)
    if sg_degree > g.sg_lobes:
        raise RuntimeError(f"sg_degree={sg_degree} but only {g.sg_lobes} spherical-Gaussian lobes were provided")
    return g


def _check_means(means3D):
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor: this rasterizer has no CPU path")


def rasterize_gaussians(means3D, means2D, sh, sg_axis, sg_sharpness, sg_color, colors_precomp, opacities, scales,
                        rotations, cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, sg_axis, sg_sharpness, sg_color, colors_precomp,
                                     opacities, scales, rotations, cov3Ds_precomp, raster_settings)


def _sum_sh_gradient_over_views(rs, means3D, d_rgb, SHM, device):
    """dL_dsh summed over the views of all ranks: all-gather (dL_drgb [P,3], campos [3]) and rebuild the rank-one
    per-view gradients locally.  12 bytes per Gaussian and view on the wire instead of all-reducing 12*SHM."""
    import torch.distributed as dist

    group = _SH_FACTORING["group"]
    campos = _dev_f32(rs.campos, device).reshape(3)
    P = means3D.shape[0]
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        world = dist.get_world_size(group)
        all_rgb = torch.empty(world, P, 3, dtype=torch.float32, device=device)
        all_cam = torch.empty(world, 3, dtype=torch.float32, device=device)
        dist.all_gather_into_tensor(all_rgb, d_rgb, group=group)
        dist.all_gather_into_tensor(all_cam, campos, group=group)
    else:
        world, all_rgb, all_cam = 1, d_rgb, campos
    d_sh = torch.empty(P, SHM, 3, dtype=torch.float32, device=device)
    _abi.check(_lib.ggs_sh_gradient_from_views(P, int(rs.sh_degree), SHM, world, _abi.ptr(means3D), _abi.ptr(all_cam),
                                               _abi.ptr(all_rgb), _abi.ptr(d_sh), _abi.stream_handle(device)),
               "ggs_sh_gradient_from_views")
    return d_sh


class _RasterizeGaussians(torch.autograd.Function):
    """Autograd node; argument and gradient order of the reference (:53-251)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, sg_axis, sg_sharpness, sg_color, colors_precomp, opacities, scales,
                rotations, cov3Ds_precomp, raster_settings):
        rs = raster_settings
        _check_means(means3D)
        if _present(cov3Ds_precomp):
            raise NotImplementedError(
                "cov3D_precomp is not supported: the reference kernel itself cannot consume precomputed "
                "covariances (render_forward.cu:344-345); pass scales and rotations")
        device = means3D.device
        P = means3D.shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        if P == 0:
            # rasterize_points.cu:95: nothing is launched, the zero-initialised outputs are returned
            ctx.empty = True
            ctx.in_shapes = [None if t is None else (t.shape, t.dtype, t.device) for t in
                             (means3D, means2D, sh, sg_axis, sg_sharpness, sg_color, colors_precomp, opacities, scales,
                              rotations)]
            z = lambda c: torch.zeros(c, H, W, dtype=torch.float32, device=device)  # noqa: E731
            radii = torch.zeros(0, dtype=torch.int32, device=device)
            ctx.mark_non_differentiable(radii)
            return z(3), radii, z(1), z(1), z(3)
        ctx.empty = False
        keep = []
        tens = [_dev_f32(t, device) if _present(t) else None
                for t in (means3D, opacities, scales, rotations, sh, sg_axis, sg_sharpness, sg_color, colors_precomp)]
        m3, op, sc, ro, sh_, sa, ss, sgc, cp = tens
        with torch.cuda.device(device):
            frame = _frame(rs, device, keep)
            g = _gaussians(m3, op, sc, ro, sh_, sa, ss, sgc, cp, int(rs.sg_degree))
            color = torch.empty(3, H, W, dtype=torch.float32, device=device)
            mdepth = torch.empty(1, H, W, dtype=torch.float32, device=device)
            alpha = torch.empty(1, H, W, dtype=torch.float32, device=device)
            normal = torch.empty(3, H, W, dtype=torch.float32, device=device)
            radii = torch.empty(P, dtype=torch.int32, device=device)
            out = _abi.GgsForwardOutputs(_abi.ptr(color), _abi.ptr(alpha), _abi.ptr(normal), _abi.ptr(mdepth),
                                         _abi.ptr(radii))
            state = _abi.GgsForwardState()
            hint_key = (device.index, P, H, W)
            state.capacity = _CAPACITY_HINTS.get(hint_key, 0) if (_USE_HINTS and not rs.debug) else 0
            arena = _abi.Arena(device)
            n = _lib.ggs_rasterize_forward(C.byref(frame), C.byref(g), arena.fn, None, C.byref(out), C.byref(state),
                                           _abi.stream_handle(device))
            arena.done()
            if n < 0:
                if arena.error is not None:
                    raise arena.error
                _abi.check(n, "ggs_rasterize_forward")
            _remember_pair_count(hint_key, int(n))
        ctx.raster_settings = rs
        ctx.num_rendered = int(n)
        ctx.capacity = int(state.capacity)
        ctx.arena = arena
        ctx.had = [_present(t) for t in (sh, sg_axis, sg_sharpness, sg_color, colors_precomp)]
        ctx.save_for_backward(m3, op, sc, ro, sh_, sa, ss, sgc, cp, alpha, normal, mdepth, radii)
        ctx.mark_non_differentiable(radii)
        return color, radii, mdepth, alpha, normal

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_mdepth, grad_alpha, grad_normal):
        if ctx.empty:
            gz = [None if (d is None or not ctx.needs_input_grad[i]) else torch.zeros(d[0], dtype=d[1], device=d[2])
                  for i, d in enumerate(ctx.in_shapes)]
            return (*gz, None, None)
        rs = ctx.raster_settings
        m3, op, sc, ro, sh_, sa, ss, sgc, cp, alpha, normal, mdepth, radii = ctx.saved_tensors
        device = m3.device
        P = m3.shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        keep = []

        def g_or_zero(gr, c):
            if gr is None:
                return torch.zeros(c, H, W, dtype=torch.float32, device=device)
            return _dev_f32(gr, device)

        gc, gd, ga, gn = g_or_zero(grad_color, 3), g_or_zero(grad_mdepth, 1), g_or_zero(grad_alpha, 1), \
            g_or_zero(grad_normal, 3)
        opts = dict(dtype=torch.float32, device=device)
        SHM = sh_.shape[1] if sh_ is not None else 0
        SGM = sgc.shape[1] if (sgc is not None and sa is not None and ss is not None) else 0
        factored = bool(_SH_FACTORING["on"]) and sh_ is not None and SGM == 0
        exchange = _SH_FACTORING["exchange"] if factored else None
        d_means2D = torch.empty(P, 3, **opts)
        d_colors = torch.empty(P, 3, **opts)
        if exchange is not None:
            # the per-Gaussian backward writes straight into this rank's region of the peer-mapped exchange buffer
            exchange.check(P, device)
            d_means3D, d_opacity, d_scales, d_rot = exchange.send_views()
            d_rgb = None
        else:
            d_means3D = torch.empty(P, 3, **opts)
            d_opacity = torch.empty(P, 1, **opts)
            d_scales = torch.empty(P, 3, **opts)
            d_rot = torch.empty(P, 4, **opts)
            d_rgb = torch.empty(P, 3, **opts) if factored else None
        d_sh = None if factored else torch.empty(P, SHM, 3, **opts)
        d_sg_axis = torch.empty(P, SGM, 3, **opts)
        d_sg_sharp = torch.empty(P, SGM, **opts)
        d_sg_color = torch.empty(P, SGM, 3, **opts)
        with torch.cuda.device(device):
            frame = _frame(rs, device, keep)
            g = _gaussians(m3, op, sc, ro, sh_, sa, ss, sgc, cp, int(rs.sg_degree))
            fwd = ctx.arena.buffers
            state = _abi.GgsForwardState(
                _abi.ptr(fwd.get(_abi.SLOT_GEOMETRY)), _abi.ptr(fwd.get(_abi.SLOT_BINNING)),
                _abi.ptr(fwd.get(_abi.SLOT_IMAGE)), _abi.ptr(fwd.get(_abi.SLOT_TILE)), ctx.num_rendered,
                ctx.capacity)
            bi = _abi.GgsBackwardInputs(_abi.ptr(alpha), _abi.ptr(normal), _abi.ptr(mdepth), _abi.ptr(radii),
                                        _abi.ptr(gc), _abi.ptr(gd), _abi.ptr(ga), _abi.ptr(gn))
            grads = _abi.GgsGradients(_abi.ptr(d_means3D), _abi.ptr(d_means2D), _abi.ptr(d_colors),
                                      _abi.ptr(d_opacity), _abi.ptr(d_scales), _abi.ptr(d_rot), _abi.ptr(d_sh),
                                      _abi.ptr(d_sg_axis), _abi.ptr(d_sg_sharp), _abi.ptr(d_sg_color), _abi.ptr(d_rgb))
            arena = _abi.Arena(device)
            if exchange is not None:
                rc = _lib.ggs_rasterize_backward_peer(C.byref(frame), C.byref(g), C.byref(state), C.byref(bi), arena.fn,
                                                      None, C.byref(grads), C.byref(exchange.px),
                                                      _abi.stream_handle(device),
                                                      C.c_void_p(exchange.side.cuda_stream))
            else:
                rc = _lib.ggs_rasterize_backward(C.byref(frame), C.byref(g), C.byref(state), C.byref(bi), arena.fn,
                                                 None, C.byref(grads), _abi.stream_handle(device))
            arena.done()
            if rc < 0:
                if arena.error is not None:
                    raise arena.error
                _abi.check(rc, "ggs_rasterize_backward")
            if exchange is not None:
                # sums over the views of all ranks, one fused launch over peer memory (parallel.PeerGradientExchange)
                d_means3D, d_opacity, d_scales, d_rot, d_sh = exchange.sum_over_views(rs, m3, SHM)
            elif factored:
                d_sh = _sum_sh_gradient_over_views(rs, m3, d_rgb, SHM, device)
        had_sh, had_sa, had_ss, had_sgc, had_cp = ctx.had
        if SGM == 0:
            # lobes absent or unused: gradients of whatever was passed are zero
            d_sg_axis = torch.zeros_like(sa) if sa is not None else None
            d_sg_sharp = torch.zeros_like(ss) if ss is not None else None
            d_sg_color = torch.zeros_like(sgc) if sgc is not None else None
        return (
            d_means3D,
            d_means2D,
            d_sh if had_sh else None,
            d_sg_axis if had_sa else None,
            d_sg_sharp if had_ss else None,
            d_sg_color if had_sgc else None,
            d_colors if had_cp else None,
            d_opacity,
            d_scales,
            d_rot,
            None,
            None,
        )


class _SampleDepth(torch.autograd.Function):
    """Autograd node of sample_depth (reference :486-655).  As in the reference, the forward runs with kernel_size
    0 (:513) and the backward's per-Gaussian chain with the settings' kernel_size (:606)."""

    @staticmethod
    def forward(ctx, points3D, means3D, opacities, scales, rotations, raster_settings):
        rs = raster_settings
        _check_means(means3D)
        if points3D.dim() < 1 or points3D.shape[-1] != 3:
            raise RuntimeError("points3D must have shape (..., 3) with last dimension == 3")
        device = means3D.device
        pts = _dev_f32(points3D, device).reshape(-1, 3)
        m3, op, sc, ro = (_dev_f32(t, device) for t in (means3D, opacities, scales, rotations))
        PN, P = pts.shape[0], m3.shape[0]
        out = torch.zeros(PN, 3, dtype=torch.float32, device=device)
        inside = torch.zeros(PN, dtype=torch.bool, device=device)
        keep = []
        state = _abi.GgsSampleState()
        arena = _abi.Arena(device)
        with torch.cuda.device(device):
            frame = _frame(rs, device, keep, kernel_size=0.0)
            g = _gaussians(m3, op, sc, ro, None, None, None, None, None, 0)
            n = _lib.ggs_sample_depth_forward(C.byref(frame), C.byref(g), _abi.ptr(pts), PN, arena.fn, None,
                                              _abi.ptr(out), _abi.ptr(inside), C.byref(state),
                                              _abi.stream_handle(device))
            arena.done()
            if n < 0:
                if arena.error is not None:
                    raise arena.error
                _abi.check(n, "ggs_sample_depth_forward")
        ctx.raster_settings = rs
        ctx.arena = arena
        ctx.counts = (int(state.num_rendered), int(state.num_evaluated), int(state.num_blocks))
        ctx.pts_shape = points3D.shape
        ctx.save_for_backward(pts, m3, op, sc, ro, inside)
        inside_out = inside.reshape(points3D.shape[:-1])
        ctx.mark_non_differentiable(inside_out)
        return out.reshape(points3D.shape), inside_out

    @staticmethod
    def backward(ctx, grad_camera_points, grad_inside):
        rs = ctx.raster_settings
        pts, m3, op, sc, ro, inside = ctx.saved_tensors
        device = m3.device
        PN, P = pts.shape[0], m3.shape[0]
        opts = dict(dtype=torch.float32, device=device)
        gout = (torch.zeros(PN, 3, **opts) if grad_camera_points is None
                else _dev_f32(grad_camera_points, device).reshape(-1, 3))
        d_means3D = torch.empty(P, 3, **opts)
        d_means2D = torch.empty(P, 3, **opts)
        d_colors = torch.empty(P, 3, **opts)
        d_opacity = torch.empty(P, 1, **opts)
        d_scales = torch.empty(P, 3, **opts)
        d_rot = torch.empty(P, 4, **opts)
        d_pts = torch.empty(PN, 3, **opts)
        keep = []
        with torch.cuda.device(device):
            frame = _frame(rs, device, keep)
            g = _gaussians(m3, op, sc, ro, None, None, None, None, None, 0)
            b = ctx.arena.buffers
            nr, ne, nb = ctx.counts
            state = _abi.GgsSampleState(
                _abi.ptr(b.get(_abi.SLOT_GEOMETRY)), _abi.ptr(b.get(_abi.SLOT_BINNING)), _abi.ptr(b.get(_abi.SLOT_POINT)),
                None, _abi.ptr(b.get(_abi.SLOT_TILE)), None, nr, ne, nb)
            grads = _abi.GgsGradients(_abi.ptr(d_means3D), _abi.ptr(d_means2D), _abi.ptr(d_colors),
                                      _abi.ptr(d_opacity), _abi.ptr(d_scales), _abi.ptr(d_rot))
            arena = _abi.Arena(device)
            rc = _lib.ggs_sample_depth_backward(C.byref(frame), C.byref(g), _abi.ptr(pts), PN, C.byref(state),
                                                _abi.ptr(inside), _abi.ptr(gout), arena.fn, None, C.byref(grads),
                                                _abi.ptr(d_pts), _abi.stream_handle(device))
            arena.done()
            if rc < 0:
                if arena.error is not None:
                    raise arena.error
                _abi.check(rc, "ggs_sample_depth_backward")
        return d_pts.reshape(ctx.pts_shape), d_means3D, d_opacity, d_scales, d_rot, None


class _ActivateGaussians(torch.autograd.Function):
    """One kernel each way for the activations the reference applies in torch before every rasterizer call
    (scene/gaussian_model.py:203-212 get_scaling_n_opacity_with_3D_filter, :158-160 get_rotation)."""

    @staticmethod
    def forward(ctx, raw_scaling, raw_opacity, raw_rotation, filter_3D):
        if not raw_scaling.is_cuda:
            raise RuntimeError("activate_gaussians needs CUDA tensors: this library has no CPU path")
        device = raw_scaling.device
        rs, ro, rq, f = (_dev_f32(t, device) for t in (raw_scaling, raw_opacity, raw_rotation, filter_3D))
        P = rs.shape[0]
        if rs.shape != (P, 3) or ro.numel() != P or rq.shape != (P, 4) or f.numel() != P:
            raise RuntimeError("expected raw_scaling [P,3], raw_opacity [P,1], raw_rotation [P,4], filter_3D [P,1]")
        scales = torch.empty(P, 3, dtype=torch.float32, device=device)
        opac = torch.empty(P, 1, dtype=torch.float32, device=device)
        rot = torch.empty(P, 4, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            _abi.check(_lib.ggs_activate_forward(P, _abi.ptr(rs), _abi.ptr(ro), _abi.ptr(rq), _abi.ptr(f),
                                                 _abi.ptr(scales), _abi.ptr(opac), _abi.ptr(rot),
                                                 _abi.stream_handle(device)), "ggs_activate_forward")
        ctx.save_for_backward(rs, ro, rq, f)
        ctx.shapes = (raw_scaling.shape, raw_opacity.shape, raw_rotation.shape)
        return scales, opac, rot

    @staticmethod
    def backward(ctx, g_scales, g_opac, g_rot):
        rs, ro, rq, f = ctx.saved_tensors
        device = rs.device
        P = rs.shape[0]
        z = lambda g, w: torch.zeros(P, w, dtype=torch.float32, device=device) if g is None else _dev_f32(g, device)  # noqa: E731
        gs, go, gq = z(g_scales, 3), z(g_opac, 1), z(g_rot, 4)
        d_rs = torch.empty(P, 3, dtype=torch.float32, device=device)
        d_ro = torch.empty(P, 1, dtype=torch.float32, device=device)
        d_rq = torch.empty(P, 4, dtype=torch.float32, device=device)
        with torch.cuda.device(device):
            _abi.check(_lib.ggs_activate_backward(P, _abi.ptr(rs), _abi.ptr(ro), _abi.ptr(rq), _abi.ptr(f), _abi.ptr(gs),
                                                  _abi.ptr(go), _abi.ptr(gq), _abi.ptr(d_rs), _abi.ptr(d_ro),
                                                  _abi.ptr(d_rq), _abi.stream_handle(device)),
                       "ggs_activate_backward")
        s0, s1, s2 = ctx.shapes
        return d_rs.view(s0), d_ro.view(s1), d_rq.view(s2), None


def activate_gaussians(raw_scaling, raw_opacity, raw_rotation, filter_3D):
    """(scales [P,3], opacities [P,1], rotations [P,4]) exactly as gaussian_renderer/__init__.py:62-63 obtains them from
    `pc.get_scaling_n_opacity_with_3D_filter` and `pc.get_rotation`, differentiable w.r.t. the three raw parameters
    (filter_3D is a buffer): one fused kernel forward and one backward instead of ~30 elementwise launches.
    Opt-in: the drop-in surface does not need it."""
    return _ActivateGaussians.apply(raw_scaling, raw_opacity, raw_rotation, filter_3D)


class PreparedView:
    """One view's Gaussians, binned once, for any number of point-query batches.

    The reference bins the Gaussians of the view inside every `integrate` / `evaluate_sdf` call
    (rasterizer_impl.cu:637-714, :861-936), i.e. once per (chunk, view) of mesh_extract_tetrahedra.py:65-87, and once
    more for every round of its binary search.  `GaussianRasterizer.prepare(...)` does that half once; `integrate` and
    `evaluate_sdf` of the returned object only bin the query points (no host synchronisation) and run the query
    kernel.  Results are those of the one-shot methods, bit for bit.  No autograd (as in the reference)."""

    def __init__(self, raster_settings, means3D, opacities, scales, rotations):
        _check_means(means3D)
        rs = raster_settings
        self.raster_settings = rs
        device = self.device = means3D.device
        with torch.no_grad():
            m3, op, sc, ro = (_dev_f32(t, device) for t in (means3D, opacities, scales, rotations))
            self._keep = []
            with torch.cuda.device(device):
                # the reference's wrappers pass kernel_size 0.0 on the point-query paths (__init__.py:382,443)
                self._frame = _frame(rs, device, self._keep, kernel_size=0.0)
                g = _gaussians(m3, op, sc, ro, None, None, None, None, None, 0)
                self._arena = arena = _abi.Arena(device)
                self._view = _abi.GgsViewState()
                n = _lib.ggs_view_prepare(C.byref(self._frame), C.byref(g), arena.fn, None, C.byref(self._view),
                                          _abi.stream_handle(device))
                arena.done()
                if n < 0:
                    if arena.error is not None:
                        raise arena.error
                    _abi.check(n, "ggs_view_prepare")
        self.num_rendered = int(n)
        self.count = int(m3.shape[0])

    def _points(self, points3D):
        if points3D.dim() != 2 or points3D.shape[1] != 3:
            raise RuntimeError("points3D must have dimensions (num_points, 3)")
        return _dev_f32(points3D, self.device)

    def integrate(self, points3D, op_counters=None):
        """(1 - transmittance [PN], inside [PN] bool), as GaussianRasterizer.integrate.  `op_counters` (measurement
        only): an int64 [3] device tensor that receives the kernel's pair counts (ggs_integrate_view_counted)."""
        device = self.device
        with torch.no_grad():
            pts = self._points(points3D)
            PN = pts.shape[0]
            transmittance = torch.zeros(PN, dtype=torch.float32, device=device)
            inside = torch.zeros(PN, dtype=torch.bool, device=device)
            with torch.cuda.device(device):
                arena = _abi.Arena(device)
                if op_counters is not None:
                    rc = _lib.ggs_integrate_view_counted(C.byref(self._frame), C.byref(self._view), _abi.ptr(pts), PN,
                                                         arena.fn, None, _abi.ptr(transmittance), _abi.ptr(inside),
                                                         _abi.ptr(op_counters), _abi.stream_handle(device))
                else:
                    rc = _lib.ggs_integrate_view(C.byref(self._frame), C.byref(self._view), _abi.ptr(pts), PN, arena.fn,
                                                 None, _abi.ptr(transmittance), _abi.ptr(inside),
                                                 _abi.stream_handle(device))
                arena.done()
                if rc < 0:
                    if arena.error is not None:
                        raise arena.error
                    _abi.check(rc, "ggs_integrate_view")
        return 1 - transmittance, inside

    def evaluate_sdf(self, points3D):
        """(median ray depth, depth minus the point's own ray depth, inside), as GaussianRasterizer.evaluate_sdf."""
        device = self.device
        with torch.no_grad():
            pts = self._points(points3D)
            PN = pts.shape[0]
            depth = torch.zeros(PN, dtype=torch.float32, device=device)
            sdf = torch.zeros(PN, dtype=torch.float32, device=device)
            inside = torch.zeros(PN, dtype=torch.bool, device=device)
            with torch.cuda.device(device):
                arena = _abi.Arena(device)
                rc = _lib.ggs_evaluate_sdf_view(C.byref(self._frame), C.byref(self._view), _abi.ptr(pts), PN, arena.fn,
                                                None, _abi.ptr(depth), _abi.ptr(sdf), _abi.ptr(inside),
                                                _abi.stream_handle(device))
                arena.done()
                if rc < 0:
                    if arena.error is not None:
                        raise arena.error
                    _abi.check(rc, "ggs_evaluate_sdf_view")
        return depth, sdf, inside


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of points in front of the near plane (reference :277-283)."""
        rs = self.raster_settings
        with torch.no_grad():
            device = positions.device
            pos = _dev_f32(positions, device)
            P = pos.shape[0]
            present = torch.zeros(P, dtype=torch.bool, device=device)
            keep = []
            with torch.cuda.device(device):
                view = _dev_f32(rs.viewmatrix, device)
                proj = _dev_f32(rs.projmatrix, device)
                keep += [view, proj]
                _abi.check(_lib.ggs_mark_visible(_abi.ptr(pos), P, _abi.ptr(view), _abi.ptr(proj),
                                                 _abi.ptr(present), _abi.stream_handle(device)), "ggs_mark_visible")
        return present

    def forward(self, means3D, means2D, opacities, shs=None, sg_axis=None, sg_sharpness=None, sg_color=None,
                colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, sg_axis, sg_sharpness, sg_color, colors_precomp, opacities,
                                   scales, rotations, cov3D_precomp, rs)

    def prepare(self, means3D, opacities, scales, rotations):
        """Bins this view's Gaussians once for repeated point queries (see PreparedView).  Not part of the reference's
        surface: `integrate` / `evaluate_sdf` below keep its one-shot signatures."""
        return PreparedView(self.raster_settings, means3D, opacities, scales, rotations)

    def integrate(self, points3D, means3D, opacities, scales=None, rotations=None, cov3D_precomp=None,
                  view2gaussian_precomp=None):
        """Opacity (1 - transmittance) of query points seen from this camera (reference :338-402)."""
        rs = self.raster_settings
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        if cov3D_precomp is not None and cov3D_precomp.numel():
            raise NotImplementedError("cov3D_precomp is not supported; pass scales and rotations")
        _check_means(means3D)
        if points3D.dim() != 2 or points3D.shape[1] != 3:
            raise RuntimeError("points3D must have dimensions (num_points, 3)")
        device = means3D.device
        with torch.no_grad():
            pts = _dev_f32(points3D, device)
            m3, op, sc, ro = (_dev_f32(t, device) for t in (means3D, opacities, scales, rotations))
            PN = pts.shape[0]
            transmittance = torch.zeros(PN, dtype=torch.float32, device=device)
            inside = torch.zeros(PN, dtype=torch.bool, device=device)
            keep = []
            with torch.cuda.device(device):
                frame = _frame(rs, device, keep, kernel_size=0.0)  # the reference passes 0.0 here (:382)
                g = _gaussians(m3, op, sc, ro, None, None, None, None, None, 0)
                arena = _abi.Arena(device)
                n = _lib.ggs_integrate_points(C.byref(frame), C.byref(g), _abi.ptr(pts), PN, arena.fn, None,
                                              _abi.ptr(transmittance), _abi.ptr(inside), _abi.stream_handle(device))
                arena.done()
                if n < 0:
                    if arena.error is not None:
                        raise arena.error
                    _abi.check(n, "ggs_integrate_points")
                # the arena's buffers are released when their last use on this stream has been enqueued;
                # torch's caching allocator keeps them stream-ordered
        return 1 - transmittance, inside

    def evaluate_sdf(self, points3D, means3D, opacities, scales=None, rotations=None, cov3D_precomp=None,
                     view2gaussian_precomp=None):
        """(median ray depth, depth minus the point's own ray depth, inside) per query point (reference :404-468)."""
        rs = self.raster_settings
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        if cov3D_precomp is not None and cov3D_precomp.numel():
            raise NotImplementedError("cov3D_precomp is not supported; pass scales and rotations")
        _check_means(means3D)
        if points3D.dim() != 2 or points3D.shape[1] != 3:
            raise RuntimeError("points3D must have dimensions (num_points, 3)")
        device = means3D.device
        with torch.no_grad():
            pts = _dev_f32(points3D, device)
            m3, op, sc, ro = (_dev_f32(t, device) for t in (means3D, opacities, scales, rotations))
            PN = pts.shape[0]
            depth = torch.zeros(PN, dtype=torch.float32, device=device)
            sdf = torch.zeros(PN, dtype=torch.float32, device=device)
            inside = torch.zeros(PN, dtype=torch.bool, device=device)
            keep = []
            with torch.cuda.device(device):
                frame = _frame(rs, device, keep, kernel_size=0.0)  # the reference passes 0.0 here (:443)
                g = _gaussians(m3, op, sc, ro, None, None, None, None, None, 0)
                arena = _abi.Arena(device)
                n = _lib.ggs_evaluate_sdf(C.byref(frame), C.byref(g), _abi.ptr(pts), PN, arena.fn, None,
                                          _abi.ptr(depth), _abi.ptr(sdf), _abi.ptr(inside), _abi.stream_handle(device))
                arena.done()
                if n < 0:
                    if arena.error is not None:
                        raise arena.error
                    _abi.check(n, "ggs_evaluate_sdf")
        return depth, sdf, inside

    def sample_depth(self, points3D, means3D, opacities, scales=None, rotations=None, cov3D_precomp=None):
        """Camera-space points at the rendered median depth of arbitrary query points (differentiable; reference
        :470-483).  Returns (camera_points [...,3], inside [...] bool)."""
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        if cov3D_precomp is not None and cov3D_precomp.numel():
            raise NotImplementedError("cov3D_precomp is not supported; pass scales and rotations")
        return _SampleDepth.apply(points3D, means3D, opacities, scales, rotations, self.raster_settings)
