"""The reference's compiled extension `_C`, as a ctypes shim over libggs_rasterizer.so.

For a maintainer who wants to keep the reference's own Python wrapper
(DGR/diff_gaussian_rasterization/__init__.py) verbatim: drop this module next to it in place of the compiled
`_C` and the wrapper's seven calls (DGR/ext.cpp:15-23) land in the C ABI of include/ggs_rasterizer.h.  Same
function names, positional arguments and return tuples as the pybind functions:

    rasterize_gaussians                RasterizeGaussiansCUDA            rasterize_points.cu:39-139
    rasterize_gaussians_backward       RasterizeGaussiansBackwardCUDA    :141-258
    mark_visible                       markVisible                       :260-277
    integrate_gaussians_to_points      IntegrateGaussiansToPointsCUDA    :279-366
    evaluate_sdf_from_signle_view      EvaluateSDFfromSingleViewCUDA     :368-457
    sample_rasterized_depth            SampleRasterizedDepthCUDA         :459-557
    sample_rasterized_depth_backward   SampleRasterizedDepthBackwardCUDA :559-634

The byte buffers the reference hands back to Python (geomBuffer, binningBuffer, ...) are this library's opaque
workspaces here; the wrapper only stores them and passes them back.  tests/test_reference_wrapper_gpu.py runs the
reference's unmodified wrapper on this module against the reference's own extension.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _abi
from ._abi import lib as _lib


def _f32(t, device):
    if t is None or t.numel() == 0:
        return None
    if t.device != device or t.dtype != torch.float32:
        t = t.to(device=device, dtype=torch.float32)
    return t.contiguous()


def _empty(device):
    return torch.empty(0, dtype=torch.uint8, device=device)


def _frame(device, keep, W, H, tanfovx, tanfovy, kernel_size, scale_modifier, viewmatrix, projmatrix, campos, bg=None,
           sh_degree=0, sg_degree=0, prefiltered=False, require_depth=True, debug=False):
    view, proj, cam = (_f32(t, device) for t in (viewmatrix, projmatrix, campos))
    bgt = _f32(bg, device) if isinstance(bg, torch.Tensor) else None
    if bgt is None:
        bgt = torch.zeros(3, dtype=torch.float32, device=device)
    keep += [view, proj, cam, bgt]
    f = _abi.GgsFrame()
    f.image_width, f.image_height = int(W), int(H)
    f.tanfovx, f.tanfovy = float(tanfovx), float(tanfovy)
    f.kernel_size, f.scale_modifier = float(kernel_size), float(scale_modifier)
    f.sh_degree, f.sg_degree = int(sh_degree), int(sg_degree)
    f.prefiltered, f.require_depth, f.debug = int(bool(prefiltered)), int(bool(require_depth)), int(bool(debug))
    f.viewmatrix, f.projmatrix, f.campos, f.background = (_abi.ptr(t) for t in (view, proj, cam, bgt))
    return f


def _gaussians(means3D, opacities, scales, rotations, sh=None, sg_axis=None, sg_sharpness=None, sg_color=None,
               colors_precomp=None):
    g = _abi.GgsGaussians()
    g.count = means3D.shape[0]
    g.means3D, g.opacities, g.scales, g.rotations = (_abi.ptr(t) for t in (means3D, opacities, scales, rotations))
    g.shs = _abi.ptr(sh)
    g.sh_coeffs = sh.shape[1] if sh is not None else 0
    has_sg = sg_axis is not None and sg_sharpness is not None and sg_color is not None
    g.sg_axis = _abi.ptr(sg_axis) if has_sg else None
    g.sg_sharpness = _abi.ptr(sg_sharpness) if has_sg else None
    g.sg_color = _abi.ptr(sg_color) if has_sg else None
    g.sg_lobes = sg_color.shape[1] if has_sg else 0
    g.colors_precomp = _abi.ptr(colors_precomp)
    return g


def _call(fn, arena, *args):
    rc = fn(*args)
    arena.done()
    if rc < 0:
        if arena.error is not None:
            raise arena.error
        raise RuntimeError(_abi.last_error())
    return rc


def _check_means(means3D):
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # AT_ERROR, rasterize_points.cu:66-68


def _no_cov(cov3D_precomp):
    if cov3D_precomp is not None and cov3D_precomp.numel():
        raise NotImplementedError("cov3D_precomp is not supported (the reference kernel cannot consume it either, "
                                  "render_forward.cu:344-345)")


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, cov3D_precomp, sh, sg_axis,
                        sg_sharpness, sg_color, sh_degree, sg_degree, scale_modifier, viewmatrix, projmatrix, tan_fovx,
                        tan_fovy, kernel_size, image_height, image_width, campos, prefiltered, require_depth, debug):
    _check_means(means3D)
    _no_cov(cov3D_precomp)
    device = means3D.device
    P, H, W = means3D.shape[0], int(image_height), int(image_width)
    opts = dict(dtype=torch.float32, device=device)
    if P == 0:  # rasterize_points.cu:95: nothing runs, zero-initialised outputs
        z = lambda c: torch.zeros(c, H, W, **opts)  # noqa: E731
        e = _empty(device)
        return 0, z(3), z(1), z(3), z(1), torch.zeros(0, dtype=torch.int32, device=device), e, e, e, e
    m3, op, sc, ro, sh_, sa, ss, sgc, cp = (_f32(t, device) for t in
                                            (means3D, opacity, scales, rotations, sh, sg_axis, sg_sharpness, sg_color,
                                             colors))
    keep = []
    with torch.cuda.device(device):
        frame = _frame(device, keep, W, H, tan_fovx, tan_fovy, kernel_size, scale_modifier, viewmatrix, projmatrix,
                       campos, background, sh_degree, sg_degree, prefiltered, require_depth, debug)
        g = _gaussians(m3, op, sc, ro, sh_, sa, ss, sgc, cp)
        color, mdepth, alpha, normal = (torch.empty(c, H, W, **opts) for c in (3, 1, 1, 3))
        radii = torch.empty(P, dtype=torch.int32, device=device)
        out = _abi.GgsForwardOutputs(_abi.ptr(color), _abi.ptr(alpha), _abi.ptr(normal), _abi.ptr(mdepth),
                                     _abi.ptr(radii))
        state = _abi.GgsForwardState()
        arena = _abi.Arena(device)
        n = _call(_lib.ggs_rasterize_forward, arena, C.byref(frame), C.byref(g), arena.fn, None, C.byref(out),
                  C.byref(state), _abi.stream_handle(device))
    b = arena.buffers
    e = _empty(device)
    return (int(n), color, alpha, normal, mdepth, radii, b.get(_abi.SLOT_GEOMETRY, e), b.get(_abi.SLOT_BINNING, e),
            b.get(_abi.SLOT_IMAGE, e), b.get(_abi.SLOT_TILE, e))


def rasterize_gaussians_backward(background, means3D, colors, opacities, scales, rotations, cov3D_precomp, sh, sg_axis,
                                 sg_sharpness, sg_color, sh_degree, sg_degree, scale_modifier, viewmatrix, projmatrix,
                                 tan_fovx, tan_fovy, kernel_size, dL_dout_color, dL_dout_mdepth, dL_dout_alpha,
                                 dL_dout_normal, alphas, normalmap, mdepth, campos, radii, geomBuffer, R, binningBuffer,
                                 imageBuffer, tileBuffer, require_depth, debug):
    _check_means(means3D)
    device = means3D.device
    P = means3D.shape[0]
    H, W = alphas.shape[-2], alphas.shape[-1]
    opts = dict(dtype=torch.float32, device=device)
    m3, op, sc, ro, sh_, sa, ss, sgc, cp = (_f32(t, device) for t in
                                            (means3D, opacities, scales, rotations, sh, sg_axis, sg_sharpness, sg_color,
                                             colors))
    SHM = sh_.shape[1] if sh_ is not None else 0
    SGM = sgc.shape[1] if (sgc is not None and sa is not None and ss is not None) else 0
    # the eleven gradient tensors of rasterize_points.cu:186-205 (every row is written by the kernels)
    d_means3D, d_means2D, d_colors = (torch.empty(P, 3, **opts) for _ in range(3))
    d_opacity = torch.empty(P, 1, **opts)
    d_cov3D = torch.zeros(P, 6, **opts)
    d_sh = torch.zeros(P, SHM, 3, **opts) if SHM == 0 else torch.empty(P, SHM, 3, **opts)
    d_sg_axis, d_sg_color = torch.empty(P, SGM, 3, **opts), torch.empty(P, SGM, 3, **opts)
    d_sg_sharp = torch.empty(P, SGM, **opts)
    d_scales, d_rot = torch.empty(P, 3, **opts), torch.empty(P, 4, **opts)
    ret = (d_means2D, d_colors, d_opacity, d_means3D, d_cov3D, d_sh, d_sg_axis, d_sg_sharp, d_sg_color, d_scales, d_rot)
    if P == 0:
        return ret
    gc, gd, ga, gn = (_f32(t, device) for t in (dL_dout_color, dL_dout_mdepth, dL_dout_alpha, dL_dout_normal))
    keep = []
    with torch.cuda.device(device):
        frame = _frame(device, keep, W, H, tan_fovx, tan_fovy, kernel_size, scale_modifier, viewmatrix, projmatrix,
                       campos, background, sh_degree, sg_degree, False, require_depth, debug)
        g = _gaussians(m3, op, sc, ro, sh_, sa, ss, sgc, cp)
        state = _abi.GgsForwardState(_abi.ptr(geomBuffer), _abi.ptr(binningBuffer), _abi.ptr(imageBuffer),
                                     _abi.ptr(tileBuffer), int(R))
        bi = _abi.GgsBackwardInputs(_abi.ptr(alphas), _abi.ptr(normalmap), _abi.ptr(mdepth), _abi.ptr(radii),
                                    _abi.ptr(gc), _abi.ptr(gd), _abi.ptr(ga), _abi.ptr(gn))
        grads = _abi.GgsGradients(_abi.ptr(d_means3D), _abi.ptr(d_means2D), _abi.ptr(d_colors), _abi.ptr(d_opacity),
                                  _abi.ptr(d_scales), _abi.ptr(d_rot), _abi.ptr(d_sh), _abi.ptr(d_sg_axis),
                                  _abi.ptr(d_sg_sharp), _abi.ptr(d_sg_color), None)
        arena = _abi.Arena(device)
        _call(_lib.ggs_rasterize_backward, arena, C.byref(frame), C.byref(g), C.byref(state), C.byref(bi), arena.fn,
              None, C.byref(grads), _abi.stream_handle(device))
    return ret


def mark_visible(means3D, viewmatrix, projmatrix):
    device = means3D.device
    pos = _f32(means3D, device)
    P = means3D.shape[0]
    present = torch.zeros(P, dtype=torch.bool, device=device)
    if P:
        view, proj = _f32(viewmatrix, device), _f32(projmatrix, device)
        with torch.cuda.device(device):
            _abi.check(_lib.ggs_mark_visible(_abi.ptr(pos), P, _abi.ptr(view), _abi.ptr(proj), _abi.ptr(present),
                                             _abi.stream_handle(device)), "ggs_mark_visible")
    return present


def _query_inputs(points3D, means3D, opacities, scales, rotations, cov3D_precomp):
    _check_means(means3D)
    _no_cov(cov3D_precomp)
    device = means3D.device
    pts = _f32(points3D, device)
    return device, pts, tuple(_f32(t, device) for t in (means3D, opacities, scales, rotations))


def integrate_gaussians_to_points(points3D, means3D, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                                  view2gaussian_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, kernel_size,
                                  image_height, image_width, campos, prefiltered, debug):
    if points3D.dim() != 2 or points3D.shape[1] != 3:
        raise RuntimeError("points3D must have dimensions (num_points, 3)")  # rasterize_points.cu:299-304
    device, pts, (m3, op, sc, ro) = _query_inputs(points3D, means3D, opacity, scales, rotations, cov3D_precomp)
    PN = points3D.shape[0]
    transmittance = torch.zeros(PN, dtype=torch.float32, device=device)
    inside = torch.zeros(PN, dtype=torch.bool, device=device)
    n = 0
    if PN and m3 is not None:
        keep = []
        with torch.cuda.device(device):
            frame = _frame(device, keep, image_width, image_height, tan_fovx, tan_fovy, kernel_size, scale_modifier,
                           viewmatrix, projmatrix, campos, prefiltered=prefiltered, debug=debug)
            g = _gaussians(m3, op, sc, ro)
            arena = _abi.Arena(device)
            n = _call(_lib.ggs_integrate_points, arena, C.byref(frame), C.byref(g), _abi.ptr(pts), PN, arena.fn, None,
                      _abi.ptr(transmittance), _abi.ptr(inside), _abi.stream_handle(device))
    return int(n), transmittance, inside


def evaluate_sdf_from_signle_view(points3D, means3D, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                                  view2gaussian_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, kernel_size,
                                  image_height, image_width, campos, prefiltered, debug):
    if points3D.dim() != 2 or points3D.shape[1] != 3:
        raise RuntimeError("points3D must have dimensions (num_points, 3)")
    device, pts, (m3, op, sc, ro) = _query_inputs(points3D, means3D, opacity, scales, rotations, cov3D_precomp)
    PN = points3D.shape[0]
    depth = torch.zeros(PN, dtype=torch.float32, device=device)
    sdf = torch.zeros(PN, dtype=torch.float32, device=device)
    inside = torch.zeros(PN, dtype=torch.bool, device=device)
    n = 0
    if PN and m3 is not None:
        keep = []
        with torch.cuda.device(device):
            frame = _frame(device, keep, image_width, image_height, tan_fovx, tan_fovy, kernel_size, scale_modifier,
                           viewmatrix, projmatrix, campos, prefiltered=prefiltered, debug=debug)
            g = _gaussians(m3, op, sc, ro)
            arena = _abi.Arena(device)
            n = _call(_lib.ggs_evaluate_sdf, arena, C.byref(frame), C.byref(g), _abi.ptr(pts), PN, arena.fn, None,
                      _abi.ptr(depth), _abi.ptr(sdf), _abi.ptr(inside), _abi.stream_handle(device))
    return int(n), depth, sdf, inside


def sample_rasterized_depth(points3D, means3D, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                            projmatrix, tan_fovx, tan_fovy, kernel_size, image_height, image_width, campos, prefiltered,
                            debug):
    if points3D.dim() < 1 or points3D.shape[-1] != 3:
        raise RuntimeError("points3D must have shape (..., 3) with last dimension == 3")
    device, pts, (m3, op, sc, ro) = _query_inputs(points3D, means3D, opacity, scales, rotations, cov3D_precomp)
    pts = pts.reshape(-1, 3) if pts is not None else None
    PN = points3D.numel() // 3
    out = torch.zeros(PN, 3, dtype=torch.float32, device=device)
    inside = torch.zeros(PN, dtype=torch.bool, device=device)
    e = _empty(device)
    bufs = {}
    n = 0
    if PN and m3 is not None:
        keep = []
        state = _abi.GgsSampleState()
        with torch.cuda.device(device):
            frame = _frame(device, keep, image_width, image_height, tan_fovx, tan_fovy, kernel_size, scale_modifier,
                           viewmatrix, projmatrix, campos, prefiltered=prefiltered, debug=debug)
            g = _gaussians(m3, op, sc, ro)
            arena = _abi.Arena(device)
            n = _call(_lib.ggs_sample_depth_forward, arena, C.byref(frame), C.byref(g), _abi.ptr(pts), PN, arena.fn,
                      None, _abi.ptr(out), _abi.ptr(inside), C.byref(state), _abi.stream_handle(device))
        bufs = arena.buffers
    # (num_rendered, num_points, num_duplicated_tiles, camera_points, inside, geomBuffer, binningBuffer, pointBuffer,
    #  point_binningBuffer, tileBuffer, duplicatedTileBuffer): rasterize_points.cu:556
    return (int(n), PN, 0, out.reshape(points3D.shape), inside.reshape(points3D.shape[:-1]),
            bufs.get(_abi.SLOT_GEOMETRY, e), bufs.get(_abi.SLOT_BINNING, e), bufs.get(_abi.SLOT_POINT, e), e,
            bufs.get(_abi.SLOT_TILE, e), e)


def sample_rasterized_depth_backward(points3D, means3D, opacities, scales, rotations, scale_modifier, cov3D_precomp,
                                     viewmatrix, projmatrix, inside, dL_doutputs, tan_fovx, tan_fovy, kernel_size,
                                     image_height, image_width, campos, geomBuffer, binningBuffer, pointBuffer,
                                     point_binningBuffer, tileBuffer, duplicatedTileBuffer, R, PN, BN, prefiltered, debug):
    device, pts, (m3, op, sc, ro) = _query_inputs(points3D, means3D, opacities, scales, rotations, cov3D_precomp)
    pts = pts.reshape(-1, 3) if pts is not None else None
    P = means3D.shape[0]
    n = points3D.numel() // 3
    opts = dict(dtype=torch.float32, device=device)
    d_means3D, d_means2D, d_colors = (torch.empty(P, 3, **opts) for _ in range(3))
    d_opacity, d_scales, d_rot = torch.empty(P, 1, **opts), torch.empty(P, 3, **opts), torch.empty(P, 4, **opts)
    d_cov3D = torch.zeros(P, 6, **opts)
    d_pts = torch.empty(n, 3, **opts)
    gout = _f32(dL_doutputs, device)
    gout = gout.reshape(-1, 3) if gout is not None else None
    ins = inside.reshape(-1).contiguous()
    keep = []
    with torch.cuda.device(device):
        frame = _frame(device, keep, image_width, image_height, tan_fovx, tan_fovy, kernel_size, scale_modifier,
                       viewmatrix, projmatrix, campos, prefiltered=prefiltered, debug=debug)
        g = _gaussians(m3, op, sc, ro) if m3 is not None else _abi.GgsGaussians()
        state = _abi.GgsSampleState(_abi.ptr(geomBuffer), _abi.ptr(binningBuffer), _abi.ptr(pointBuffer), None,
                                    _abi.ptr(tileBuffer), None, int(R), int(PN), 0)
        grads = _abi.GgsGradients(_abi.ptr(d_means3D), _abi.ptr(d_means2D), _abi.ptr(d_colors), _abi.ptr(d_opacity),
                                  _abi.ptr(d_scales), _abi.ptr(d_rot))
        arena = _abi.Arena(device)
        _call(_lib.ggs_sample_depth_backward, arena, C.byref(frame), C.byref(g), _abi.ptr(pts), n, C.byref(state),
              _abi.ptr(ins), _abi.ptr(gout), arena.fn, None, C.byref(grads), _abi.ptr(d_pts),
              _abi.stream_handle(device))
    # (dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dscales, dL_drotations, dL_dpoints3D): rasterize_points.cu:633
    return d_opacity, d_means3D, d_cov3D, d_scales, d_rot, d_pts.reshape(points3D.shape)
