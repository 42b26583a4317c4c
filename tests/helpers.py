"""Shared test plumbing: run the product (C ABI via the drop-in package), the compiled reference
(Oracle A, GPU only) and the C restatement (Oracle B, CPU) on a `scenes.Scene`, and compare.

Tolerances (BASELINE.json north_star):
    integers (radii, tiles_touched, sorted keys, point list, tile ranges, n_contrib) bit-exact;
    forward images  max |a-b| / max(|b|, 1e-3 * max|b|) <= 1e-5;
    gradients       ||a-b||_inf / ||b||_inf <= 1e-4 per tensor.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

IMG_TOL = 1e-5
GRAD_TOL = 1e-4

GRAD_NAMES = ("means3D", "means2D", "shs", "opacities", "scales", "rotations", "sg_axis", "sg_sharpness", "sg_color")


def settings_for(mod, sc, device, require_depth=True, debug=False):
    c = sc.cam
    return mod.GaussianRasterizationSettings(
        image_height=c.H, image_width=c.W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, kernel_size=sc.kernel_size,
        bg=sc.bg.to(device), scale_modifier=1.0, viewmatrix=c.viewmatrix.to(device),
        projmatrix=c.projmatrix.to(device), sh_degree=sc.sh_degree, sg_degree=sc.sg_degree,
        campos=c.campos.to(device), prefiltered=False, require_depth=require_depth, debug=debug)


def leaf_inputs(sc, device):
    d = {}
    d["means3D"] = sc.means3D.to(device).requires_grad_(True)
    d["means2D"] = torch.zeros(sc.P, 3, device=device, requires_grad=True)
    d["opacities"] = sc.opacities.to(device).requires_grad_(True)
    d["scales"] = sc.scales.to(device).requires_grad_(True)
    d["rotations"] = sc.rotations.to(device).requires_grad_(True)
    d["shs"] = sc.shs.to(device).requires_grad_(True)
    d["sg_axis"] = sc.sg_axis.to(device).requires_grad_(True)
    d["sg_sharpness"] = sc.sg_sharpness.to(device).requires_grad_(True)
    d["sg_color"] = sc.sg_color.to(device).requires_grad_(True)
    return d


def run_module(mod, sc, device, require_depth=True, backward=True, mask_depth_grad=False):
    """forward (+ backward with the scene's fixed upstream gradients) through a module exposing the
    reference's Python surface.  Returns (outputs dict, grads dict)."""
    rs = settings_for(mod, sc, device, require_depth)
    x = leaf_inputs(sc, device)
    rast = mod.GaussianRasterizer(rs)
    color, radii, mdepth, alpha, normal = rast(
        means3D=x["means3D"], means2D=x["means2D"], opacities=x["opacities"], shs=x["shs"], sg_axis=x["sg_axis"],
        sg_sharpness=x["sg_sharpness"], sg_color=x["sg_color"], colors_precomp=None, scales=x["scales"],
        rotations=x["rotations"], cov3D_precomp=None)
    out = dict(color=color.detach(), radii=radii.detach(), mdepth=mdepth.detach(), alpha=alpha.detach(),
               normal=normal.detach())
    grads = {}
    if backward:
        g = {k: v.to(device) for k, v in sc.grads.items()}
        gd = g["mdepth"]
        if mask_depth_grad:
            gd = gd * (mdepth.detach() > 0)
        loss = (color * g["color"]).sum() + (alpha * g["alpha"]).sum()
        if require_depth:
            loss = loss + (mdepth * gd).sum() + (normal * g["normal"]).sum()
        loss.backward()
        for k in GRAD_NAMES:
            t = x[k]
            grads[k] = t.grad.detach().clone() if t.grad is not None else torch.zeros_like(t)
    torch.cuda.synchronize(device)
    return out, grads


def rasterize_with(mod, sc, device, scales, opacities, rotations, require_depth=True):
    """Rasterizer forward with caller-provided (differentiable) scales / opacities / rotations; the other inputs are
    the scene's.  Returns (color, mdepth, alpha, normal) attached to the graph."""
    rs = settings_for(mod, sc, device, require_depth)
    x = leaf_inputs(sc, device)
    rast = mod.GaussianRasterizer(rs)
    color, _radii, mdepth, alpha, normal = rast(
        means3D=x["means3D"], means2D=x["means2D"], opacities=opacities, shs=x["shs"], sg_axis=x["sg_axis"],
        sg_sharpness=x["sg_sharpness"], sg_color=x["sg_color"], colors_precomp=None, scales=scales,
        rotations=rotations, cov3D_precomp=None)
    return color, mdepth, alpha, normal


# ---- intermediates -------------------------------------------------------------------------------

def _bytes_view(buf, offset, dtype, count):
    itemsize = torch.empty((), dtype=dtype).element_size()
    return buf[offset:offset + count * itemsize].view(dtype)


def ours_forward_raw(sc, device, require_depth=True, want_ctx=False, capacity_hint=0):
    """Forward through the C ABI keeping the workspaces; returns outputs + bit-exact intermediates (+ the call's frame /
    Gaussians / state structures with want_ctx, for entry points that take a finished forward)."""
    import ctypes as C

    import diff_gaussian_rasterization as ours
    from diff_gaussian_rasterization import _abi

    rs = settings_for(ours, sc, device, require_depth)
    keep = []
    P, H, W = sc.P, sc.cam.H, sc.cam.W
    t = {k: getattr(sc, k).to(device).contiguous() for k in
         ("means3D", "opacities", "scales", "rotations", "shs", "sg_axis", "sg_sharpness", "sg_color")}
    frame = ours._frame(rs, torch.device(device), keep)
    g = ours._gaussians(t["means3D"], t["opacities"], t["scales"], t["rotations"], t["shs"], t["sg_axis"],
                        t["sg_sharpness"], t["sg_color"], None, sc.sg_degree)
    color = torch.empty(3, H, W, device=device)
    mdepth = torch.empty(1, H, W, device=device)
    alpha = torch.empty(1, H, W, device=device)
    normal = torch.empty(3, H, W, device=device)
    radii = torch.empty(P, dtype=torch.int32, device=device)
    out = _abi.GgsForwardOutputs(_abi.ptr(color), _abi.ptr(alpha), _abi.ptr(normal), _abi.ptr(mdepth),
                                 _abi.ptr(radii))
    state = _abi.GgsForwardState()
    state.capacity = int(capacity_hint)  # 0: the forward blocks on the pair count, as the reference does
    arena = _abi.Arena(torch.device(device))
    n = _abi.lib.ggs_rasterize_forward(C.byref(frame), C.byref(g), arena.fn, None, C.byref(out), C.byref(state),
                                       _abi.stream_handle(torch.device(device)))
    _abi.check(n, "ggs_rasterize_forward")
    torch.cuda.synchronize(device)
    D = int(n)
    assert int(state.num_rendered) == D and int(state.capacity) >= D
    # the reference's 64-bit keys are not formed on the hot path: write them into the workspace for the comparisons
    _abi.check(_abi.lib.ggs_materialize_sorted_keys(P, W, H, C.byref(state), _abi.stream_handle(torch.device(device))),
               "ggs_materialize_sorted_keys")
    torch.cuda.synchronize(device)
    L = _abi.describe_layout(P, W, H, int(state.capacity))
    geom, binb, img, tile = (arena.buffers[s] for s in (_abi.SLOT_GEOMETRY, _abi.SLOT_BINNING, _abi.SLOT_IMAGE,
                                                       _abi.SLOT_TILE))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    rec = _bytes_view(geom, L.records, torch.float32, P * 16).view(P, 16)
    inter = dict(
        num_rendered=D,
        capacity=int(state.capacity),
        tiles_touched=_bytes_view(geom, L.tiles_touched, torch.int32, P).clone(),
        point_offsets=_bytes_view(geom, L.point_offsets, torch.int32, P).clone(),  # scan in depth order
        depth_order=_bytes_view(geom, L.depth_order, torch.int32, P).clone(),
        depths=_bytes_view(geom, L.depths, torch.float32, P).clone(),
        clamped=_bytes_view(geom, L.clamped, torch.uint8, P).clone(),
        records=rec.clone(),
        sorted_keys=_bytes_view(binb, L.sorted_keys, torch.int64, D).clone(),
        point_list=_bytes_view(binb, L.point_list, torch.int32, D).clone(),
        ranges=_bytes_view(tile, L.ranges, torch.int32, tiles * 2).view(tiles, 2).clone(),
        max_contributor=_bytes_view(tile, L.max_contributor, torch.int32, tiles).clone(),
        n_contrib=_bytes_view(img, L.n_contrib, torch.int32, H * W).view(H, W).clone(),
        sort_bits=int(L.sort_bits),
    )
    outs = dict(color=color, radii=radii, mdepth=mdepth, alpha=alpha, normal=normal)
    if want_ctx:
        return outs, inter, dict(frame=frame, g=g, state=state, keep=(keep, t, arena))
    return outs, inter


def _align(x, a=128):
    return (x + a - 1) // a * a


def ref_forward_raw(ref, sc, device, require_depth=True):
    """Forward through the reference's pybind entry point; intermediates are cut out of its byte
    buffers by replaying the `obtain` layout (rasterizer_impl.h:20-25, rasterizer_impl.cu:199-281)."""
    P, H, W = sc.P, sc.cam.H, sc.cam.W
    c = sc.cam
    dev = torch.device(device)
    e = torch.Tensor([])
    args = (sc.bg.to(dev), sc.means3D.to(dev), e, sc.opacities.to(dev), sc.scales.to(dev), sc.rotations.to(dev), e,
            sc.shs.to(dev), sc.sg_axis.to(dev), sc.sg_sharpness.to(dev), sc.sg_color.to(dev), sc.sh_degree,
            sc.sg_degree, 1.0, c.viewmatrix.to(dev), c.projmatrix.to(dev), c.tanfovx, c.tanfovy, sc.kernel_size, H, W,
            c.campos.to(dev), False, require_depth, False)
    D, color, alpha, normal, mdepth, radii, geom, binb, img, tile = ref._C.rasterize_gaussians(*args)
    torch.cuda.synchronize(device)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def cut(buf, fields):
        base = buf.data_ptr()
        p = base
        out = {}
        for name, dtype, count in fields:
            p = _align(p)
            if dtype is not None:
                out[name] = _bytes_view(buf, p - base, dtype, count).clone()
                p += count * torch.empty((), dtype=dtype).element_size()
        return out

    f32, i32, u8, i64 = torch.float32, torch.int32, torch.uint8, torch.int64
    g = cut(geom, [("depths", f32, P), ("ray_planes", f32, 4 * P), ("normals", f32, 3 * P), ("clamped", u8, 3 * P),
                   ("internal_radii", i32, P), ("means2D", f32, 2 * P), ("conic_opacity", f32, 4 * P),
                   ("rgb", f32, 3 * P), ("tiles_touched", i32, P)])
    b = cut(binb, [("point_list", i32, D), ("point_list_unsorted", i32, D), ("sorted_keys", i64, D)]) if D else \
        dict(point_list=torch.zeros(0, dtype=i32), sorted_keys=torch.zeros(0, dtype=i64))
    t = cut(tile, [("ranges", i32, 2 * tiles), ("max_contributor", i32, tiles)])
    i = cut(img, [("n_contrib", i32, H * W)])
    rec = torch.cat([g["means2D"].view(P, 2), g["conic_opacity"].view(P, 4), g["ray_planes"].view(P, 4),
                     g["rgb"].view(P, 3), g["normals"].view(P, 3)], 1)
    cl = g["clamped"].view(P, 3)
    inter = dict(num_rendered=int(D), tiles_touched=g["tiles_touched"],
                 point_offsets=torch.cumsum(g["tiles_touched"].to(torch.int64), 0).to(i32), depths=g["depths"],
                 clamped=(cl[:, 0] | (cl[:, 1] << 1) | (cl[:, 2] << 2)).to(u8), records=rec,
                 sorted_keys=b["sorted_keys"], point_list=b["point_list"], ranges=t["ranges"].view(tiles, 2),
                 max_contributor=t["max_contributor"], n_contrib=i["n_contrib"].view(H, W))
    outs = dict(color=color, radii=radii, mdepth=mdepth, alpha=alpha, normal=normal)
    return outs, inter


def check_depth_order(it, radii):
    """Ours only: depth_order is the STABLE ascending order of the visible Gaussians' depth bits (culled ones behind
    them, any order), and point_offsets is the inclusive scan of tiles_touched taken in that order."""
    order = it["depth_order"].to(torch.int64)
    P = order.numel()
    assert torch.equal(torch.sort(order).values, torch.arange(P, device=order.device)), "not a permutation"
    vis = radii.to(order.device) > 0
    V = int(vis.sum())
    assert bool(vis[order[:V]].all()) and not bool(vis[order[V:]].any()), "culled Gaussians must come last"
    bits = it["depths"].view(torch.int32).to(torch.int64)[order[:V]]
    key = bits * P + order[:V]  # stable: ties in depth keep index order
    assert bool((key[1:] > key[:-1]).all()), "depth order is not the stable sort of the depth bits"
    tt = it["tiles_touched"].to(torch.int64)[order]
    assert torch.equal(torch.cumsum(tt, 0).to(torch.int32), it["point_offsets"])


# ---- metrics ---------------------------------------------------------------------------------------

def image_rel_err(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    if b.numel() == 0:
        return 0.0, 0
    eps = 1e-3 * float(b.abs().max()) + 1e-30
    rel = (a - b).abs() / torch.clamp(b.abs(), min=eps)
    return float(rel.max()), int((rel > IMG_TOL).sum())


def grad_rel_err(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    if b.numel() == 0:
        return 0.0
    den = float(b.abs().max())
    if den == 0.0:
        return float((a - b).abs().max())
    return float((a - b).abs().max()) / den


def ours_backward_raw(sc, device, require_depth=True):
    """Forward + backward through the C ABI with the scene's fixed upstream gradients, keeping the backward scratch.
    Returns (grads dict, accumulators double[P,18], ctx) where ctx can be passed to `chain_from_accumulators`."""
    import ctypes as C

    import diff_gaussian_rasterization as ours
    from diff_gaussian_rasterization import _abi

    dev = torch.device(device)
    rs = settings_for(ours, sc, device, require_depth)
    keep = []
    P, Hh, W = sc.P, sc.cam.H, sc.cam.W
    t = {k: getattr(sc, k).to(device).contiguous() for k in
         ("means3D", "opacities", "scales", "rotations", "shs", "sg_axis", "sg_sharpness", "sg_color")}
    frame = ours._frame(rs, dev, keep)
    g = ours._gaussians(t["means3D"], t["opacities"], t["scales"], t["rotations"], t["shs"], t["sg_axis"],
                        t["sg_sharpness"], t["sg_color"], None, sc.sg_degree)
    color = torch.empty(3, Hh, W, device=device)
    mdepth = torch.zeros(1, Hh, W, device=device)
    alpha = torch.empty(1, Hh, W, device=device)
    normal = torch.zeros(3, Hh, W, device=device)
    radii = torch.empty(P, dtype=torch.int32, device=device)
    out = _abi.GgsForwardOutputs(_abi.ptr(color), _abi.ptr(alpha), _abi.ptr(normal), _abi.ptr(mdepth),
                                 _abi.ptr(radii))
    state = _abi.GgsForwardState()
    arena = _abi.Arena(dev)
    stream = _abi.stream_handle(dev)
    n = _abi.lib.ggs_rasterize_forward(C.byref(frame), C.byref(g), arena.fn, None, C.byref(out), C.byref(state), stream)
    _abi.check(n, "ggs_rasterize_forward")
    up = {k: v.to(device).contiguous() for k, v in sc.grads.items()}
    if not require_depth:
        up["mdepth"] = torch.zeros_like(up["mdepth"])
        up["normal"] = torch.zeros_like(up["normal"])
    bi = _abi.GgsBackwardInputs(_abi.ptr(alpha), _abi.ptr(normal), _abi.ptr(mdepth), _abi.ptr(radii),
                                _abi.ptr(up["color"]), _abi.ptr(up["mdepth"]), _abi.ptr(up["alpha"]),
                                _abi.ptr(up["normal"]))

    def alloc_grads():
        SHM, SGM = t["shs"].shape[1], t["sg_color"].shape[1]
        d = dict(means3D=torch.empty(P, 3, device=device), means2D=torch.empty(P, 3, device=device),
                 colors=torch.empty(P, 3, device=device), opacities=torch.empty(P, 1, device=device),
                 scales=torch.empty(P, 3, device=device), rotations=torch.empty(P, 4, device=device),
                 shs=torch.empty(P, SHM, 3, device=device), sg_axis=torch.empty(P, SGM, 3, device=device),
                 sg_sharpness=torch.empty(P, SGM, device=device), sg_color=torch.empty(P, SGM, 3, device=device),
                 rgb=torch.empty(P, 3, device=device))
        st = _abi.GgsGradients(*[_abi.ptr(d[k]) for k in ("means3D", "means2D", "colors", "opacities", "scales",
                                                           "rotations", "shs", "sg_axis", "sg_sharpness", "sg_color",
                                                           "rgb")])
        return d, st

    grads, gst = alloc_grads()
    barena = _abi.Arena(dev)
    rc = _abi.lib.ggs_rasterize_backward(C.byref(frame), C.byref(g), C.byref(state), C.byref(bi), barena.fn, None,
                                         C.byref(gst), stream)
    _abi.check(rc, "ggs_rasterize_backward")
    torch.cuda.synchronize(device)
    acc = barena.buffers[_abi.SLOT_BACKWARD][:P * 18 * 8].view(torch.float64).view(P, 18).clone()
    ctx = dict(frame=frame, g=g, state=state, radii=radii, alloc_grads=alloc_grads, stream=stream,
               keep=(keep, t, arena, barena, color, mdepth, alpha, normal, up))
    return grads, acc, ctx


def chain_from_accumulators(ctx, acc):
    """Per-Gaussian chain rule on a caller-provided accumulator array (ggs_backward_from_accumulators)."""
    import ctypes as C

    from diff_gaussian_rasterization import _abi

    grads, gst = ctx["alloc_grads"]()
    acc = acc.contiguous()
    rc = _abi.lib.ggs_backward_from_accumulators(C.byref(ctx["frame"]), C.byref(ctx["g"]), C.byref(ctx["state"]),
                                                 _abi.ptr(ctx["radii"]), _abi.ptr(acc), C.byref(gst), ctx["stream"])
    _abi.check(rc, "ggs_backward_from_accumulators")
    torch.cuda.synchronize()
    return grads


def grad_cancellation(mod, sc, device, require_depth=True):
    """Conditioning of the per-Gaussian chain rule with respect to its 17 accumulated sums.

    The chain (render_backward.cu:249-713) is linear in the accumulators a_j, so grad = sum_j chain(a_j e_j) and
    A = sum_j |chain(a_j e_j)| is the magnitude a relative perturbation of the a_j is multiplied with.  The reference
    accumulates the a_j with float atomics (relative error of a few eps32 each, in a timing-dependent order); on a
    thin, strongly anisotropic splat the chain multiplies that by (s_max / s_min)^2 and the result moves by more
    than 1e-4 although both implementations are correct.  Returns A per gradient tensor (CPU, double)."""
    _, acc, ctx = ours_backward_raw(sc, device, require_depth)
    A = None
    for j in range(17):
        a = torch.zeros_like(acc)
        a[:, j] = acc[:, j]
        gr = chain_from_accumulators(ctx, a)
        if A is None:
            A = {n: v.abs().double().cpu() for n, v in gr.items()}
        else:
            for n, v in gr.items():
                A[n] += v.abs().double().cpu()
    return A


def grad_excess(a, b, cancel, tol=None, k_eps=32.0):
    """max over elements of |a-b| - (tol * max|b| + k_eps * eps32 * cancel): <= 0 means within tolerance.
    k_eps = 32: a float-atomic sum of a few dozen partial sums is off by a few eps32, and the chain's own float
    arithmetic (identical in both implementations up to instruction order) adds a few more."""
    tol = GRAD_TOL if tol is None else tol
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    if b.numel() == 0:
        return 0.0
    bound = tol * float(b.abs().max()) + k_eps * 5.96e-8 * cancel
    return float(((a - b).abs() - bound).max())


def int_mismatch(a, b):
    a = a.detach().cpu()
    b = b.detach().cpu()
    if a.shape != b.shape:
        return -1
    return int((a != b).sum())


def to_numpy(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
