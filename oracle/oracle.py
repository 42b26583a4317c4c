"""ctypes front-end of the C oracle (oracle/ggs_oracle.c).  TEST INFRASTRUCTURE ONLY.

Mirrors the reference call surface: ``forward`` ~ RasterizeGaussiansCUDA
(DGR/rasterize_points.cu:39-139), ``backward`` ~ RasterizeGaussiansBackwardCUDA (:141-258),
``integrate`` ~ IntegrateGaussiansToPointsCUDA (:279-366).  Inputs/outputs are CPU torch tensors.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libggs_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("ggs_oracle.c", "ggs_oracle_backward.inc", "ggs_oracle_points.inc",
                                             "ggs_oracle_ncc.inc")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "all"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.ggs_oracle_create.restype = C.c_void_p
        L.ggs_oracle_destroy.argtypes = [C.c_void_p]
        L.ggs_oracle_forward.restype = C.c_int64
        L.ggs_oracle_integrate.restype = C.c_int64
        L.ggs_oracle_num_rendered.restype = C.c_int64
        L.ggs_oracle_num_rendered.argtypes = [C.c_void_p]
        L.ggs_oracle_higher_msb.restype = C.c_uint32
        L.ggs_oracle_higher_msb.argtypes = [C.c_uint32]
        for n in ("tiles_touched", "point_offsets", "sorted_keys", "point_list", "ranges", "n_contrib",
                  "max_contributor", "depths", "means2D", "conic_opacity", "rgb", "ray_planes", "normals", "clamped"):
            f = getattr(L, "ggs_oracle_" + n)
            f.restype = C.c_void_p
            f.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _f(t):
    """float32 contiguous CPU tensor -> (keepalive, pointer); empty/None -> NULL."""
    if t is None or t.numel() == 0:
        return None, None
    t = t.detach().to("cpu", torch.float32).contiguous()
    return t, C.c_void_p(t.data_ptr())


def _ptr(t):
    return C.c_void_p(t.data_ptr())


class Oracle:
    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p(self.L.ggs_oracle_create())
        self._meta = None

    def __del__(self):
        try:
            self.L.ggs_oracle_destroy(self.h)
        except Exception:
            pass

    # ------------------------------------------------------------------ forward
    def forward(self, *, bg, means3D, opacities, scales, rotations, shs=None, sg_axis=None, sg_sharpness=None,
                sg_color=None, colors_precomp=None, sh_degree=0, sg_degree=0, scale_modifier=1.0, viewmatrix=None,
                projmatrix=None, campos=None, tanfovx=1.0, tanfovy=1.0, kernel_size=0.0, image_height=0,
                image_width=0, require_depth=True):
        P, H, W = means3D.shape[0], int(image_height), int(image_width)
        SHM = shs.shape[1] if shs is not None and shs.numel() else 0
        SGM = sg_color.shape[1] if sg_color is not None and sg_color.numel() else 0
        keep = [_f(x) for x in (bg, means3D, colors_precomp, opacities, scales, rotations, shs, sg_axis, sg_sharpness,
                                sg_color, viewmatrix, projmatrix, campos)]
        (bg_, me, cp, op, sc, ro, sh, sa, ss, sgc, vm, pm, cpos) = [k[1] for k in keep]
        color = torch.zeros(3, H, W)
        mdepth = torch.zeros(1, H, W)
        alpha = torch.zeros(1, H, W)
        normal = torch.zeros(3, H, W)
        radii = torch.zeros(P, dtype=torch.int32)
        D = self.L.ggs_oracle_forward(
            self.h, C.c_int(P), C.c_int(sh_degree), C.c_int(SHM), C.c_int(sg_degree), C.c_int(SGM), bg_, C.c_int(W),
            C.c_int(H), me, cp, op, sc, ro, sh, sa, ss, sgc, C.c_float(scale_modifier), vm, pm, cpos,
            C.c_float(tanfovx), C.c_float(tanfovy), C.c_float(kernel_size), C.c_int(1 if require_depth else 0),
            _ptr(color), _ptr(mdepth), _ptr(alpha), _ptr(normal), _ptr(radii))
        self._meta = dict(P=P, H=H, W=W, D=int(D), SHM=SHM, SGM=SGM, keep=keep, alpha=alpha, normal=normal,
                          mdepth=mdepth)
        return dict(num_rendered=int(D), color=color, radii=radii, mdepth=mdepth, alpha=alpha, normal=normal)

    def _arr(self, name, dtype, n):
        p = getattr(self.L, "ggs_oracle_" + name)(self.h)
        if not p or n == 0:
            return np.zeros(0, dtype=dtype)
        buf = (C.c_char * (np.dtype(dtype).itemsize * n)).from_address(p)
        return np.frombuffer(buf, dtype=dtype).copy()

    def intermediates(self):
        m = self._meta
        P, D, W, H = m["P"], m["D"], m["W"], m["H"]
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        return dict(
            tiles_touched=self._arr("tiles_touched", np.uint32, P),
            point_offsets=self._arr("point_offsets", np.uint32, P),
            sorted_keys=self._arr("sorted_keys", np.uint64, D),
            point_list=self._arr("point_list", np.uint32, D),
            ranges=self._arr("ranges", np.uint32, 2 * tiles).reshape(tiles, 2),
            n_contrib=self._arr("n_contrib", np.uint32, W * H).reshape(H, W),
            max_contributor=self._arr("max_contributor", np.uint32, tiles),
            depths=self._arr("depths", np.float32, P),
            means2D=self._arr("means2D", np.float32, 2 * P).reshape(P, 2),
            conic_opacity=self._arr("conic_opacity", np.float32, 4 * P).reshape(P, 4),
            rgb=self._arr("rgb", np.float32, 3 * P).reshape(P, 3),
            ray_planes=self._arr("ray_planes", np.float32, 4 * P).reshape(P, 4),
            normals=self._arr("normals", np.float32, 3 * P).reshape(P, 3),
            clamped=self._arr("clamped", np.uint8, 3 * P).reshape(P, 3),
        )

    # ----------------------------------------------------------------- backward
    def backward(self, grad_color, grad_mdepth, grad_alpha, grad_normal):
        m = self._meta
        P, SHM, SGM = m["P"], m["SHM"], m["SGM"]
        (bg_, me, cp, op, sc, ro, sh, sa, ss, sgc, vm, pm, cpos) = [k[1] for k in m["keep"]]
        gk = [_f(x) for x in (grad_color, grad_mdepth, grad_alpha, grad_normal)]
        out = dict(
            means3D=torch.zeros(P, 3), means2D=torch.zeros(P, 3), colors=torch.zeros(P, 3), opacities=torch.zeros(P, 1),
            scales=torch.zeros(P, 3), rotations=torch.zeros(P, 4), shs=torch.zeros(P, SHM, 3),
            sg_axis=torch.zeros(P, SGM, 3), sg_sharpness=torch.zeros(P, SGM), sg_color=torch.zeros(P, SGM, 3))
        self.L.ggs_oracle_backward(
            self.h, me, cp, op, sc, ro, sh, sa, ss, sgc, _ptr(m["alpha"]), _ptr(m["normal"]), _ptr(m["mdepth"]),
            gk[0][1], gk[1][1], gk[2][1], gk[3][1], _ptr(out["means3D"]), _ptr(out["means2D"]), _ptr(out["colors"]),
            _ptr(out["opacities"]), _ptr(out["scales"]), _ptr(out["rotations"]), _ptr(out["shs"]),
            _ptr(out["sg_axis"]), _ptr(out["sg_sharpness"]), _ptr(out["sg_color"]))
        return out

    # ---------------------------------------------------------------- integrate
    def integrate(self, *, points3D, means3D, opacities, scales, rotations, scale_modifier=1.0, viewmatrix=None,
                  projmatrix=None, campos=None, tanfovx=1.0, tanfovy=1.0, kernel_size=0.0, image_height=0,
                  image_width=0):
        PN, P = points3D.shape[0], means3D.shape[0]
        keep = [_f(x) for x in (points3D, means3D, opacities, scales, rotations, viewmatrix, projmatrix, campos)]
        pts, me, op, sc, ro, vm, pm, cpos = [k[1] for k in keep]
        T = torch.zeros(PN)
        inside = torch.zeros(PN, dtype=torch.uint8)
        D = self.L.ggs_oracle_integrate(
            self.h, C.c_int(PN), C.c_int(P), C.c_int(int(image_width)), C.c_int(int(image_height)), pts, me, op, sc,
            C.c_float(scale_modifier), ro, vm, pm, cpos, C.c_float(tanfovx), C.c_float(tanfovy),
            C.c_float(kernel_size), _ptr(T), _ptr(inside))
        return dict(num_rendered=int(D), alpha=1 - T, transmittance=T, inside=inside.bool())


    # ---------------------------------------------------------------- sample_depth / evaluate_sdf (forward)
    def _sample(self, sdf, *, points3D, means3D, opacities, scales, rotations, scale_modifier=1.0, viewmatrix=None,
                projmatrix=None, campos=None, tanfovx=1.0, tanfovy=1.0, kernel_size=0.0, image_height=0,
                image_width=0):
        PN, P = points3D.shape[0], means3D.shape[0]
        keep = [_f(x) for x in (points3D, means3D, opacities, scales, rotations, viewmatrix, projmatrix, campos)]
        pts, me, op, sc, ro, vm, pm, cpos = [k[1] for k in keep]
        out3 = torch.zeros(PN, 3)
        depth = torch.zeros(PN)
        sdfv = torch.zeros(PN)
        inside = torch.zeros(PN, dtype=torch.uint8)
        self.L.ggs_oracle_sample_depth.restype = C.c_int64
        D = self.L.ggs_oracle_sample_depth(
            self.h, C.c_int(PN), C.c_int(P), C.c_int(int(image_width)), C.c_int(int(image_height)), pts, me, op, sc,
            C.c_float(scale_modifier), ro, vm, pm, cpos, C.c_float(tanfovx), C.c_float(tanfovy),
            C.c_float(kernel_size), C.c_int(1 if sdf else 0), _ptr(out3), _ptr(depth), _ptr(sdfv), _ptr(inside))
        return dict(num_rendered=int(D), camera_points=out3, depth=depth, sdf=sdfv, inside=inside.bool())

    def sample_depth(self, **kw):
        """Forward of GaussianRasterizer.sample_depth for [PN,3] query points (the reference passes kernel_size 0)."""
        return self._sample(False, **kw)

    def evaluate_sdf(self, **kw):
        return self._sample(True, **kw)

    def sample_depth_backward(self, dL_dcamera_points, *, points3D, means3D, opacities, scales, rotations,
                              scale_modifier=1.0, viewmatrix=None, projmatrix=None, campos=None, tanfovx=1.0,
                              tanfovy=1.0, kernel_size=0.0, image_height=0, image_width=0):
        """Backward of GaussianRasterizer.sample_depth: gradients of sum(camera_points * dL_dcamera_points).  The
        forward is recomputed with kernel size 0 and the per-Gaussian chain uses `kernel_size`, as the reference's
        autograd node does (__init__.py:513,606)."""
        PN, P = points3D.shape[0], means3D.shape[0]
        keep = [_f(x) for x in (points3D, means3D, opacities, scales, rotations, viewmatrix, projmatrix, campos,
                                dL_dcamera_points)]
        pts, me, op, sc, ro, vm, pm, cpos, up = [k[1] for k in keep]
        g = dict(points3D=torch.zeros(PN, 3), means3D=torch.zeros(P, 3), opacities=torch.zeros(P, 1),
                 scales=torch.zeros(P, 3), rotations=torch.zeros(P, 4))
        self.L.ggs_oracle_sample_depth_backward.restype = None
        self.L.ggs_oracle_sample_depth_backward(
            self.h, C.c_int(PN), C.c_int(P), C.c_int(int(image_width)), C.c_int(int(image_height)), pts, me, op, sc,
            C.c_float(scale_modifier), ro, vm, pm, cpos, C.c_float(tanfovx), C.c_float(tanfovy), C.c_float(0.0),
            C.c_float(kernel_size), up, _ptr(g["points3D"]), _ptr(g["means3D"]), _ptr(g["opacities"]),
            _ptr(g["scales"]), _ptr(g["rotations"]))
        return g


def warp_patch_ncc(depths, normals, uvs, R, T, image_r, image_n, fx_r, fy_r, cx_r, cy_r, fx_n, fy_n, cx_n, cy_n,
                   debug=False):
    """The reference's `_C.warp_patch_ncc` (submodules/warp-patch-ncc/warp_patch_ncc.cu:5-57) on CPU tensors:
    (ncc [P], grad_depths [P], grad_normals [P,3], valid [P] bool).  R is read as the reference reads it: 9 floats
    in memory order, columns of the rotation."""
    L = lib()
    P = int(depths.shape[0])
    keep = [_f(x) for x in (depths, normals, R, T, image_r, image_n)]
    d, n, r, t, ir, im = [k[1] for k in keep]
    uv = uvs.detach().to("cpu", torch.int32).contiguous()
    ncc, gd, gn = torch.zeros(P), torch.zeros(P), torch.zeros(P, 3)
    valid = torch.zeros(P, dtype=torch.uint8)
    L.ggs_oracle_warp_patch_ncc.restype = None
    if P:
        L.ggs_oracle_warp_patch_ncc(
            C.c_int(P), d, n, _ptr(uv), r, t, ir, im, *[C.c_float(float(v)) for v in (fx_r, fy_r, cx_r, cy_r, fx_n, fy_n,
                                                                                    cx_n, cy_n)],
            C.c_int(int(image_r.shape[0])), C.c_int(int(image_r.shape[1])), C.c_int(int(image_n.shape[0])),
            C.c_int(int(image_n.shape[1])), _ptr(ncc), _ptr(gd), _ptr(gn), _ptr(valid))
    return ncc, gd, gn, valid.bool()


def scene_kwargs(sc):
    """Keyword arguments of Oracle.forward for a scenes.Scene."""
    return dict(bg=sc.bg, means3D=sc.means3D, opacities=sc.opacities, scales=sc.scales, rotations=sc.rotations,
                shs=sc.shs, sg_axis=sc.sg_axis, sg_sharpness=sc.sg_sharpness, sg_color=sc.sg_color,
                sh_degree=sc.sh_degree, sg_degree=sc.sg_degree, viewmatrix=sc.cam.viewmatrix,
                projmatrix=sc.cam.projmatrix, campos=sc.cam.campos, tanfovx=sc.cam.tanfovx, tanfovy=sc.cam.tanfovy,
                kernel_size=sc.kernel_size, image_height=sc.cam.H, image_width=sc.cam.W)
