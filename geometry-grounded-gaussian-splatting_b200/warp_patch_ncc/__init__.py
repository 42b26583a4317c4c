"""Drop-in for the reference's `warp_patch_ncc` module (submodules/warp-patch-ncc/warp_patch_ncc/__init__.py:1-84):
same function, argument order, outputs and autograd behaviour, on top of libggs_patch_ncc.so
(C ABI: include/ggs_patch_ncc.h).  PyTorch owns the memory and the stream; the kernel is ours.  There is no
fallback: if the library is missing or does not load, importing this module fails.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import NamedTuple

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# GGS_PATCH_NCC_LIB: a differently built copy of the same library (A/B timing of kernel variants); no fallback either way
LIB_PATH = os.environ.get("GGS_PATCH_NCC_LIB") or os.path.join(_HERE, "libggs_patch_ncc.so")
EXPORTS = ("ggs_ncc_abi_version", "ggs_ncc_last_error", "ggs_ncc_warp_patch")
ABI_VERSION = 1


class GgsNccPair(C.Structure):
    _fields_ = [("R", C.c_void_p), ("T", C.c_void_p)] + [(n, C.c_float) for n in (
        "fx_r", "fy_r", "cx_r", "cy_r", "fx_n", "fy_n", "cx_n", "cy_n")] + [(n, C.c_int32) for n in (
        "height_r", "width_r", "height_n", "width_n")]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C geometry-grounded-gaussian-splatting_b200/csrc` "
                          "(or __graft_entry__.build()); there is no CPU or PyTorch fallback for this kernel")
    lib = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise ImportError(f"{LIB_PATH} does not export {name}")
    lib.ggs_ncc_abi_version.restype = C.c_int
    lib.ggs_ncc_last_error.restype = C.c_char_p
    lib.ggs_ncc_warp_patch.restype = C.c_int
    lib.ggs_ncc_warp_patch.argtypes = [C.POINTER(GgsNccPair), C.c_int64] + [C.c_void_p] * 10
    if lib.ggs_ncc_abi_version() != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {lib.ggs_ncc_abi_version()}, this module expects {ABI_VERSION}")
    return lib


_lib = _load()


class WarpParams(NamedTuple):
    R: torch.Tensor
    T: torch.Tensor
    fx_r: float
    fy_r: float
    cx_r: float
    cy_r: float
    fx_n: float
    fy_n: float
    cx_n: float
    cy_n: float
    debug: bool


def warp_patch_ncc(depths, normals, uvs, R, T, image_r, image_n, fx_r, fy_r, cx_r, cy_r, fx_n, fy_n, cx_n, cy_n,
                   debug):
    """(ncc [P], valid [P] bool): squared NCC of the 7x7 half-pixel patch around each reference pixel `uvs` and its
    warp into the neighbour image through the plane (depth, normal); differentiable in `depths` and `normals`."""
    params = WarpParams(R.contiguous(), T.contiguous(), fx_r, fy_r, cx_r, cy_r, fx_n, fy_n, cx_n, cy_n, debug)
    ncc, valid = _WarpPatchNCC.apply(depths, normals, uvs, image_r, image_n, params)
    return ncc, valid


def _f32(t, device):
    return t.detach().to(device=device, dtype=torch.float32).contiguous()


def _run(depths, normals, uvs, image_r, image_n, p: WarpParams):
    """The reference's `_C.warp_patch_ncc` (warp_patch_ncc.cu:5-57): (ncc, grad_depths, grad_normals, valid)."""
    if normals.dim() != 2 or normals.shape[1] != 3:
        raise RuntimeError("normals must have dimensions (num_points, 3)")  # warp_patch_ncc.cu:18-20
    if not depths.is_cuda:
        raise RuntimeError("warp_patch_ncc: tensors must live on a CUDA device (this module has no CPU path)")
    dev = depths.device
    P = int(depths.shape[0])
    d, n, ir, im = _f32(depths, dev), _f32(normals, dev), _f32(image_r, dev), _f32(image_n, dev)
    uv = uvs.detach().to(device=dev, dtype=torch.int32).contiguous()
    if uv.numel() != 2 * P or d.numel() != P or n.shape[0] != P:
        raise RuntimeError("warp_patch_ncc: depths, normals and uvs disagree on the number of points")
    if ir.dim() != 2 or im.dim() != 2:
        raise RuntimeError("warp_patch_ncc: images must be [H, W] grey images")
    pair = GgsNccPair()
    Rd, Td = _f32(p.R, dev), _f32(p.T, dev)  # stay on the device: no host synchronisation in a training step
    if Rd.numel() != 9 or Td.numel() != 3:
        raise RuntimeError("warp_patch_ncc: R must have 9 elements and T 3")
    pair.R, pair.T = Rd.data_ptr(), Td.data_ptr()
    for k in ("fx_r", "fy_r", "cx_r", "cy_r", "fx_n", "fy_n", "cx_n", "cy_n"):
        setattr(pair, k, float(getattr(p, k)))
    pair.height_r, pair.width_r = int(ir.shape[0]), int(ir.shape[1])
    pair.height_n, pair.width_n = int(im.shape[0]), int(im.shape[1])
    ncc = torch.empty(P, device=dev, dtype=torch.float32)
    g_d = torch.empty_like(d)
    g_n = torch.empty(P, 3, device=dev, dtype=torch.float32)
    valid = torch.empty(P, device=dev, dtype=torch.uint8)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = _lib.ggs_ncc_warp_patch(C.byref(pair), P, d.data_ptr(), n.data_ptr(), uv.data_ptr(), ir.data_ptr(),
                                     im.data_ptr(), ncc.data_ptr(), g_d.data_ptr(), g_n.data_ptr(), valid.data_ptr(),
                                     C.c_void_p(stream))
        if rc != 0:
            raise RuntimeError(_lib.ggs_ncc_last_error().decode())
        if p.debug:
            torch.cuda.synchronize(dev)  # the reference's CHECK_CUDA(..., debug)
    return ncc, g_d, g_n, valid.bool()


class _WarpPatchNCC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depths, normals, uvs, image_r, image_n, params):
        ncc, grad_depths, grad_normals, valid = _run(depths, normals, uvs, image_r, image_n, params)
        ctx.save_for_backward(grad_depths, grad_normals)
        ctx.mark_non_differentiable(valid)
        return ncc, valid

    @staticmethod
    def backward(ctx, grad_ncc, grad_valid):
        grad_depths, grad_normals = ctx.saved_tensors
        return grad_ncc * grad_depths, grad_ncc.unsqueeze(-1) * grad_normals, None, None, None, None
