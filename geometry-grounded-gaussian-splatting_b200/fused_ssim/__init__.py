"""Drop-in for the third-party `fused_ssim` package the reference imports (utils/loss_utils.py:18,48-49:
`fused_ssim(img1, img2, padding="valid")`; rahul-goel/fused-ssim, README.md:37): same function and arguments, the mean
SSIM of two [B,C,H,W] images as a scalar, differentiable in `img1`.  Kernels: libggs_ssim.so (C ABI:
include/ggs_ssim.h).  What is fused beyond that package: the mean is reduced inside the forward (the map is never
written) and "valid" padding computes only the cropped region; the backward reads the scalar gradient from device
memory.  There is no fallback: if the library is missing or does not load, importing this module fails.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libggs_ssim.so")
EXPORTS = ("ggs_ssim_abi_version", "ggs_ssim_last_error", "ggs_ssim_partials", "ggs_ssim_forward",
           "ggs_ssim_backward")
ABI_VERSION = 1
TAPS = 11
allowed_padding = ["same", "valid"]


class GgsSsimPlan(C.Structure):
    _fields_ = [("planes", C.c_int32), ("height", C.c_int32), ("width", C.c_int32), ("valid", C.c_int32),
                ("c1", C.c_float), ("c2", C.c_float), ("window", C.c_float * TAPS)]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C geometry-grounded-gaussian-splatting_b200/csrc` "
                          "(or __graft_entry__.build()); there is no CPU or PyTorch fallback for this kernel")
    lib = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise ImportError(f"{LIB_PATH} does not export {name}")
    lib.ggs_ssim_abi_version.restype = C.c_int
    lib.ggs_ssim_last_error.restype = C.c_char_p
    lib.ggs_ssim_partials.restype = C.c_int64
    lib.ggs_ssim_partials.argtypes = [C.POINTER(GgsSsimPlan)]
    lib.ggs_ssim_forward.restype = C.c_int
    lib.ggs_ssim_forward.argtypes = [C.POINTER(GgsSsimPlan)] + [C.c_void_p] * 8
    lib.ggs_ssim_backward.restype = C.c_int
    lib.ggs_ssim_backward.argtypes = [C.POINTER(GgsSsimPlan)] + [C.c_void_p] * 8
    if lib.ggs_ssim_abi_version() != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {lib.ggs_ssim_abi_version()}, this module expects {ABI_VERSION}")
    return lib


_lib = _load()


def gaussian_window(size: int = TAPS, sigma: float = 1.5):
    """utils/loss_utils.py:34-37, in float32 like the reference's torch.Tensor."""
    g = torch.tensor([math.exp(-((x - size // 2) ** 2) / float(2 * sigma ** 2)) for x in range(size)],
                     dtype=torch.float32)
    return g / g.sum()


def make_plan(shape, padding, C1, C2) -> GgsSsimPlan:
    B, Ch, H, W = shape
    plan = GgsSsimPlan()
    plan.planes, plan.height, plan.width = int(B * Ch), int(H), int(W)
    plan.valid = 1 if padding == "valid" else 0
    plan.c1, plan.c2 = float(C1), float(C2)
    plan.window[:] = gaussian_window().tolist()
    return plan


def _check(rc):
    if rc != 0:
        raise RuntimeError(_lib.ggs_ssim_last_error().decode())


class _FusedSSIM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding, train):
        if img1.dim() != 4 or img1.shape != img2.shape:
            raise RuntimeError("fused_ssim: img1 and img2 must be [B, C, H, W] tensors of the same shape")
        if not img1.is_cuda:
            raise RuntimeError("fused_ssim: tensors must live on a CUDA device (this module has no CPU path)")
        dev = img1.device
        a = img1.detach().to(torch.float32).contiguous()
        b = img2.detach().to(device=dev, dtype=torch.float32).contiguous()
        plan = make_plan(a.shape, padding, C1, C2)
        n = _lib.ggs_ssim_partials(C.byref(plan))
        if n < 0:
            raise RuntimeError(_lib.ggs_ssim_last_error().decode())
        partials = torch.empty(int(n), device=dev, dtype=torch.float32)
        mean = torch.empty((), device=dev, dtype=torch.float32)
        maps = [torch.empty_like(a) for _ in range(3)] if train else [None] * 3
        ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            _check(_lib.ggs_ssim_forward(C.byref(plan), a.data_ptr(), b.data_ptr(), ptr(maps[0]), ptr(maps[1]),
                                         ptr(maps[2]), partials.data_ptr(), mean.data_ptr(), C.c_void_p(stream)))
        ctx.plan = plan
        ctx.train = bool(train)
        if train:
            ctx.save_for_backward(a, b, *maps)
        return mean

    @staticmethod
    def backward(ctx, grad_mean):
        if not ctx.train:
            raise RuntimeError("fused_ssim: called with train=False, no derivative maps were kept")
        a, b, m0, m1, m2 = ctx.saved_tensors
        g = grad_mean.detach().to(device=a.device, dtype=torch.float32).reshape(1).contiguous()
        grad = torch.empty_like(a)
        with torch.cuda.device(a.device):
            stream = torch.cuda.current_stream(a.device).cuda_stream
            _check(_lib.ggs_ssim_backward(C.byref(ctx.plan), a.data_ptr(), b.data_ptr(), m0.data_ptr(), m1.data_ptr(),
                                          m2.data_ptr(), g.data_ptr(), grad.data_ptr(), C.c_void_p(stream)))
        return None, None, grad, None, None, None


def fused_ssim(img1, img2, padding="same", train=True):
    """Mean SSIM of img1 vs img2 ([B,C,H,W]); gradient flows to img1 only, as in the package this replaces."""
    C1 = 0.01 ** 2
    C2 = 0.03 ** 2
    assert padding in allowed_padding
    return _FusedSSIM.apply(C1, C2, img1, img2, padding, train)
