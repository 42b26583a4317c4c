"""ctypes binding of libggs_rasterizer.so (C ABI declared in include/ggs_rasterizer.h).

PyTorch is used for device memory and streams only; every kernel lives in the shared library.
There is no fallback: if the library is missing or does not load, importing this module fails.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# GGS_RASTERIZER_LIB: a differently built copy of the same library (A/B timing of kernel variants, tools/variants.sh);
# there is still no fallback -- whatever path is named must exist and export the whole ABI.
LIB_PATH = os.environ.get("GGS_RASTERIZER_LIB") or os.path.join(_HERE, "libggs_rasterizer.so")

ABI_VERSION = 2  # GGS_ABI_VERSION of include/ggs_rasterizer.h this binding was written against

(SLOT_GEOMETRY, SLOT_BINNING, SLOT_IMAGE, SLOT_TILE, SLOT_POINT, SLOT_POINT_BINNING, SLOT_BACKWARD,
 SLOT_BLOCKS) = range(8)

_f32p = C.c_void_p


class GgsFrame(C.Structure):
    _fields_ = [
        ("image_width", C.c_int32), ("image_height", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float),
        ("kernel_size", C.c_float), ("scale_modifier", C.c_float),
        ("sh_degree", C.c_int32), ("sg_degree", C.c_int32),
        ("prefiltered", C.c_int32), ("require_depth", C.c_int32), ("debug", C.c_int32),
        ("viewmatrix", _f32p), ("projmatrix", _f32p), ("campos", _f32p), ("background", _f32p),
    ]


class GgsGaussians(C.Structure):
    _fields_ = [
        ("count", C.c_int64),
        ("means3D", _f32p), ("opacities", _f32p), ("scales", _f32p), ("rotations", _f32p),
        ("shs", _f32p), ("sh_coeffs", C.c_int32),
        ("sg_axis", _f32p), ("sg_sharpness", _f32p), ("sg_color", _f32p), ("sg_lobes", C.c_int32),
        ("colors_precomp", _f32p),
    ]


class GgsForwardOutputs(C.Structure):
    _fields_ = [("color", _f32p), ("alpha", _f32p), ("normal", _f32p), ("mdepth", _f32p), ("radii", C.c_void_p)]


class GgsForwardState(C.Structure):
    # capacity: in = hint (0: none, the forward blocks on the pair count), out = pairs the binning workspace is laid out for
    _fields_ = [("geometry", C.c_void_p), ("binning", C.c_void_p), ("image", C.c_void_p), ("tile", C.c_void_p),
                ("num_rendered", C.c_int64), ("capacity", C.c_int64)]


class GgsSampleState(C.Structure):
    _fields_ = [("geometry", C.c_void_p), ("binning", C.c_void_p), ("point", C.c_void_p),
                ("point_binning", C.c_void_p), ("tile", C.c_void_p), ("blocks", C.c_void_p),
                ("num_rendered", C.c_int64), ("num_evaluated", C.c_int64), ("num_blocks", C.c_int64)]


class GgsViewState(C.Structure):
    _fields_ = [("geometry", C.c_void_p), ("binning", C.c_void_p), ("tile", C.c_void_p),
                ("num_rendered", C.c_int64), ("count", C.c_int64)]


class GgsBackwardInputs(C.Structure):
    _fields_ = [("alpha", _f32p), ("normal", _f32p), ("mdepth", _f32p), ("radii", C.c_void_p),
                ("dL_dcolor", _f32p), ("dL_dmdepth", _f32p), ("dL_dalpha", _f32p), ("dL_dnormal", _f32p)]


class GgsGradients(C.Structure):
    _fields_ = [("dL_dmeans3D", _f32p), ("dL_dmeans2D", _f32p), ("dL_dcolors", _f32p), ("dL_dopacity", _f32p),
                ("dL_dscales", _f32p), ("dL_drotations", _f32p), ("dL_dsh", _f32p), ("dL_dsg_axis", _f32p),
                ("dL_dsg_sharpness", _f32p), ("dL_dsg_color", _f32p), ("dL_drgb", _f32p)]


MAX_CHUNKS = 8  # GGS_MAX_CHUNKS
SORT_STATIC_MAX_PAIRS = 1 << 18  # GGS_SORT_STATIC_MAX_PAIRS


class GgsPeerExchange(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("peer_buffers", C.POINTER(C.c_uint64)),
                ("multicast_base", C.c_uint64)] + [(n, C.c_int64) for n in (
        "send_offset", "recv_offset", "reduce_floats", "rgb_stage_offset", "rgb_stage_stride",
        "campos_stage_offset", "reduce_stage_offset", "reduce_stage_stride")] + [
        ("pull_reduce", C.c_int32), ("chunks", C.c_int32), ("chunk_end", C.c_int64 * MAX_CHUNKS),
        ("chunk_events", C.c_uint64 * MAX_CHUNKS), ("fuse_rgb_push", C.c_int32)]


class GgsLayout(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "geometry_bytes", "records", "depths", "tiles_touched", "point_offsets", "clamped", "rects",
        "binning_bytes", "sorted_keys", "point_list", "unsorted_keys", "unsorted_list",
        "tile_bytes", "ranges", "max_contributor", "image_bytes", "n_contrib")] + [
        ("sort_bits", C.c_uint32), ("sort_passes", C.c_uint32), ("depth_order", C.c_uint64)]


class GgsStageTimes(C.Structure):
    _fields_ = [("have_forward", C.c_int32), ("have_backward", C.c_int32)] + [(n, C.c_float) for n in (
        "preprocess_ms", "scan_ms", "duplicate_ms", "sort_ms", "ranges_ms", "render_forward_ms",
        "render_backward_ms", "gaussian_backward_ms")] + [("sort_passes", C.c_int32), ("depth_order_ms", C.c_float),
                                                          ("have_query", C.c_int32), ("query_binning_ms", C.c_float),
                                                          ("query_kernel_ms", C.c_float)]


ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_int, C.c_size_t, C.c_void_p)

EXPORTS = ("ggs_abi_version", "ggs_last_error", "ggs_rasterize_forward", "ggs_rasterize_backward",
           "ggs_backward_from_accumulators", "ggs_sh_gradient_from_views", "ggs_sample_depth_forward",
           "ggs_sample_depth_backward", "ggs_evaluate_sdf",
           "ggs_integrate_points", "ggs_view_prepare", "ggs_integrate_view", "ggs_integrate_view_counted", "ggs_evaluate_sdf_view", "ggs_mark_visible", "ggs_count_blend_ops", "ggs_describe_layout", "ggs_materialize_sorted_keys", "ggs_sort_scratch_bytes",
           "ggs_sort_pairs", "ggs_sort_pairs_static", "ggs_profile_enable", "ggs_profile_read", "ggs_view_sum_peer",
           "ggs_peer_sum_ranges", "ggs_rasterize_backward_peer", "ggs_activate_forward", "ggs_activate_backward")


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C geometry-grounded-gaussian-splatting_b200/csrc` "
            "(or __graft_entry__.build()).  There is no CPU or PyTorch fallback for the rasterizer.")
    lib = C.CDLL(LIB_PATH)
    lib.ggs_abi_version.restype = C.c_int
    lib.ggs_last_error.restype = C.c_char_p
    lib.ggs_rasterize_forward.restype = C.c_int64
    lib.ggs_rasterize_forward.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), ALLOC_FN, C.c_void_p,
                                          C.POINTER(GgsForwardOutputs), C.POINTER(GgsForwardState), C.c_void_p]
    lib.ggs_rasterize_backward.restype = C.c_int
    lib.ggs_rasterize_backward.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), C.POINTER(GgsForwardState),
                                           C.POINTER(GgsBackwardInputs), ALLOC_FN, C.c_void_p,
                                           C.POINTER(GgsGradients), C.c_void_p]
    lib.ggs_backward_from_accumulators.restype = C.c_int
    lib.ggs_backward_from_accumulators.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians),
                                                   C.POINTER(GgsForwardState), C.c_void_p, C.c_void_p,
                                                   C.POINTER(GgsGradients), C.c_void_p]
    lib.ggs_sh_gradient_from_views.restype = C.c_int
    lib.ggs_sh_gradient_from_views.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ggs_sample_depth_forward.restype = C.c_int64
    lib.ggs_sample_depth_forward.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), C.c_void_p, C.c_int64,
                                             ALLOC_FN, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(GgsSampleState),
                                             C.c_void_p]
    lib.ggs_evaluate_sdf.restype = C.c_int64
    lib.ggs_evaluate_sdf.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), C.c_void_p, C.c_int64, ALLOC_FN,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ggs_sample_depth_backward.restype = C.c_int
    lib.ggs_sample_depth_backward.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), C.c_void_p, C.c_int64,
                                              C.POINTER(GgsSampleState), C.c_void_p, C.c_void_p, ALLOC_FN, C.c_void_p,
                                              C.POINTER(GgsGradients), C.c_void_p, C.c_void_p]
    lib.ggs_integrate_points.restype = C.c_int64
    lib.ggs_integrate_points.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), C.c_void_p, C.c_int64,
                                         ALLOC_FN, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ggs_view_prepare.restype = C.c_int64
    lib.ggs_view_prepare.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), ALLOC_FN, C.c_void_p,
                                     C.POINTER(GgsViewState), C.c_void_p]
    lib.ggs_integrate_view.restype = C.c_int
    lib.ggs_integrate_view.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsViewState), C.c_void_p, C.c_int64, ALLOC_FN,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ggs_integrate_view_counted.restype = C.c_int
    lib.ggs_integrate_view_counted.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsViewState), C.c_void_p, C.c_int64,
                                               ALLOC_FN, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ggs_evaluate_sdf_view.restype = C.c_int
    lib.ggs_evaluate_sdf_view.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsViewState), C.c_void_p, C.c_int64,
                                          ALLOC_FN, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ggs_mark_visible.restype = C.c_int
    lib.ggs_mark_visible.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ggs_count_blend_ops.restype = C.c_int
    lib.ggs_count_blend_ops.argtypes = [C.POINTER(GgsFrame), C.POINTER(GgsGaussians), C.POINTER(GgsForwardState),
                                        C.c_void_p, C.c_void_p]
    lib.ggs_describe_layout.restype = C.c_int
    lib.ggs_describe_layout.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int64, C.POINTER(GgsLayout)]
    lib.ggs_sort_scratch_bytes.restype = C.c_size_t
    lib.ggs_sort_scratch_bytes.argtypes = [C.c_int64, C.c_int32, C.c_int32]
    lib.ggs_sort_pairs.restype = C.c_int
    lib.ggs_sort_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                   C.c_void_p, C.c_void_p]
    # no allocation anywhere: the scratch is a static device array inside the library (include/ggs_rasterizer.h)
    lib.ggs_sort_pairs_static.restype = C.c_int
    lib.ggs_sort_pairs_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                                          C.c_int32, C.c_void_p]
    lib.ggs_profile_enable.restype = C.c_int
    lib.ggs_profile_enable.argtypes = [C.c_int]
    lib.ggs_materialize_sorted_keys.restype = C.c_int
    lib.ggs_materialize_sorted_keys.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.POINTER(GgsForwardState), C.c_void_p]
    lib.ggs_peer_sum_ranges.restype = C.c_int
    lib.ggs_peer_sum_ranges.argtypes = [C.POINTER(GgsPeerExchange), C.c_int32, C.POINTER(C.c_int64),
                                        C.POINTER(C.c_int64), C.c_void_p]
    lib.ggs_view_sum_peer.restype = C.c_int
    lib.ggs_view_sum_peer.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.POINTER(GgsPeerExchange),
                                      C.c_void_p, C.c_int32, C.c_void_p]
    lib.ggs_rasterize_backward_peer.restype = C.c_int
    lib.ggs_rasterize_backward_peer.argtypes = list(lib.ggs_rasterize_backward.argtypes[:-1]) + [
        C.POINTER(GgsPeerExchange), C.c_void_p, C.c_void_p]
    lib.ggs_activate_forward.restype = C.c_int
    lib.ggs_activate_forward.argtypes = [C.c_int64] + [C.c_void_p] * 8
    lib.ggs_activate_backward.restype = C.c_int
    lib.ggs_activate_backward.argtypes = [C.c_int64] + [C.c_void_p] * 11
    lib.ggs_profile_read.restype = C.c_int
    lib.ggs_profile_read.argtypes = [C.POINTER(GgsStageTimes)]
    if lib.ggs_abi_version() != ABI_VERSION:
        raise ImportError("libggs_rasterizer.so has an unexpected ABI version")
    return lib


lib = _load()


def last_error() -> str:
    return (lib.ggs_last_error() or b"").decode()


def check(rc: int, what: str) -> int:
    if rc < 0:
        raise RuntimeError(f"{what} failed ({rc}): {last_error()}")
    return rc


def ptr(t):
    """Device pointer of a tensor, NULL for None / empty tensors."""
    if t is None or t.numel() == 0:
        return None
    return C.c_void_p(t.data_ptr())


class Arena:
    """Hands torch-owned device buffers to the library (the reference's resize lambdas,
    rasterize_points.cu:27-37).  The tensors stay referenced in `buffers` until the arena is dropped.

    The C callback closes over a plain dict, not over the arena, and `done()` drops it after the call: no
    reference cycle, so the (hundreds of MB of) workspaces are released by reference counting as soon as the
    autograd context dies instead of waiting for Python's cycle collector."""

    def __init__(self, device):
        self.device = device
        self.buffers = buffers = {}
        self._err = err = []

        def _alloc(slot, nbytes, _user):
            try:
                t = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
                buffers[int(slot)] = t
                return t.data_ptr()
            except Exception as ex:  # surfaces as GGS_ERR_ALLOC
                err.append(ex)
                return 0

        self.fn = ALLOC_FN(_alloc)

    @property
    def error(self):
        return self._err[0] if self._err else None

    def done(self):
        """Call after the library call returned: the callback is not needed any more."""
        self.fn = None


def stream_handle(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def profile_enable(on: bool) -> None:
    check(lib.ggs_profile_enable(1 if on else 0), "ggs_profile_enable")


def profile_read() -> dict:
    t = GgsStageTimes()
    check(lib.ggs_profile_read(C.byref(t)), "ggs_profile_read")
    return {n: getattr(t, n) for n, _ in GgsStageTimes._fields_}


def describe_layout(P: int, W: int, H: int, num_rendered: int) -> GgsLayout:
    # num_rendered: the CAPACITY the binning workspace is laid out for (GgsForwardState.capacity; == the pair count
    # when the forward ran without a hint)
    L = GgsLayout()
    check(lib.ggs_describe_layout(P, W, H, num_rendered, C.byref(L)), "ggs_describe_layout")
    return L
