"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/ggs_rasterizer.h declares, layout queries work without a GPU, and the Python surface keeps the
reference's names, field order and argument errors (DGR/diff_gaussian_rasterization/__init__.py:254-336)."""
import ctypes as C
import inspect
import os
import re

import pytest
import torch

import helpers as H


def _declared():
    hdr = open(os.path.join(H.ROOT, "include", "ggs_rasterizer.h")).read()
    return sorted(set(re.findall(r"GGS_API\s+[\w\s\*]+?\b(ggs_\w+)\s*\(", hdr)))


def test_header_symbols_exported():
    from diff_gaussian_rasterization import _abi

    names = _declared()
    assert len(names) >= 11
    lib = C.CDLL(_abi.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ggs_rasterizer.h but not exported"
    assert set(names) == set(_abi.EXPORTS)
    assert lib.ggs_abi_version() == 2


def test_no_torch_in_abi_and_arch():
    """The shared library is plain C ABI (no libtorch / libc10 dependency)."""
    from diff_gaussian_rasterization import _abi
    import subprocess

    out = subprocess.run(["ldd", _abi.LIB_PATH], capture_output=True, text=True).stdout
    libs = [ln.split()[0] for ln in out.splitlines() if ln.strip()]  # names only: load addresses are random hex
    assert not any("torch" in n or "c10" in n for n in libs), libs


def test_layout_query_on_cpu():
    from diff_gaussian_rasterization import _abi

    L = _abi.describe_layout(1000, 1237, 822, 12345)
    assert L.sort_bits == 44 and L.sort_passes == 2          # 4056 tiles -> getHigherMsb = 12; pair passes = tile bits only
    L = _abi.describe_layout(10, 256, 256, 0)
    assert L.sort_bits == 41                                  # 256 tiles -> 9 (NOT ceil(log2))
    L = _abi.describe_layout(10, 800, 600, 5)
    assert L.sort_bits == 43
    assert L.records % 256 == 0 and L.point_list % 256 == 0 and L.ranges % 256 == 0
    assert _abi.lib.ggs_sort_scratch_bytes(0, 0, 44) >= 0
    assert _abi.lib.ggs_sort_scratch_bytes(1 << 20, 0, 44) > (1 << 20) * 24
    bad = _abi.GgsLayout()
    assert _abi.lib.ggs_describe_layout(-1, 10, 10, 0, C.byref(bad)) < 0
    assert b"layout" in _abi.lib.ggs_last_error()


def test_argument_errors_without_gpu():
    from diff_gaussian_rasterization import _abi

    # null frame -> invalid argument, message retrievable; nothing touches CUDA
    rc = _abi.lib.ggs_rasterize_forward(None, None, _abi.ALLOC_FN(lambda a, b, c: 0), None, None, None, None)
    assert rc == -1 and _abi.last_error()
    rc = _abi.lib.ggs_sort_pairs(None, None, None, None, -5, 0, 8, None, None)
    assert rc == -1
    rc = _abi.lib.ggs_sort_pairs(None, None, None, None, 10, 8, 8, None, None)
    assert rc == -1


def test_static_sort_rejects_bad_sizes_without_gpu():
    """ggs_sort_pairs_static (no allocation; static device scratch): every refusal happens before any CUDA call."""
    from diff_gaussian_rasterization import _abi

    f = _abi.lib.ggs_sort_pairs_static
    one = C.c_void_p(256)  # never dereferenced: the size check comes first
    assert f(one, one, one, one, _abi.SORT_STATIC_MAX_PAIRS + 1, 0, 44, None) == -1
    msg = _abi.last_error()
    assert "GGS_SORT_STATIC_MAX_PAIRS" in msg and str(_abi.SORT_STATIC_MAX_PAIRS) in msg
    assert f(one, one, one, one, -1, 0, 44, None) == -1
    assert f(one, one, one, one, 10, 44, 44, None) == -1
    assert f(one, one, one, one, 10, 0, 65, None) == -1
    assert f(None, None, None, None, 10, 0, 8, None) == -1
    assert f(None, None, None, None, 0, 0, 8, None) == 0  # nothing to sort, nothing touched
    hdr = open(os.path.join(H.ROOT, "include", "ggs_rasterizer.h")).read()
    assert f"#define GGS_SORT_STATIC_MAX_PAIRS (1 << {_abi.SORT_STATIC_MAX_PAIRS.bit_length() - 1})" in hdr


def test_python_surface_matches_reference_names():
    import diff_gaussian_rasterization as dgr

    fields = dgr.GaussianRasterizationSettings._fields
    assert fields == ("image_height", "image_width", "tanfovx", "tanfovy", "kernel_size", "bg", "scale_modifier",
                      "viewmatrix", "projmatrix", "sh_degree", "sg_degree", "campos", "prefiltered",
                      "require_depth", "debug")
    sig = list(inspect.signature(dgr.GaussianRasterizer.forward).parameters)
    assert sig == ["self", "means3D", "means2D", "opacities", "shs", "sg_axis", "sg_sharpness", "sg_color",
                   "colors_precomp", "scales", "rotations", "cov3D_precomp"]
    sig = list(inspect.signature(dgr.GaussianRasterizer.integrate).parameters)
    assert sig[:6] == ["self", "points3D", "means3D", "opacities", "scales", "rotations"]
    for name in ("markVisible", "integrate", "evaluate_sdf", "sample_depth"):
        assert hasattr(dgr.GaussianRasterizer, name)


def test_either_or_argument_rules():
    import diff_gaussian_rasterization as dgr

    rs = dgr.GaussianRasterizationSettings(16, 16, 1.0, 1.0, 0.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                           0, torch.zeros(3), False, True, False)
    r = dgr.GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), shs=None, colors_precomp=None, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=None, rotations=None, cov3D_precomp=None)
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r.integrate(m, m, torch.zeros(4, 1), scales=None, rotations=None)
    # CPU tensors are refused loudly: there is no CPU path
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="dimensions"):
        r(torch.zeros(4), m, torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), scales=m, rotations=torch.zeros(4, 4))


def test_product_never_imports_oracle():
    """The product package must not reference the oracle (test infrastructure) or any CPU fallback."""
    for name in ("diff_gaussian_rasterization", "warp_patch_ncc", "fused_ssim"):
        pkg = os.path.join(H.PKG, name)
        for fn in os.listdir(pkg):
            if fn.endswith(".py"):
                src = open(os.path.join(pkg, fn)).read()
                assert "oracle" not in src.replace("Oracle A", ""), fn
    for fn in os.listdir(os.path.join(H.PKG, "csrc")):
        if fn.endswith((".cu", ".cuh")):
            assert "oracle" not in open(os.path.join(H.PKG, "csrc", fn)).read(), fn


def test_public_headers_are_plain_c99():
    """The boundary is a C ABI: every header under include/ must compile as C99 with no C++ or CUDA in it."""
    import glob
    import subprocess

    headers = sorted(glob.glob(os.path.join(H.ROOT, "include", "*.h")))
    assert len(headers) >= 3
    for h in headers:
        r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", "-"],
                           input=f'#include "{h}"\n', capture_output=True, text=True)
        assert r.returncode == 0, (h, r.stderr[-800:])
