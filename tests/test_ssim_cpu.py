"""Fused SSIM without a GPU: the kernels' SOURCE run on the host under tests/native/cuda_emu.h
against the reference's torch statement of SSIM in float64, and the C ABI of libggs_ssim.so (symbols and host-side
queries only: no compute without a GPU)."""
import ctypes as C
import os
import re
import subprocess

import pytest
import torch

import helpers as H
import ssim_reference as R

CSRC = os.path.join(H.ROOT, "geometry-grounded-gaussian-splatting_b200", "csrc")


@pytest.fixture(scope="module")
def emulated(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("ssim_emu") / "ssim_emulated.so")
    native = os.path.join(H.ROOT, "tests", "native")
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-pthread", "-shared", "-fPIC", "-I" + CSRC, "-I" + native,
                           "-o", so, os.path.join(native, "ssim_emulated.cpp")])
    return C.CDLL(so)


CASES = [  # (B, C, H, W, padding): several tiles, partial tiles, several planes, the smallest valid image
    (1, 3, 40, 50, "valid"), (1, 3, 40, 50, "same"), (2, 1, 33, 64, "valid"), (1, 1, 75, 70, "same"),
    (1, 2, 12, 11, "valid"), (1, 1, 5, 7, "same"),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "x".join(map(str, c)))
def test_kernel_sources_under_emulation_match_torch_float64(case, emulated):
    B, Ch, Hh, W, padding = case
    img1, img2 = R.image_pair((B, Ch, Hh, W), seed=3 + Hh)
    want, want_grad = R.value_and_grad(img1, img2, padding, upstream=0.7)
    planes, valid = B * Ch, 1 if padding == "valid" else 0
    rh, rw = (Hh - 10, W - 10) if valid else (Hh, W)
    tiles = ((rh + 31) // 32) * ((rw + 31) // 32) * planes
    maps = [torch.full((planes, Hh, W), float("nan")) for _ in range(3)]
    partials, mean = torch.zeros(tiles), torch.zeros(1)
    window = R.gaussian().float()
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    emulated.ggs_test_ssim_forward(planes, Hh, W, valid, p(window), p(img1), p(img2), p(maps[0]), p(maps[1]),
                                   p(maps[2]), p(partials), p(mean))
    assert abs(float(mean) - float(want)) <= 1e-6 * abs(float(want))
    inside = torch.zeros(Hh, W, dtype=torch.bool)
    inside[(5 if valid else 0):(Hh - 5 if valid else Hh), (5 if valid else 0):(W - 5 if valid else W)] = True
    for m in maps:  # written exactly on the map region
        assert not bool(torch.isnan(m[:, inside]).any()) and bool(torch.isnan(m[:, ~inside]).all())
    grad = torch.full((planes, Hh, W), float("nan"))
    upstream = torch.tensor([0.7])
    emulated.ggs_test_ssim_backward(planes, Hh, W, valid, p(window), p(img1), p(img2), p(maps[0]), p(maps[1]),
                                    p(maps[2]), p(upstream), p(grad))
    assert not bool(torch.isnan(grad).any())
    err = float((grad.reshape(B, Ch, Hh, W).double() - want_grad).abs().max() / want_grad.abs().max())
    assert err <= 1e-5, err


def test_library_exports_what_the_header_declares():
    import fused_ssim as ours

    with open(os.path.join(H.ROOT, "include", "ggs_ssim.h")) as fh:
        text = re.sub(r"/\*.*?\*/", "", fh.read(), flags=re.S)
    declared = set(re.findall(r"\b(ggs_ssim_\w+)\s*\(", text))
    assert declared == set(ours.EXPORTS), declared ^ set(ours.EXPORTS)
    lib = C.CDLL(ours.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ggs_ssim_abi_version() == ours.ABI_VERSION
    assert C.sizeof(ours.GgsSsimPlan) == 4 * 4 + 2 * 4 + 11 * 4
    # host-side query: tiles of the map region times planes
    plan = ours.make_plan((1, 3, 822, 1237), "valid", 0.01 ** 2, 0.03 ** 2)
    assert ours._lib.ggs_ssim_partials(C.byref(plan)) == 3 * ((812 + 31) // 32) * ((1227 + 31) // 32)
    plan = ours.make_plan((2, 3, 64, 64), "same", 0.01 ** 2, 0.03 ** 2)
    assert ours._lib.ggs_ssim_partials(C.byref(plan)) == 6 * 2 * 2
    plan = ours.make_plan((1, 1, 10, 64), "valid", 0.01 ** 2, 0.03 ** 2)  # no map left
    assert ours._lib.ggs_ssim_partials(C.byref(plan)) < 0
    assert torch.equal(ours.gaussian_window(), R.gaussian().float())
    assert abs(float(sum(plan.window)) - 1.0) < 1e-6
    img1, img2 = R.image_pair((1, 3, 20, 20), seed=1)
    with pytest.raises(RuntimeError, match="CUDA"):
        ours.fused_ssim(img1, img2, padding="valid")
    with pytest.raises(AssertionError):
        ours.fused_ssim(img1, img2, padding="reflect")


def test_smooth_images_with_flat_regions_are_as_accurate_as_torch_float32(emulated):
    """Rendered images are smooth and have flat areas, where sigma^2 = E[x^2] - mu^2 cancels and the map divides by
    ~C2: the fused kernels must not lose more there than the reference's own float32 formulation does."""
    import ncc_scenes

    tex = ncc_scenes.make_pair(height=120, width=160, n=4)["image_r"]
    img1 = tex[None, None].repeat(1, 3, 1, 1).contiguous()
    img1[:, :, :40, :60] = 0.3
    img2 = (img1 + 0.01 * torch.randn(img1.shape, generator=torch.Generator().manual_seed(0))).clamp(0, 1).contiguous()
    img2[:, :, :40, :60] = 0.31
    B, Ch, Hh, W = img1.shape
    want, want_grad = R.value_and_grad(img1, img2, "valid")
    v32, g32 = R.value_and_grad(img1, img2, "valid", dtype=torch.float32)
    torch_err = float((g32.double() - want_grad).abs().max() / want_grad.abs().max())
    planes = B * Ch
    tiles = (((Hh - 10) + 31) // 32) * (((W - 10) + 31) // 32) * planes
    maps = [torch.zeros(planes, Hh, W) for _ in range(3)]
    partials, mean, grad = torch.zeros(tiles), torch.zeros(1), torch.zeros(planes, Hh, W)
    window, one = R.gaussian().float(), torch.tensor([1.0])
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    emulated.ggs_test_ssim_forward(planes, Hh, W, 1, p(window), p(img1), p(img2), p(maps[0]), p(maps[1]), p(maps[2]),
                                   p(partials), p(mean))
    emulated.ggs_test_ssim_backward(planes, Hh, W, 1, p(window), p(img1), p(img2), p(maps[0]), p(maps[1]), p(maps[2]),
                                    p(one), p(grad))
    ours_err = float((grad.reshape(B, Ch, Hh, W).double() - want_grad).abs().max() / want_grad.abs().max())
    assert abs(float(mean) - float(want)) <= 1e-5 * abs(float(want))
    assert ours_err <= 1e-4 and ours_err <= 2.0 * torch_err, (ours_err, torch_err)


def test_emulated_kernels_have_no_out_of_bounds_access():
    """tools/emu_asan.sh: the NCC and SSIM kernel sources under the host shim, built with AddressSanitizer, on
    exactly-sized buffers (global memory and the static "shared" arrays get redzones)."""
    asan = subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(asan) or not os.path.exists(asan):
        pytest.skip("libasan not installed")
    r = subprocess.run(["bash", os.path.join(H.ROOT, "tools", "emu_asan.sh")], capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0 and "no out-of-bounds access" in r.stdout, (r.stdout[-500:], r.stderr[-1500:])


def test_reference_golden_pins_the_checker_and_the_kernels(emulated):
    """tests/golden/ssim/same.npz was produced by executing the reference's own `_ssim` (utils/loss_utils.py:52-73,
    tests/golden/make_ssim_golden.py): it pins tests/ssim_reference.py, the checker of every SSIM test, and the
    kernels (run from source on the host) for the padding the reference's torch code has."""
    import numpy as np

    gold = np.load(os.path.join(H.ROOT, "tests", "golden", "ssim", "same.npz"))
    img1, img2 = torch.from_numpy(gold["img1"]), torch.from_numpy(gold["img2"])
    val, grad = R.value_and_grad(img1, img2, "same")
    # the reference rounds the 2-D window's products to float32 (create_window, :41-43); the checker and the separable
    # kernels use the products of the float32 taps unrounded: agreement to 1e-8, not to the last bit
    assert abs(float(val) - float(gold["value_f64"])) < 1e-8
    want64 = torch.from_numpy(gold["grad_f64"])
    assert float((grad - want64).abs().max()) < 1e-7 * float(want64.abs().max())
    B, Ch, Hh, W = img1.shape
    planes = B * Ch
    tiles = ((Hh + 31) // 32) * ((W + 31) // 32) * planes
    maps = [torch.zeros(planes, Hh, W) for _ in range(3)]
    partials, mean, out = torch.zeros(tiles), torch.zeros(1), torch.zeros(planes, Hh, W)
    window, one = R.gaussian().float(), torch.tensor([1.0])
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    emulated.ggs_test_ssim_forward(planes, Hh, W, 0, p(window), p(img1), p(img2), p(maps[0]), p(maps[1]), p(maps[2]),
                                   p(partials), p(mean))
    emulated.ggs_test_ssim_backward(planes, Hh, W, 0, p(window), p(img1), p(img2), p(maps[0]), p(maps[1]), p(maps[2]),
                                    p(one), p(out))
    want_grad = torch.from_numpy(gold["grad_f64"])
    assert abs(float(mean) - float(gold["value_f64"])) <= 1e-6 * float(gold["value_f64"])
    err = float((out.reshape(B, Ch, Hh, W).double() - want_grad).abs().max() / want_grad.abs().max())
    ref32 = float((torch.from_numpy(gold["grad_f32"]).double() - want_grad).abs().max() / want_grad.abs().max())
    assert err <= 1e-5 and err <= 4.0 * max(ref32, 1e-7), (err, ref32)
