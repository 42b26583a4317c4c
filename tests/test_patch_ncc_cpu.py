"""Patch NCC without a GPU: the C oracle against a float64 model with numerical derivatives, the
kernel's own per-point routine (ggs_patch_ncc_point.cuh compiled for the host by tests/native/patch_ncc_host.cpp)
against the oracle, and the C ABI of libggs_patch_ncc.so (symbols only: no compute without a GPU)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest
import torch

import helpers as H
import ncc_scenes

CSRC = os.path.join(H.ROOT, "geometry-grounded-gaussian-splatting_b200", "csrc")
INTR = ("fx_r", "fy_r", "cx_r", "cy_r", "fx_n", "fy_n", "cx_n", "cy_n")


def _bilinear(img, x, y):
    x0, y0 = np.floor(x).astype(int), np.floor(y).astype(int)
    x0 = np.clip(x0, 0, img.shape[1] - 2)
    y0 = np.clip(y0, 0, img.shape[0] - 2)
    ax, ay = x - x0, y - y0
    return ((1 - ay) * ((1 - ax) * img[y0, x0] + ax * img[y0, x0 + 1])
            + ay * ((1 - ax) * img[y0 + 1, x0] + ax * img[y0 + 1, x0 + 1]))


def model_ncc(sc, depths, normals):
    """float64, straight from the definition: x_n ~ K_n (R - T n^T / d) K_r^-1 x_r, d = -(ray . n) depth; squared NCC of
    the 7x7 half-pixel patch.  No validity logic: callers look at points every implementation calls valid."""
    f = {k: float(sc[k]) for k in INTR}
    R = sc["R"].double().numpy().T  # the tensor holds the columns of R in memory order
    T = sc["T"].double().numpy()
    ir, im = sc["image_r"].double().numpy(), sc["image_n"].double().numpy()
    uv = sc["uvs"].numpy().astype(np.float64)
    Kn = np.array([[f["fx_n"], 0, f["cx_n"]], [0, f["fy_n"], f["cy_n"]], [0, 0, 1]])
    Kr_inv = np.linalg.inv(np.array([[f["fx_r"], 0, f["cx_r"]], [0, f["fy_r"], f["cy_r"]], [0, 0, 1]]))
    out = np.zeros(len(uv))
    offs = 0.5 * np.arange(-3, 4)
    dv, du = np.meshgrid(offs, offs, indexing="ij")
    for i in range(len(uv)):
        ray = Kr_inv @ np.array([uv[i, 0], uv[i, 1], 1.0])
        dist = -(ray @ normals[i]) * depths[i]
        Hm = Kn @ (R - np.outer(T, normals[i]) / dist) @ Kr_inv
        xr, yr = uv[i, 0] + du.ravel(), uv[i, 1] + dv.ravel()
        p = Hm @ np.stack([xr, yr, np.ones_like(xr)])
        c_r = _bilinear(ir, xr, yr)
        c_n = _bilinear(im, p[0] / p[2], p[1] / p[2])
        cross = (c_r * c_n).sum() - c_r.sum() * c_n.sum() / 49
        vr = (c_r ** 2).sum() - c_r.sum() ** 2 / 49
        vn = (c_n ** 2).sum() - c_n.sum() ** 2 / 49
        out[i] = cross * cross / (vr * vn + 1e-8)
    return out


@pytest.fixture(scope="module")
def host_routine(tmp_path_factory):
    """The kernel's per-point routine compiled for the host (IEEE, no contraction, like the oracle)."""
    so = str(tmp_path_factory.mktemp("ncc_host") / "patch_ncc_host.so")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I" + CSRC, "-o", so,
                           os.path.join(H.ROOT, "tests", "native", "patch_ncc_host.cpp")])
    lib = C.CDLL(so)

    def run(sc):
        P = int(sc["depths"].shape[0])
        o = [torch.zeros(P), torch.zeros(P), torch.zeros(P, 3), torch.zeros(P, dtype=torch.uint8)]
        intr = torch.tensor([sc[k] for k in INTR], dtype=torch.float32)
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        lib.ggs_test_ncc_points(P, p(sc["depths"]), p(sc["normals"]), p(sc["uvs"]), p(sc["R"]), p(sc["T"]),
                                p(sc["image_r"]), p(sc["image_n"]), p(intr), *[int(v) for v in sc["image_r"].shape],
                                *[int(v) for v in sc["image_n"].shape], p(o[0]), p(o[1]), p(o[2]), p(o[3]))
        return o[0], o[1], o[2], o[3].bool()

    return run


SCENES = {
    "same_size": dict(),
    "other_size": dict(height=96, width=128, neighbour_size=(110, 150), seed=11, shift=(12.0, 6.0)),
    "big_shift": dict(seed=13, shift=(-40.0, 25.0), n=3000),
}


@pytest.mark.parametrize("tag", sorted(SCENES))
def test_kernel_routine_matches_oracle_bit_for_bit(tag, host_routine):
    """Same IEEE arithmetic, two independently written restatements: the kernel's routine (register patch, early
    exits) and the oracle (plain bilinear, all 49 samples, zeroed at the end) must agree exactly."""
    from oracle import oracle as O

    sc = ncc_scenes.make_pair(**SCENES[tag])
    want = O.warp_patch_ncc(*ncc_scenes.args_of(sc))
    got = host_routine(sc)
    assert int(want[3].sum()) > len(want[3]) // 2 and int((~want[3]).sum()) > 20, "scene must hold both kinds"
    assert torch.equal(got[3], want[3])
    for a, b, name in zip(got[:3], want[:3], ("ncc", "grad_depths", "grad_normals")):
        assert torch.equal(a, b), name


def test_invalid_patches_are_zero_and_follow_the_rules():
    from oracle import oracle as O

    sc = ncc_scenes.make_pair()
    ncc, gd, gn, valid = O.warp_patch_ncc(*ncc_scenes.args_of(sc))
    h, w = sc["image_r"].shape
    u, v = sc["uvs"][:, 0], sc["uvs"][:, 1]
    border = (u < 2) | (u > w - 3) | (v < 2) | (v > h - 3)  # u - 1.5 > 0 and u + 1.5 < w - 1
    assert int(border.sum()) > 0 and not bool(valid[border].any())
    band = (v >= h // 2 - 4) & (v < h // 2 + 4) & (u >= 2) & (u < w // 3 - 2)  # 5x5 pixels all equal: zero variance
    assert int(band.sum()) > 0 and not bool(valid[band].any())
    for t in (ncc, gd, gn):
        assert float(t[~valid].abs().max()) == 0.0
    assert bool((ncc[valid] >= 0).all()) and bool((ncc[valid] <= 1.0 + 1e-4).all())  # squared correlation


def test_oracle_value_and_derivatives_against_float64_model():
    """The forward-mode derivative the reference accumulates (d ncc / d depth, d ncc / d normal) against central
    differences of a float64 model written from the definition."""
    from oracle import oracle as O

    sc = ncc_scenes.make_pair(n=600, seed=5)
    ncc, gd, gn, valid = O.warp_patch_ncc(*ncc_scenes.args_of(sc))
    d0, n0 = sc["depths"].double().numpy(), sc["normals"].double().numpy()
    base = model_ncc(sc, d0, n0)
    ok = valid.numpy()
    # float32 sums of 49 terms, then variance = sum of squares - square of sum / 49: the cancellation costs a few 1e-4
    # on low-contrast patches (measured: median 9e-6, worst 5.6e-4)
    dv = np.abs(base[ok] - ncc.numpy()[ok])
    assert np.median(dv) < 5e-5 and dv.max() < 3e-3, (np.median(dv), dv.max())
    step = 1e-6
    fd_d = (model_ncc(sc, d0 + step, n0) - model_ncc(sc, d0 - step, n0)) / (2 * step)
    scale = np.abs(fd_d[ok]).max()
    err = np.abs(fd_d[ok] - gd.numpy()[ok]) / scale
    assert np.median(err) < 1e-4 and np.quantile(err, 0.95) < 2e-3, (np.median(err), np.quantile(err, 0.95))
    for c in range(3):
        e = np.zeros(3)
        e[c] = step
        fd = (model_ncc(sc, d0, n0 + e) - model_ncc(sc, d0, n0 - e)) / (2 * step)
        scale = np.abs(fd[ok]).max()
        err = np.abs(fd[ok] - gn.numpy()[ok, c]) / scale
        assert np.median(err) < 1e-4 and np.quantile(err, 0.95) < 2e-3, (c, np.median(err), np.quantile(err, 0.95))


def test_library_exports_what_the_header_declares():
    """C ABI of libggs_patch_ncc.so: every function include/ggs_patch_ncc.h declares is exported, the Python module
    loads it, and it refuses to compute without a CUDA device instead of falling back."""
    import warp_patch_ncc as ours

    with open(os.path.join(H.ROOT, "include", "ggs_patch_ncc.h")) as fh:
        text = re.sub(r"/\*.*?\*/", "", fh.read(), flags=re.S)
    declared = set(re.findall(r"\b(ggs_ncc_\w+)\s*\(", text))
    assert declared == set(ours.EXPORTS), declared ^ set(ours.EXPORTS)
    lib = C.CDLL(ours.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ggs_ncc_abi_version() == ours.ABI_VERSION
    assert C.sizeof(ours.GgsNccPair) == 8 + 8 + 8 * 4 + 4 * 4
    sc = ncc_scenes.make_pair(n=8)
    with pytest.raises(RuntimeError, match="CUDA"):
        ours.warp_patch_ncc(*ncc_scenes.args_of(sc), False)
    with pytest.raises(RuntimeError, match="normals must have dimensions"):
        ours.warp_patch_ncc(sc["depths"], sc["normals"].reshape(-1), *ncc_scenes.args_of(sc)[2:], False)


@pytest.mark.parametrize("n", [1, 127, 128, 129, 700])
def test_kernel_source_under_emulation_matches_oracle(n, tmp_path_factory):
    """warp_patch_kernel's SOURCE run on the host (tests/native/cuda_emu.h: one OS thread per CUDA thread, real
    barriers), launched as ggs_ncc_warp_patch launches it: shared-memory staging of the [P,3] rows, partial last
    block, every output element written."""
    from oracle import oracle as O

    so = str(tmp_path_factory.mktemp("ncc_emu") / "patch_ncc_emulated.so")
    native = os.path.join(H.ROOT, "tests", "native")
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-pthread", "-shared", "-fPIC", "-I" + CSRC, "-I" + native,
                           "-o", so, os.path.join(native, "patch_ncc_emulated.cpp")])
    lib = C.CDLL(so)
    sc = ncc_scenes.make_pair(n=n, seed=31)
    want = O.warp_patch_ncc(*ncc_scenes.args_of(sc))
    out = [torch.full((n,), 7.0), torch.full((n,), 7.0), torch.full((n, 3), 7.0), torch.full((n,), 9, dtype=torch.uint8)]
    intr = torch.tensor([sc[k] for k in INTR], dtype=torch.float32)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    lib.ggs_test_ncc_kernel(C.c_longlong(n), p(sc["depths"]), p(sc["normals"]), p(sc["uvs"]), p(sc["R"]), p(sc["T"]),
                            p(sc["image_r"]), p(sc["image_n"]), p(intr), *[int(v) for v in sc["image_r"].shape],
                            *[int(v) for v in sc["image_n"].shape], p(out[0]), p(out[1]), p(out[2]), p(out[3]))
    assert int(out[3].max()) <= 1
    assert torch.equal(out[3].bool(), want[3])
    for a, b, name in zip(out[:3], want[:3], ("ncc", "grad_depths", "grad_normals")):
        assert torch.equal(a, b), name


NCC_GOLDEN = {
    "same_size": dict(n=4000),
    "other_size": dict(height=96, width=128, neighbour_size=(110, 150), seed=11, shift=(12.0, 6.0), n=4000),
    "big_shift": dict(seed=13, shift=(-40.0, 25.0), n=3000),
}


@pytest.mark.parametrize("tag", sorted(NCC_GOLDEN))
def test_oracle_and_kernel_routine_match_the_reference_kernel_source_bit_for_bit(tag, host_routine):
    """tests/golden/ncc/*.npz are outputs of the reference's OWN kernel source executed on the host under the CUDA
    shim (tests/golden/make_ncc_golden.py; IEEE arithmetic, i.e. the reference's algorithm without fast-math).  The C
    oracle and the kernel's routine, both IEEE on the host, must reproduce them exactly: this is what pins them."""
    from oracle import oracle as O

    gold = np.load(os.path.join(H.ROOT, "tests", "golden", "ncc", tag + ".npz"))
    assert str(gold["scene"]) == repr(sorted(NCC_GOLDEN[tag].items()))
    sc = ncc_scenes.make_pair(**NCC_GOLDEN[tag])
    for arm, got in (("oracle", O.warp_patch_ncc(*ncc_scenes.args_of(sc))), ("kernel routine", host_routine(sc))):
        assert np.array_equal(got[3].numpy(), gold["valid"]), arm
        assert int(gold["valid"].sum()) > len(gold["valid"]) // 2
        for t, key in zip(got[:3], ("ncc", "grad_depths", "grad_normals")):
            assert np.array_equal(t.numpy(), gold[key]), (arm, key)
