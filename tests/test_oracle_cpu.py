"""The C oracle (oracle/ggs_oracle.c) against the golden vectors produced by the reference CUDA rasterizer
itself (tests/golden/*.npz, generated on a B200 by tests/golden/make_golden.py), plus internal consistency.

The oracle uses IEEE float arithmetic, the reference was compiled with --use_fast_math: integers must agree
exactly on these fixtures, floats to ~1e-3 of the image / gradient scale (see DESIGN.md "oracle pinning")."""
import glob
import os

import numpy as np
import pytest
import torch

import helpers as H
import scenes
from oracle.oracle import Oracle, lib, scene_kwargs
from golden.make_golden import SCENES

GOLDEN = sorted(glob.glob(os.path.join(H.ROOT, "tests", "golden", "*.npz")))


def test_golden_present():
    assert len(GOLDEN) >= 2, "golden vectors missing: run tests/golden/make_golden.py on a GPU box"


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_matches_reference_golden(path):
    tag = os.path.basename(path)[:-4]
    gold = np.load(path)
    sc = scenes.make_scene(**SCENES[tag])
    orc = Oracle()
    out = orc.forward(**scene_kwargs(sc))
    it = orc.intermediates()
    assert out["num_rendered"] == int(gold["meta_D"])
    assert np.array_equal(out["radii"].numpy(), gold["out_radii"])
    for k in ("tiles_touched", "point_offsets", "point_list", "max_contributor"):
        assert np.array_equal(it[k].astype(np.int64), gold["int_" + k].astype(np.int64)), k
    # keys: the tile half is exact; the depth half is norm3df (fast-math) vs IEEE sqrt: at most 2 ulp apart
    ok, gk = it["sorted_keys"].astype(np.uint64), gold["int_sorted_keys"].astype(np.uint64)
    assert np.array_equal(ok >> np.uint64(32), gk >> np.uint64(32))
    lo = (ok & np.uint64(0xFFFFFFFF)).astype(np.int64) - (gk & np.uint64(0xFFFFFFFF)).astype(np.int64)
    assert np.abs(lo).max() <= 2
    assert np.array_equal(it["ranges"].astype(np.int64), gold["int_ranges"].astype(np.int64))
    bad = int((it["n_contrib"].astype(np.int64) != gold["int_n_contrib"].astype(np.int64)).sum())
    assert bad <= 2, f"n_contrib differs on {bad} pixels"
    for k, tol in (("color", 2e-3), ("alpha", 2e-3), ("mdepth", 2e-3), ("normal", 2e-2)):
        err, _ = H.image_rel_err(out[k], torch.from_numpy(gold["out_" + k]))
        assert err < tol, (k, err)
    g = sc.grads
    gr = orc.backward(g["color"], g["mdepth"], g["alpha"], g["normal"])
    for k in H.GRAD_NAMES:
        ref = torch.from_numpy(gold["grad_" + k])
        err = H.grad_rel_err(gr[k].reshape(ref.shape), ref)
        assert err < 2e-2, (k, err)


def test_higher_msb_is_the_reference_formula():
    L = lib()
    # rasterizer_impl.cu:37-50: 256 tiles -> 9, 1900 -> 11, 4056 -> 12, 2170 -> 12
    assert [L.ggs_oracle_higher_msb(n) for n in (256, 1900, 4056, 2170, 1, 2, 3)] == [9, 11, 12, 12, 1, 2, 2]


def test_oracle_binning_invariants():
    sc = scenes.make_scene("c1")
    orc = Oracle()
    out = orc.forward(**scene_kwargs(sc))
    it = orc.intermediates()
    D = out["num_rendered"]
    assert D == int(it["tiles_touched"].sum()) == int(it["point_offsets"][-1])
    assert np.array_equal(np.cumsum(it["tiles_touched"], dtype=np.uint64).astype(np.uint32), it["point_offsets"])
    keys = it["sorted_keys"]
    assert np.all(keys[1:] >= keys[:-1])
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    for t in np.unique(tiles)[:50]:
        lo, hi = it["ranges"][t]
        assert np.all(tiles[lo:hi] == t) and (lo == 0 or tiles[lo - 1] != t) and (hi == D or tiles[hi] != t)
    # radii > 0 exactly where the Gaussian touches tiles
    assert np.array_equal(out["radii"].numpy() > 0, it["tiles_touched"] > 0)
    # alpha in [0,1), depth zero wherever the pixel is not opaque enough for the median search
    a = out["alpha"].numpy()
    assert a.min() >= 0 and a.max() < 1
    assert np.all(out["mdepth"].numpy()[a < 0.55 - 1e-6] == 0)


def test_oracle_edge_cases():
    orc = Oracle()
    sc = scenes.make_scene("tiny", P=64)
    # everything behind the camera: nothing rendered, background shows through
    sc.means3D[:, 2] = -10.0
    sc.bg = torch.tensor([0.25, 0.5, 0.75])
    out = orc.forward(**scene_kwargs(sc))
    assert out["num_rendered"] == 0 and int(out["radii"].abs().sum()) == 0
    assert torch.allclose(out["color"][1], torch.full_like(out["color"][1], 0.5))
    assert float(out["alpha"].abs().max()) == 0
    # require_depth=False leaves depth and normal at zero
    sc2 = scenes.make_scene("tiny")
    o2 = Oracle().forward(require_depth=False, **scene_kwargs(sc2))
    assert float(o2["mdepth"].abs().max()) == 0 and float(o2["normal"].abs().max()) == 0
    o3 = Oracle().forward(**scene_kwargs(sc2))
    assert torch.equal(o2["color"], o3["color"]) and torch.equal(o2["alpha"], o3["alpha"])
