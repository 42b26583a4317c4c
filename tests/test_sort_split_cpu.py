"""Host-side statement of the invariant the binning relies on: emitting the (tile, Gaussian) pairs in
the stable depth order of the Gaussians and then sorting them STABLY on the tile alone gives exactly the order of the
reference's stable 64-bit sort of (tile << 32 | depth bits) over pairs emitted in Gaussian order
(rasterizer_impl.cu:70-107, 406-412).  Pure numpy; the CUDA path is checked against the compiled reference in
test_parity_gpu.py."""
import numpy as np
import pytest


def _reference_order(rect_lo, rect_hi, depth_bits, gx):
    keys, vals = [], []
    for g in range(len(depth_bits)):
        for y in range(rect_lo[g, 1], rect_hi[g, 1]):
            for x in range(rect_lo[g, 0], rect_hi[g, 0]):
                keys.append(((y * gx + x) << 32) | int(depth_bits[g]))
                vals.append(g)
    keys = np.array(keys, dtype=np.uint64)
    vals = np.array(vals, dtype=np.uint32)
    order = np.argsort(keys, kind="stable")  # SortPairs: stable, ascending
    return keys[order], vals[order]


def _split_order(rect_lo, rect_hi, depth_bits, gx):
    touched = (rect_hi[:, 0] - rect_lo[:, 0]) * (rect_hi[:, 1] - rect_lo[:, 1])
    pre = np.where(touched > 0, depth_bits.astype(np.uint64), np.uint64(0xFFFFFFFF))  # culled Gaussians last
    depth_order = np.argsort(pre, kind="stable")
    tiles, vals = [], []
    for g in depth_order:
        for y in range(rect_lo[g, 1], rect_hi[g, 1]):
            for x in range(rect_lo[g, 0], rect_hi[g, 0]):
                tiles.append(y * gx + x)
                vals.append(g)
    tiles = np.array(tiles, dtype=np.uint32)
    vals = np.array(vals, dtype=np.uint32)
    order = np.argsort(tiles, kind="stable")  # the tile passes
    tiles, vals = tiles[order], vals[order]
    keys = (tiles.astype(np.uint64) << np.uint64(32)) | depth_bits[vals].astype(np.uint64)  # what the last pass writes
    return keys, vals


@pytest.mark.parametrize("seed", range(6))
def test_depth_presort_plus_stable_tile_sort_equals_the_64_bit_sort(seed):
    rng = np.random.default_rng(seed)
    P, gx, gy = 400, 13, 9
    lo = np.stack([rng.integers(0, gx, P), rng.integers(0, gy, P)], 1)
    ext = np.stack([rng.integers(0, 4, P), rng.integers(0, 4, P)], 1)  # zero extent = culled
    hi = np.minimum(lo + ext, [gx, gy])
    # few distinct depths: plenty of ties, which must keep Gaussian-index order
    depth = rng.choice(np.float32([0.25, 0.5, 1.0, 1.5, 3.0, 3.0000002, 7.5, 90.0]), P).astype(np.float32)
    bits = depth.view(np.uint32)
    k_ref, v_ref = _reference_order(lo, hi, bits, gx)
    k_new, v_new = _split_order(lo, hi, bits, gx)
    assert np.array_equal(k_ref, k_new)
    assert np.array_equal(v_ref, v_new)
