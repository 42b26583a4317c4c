"""The hand-written onesweep radix sort (ggs_sort_pairs, replacement of cub::DeviceRadixSort::SortPairs at
rasterizer_impl.cu:406-412): stable, ascending, restricted to bits [begin, end)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _sort(keys, vals, b0, b1):
    from diff_gaussian_rasterization import _abi

    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    nbytes = _abi.lib.ggs_sort_scratch_bytes(n, b0, b1)
    scratch = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=keys.device)
    rc = _abi.lib.ggs_sort_pairs(_abi.ptr(keys), _abi.ptr(vals), _abi.ptr(ko), _abi.ptr(vo), n, b0, b1,
                                 C.c_void_p(scratch.data_ptr()), _abi.stream_handle(keys.device))
    _abi.check(rc, "ggs_sort_pairs")
    torch.cuda.synchronize()
    return ko, vo


@pytest.mark.parametrize("n", [0, 1, 2, 255, 256, 3071, 3072, 3073, 100_003, 4_000_001])
@pytest.mark.parametrize("bits", [(0, 44), (0, 41), (32, 44), (0, 64), (0, 8), (5, 23)])
def test_sort_pairs_matches_stable_sort(n, bits, device):
    b0, b1 = bits
    g = torch.Generator(device="cpu").manual_seed(n * 131 + b0 * 7 + b1)
    keys = torch.randint(-(2 ** 63), 2 ** 63 - 1, (n,), generator=g, dtype=torch.int64)
    if n > 10:
        keys[n // 2:] = keys[: n - n // 2] & ~((1 << 20) - 1)  # many ties on the low bits
        keys[:: 7] = keys[0]
    keys = keys.to(device)
    vals = torch.arange(n, dtype=torch.int32, device=device)
    ko, vo = _sort(keys, vals, b0, b1)
    width = b1 - b0
    field = (keys >> b0) & ((1 << width) - 1) if width < 64 else keys ^ (-(2 ** 63))
    if width == 64:
        # unsigned order of the full 64 bits == signed order after flipping the sign bit
        order = torch.sort(field, stable=True).indices
    else:
        order = torch.sort(field, stable=True).indices
    assert torch.equal(ko, keys[order])
    assert torch.equal(vo, vals[order])


def test_sort_all_equal_and_presorted(device):
    n = 50_000
    keys = torch.full((n,), 0x0000_0123_4567_89AB, dtype=torch.int64, device=device)
    vals = torch.arange(n, dtype=torch.int32, device=device)
    ko, vo = _sort(keys, vals, 0, 44)
    assert torch.equal(vo, vals) and torch.equal(ko, keys)
    keys = torch.arange(n, dtype=torch.int64, device=device).flip(0)
    ko, vo = _sort(keys, vals, 0, 44)
    assert torch.equal(ko, keys.flip(0)) and torch.equal(vo, vals.flip(0))
