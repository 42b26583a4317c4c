"""ggs_sort_pairs_static: the stand-alone radix sort with its scratch in a static device array inside the library
(include/ggs_rasterizer.h; no counterpart in the reference -- added for callers that cannot call cudaMalloc).

Collected last on purpose.  The checks: the same bytes as ggs_sort_pairs (bit-exact,
the tolerance of tests/test_sort_gpu.py) on calls in a row that share the array -- a long sort, then a short one over a
different bit range, then the long one again, so that stale histograms / look-back descriptors would show -- no device
allocation during the calls, and a clear refusal beyond the stated maximum."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _pairs(n, seed, device):
    g = torch.Generator(device="cpu").manual_seed(seed)
    keys = torch.randint(-(2 ** 63), 2 ** 63 - 1, (n,), generator=g, dtype=torch.int64)
    if n > 10:
        keys[n // 2:] = keys[: n - n // 2] & ~((1 << 20) - 1)  # many ties on the low bits: stability matters
        keys[::7] = keys[0]
    return keys.to(device), torch.arange(n, dtype=torch.int32, device=device)


def _sort_with_caller_scratch(keys, vals, b0, b1):
    from diff_gaussian_rasterization import _abi

    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    scratch = torch.empty(max(_abi.lib.ggs_sort_scratch_bytes(n, b0, b1), 256), dtype=torch.uint8, device=keys.device)
    rc = _abi.lib.ggs_sort_pairs(_abi.ptr(keys), _abi.ptr(vals), _abi.ptr(ko), _abi.ptr(vo), n, b0, b1,
                                 C.c_void_p(scratch.data_ptr()), _abi.stream_handle(keys.device))
    _abi.check(rc, "ggs_sort_pairs")
    torch.cuda.synchronize()
    return ko, vo


def _sort_static(keys, vals, ko, vo, b0, b1):
    """Outputs are the caller's, made beforehand: nothing is allocated between the two memory readings."""
    from diff_gaussian_rasterization import _abi

    stream = _abi.stream_handle(keys.device)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(keys.device)["allocation.all.allocated"]
    free_before, _ = torch.cuda.mem_get_info(keys.device)
    rc = _abi.lib.ggs_sort_pairs_static(_abi.ptr(keys), _abi.ptr(vals), _abi.ptr(ko), _abi.ptr(vo), keys.numel(), b0, b1,
                                        stream)
    _abi.check(rc, "ggs_sort_pairs_static")
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(keys.device)
    assert torch.cuda.memory_stats(keys.device)["allocation.all.allocated"] == before  # nothing through torch
    return free_before - free_after


def test_static_sort_matches_sort_pairs_on_calls_in_a_row(device):
    from diff_gaussian_rasterization import _abi

    n_max = _abi.SORT_STATIC_MAX_PAIRS
    cases = [(n_max, (0, 64), 1), (3073, (32, 44), 2), (n_max, (0, 64), 3), (100_003, (5, 23), 4), (1, (0, 8), 5),
             (n_max - 1, (0, 44), 6)]
    inputs = [(_pairs(n, seed, device), bits) for n, bits, seed in cases]
    expected = [_sort_with_caller_scratch(k, v, *bits) for (k, v), bits in inputs]
    outputs = [(torch.empty_like(k), torch.empty_like(v)) for (k, v), _ in inputs]
    # the first call on a device loads the kernels and asks for the array's address: driver set-up, not counted
    drops = [_sort_static(k, v, ko, vo, *bits) for ((k, v), bits), (ko, vo) in zip(inputs, outputs)]
    for (ko, vo), (ke, ve), case in zip(outputs, expected, cases):
        assert torch.equal(ko, ke), case
        assert torch.equal(vo, ve), case
    assert all(d <= 0 for d in drops[1:]), drops  # device memory in use did not grow over the later calls
    # and once more over the first input, after all the others have been through the array
    (k, v), bits = inputs[0]
    ko, vo = torch.empty_like(k), torch.empty_like(v)
    assert _sort_static(k, v, ko, vo, *bits) <= 0
    assert torch.equal(ko, expected[0][0]) and torch.equal(vo, expected[0][1])


def test_static_sort_empty_and_limits(device):
    from diff_gaussian_rasterization import _abi

    stream = _abi.stream_handle(device)
    assert _abi.lib.ggs_sort_pairs_static(None, None, None, None, 0, 0, 44, stream) == 0
    n = _abi.SORT_STATIC_MAX_PAIRS + 1
    keys, vals = _pairs(n, 9, device)
    ko, vo = torch.full_like(keys, -7), torch.full_like(vals, -7)
    rc = _abi.lib.ggs_sort_pairs_static(_abi.ptr(keys), _abi.ptr(vals), _abi.ptr(ko), _abi.ptr(vo), n, 0, 44, stream)
    assert rc == -1
    msg = _abi.last_error()
    assert "GGS_SORT_STATIC_MAX_PAIRS" in msg and str(_abi.SORT_STATIC_MAX_PAIRS) in msg
    torch.cuda.synchronize()
    assert bool((ko == -7).all()) and bool((vo == -7).all())  # refused before anything was enqueued
    assert _abi.lib.ggs_sort_pairs_static(_abi.ptr(keys), _abi.ptr(vals), _abi.ptr(ko), _abi.ptr(vo), 10, 8, 8,
                                          stream) == -1
    assert _abi.lib.ggs_sort_pairs_static(None, None, None, None, 10, 0, 8, stream) == -1
