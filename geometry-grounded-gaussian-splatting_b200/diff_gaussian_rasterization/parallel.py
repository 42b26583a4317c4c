"""Multi-GPU plumbing for the two places the path shards; the reference has no
distributed code at all, so nothing here replaces a reference file.

* view-parallel batch: every rank rasterizes its own view of the replicated Gaussian model; the parameter
  gradients are summed with ONE all-reduce over a flat bucket (`GradientBucket`).
* mesh-extraction queries: `integrate` is independent per query point, so points are split into contiguous
  ranges, one per rank, with no data-path collective; `integrate_sharded` gathers the per-range results.

One process per GPU (`torchrun`), `torch.distributed` with the NCCL backend on the box; the same code runs on
`gloo` with CPU tensors in the tests (the collective plumbing only -- the rasterizer itself has no CPU path).
"""
from __future__ import annotations

import contextlib
from typing import Callable, Sequence

import torch
import torch.distributed as dist


@contextlib.contextmanager
def factored_sh_gradients(group=None):
    """Inside this context the rasterizer's backward returns the SH-coefficient gradient already SUMMED over the
    views of all ranks of `group`: instead of all-reducing [P, SHM, 3] (81 % of the parameter-gradient bytes at SH3)
    every rank all-gathers the 3-float colour gradient per Gaussian and view and rebuilds the rank-one per-view SH
    gradients locally.  Reduce the remaining parameters with `GradientBucket` as usual and leave `shs` out of it."""
    from . import _SH_FACTORING

    prev = dict(_SH_FACTORING)
    _SH_FACTORING.update(on=True, group=group)
    try:
        yield
    finally:
        _SH_FACTORING.update(prev)


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous range [lo, hi) of `n` items owned by `rank`; sizes differ by at most one and cover [0, n)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(int(n), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class GradientBucket:
    """Flat fp32 bucket holding the gradients of a fixed list of parameters; `allreduce` sums it over the
    group in one collective and scatters the result back into the `.grad` tensors (views into the bucket, so no
    copy on the way back)."""

    def __init__(self, params: Sequence[torch.Tensor]):
        self.params = list(params)
        if not self.params:
            raise ValueError("no parameters")
        dev, dt = self.params[0].device, self.params[0].dtype
        self.sizes = [p.numel() for p in self.params]
        self.flat = torch.zeros(sum(self.sizes), dtype=dt, device=dev)
        self.views = []
        o = 0
        for p, n in zip(self.params, self.sizes):
            self.views.append(self.flat[o:o + n].view(p.shape))
            o += n

    def pack(self) -> torch.Tensor:
        """Copy the current `.grad`s into the bucket (missing gradients count as zero)."""
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                v.copy_(p.grad)
        return self.flat

    def allreduce(self, group=None, average: bool = False) -> torch.Tensor:
        self.pack()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            if average:
                self.flat.div_(dist.get_world_size(group))
        for p, v in zip(self.params, self.views):
            p.grad = v
        return self.flat


def integrate_sharded(integrate: Callable[[torch.Tensor], tuple[torch.Tensor, torch.Tensor]],
                      points3D: torch.Tensor, group=None) -> tuple[torch.Tensor, torch.Tensor]:
    """Run `integrate(points[lo:hi])` on this rank's contiguous point range and all-gather `(alpha, inside)` for
    all points.  `integrate` is e.g. `lambda p: rasterizer.integrate(p, means3D, opacities, scales, rotations)`
    (mesh_extract_tetrahedra.py:75).  With no process group it is the plain call."""
    n = points3D.shape[0]
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return integrate(points3D)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_range(n, rank, world)
    alpha, inside = integrate(points3D[lo:hi])
    cap = (n + world - 1) // world
    # equal-sized slots so one all_gather_into_tensor suffices; 5 bytes per point on the wire
    a_pad = torch.zeros(cap, dtype=alpha.dtype, device=alpha.device)
    i_pad = torch.zeros(cap, dtype=torch.uint8, device=alpha.device)
    a_pad[:hi - lo] = alpha
    i_pad[:hi - lo] = inside.to(torch.uint8)
    a_all = torch.empty(world * cap, dtype=alpha.dtype, device=alpha.device)
    i_all = torch.empty(world * cap, dtype=torch.uint8, device=alpha.device)
    dist.all_gather_into_tensor(a_all, a_pad, group=group)
    dist.all_gather_into_tensor(i_all, i_pad, group=group)
    out_a = torch.empty(n, dtype=alpha.dtype, device=alpha.device)
    out_i = torch.empty(n, dtype=torch.bool, device=alpha.device)
    for r in range(world):
        rlo, rhi = shard_range(n, r, world)
        out_a[rlo:rhi] = a_all[r * cap:r * cap + (rhi - rlo)]
        out_i[rlo:rhi] = i_all[r * cap:r * cap + (rhi - rlo)].bool()
    return out_a, out_i
