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
import os
from typing import Callable, Sequence

import torch
import torch.distributed as dist


@contextlib.contextmanager
def factored_sh_gradients(group=None, exchange=None):
    """Inside this context the rasterizer's backward returns the SH-coefficient gradient already SUMMED over the
    views of all ranks of `group`: instead of all-reducing [P, SHM, 3] (81 % of the parameter-gradient bytes at SH3)
    every rank all-gathers the 3-float colour gradient per Gaussian and view and rebuilds the rank-one per-view SH
    gradients locally.  Reduce the remaining parameters with `GradientBucket` as usual and leave `shs` out of it.

    With `exchange` (a `PeerGradientExchange`) nothing is left to reduce: the backward returns EVERY parameter
    gradient summed over the views, formed by one fused launch over peer memory (no NCCL call on the data path)."""
    from . import _SH_FACTORING

    prev = dict(_SH_FACTORING)
    _SH_FACTORING.update(on=True, group=group, exchange=exchange)
    try:
        yield
    finally:
        _SH_FACTORING.update(prev)


def exchange_layout(P: int, world: int) -> dict:
    """Float offsets of the regions of a `PeerGradientExchange` buffer (`GgsPeerExchange` in the C ABI).  Every
    region starts on a 256-byte boundary; rotations come first because their rows are stored as float4."""
    pad = lambda n: (int(n) + 63) // 64 * 64  # noqa: E731
    L, o = {}, 0
    for name, width in (("rotations", 4), ("means3D", 3), ("scales", 3), ("opacities", 1)):
        L[name] = (o, width)
        o += pad(width * P)
    L["send"], L["reduce"] = 0, o
    L["recv"] = o
    o += L["reduce"]
    L["rgb_stage"], L["rgb_stride"] = o, pad(3 * P)
    o += world * L["rgb_stride"]
    L["campos_stage"] = o
    o += pad(4 * world)
    L["reduce_stride"] = pad((L["reduce"] // 4 + world - 1) // world * 4)
    L["reduce_stage"] = o
    o += world * L["reduce_stride"]
    L["total"] = o
    return L


class ResultLease:
    """Guards results handed out as views into persistent buffers: `check()` raises if tensors of the previous hand-out
    are still alive when the buffers are about to be overwritten (autograd's AccumulateGrad would then add the new sum
    to a `.grad` that aliases it: twice the new gradient, the old one lost)."""

    def __init__(self, what: str):
        import weakref

        self._weakref = weakref
        self.what = what
        self._refs = []

    def alive(self) -> int:
        return sum(1 for r in self._refs if r() is not None)

    def check(self):
        n = self.alive()
        if n:
            raise RuntimeError(
                f"{self.what}: {n} gradient tensor(s) of the previous backward are still referenced and alias the "
                "exchange buffers the next backward overwrites.  Drop them first (optimizer.zero_grad(set_to_none=True) "
                "/ p.grad = None), or build the exchange with zero_copy=False (results are then cloned out of the buffers).")

    def hand_out(self, tensors):
        self._refs = [self._weakref.ref(t) for t in tensors]
        return tensors


class PeerGradientExchange:
    """View-parallel gradient sums through peer memory (NVLink / NVSwitch) instead of NCCL collectives.

    Every rank allocates one symmetric buffer (`torch.distributed._symmetric_memory`: cuMem allocations whose handles
    are exchanged once, here) laid out by `exchange_layout`.  Per step, inside the rasterizer's backward:
      1. `ggs_rasterize_backward_peer`: the per-Gaussian backward writes this rank's gradients of means3D / opacities /
         scales / rotations straight into the buffer; the masked colour gradient of the view is pushed into every
         rank's colour stage while that kernel still runs, and each 1/world slice of the gradient region is pushed to
         the rank that owns it -- all posted stores, one multimem.st per 16 bytes where the switch can replicate;
      2. node barrier; `ggs_view_sum_peer` (all reads local): sums the slots of this rank's slice and stores the sum
         into every rank's result region, and forms the SH gradient summed over views from the staged colour
         gradients; node barrier.
    Wire bytes received per rank and step: 12 B x P x (world - 1) colour gradients + 2 x 44 B x P x (world - 1) / world."""

    def __init__(self, P: int, device, group=None, multicast: bool | None = None, pull_reduce: bool | None = None,
                 zero_copy: bool = False, chunks: int | None = None):
        """`zero_copy=False` (default): the summed gradients are cloned out of the exchange buffers (236 B per
        Gaussian at SH3), so they may be kept, accumulated into, or outlive the next backward like any gradient.
        `zero_copy=True`: they are views into the buffers, valid until the next backward through this exchange; that
        backward raises if they are still referenced (`ResultLease`).  The buffers are sized for `P` Gaussians: after
        densification / pruning build a new exchange (`resized`)."""
        import ctypes as C

        import torch.distributed._symmetric_memory as symm

        from . import _abi

        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("PeerGradientExchange needs an initialised process group")
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        if self.world > 16:
            raise RuntimeError("at most 16 ranks (one node)")
        self.P, self.device = int(P), torch.device(device)
        self.zero_copy = bool(zero_copy)
        self.sh_after_backward = os.environ.get("GGS_PEER_SH_LATE", "1") == "1"
        self._want_fused_push = os.environ.get("GGS_PEER_FUSE_RGB", "1") == "1"
        self._lease = ResultLease("PeerGradientExchange(zero_copy=True)")
        self._args = dict(multicast=multicast, pull_reduce=pull_reduce, zero_copy=zero_copy, chunks=chunks)
        L = self.layout = exchange_layout(self.P, self.world)
        self.buf = symm.empty(L["total"], dtype=torch.float32, device=self.device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, self.group)
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        if len(ptrs) != self.world or any(p == 0 for p in ptrs):
            raise RuntimeError("symmetric-memory rendezvous returned no peer pointers")
        self._ptrs = (C.c_uint64 * self.world)(*ptrs)
        mc = 0
        try:
            if (multicast is None or multicast) and self.hdl.has_multicast_support:
                mc = int(self.hdl.multicast_ptr)
        except Exception:  # noqa: BLE001  (no multicast attributes)
            mc = 0
        if multicast and not mc:
            raise RuntimeError("no NVSwitch multicast mapping for the exchange buffer")
        self.multicast_ptr = mc
        # slices summed inside the switch (multimem.ld_reduce) where the buffer has a multicast mapping: no slice push,
        # bit-identical to the NCCL sums at 2 and 8 ranks (profiles/r02_peer_check_n8_c3.log); posted stores otherwise
        if pull_reduce is None:
            pull_reduce = bool(mc)
        self.pull_reduce = bool(pull_reduce)
        self.px = _abi.GgsPeerExchange(self.world, self.rank, C.cast(self._ptrs, C.POINTER(C.c_uint64)), mc, L["send"],
                                       L["recv"], L["reduce"], L["rgb_stage"], L["rgb_stride"], L["campos_stage"],
                                       L["reduce_stage"], L["reduce_stride"], int(bool(pull_reduce)))
        # the per-Gaussian backward pushes the colour gradient itself (no second pass over the accumulators next to
        # it) where the SH sum waits for that kernel anyway and the switch can replicate the stores
        self.fused_push = bool(self._want_fused_push and self.sh_after_backward and mc)
        self.px.fuse_rgb_push = int(self.fused_push)
        # Pipelining: with the switch-reduced sum the per-Gaussian backward runs in pieces and the links sum piece k
        # while the SMs compute piece k + 1 (GGS_PEER_CHUNKS).  Default 1 = one piece, summed after the backward: at 8 GPUs two
        # pieces tie and four lose (C3 2656 / 2648 / 2617 iters/s, C4 6.64x / 6.71x / 6.59x for 1 / 2 / 4 pieces) -- every piece
        # pays a node barrier of its own and the switch sums 44 MB in ~110 us whatever the pieces
        if chunks is None:
            chunks = int(os.environ.get("GGS_PEER_CHUNKS", "1"))
        self.chunks = max(1, min(int(chunks), _abi.MAX_CHUNKS)) if (self.pull_reduce and mc and self.P >= 128 * 8) else 1
        self._chunk_events, self._chunk_ranges = [], []
        if self.chunks > 1:
            ends, self._chunk_ranges = chunk_ranges(self.P, self.world, self.rank, self.chunks, L)
            with torch.cuda.device(self.device):
                for k in range(self.chunks):
                    ev = torch.cuda.Event()
                    ev.record()  # creates the underlying cudaEvent_t
                    self._chunk_events.append(ev)
                    self.px.chunk_end[k] = ends[k]
                    self.px.chunk_events[k] = int(ev.cuda_event)
            self.px.chunks = self.chunks
            self.sum_stream = torch.cuda.Stream(self.device)
        # the colour push and the SH half of the sum run on this stream, concurrent with the per-Gaussian backward
        self.side = torch.cuda.Stream(self.device)
        self._d_sh = None  # persistent: written on the side stream, so it must not come from the caching allocator
        torch.cuda.synchronize(self.device)
        self.hdl.barrier(channel=0)

    def check(self, P, device):
        """Called when a backward starts: the buffers must fit, and nothing may still alias what is about to be
        overwritten."""
        if int(P) != self.P or torch.device(device) != self.device:
            raise RuntimeError(f"PeerGradientExchange was built for {self.P} Gaussians on {self.device}, the backward has "
                               f"{int(P)} on {torch.device(device)}: after densification / pruning build a new one "
                               "(exchange.resized(P)) on every rank")
        if self.zero_copy:
            self._lease.check()

    def resized(self, P: int) -> "PeerGradientExchange":
        """A new exchange for `P` Gaussians with the same options (collective: call it on every rank of the group)."""
        return PeerGradientExchange(P, self.device, self.group, **self._args)

    def _view(self, base, name):
        o, w = self.layout[name]
        return self.buf[base + o:base + o + w * self.P].view(self.P, w)

    def send_views(self):
        """(d_means3D, d_opacity, d_scales, d_rotations) of THIS rank's view, inside the exchange buffer."""
        s = self.layout["send"]
        return (self._view(s, "means3D"), self._view(s, "opacities"), self._view(s, "scales"),
                self._view(s, "rotations"))

    def sum_over_views(self, rs, means3D, SHM):
        """Call after `ggs_rasterize_backward_peer(..., side_stream=self.side)` has been enqueued on the current stream.
        Returns the gradients of (means3D, opacities, scales, rotations, shs) summed over the views of all ranks: clones,
        or with zero_copy=True views into buffers of this object that are valid until the next backward."""
        import ctypes as C

        from . import _abi

        if self._d_sh is None or self._d_sh.shape[1] != SHM:
            self._d_sh = torch.empty(self.P, SHM, 3, dtype=torch.float32, device=self.device)
        d_sh = self._d_sh
        main = torch.cuda.current_stream(self.device)

        def launch(parts, stream):
            _abi.check(_abi.lib.ggs_view_sum_peer(self.P, int(rs.sh_degree), int(SHM), _abi.ptr(means3D),
                                                  C.byref(self.px), _abi.ptr(d_sh), parts,
                                                  C.c_void_p(stream.cuda_stream)), "ggs_view_sum_peer")

        # side stream, behind the colour push and concurrent with the per-Gaussian backward: once every rank's colour
        # gradients have landed, the SH gradient summed over views (all reads local)
        backward_done = main.record_event() if self.sh_after_backward else None
        with torch.cuda.stream(self.side):
            if self.fused_push:
                self.side.wait_event(backward_done)  # the backward kernel itself pushed this view's colour gradient
                self.hdl.barrier(channel=2)          # ... and so did every other rank's
            else:
                self.hdl.barrier(channel=2)
                if backward_done is not None:
                    # both are HBM-bound: rather than share the memory system with the per-Gaussian backward, the SH
                    # sum runs next to the slice sum, which is bound by the links
                    self.side.wait_event(backward_done)
            launch(2, self.side)  # GGS_VIEW_SUM_SH
            sh_done = self.side.record_event()
        if self.chunks > 1:
            # the pieces of the per-Gaussian backward are summed as they finish: node barrier (every rank has written
            # piece k) + this rank's share of piece k through the switch, on a stream of their own
            with torch.cuda.stream(self.sum_stream):
                for k in range(self.chunks):
                    self.sum_stream.wait_event(self._chunk_events[k])
                    self.hdl.barrier(channel=3)
                    mine = self._chunk_ranges[k]
                    lo = (C.c_int64 * len(mine))(*[a for a, _ in mine])
                    hi = (C.c_int64 * len(mine))(*[b for _, b in mine])
                    _abi.check(_abi.lib.ggs_peer_sum_ranges(C.byref(self.px), len(mine), lo, hi,
                                                            C.c_void_p(self.sum_stream.cuda_stream)),
                               "ggs_peer_sum_ranges")
                main.wait_event(self.sum_stream.record_event())
        else:
            # main stream, behind the per-Gaussian backward and the slice push
            self.hdl.barrier(channel=0)  # every rank's slices have landed
            launch(1, main)  # GGS_VIEW_SUM_SLICE
        main.wait_event(sh_done)
        self.hdl.barrier(channel=1)  # result regions complete everywhere; every stage may be overwritten
        r = self.layout["recv"]
        # fresh view objects: autograd's AccumulateGrad adopts a gradient it holds the only reference to, and clones
        # (192 MB for the SH gradient) one that something else still references
        out = (self._view(r, "means3D"), self._view(r, "opacities"), self._view(r, "scales"),
               self._view(r, "rotations"), d_sh.view(self.P, SHM, 3))
        if self.zero_copy:
            return self._lease.hand_out(out)
        return tuple(t.clone() for t in out)


def chunk_ranges(P: int, world: int, rank: int, chunks: int, layout: dict) -> tuple[list, list]:
    """Pipelined sum (GgsPeerExchange.chunks): the per-Gaussian backward runs in `chunks` pieces of the model; piece k =
    Gaussians [ends[k-1], ends[k]), every boundary but the last a multiple of 128.  Returns (ends, ranges) where
    ranges[k] = this rank's share of the 16-byte units of the gradient region that piece k writes -- per parameter array
    one contiguous run split evenly over the ranks -- as [(lo4, hi4), ...] in units of 4 floats from the region's base.
    Over the ranks and pieces every unit of the (padded) arrays is covered exactly once."""
    blocks = (int(P) + 127) // 128
    ends = [min(int(P), (blocks * (k + 1) // chunks) * 128) for k in range(chunks)]
    ends[-1] = int(P)
    ranges, first = [], 0
    for k, last in enumerate(ends):
        mine = []
        for name in ("rotations", "means3D", "scales", "opacities"):
            o, w = layout[name]
            lo = (o + w * first) // 4  # first is a multiple of 128 and o of 64: exact
            # the last piece runs to the end of the array's padded extent (the pad is zero on every rank)
            hi = (o + w * last + 3) // 4 if k + 1 < chunks else (o + (w * int(P) + 63) // 64 * 64) // 4
            n = hi - lo
            a, b = lo + n * rank // world, lo + n * (rank + 1) // world
            if b > a:
                mine.append((a, b))
        ranges.append(mine)
        first = last
    return ends, ranges


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


def prepare_views(views: Sequence, means3D, opacities, scales, rotations, prepare: Callable | None = None) -> list:
    """Bins the Gaussians of every view once (`GaussianRasterizer.prepare`) for any number of `alpha_cull` rounds over
    the same model: the tetrahedral extraction evaluates its vertices, then refines the level set with a binary search
    -- every round queries the same views (mesh_extract_tetrahedra.py).  About 0.3 KB per (view, Gaussian) + 24 B per
    (tile, Gaussian) pair of device memory stay allocated per view."""
    if prepare is None:
        from . import GaussianRasterizer

        def prepare(rs):
            return GaussianRasterizer(rs).prepare(means3D, opacities, scales, rotations)

    return [prepare(rs) for rs in views]


def alpha_cull(points3D: torch.Tensor, views: Sequence, means3D, opacities, scales, rotations, *,
               chunk_size: int = 10_000_000, validate: Callable | None = None, group=None,
               gather: bool = True, prepare: Callable | None = None,
               prepared: Sequence | None = None) -> tuple[torch.Tensor, torch.Tensor]:
    """The opacity test of the tetrahedral mesh extraction over all views (`evaluage_alpha_cull`,
    mesh_extract_tetrahedra.py:65-87): per point the minimum over the views that see it of the integrated opacity.
    Returns `(sdf = 0.5 - min_alpha  [N], any_valid [N] bool)`.

    `views`: one `GaussianRasterizationSettings` per camera.  `validate(view_index, points, inside) -> bool [n]` is the
    reference's `evaluation_validation` (its ground-truth-mask test); None keeps `inside`.

    What differs from the reference's loop, not from its result:
      * view-major instead of chunk-major: the Gaussians of a view are binned ONCE (`GaussianRasterizer.prepare`) and
        every chunk is queried against that state, where the reference re-runs preprocess + sort for every
        (chunk, view) pair (rasterizer_impl.cu:637-714).  min() over views and or() are order independent and exact;
      * with a process group, rank r owns the contiguous point range `shard_range(N, r, world)` (queries are
        independent: no data-path collective) and the two result arrays are all-gathered at the end, 5 bytes per
        point (`gather=False` returns the rank's own range instead).
    `prepare(settings) -> object with .integrate(points)` replaces `GaussianRasterizer(settings).prepare(...)` (tests of
    the sharding logic on CPU tensors).  `prepared` (from `prepare_views`): the views' binned Gaussians from an earlier
    round over the same model; nothing is binned again."""
    if prepared is not None and len(prepared) != len(views):
        raise ValueError("`prepared` must hold one prepared view per entry of `views`")
    if prepare is None:
        from . import GaussianRasterizer

        def prepare(rs):
            return GaussianRasterizer(rs).prepare(means3D, opacities, scales, rotations)

    n = points3D.shape[0]
    sharded = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    world, rank = (dist.get_world_size(group), dist.get_rank(group)) if sharded else (1, 0)
    lo, hi = shard_range(n, rank, world)
    mine = points3D[lo:hi]
    m = hi - lo
    device = points3D.device
    weight = torch.ones(m, dtype=torch.float32, device=device)
    any_valid = torch.zeros(m, dtype=torch.bool, device=device)
    chunks = [(a, min(a + chunk_size, m)) for a in range(0, m, chunk_size)]
    for vi, rs in enumerate(views):
        view = prepared[vi] if prepared is not None else prepare(rs)
        for a, b in chunks:
            pts = mine[a:b]
            alpha, inside = view.integrate(pts)
            valid = inside if validate is None else validate(vi, pts, inside)
            any_valid[a:b] |= valid
            w = weight[a:b]
            weight[a:b] = torch.where(valid, torch.min(alpha, w), w)
        del view
    weight[~any_valid] = 0
    sdf = 0.5 - weight
    if not sharded or not gather:
        return sdf, any_valid
    cap = (n + world - 1) // world
    s_pad = torch.zeros(cap, dtype=torch.float32, device=device)
    v_pad = torch.zeros(cap, dtype=torch.uint8, device=device)
    s_pad[:m] = sdf
    v_pad[:m] = any_valid.to(torch.uint8)
    s_all = torch.empty(world * cap, dtype=torch.float32, device=device)
    v_all = torch.empty(world * cap, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(s_all, s_pad, group=group)
    dist.all_gather_into_tensor(v_all, v_pad, group=group)
    out_s = torch.empty(n, dtype=torch.float32, device=device)
    out_v = torch.empty(n, dtype=torch.bool, device=device)
    for r in range(world):
        rlo, rhi = shard_range(n, r, world)
        out_s[rlo:rhi] = s_all[r * cap:r * cap + (rhi - rlo)]
        out_v[rlo:rhi] = v_all[r * cap:r * cap + (rhi - rlo)].bool()
    return out_s, out_v
