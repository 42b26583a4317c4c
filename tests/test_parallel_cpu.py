"""World-size-2 gloo tests of the multi-GPU plumbing (view-parallel gradient all-reduce, point-range sharding
of integrate).  Only the collectives run here; the rasterizer itself needs a GPU."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers  # noqa: F401  (sys.path)
from diff_gaussian_rasterization import parallel


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 8, 9, 50_000_001):
        for world in (1, 2, 3, 8):
            spans = [parallel.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        parallel.shard_range(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # ---- gradient bucket: rank r holds gradient (r+1) * pattern; the sum is 3 * pattern at world 2 ----
        g = torch.Generator().manual_seed(0)
        shapes = [(100, 3), (100, 1), (100, 16, 3), (100, 4)]
        pattern = [torch.randn(*s, generator=g) for s in shapes]
        params = [torch.zeros(*s, requires_grad=True) for s in shapes]
        for p, pat in zip(params, pattern):
            p.grad = (rank + 1) * pat.clone()
        params[1].grad = None  # a missing gradient counts as zero
        bucket = parallel.GradientBucket(params)
        flat = bucket.allreduce()
        total = sum(r + 1 for r in range(world))
        ok = flat.numel() == sum(p.numel() for p in params)
        for i, (p, pat) in enumerate(zip(params, pattern)):
            want = torch.zeros_like(pat) if i == 1 else total * pat
            ok = ok and torch.allclose(p.grad, want, rtol=1e-6, atol=1e-6)
        # a second reduction reuses the views without copying
        for p, pat in zip(params, pattern):
            p.grad.copy_((rank + 1) * pat)
        bucket.allreduce(average=True)
        ok = ok and torch.allclose(params[0].grad, total * pattern[0] / world, rtol=1e-6, atol=1e-6)

        # ---- integrate sharding: a stand-in for rasterizer.integrate that is a pure function of the point ----
        calls = []

        def fake_integrate(pts):
            calls.append(pts.shape[0])
            return pts.sum(dim=1) * 0.5, pts[:, 0] > 0

        pts = torch.randn(1001, 3, generator=torch.Generator().manual_seed(5))
        a, ins = parallel.integrate_sharded(fake_integrate, pts)
        lo, hi = parallel.shard_range(1001, rank, world)
        ok = ok and calls == [hi - lo]
        ok = ok and torch.equal(a, pts.sum(dim=1) * 0.5) and torch.equal(ins, pts[:, 0] > 0)
        # fewer points than ranks, and zero points
        a1, i1 = parallel.integrate_sharded(fake_integrate, pts[:1])
        ok = ok and torch.equal(a1, pts[:1].sum(dim=1) * 0.5) and i1.shape == (1,)
        a0, i0 = parallel.integrate_sharded(fake_integrate, pts[:0])
        ok = ok and a0.numel() == 0 and i0.numel() == 0
        # ---- alpha_cull: view-major loop over prepared views, point-range sharding, final all-gather ----
        prepared = []

        class FakeView:
            def __init__(self, k):
                self.k = k
                prepared.append(k)

            def integrate(self, p):
                # "opacity" depends on the view and the point; a view sees the point iff a coordinate is positive
                return (p[:, 0].abs() * (self.k + 1)).clamp(max=1.0) * 0.5, p[:, self.k % 3] > 0

        views = [0, 1, 2, 3]
        sdf, valid = parallel.alpha_cull(pts, views, None, None, None, None, chunk_size=170, prepare=FakeView)
        want_w = torch.ones(1001)
        want_v = torch.zeros(1001, dtype=torch.bool)
        for k in views:
            al, ins = FakeView(k).integrate(pts)
            want_v |= ins
            want_w = torch.where(ins, torch.min(al, want_w), want_w)
        want_w[~want_v] = 0
        ok = ok and torch.equal(sdf, 0.5 - want_w) and torch.equal(valid, want_v)
        ok = ok and prepared[:4] == [0, 1, 2, 3]  # every view prepared once per rank, whatever the number of chunks
        s_own, v_own = parallel.alpha_cull(pts, views, None, None, None, None, chunk_size=170, prepare=FakeView,
                                           gather=False)
        ok = ok and torch.equal(s_own, sdf[lo:hi]) and torch.equal(v_own, valid[lo:hi])
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_gloo_world2_allreduce_and_point_sharding():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert dict(ret) == {0: True, 1: True}


def test_single_process_paths_need_no_group():
    pts = torch.randn(10, 3)
    a, i = parallel.integrate_sharded(lambda p: (p[:, 0], p[:, 1] > 0), pts)
    assert torch.equal(a, pts[:, 0]) and torch.equal(i, pts[:, 1] > 0)
    p = torch.zeros(4, requires_grad=True)
    p.grad = torch.ones(4)
    b = parallel.GradientBucket([p])
    b.allreduce()
    assert torch.equal(p.grad, torch.ones(4))


def test_exchange_layout_regions_are_disjoint_and_aligned():
    """Layout of the peer-memory exchange buffer (GgsPeerExchange): 256-byte aligned regions that do not overlap,
    reduce-stage slots large enough for the longest 1/world slice."""
    from diff_gaussian_rasterization.parallel import exchange_layout

    for P in (1, 10, 1000, 200_003, 1_000_000):
        for world in (1, 2, 3, 8, 16):
            L = exchange_layout(P, world)
            spans = []
            for name, width in (("rotations", 4), ("means3D", 3), ("scales", 3), ("opacities", 1)):
                o, w = L[name]
                assert w == width and o % 64 == 0
                spans.append((L["send"] + o, L["send"] + o + w * P))
                spans.append((L["recv"] + o, L["recv"] + o + w * P))
            assert spans[-2][1] <= L["reduce"]
            for r in range(world):
                spans.append((L["rgb_stage"] + r * L["rgb_stride"], L["rgb_stage"] + r * L["rgb_stride"] + 3 * P))
                spans.append((L["campos_stage"] + 4 * r, L["campos_stage"] + 4 * r + 3))
                spans.append((L["reduce_stage"] + r * L["reduce_stride"],
                              L["reduce_stage"] + (r + 1) * L["reduce_stride"]))
            spans.sort()
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 <= b0, (P, world, (a0, a1), (b0, b1))
            assert spans[-1][1] <= L["total"]
            for k in ("send", "recv", "reduce", "rgb_stage", "rgb_stride", "campos_stage", "reduce_stage",
                      "reduce_stride"):
                assert L[k] % 4 == 0
            n4 = L["reduce"] // 4
            longest = max(n4 * (s + 1) // world - n4 * s // world for s in range(world))
            assert L["reduce_stride"] // 4 >= longest


def test_bench_watchdog_fails_fast_instead_of_hanging():
    """bench.py's view-parallel runs hold node barriers in every step; a stuck rank must end the run, not hang it."""
    import subprocess
    import sys

    code = ("import sys, time; sys.argv=['bench.py']; import bench; bench.watchdog(0.3, 0); time.sleep(30); "
            "print('still here')")
    r = subprocess.run([sys.executable, "-c", code], cwd=helpers.ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 3, (r.returncode, r.stderr[-400:])
    assert "still here" not in r.stdout and "stuck" in r.stderr
    # a cancelled guard does nothing
    code = "import sys, time; sys.argv=['bench.py']; import bench; g = bench.watchdog(0.2, 0); g.cancel(); time.sleep(0.6)"
    r = subprocess.run([sys.executable, "-c", code], cwd=helpers.ROOT, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-400:]


def test_bench_reference_arm_runs_on_rank_zero_only_and_needs_a_gpu():
    """Contract of `bench.py --impl reference` under torchrun: ranks other than 0 exit 0 without work; and no arm
    pretends to run without a CUDA device."""
    import subprocess
    import sys

    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1"],
                       cwd=helpers.ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip() == "", (r.returncode, r.stdout[-200:], r.stderr[-200:])
    if not torch.cuda.is_available():
        env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
        r = subprocess.run([sys.executable, "bench.py", "--steps", "1"], cwd=helpers.ROOT, env=env,
                           capture_output=True, text=True, timeout=300)
        assert r.returncode != 0 and "no CPU path" in (r.stderr + r.stdout)


def _agree_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import sys

    sys.argv = ["bench.py"]
    sys.path.insert(0, helpers.ROOT)
    import bench

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # rank r alone would stop after 3 + 2 r steps; every step holds a barrier, as a view-parallel step does
        steps = 0
        while bench.agreed(steps < 3 + 2 * rank, "cpu"):
            dist.barrier()
            steps += 1
        ret[rank] = steps
    finally:
        dist.destroy_process_group()


def test_gloo_world2_ranks_agree_on_the_number_of_pre_run_steps():
    """bench.py's untimed pre-run is time-based; with a node barrier inside every step the ranks must stop after the
    same step (a rank that went one step further hung two 8-GPU runs): the decision is a MAX all-reduce."""
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_agree_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert dict(ret) == {0: 5, 1: 5}


def test_result_lease_detects_aliasing_gradients():
    """results handed out as views into persistent buffers must not survive into the next
    backward unnoticed."""
    import gc

    import pytest
    import torch

    from diff_gaussian_rasterization.parallel import ResultLease

    buf = torch.zeros(8)
    lease = ResultLease("test")
    lease.check()  # nothing handed out yet
    views = lease.hand_out((buf[:4].view(2, 2), buf[4:].view(2, 2)))
    assert lease.alive() == 2
    with pytest.raises(RuntimeError, match="still referenced"):
        lease.check()
    holder = views[0]  # e.g. a parameter's .grad
    del views
    gc.collect()
    assert lease.alive() == 1
    with pytest.raises(RuntimeError):
        lease.check()
    del holder
    gc.collect()
    lease.check()


def test_alpha_cull_with_prepared_views_matches():
    """parallel.prepare_views: later rounds reuse the prepared views; same result as preparing inside the call."""
    import torch

    from diff_gaussian_rasterization import parallel

    torch.manual_seed(3)
    pts = torch.rand(1000, 3)
    views = [0.3, 0.5, 0.9]
    made = []

    class View:
        def __init__(self, thr):
            self.thr = thr
            made.append(thr)

        def integrate(self, p):
            return (p[:, 0] * self.thr).float(), p[:, 1] < self.thr

    a = parallel.alpha_cull(pts, views, None, None, None, None, prepare=View, chunk_size=300)
    assert made == views
    prepared = parallel.prepare_views(views, None, None, None, None, prepare=View)
    n_made = len(made)
    b = parallel.alpha_cull(pts, views, None, None, None, None, prepare=View, prepared=prepared, chunk_size=170)
    assert len(made) == n_made, "nothing may be prepared again"
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_chunk_ranges_cover_the_region_once():
    """Pipelined sum of the view-parallel exchange: over ranks and pieces every 16-byte unit of the padded gradient
    region is summed exactly once, piece boundaries are multiples of 128 Gaussians."""
    from diff_gaussian_rasterization import parallel

    for P in (1024, 1025, 12800, 99_999):
        for world in (1, 2, 3, 8):
            L = parallel.exchange_layout(P, world)
            for chunks in (1, 2, 4, 8):
                cover = [0] * (L["reduce"] // 4)
                for r in range(world):
                    ends, ranges = parallel.chunk_ranges(P, world, r, chunks, L)
                    assert ends[-1] == P and ends == sorted(ends) and all(e % 128 == 0 for e in ends[:-1])
                    assert len(ranges) == chunks and all(len(m) <= 4 for m in ranges)
                    for mine in ranges:
                        for a, b in mine:
                            for i in range(a, b):
                                cover[i] += 1
                assert all(c == 1 for c in cover), (P, world, chunks)
