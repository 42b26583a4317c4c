"""torchrun -N ranks: device time of the parts of the view-parallel gradient exchange on static data, each alone:
node barrier, the SH half of ggs_view_sum_peer (parts=2), the slice half (parts=1), whole sum_over_views, and a whole
step.  PEER_WORKLOAD=c3|c4."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"))
import torch
import torch.distributed as dist


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import bench
    import diff_gaussian_rasterization as dgr
    from diff_gaussian_rasterization import _abi

    workload = os.environ.get("PEER_WORKLOAD", "c3")
    sc = bench.make_view(workload, rank, world)
    step = bench.Step(dgr, sc, dev, allreduce=True)
    for _ in range(5):
        step.resident()
    torch.cuda.synchronize()
    ex = step.exchange
    assert ex is not None
    P, SHM = sc.P, step.params["shs"].shape[1]
    d_sh = torch.empty(P, SHM, 3, device=dev)
    means = step.params["means3D"].detach()
    cur = torch.cuda.current_stream()

    def view_sum(parts):
        _abi.check(_abi.lib.ggs_view_sum_peer(P, sc.sh_degree, SHM, _abi.ptr(means), C.byref(ex.px), _abi.ptr(d_sh), parts,
                                              C.c_void_p(cur.cuda_stream)), "ggs_view_sum_peer")

    class RS:
        sh_degree = sc.sh_degree

    def timed(fn, n=20):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / n * 1e3], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    res = {"workload": workload, "world": world, "P": P, "exchange": step.exchange_note}
    res["barrier_us"] = timed(lambda: ex.hdl.barrier(channel=0))
    res["sh_sum_alone_us"] = timed(lambda: view_sum(2))

    def slice_with_barriers():
        ex.hdl.barrier(channel=0)
        view_sum(1)
        ex.hdl.barrier(channel=1)

    res["slice_sum_with_2_barriers_us"] = timed(slice_with_barriers)
    res["sum_over_views_us"] = timed(lambda: ex.sum_over_views(RS, means, SHM))
    res["step_us"] = timed(step.resident)
    # the per-Gaussian part of a one-GPU step for scale
    if rank == 0:
        print(res, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
