"""torchrun -N ranks: view-parallel fwd+bwd with the gradient sums through peer memory (PeerGradientExchange:
ggs_rasterize_backward_peer + ggs_view_sum_peer) against the same step with NCCL collectives (all-gather of the colour
gradients + bucket all-reduce), on the same views.  Exit code 0 = the gradients agree.
  PEER_WORKLOAD=c1|c3 (default c1), PEER_MULTICAST=0|1 (default: use it when the box has it)."""
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

    workload = os.environ.get("PEER_WORKLOAD", "c1")
    sc = bench.make_view(workload, rank, world)
    names = ("means3D", "opacities", "scales", "rotations", "shs")

    def run(peer, multicast=None, reps=1, pull=False, extra=None):
        for k in ("GGS_PEER_FUSE_RGB", "GGS_PEER_SH_LATE", "GGS_PEER_CHUNKS"):
            os.environ.pop(k, None)
        os.environ.update(extra or {})
        os.environ["GGS_PEER"] = "1" if peer else "0"
        os.environ["GGS_PEER_PULL"] = "1" if pull else "0"
        os.environ["GGS_MULTICAST"] = "" if multicast is None else ("1" if multicast else "0")
        step = bench.Step(dgr, sc, dev, allreduce=True)
        for _ in range(reps):
            step.resident()
        torch.cuda.synchronize()
        grads = {k: step.params[k].grad.detach().clone() for k in names}
        return step, grads

    _, ref = run(False)
    if rank == 0:
        print(f"nccl reference done ({workload}, P={sc.P}, world={world})", flush=True)
    rc = 0
    # the last four are the switch-reduced default (colour gradients pushed by the per-Gaussian backward itself, SH sum
    # behind the backward) and each of its parts switched off / the pipelined sum switched on
    variants = ((False, False, {}), (True, False, {}), (True, True, {}), (True, True, {"GGS_PEER_FUSE_RGB": "0"}),
                (True, True, {"GGS_PEER_SH_LATE": "0"}), (True, True, {"GGS_PEER_CHUNKS": "2"}))
    for mc, pull, extra in variants:
        try:
            step, got = run(True, mc, reps=3, pull=pull, extra=extra)
        except Exception as e:  # noqa: BLE001
            print(f"rank {rank}: exchange(multicast={mc}) failed: {type(e).__name__}: {e}", flush=True)
            if mc:
                continue
            rc = 2
            break
        if step.exchange is None:
            print(f"rank {rank}: exchange(multicast={mc}) unavailable", flush=True)
            if not mc:
                rc = 2
            continue
        errs = {k: float((got[k] - ref[k]).abs().max() / ref[k].abs().max().clamp_min(1e-30)) for k in names}
        ok = max(errs.values()) < 1e-5
        msg = "  ".join(f"{k} {v:.1e}" for k, v in errs.items())
        ex = step.exchange
        flags = f"fused_push={int(ex.fused_push)} sh_late={int(ex.sh_after_backward)} chunks={ex.chunks}"
        print(f"rank {rank}: {step.exchange_note} [{flags}]: {'OK' if ok else 'MISMATCH'}  {msg}", flush=True)
        if not ok:
            rc = 1
        # the exchange alone: barrier + view-sum launch + barrier, on static data
        class RS:
            sh_degree = sc.sh_degree

        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ex = step.exchange
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            ex.sum_over_views(RS, step.params["means3D"], step.params["shs"].shape[1])
        e1.record()
        torch.cuda.synchronize()
        # and whole steps
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize()
        f0.record()
        for _ in range(20):
            step.resident()
        f1.record()
        torch.cuda.synchronize()
        print(f"rank {rank}: barrier+view_sum+barrier {e0.elapsed_time(e1) / 20 * 1e3:.0f} us, "
              f"step {f0.elapsed_time(f1) / 20:.3f} ms", flush=True)
        del step
    step, _ = run(False, reps=3)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    f0.record()
    for _ in range(20):
        step.resident()
    f1.record()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"nccl step {f0.elapsed_time(f1) / 20:.3f} ms", flush=True)
    dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
