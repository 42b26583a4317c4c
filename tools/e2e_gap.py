"""One GPU: where the end-to-end step of bench.py loses time against the resident step.  Times Step.resident and
variants of Step.end_to_end with pieces switched off (H2D prefetch, colour copy-out + D2H, checksum).
    python tools/e2e_gap.py [workload] [steps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"))
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    import diff_gaussian_rasterization as mod

    workload = sys.argv[1] if len(sys.argv) > 1 else "c3"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    dev = torch.device("cuda", 0)
    sc = bench.make_view(workload, 0, 1)
    step = bench.Step(mod, sc, dev, allreduce=False)
    for _ in range(5):
        step.resident()
    out = {}
    out["resident"] = bench.timed(step.resident, steps, 5, 1, None)[0] / steps
    out["end_to_end"] = bench.timed(step.end_to_end, steps, 5, 1, None)[0] / steps

    cur = torch.cuda.current_stream()

    def no_h2d():  # same slot tensors, no prefetch copies
        slot = step.slots[step.pipe_i & 1]
        if slot["d2h"] is not None:
            slot["d2h"].synchronize()
        color = step._fwd_bwd(slot["cam"], slot["g"])
        slot["out"].copy_(color.detach())
        torch.sum(step.params["means3D"].grad.abs(), dim=None, out=slot["chk"].view(()))
        slot["free"].record(cur)
        d2h = torch.cuda.Event()
        with torch.cuda.stream(step.copy_stream):
            step.copy_stream.wait_event(slot["free"])
            slot["color_host"].copy_(slot["out"], non_blocking=True)
            slot["check_host"].copy_(slot["chk"], non_blocking=True)
            d2h.record(step.copy_stream)
        slot["d2h"] = d2h
        step.pipe_i += 1

    def no_d2h():
        slot = step.slots[step.pipe_i & 1]
        step._prefetch(step.slots[(step.pipe_i + 1) & 1])
        cur.wait_event(slot["ready"])
        color = step._fwd_bwd(slot["cam"], slot["g"])
        slot["out"].copy_(color.detach())
        torch.sum(step.params["means3D"].grad.abs(), dim=None, out=slot["chk"].view(()))
        slot["free"].record(cur)
        step.pipe_i += 1

    def slot_tensors_only():  # resident step, but on the slot's tensors (fresh settings objects per step either way)
        slot = step.slots[step.pipe_i & 1]
        step._fwd_bwd(slot["cam"], slot["g"])
        step.pipe_i += 1

    def resident_plus_outputs():
        slot = step.slots[0]
        color = step._fwd_bwd(step.cam_dev, step.grads_dev)
        slot["out"].copy_(color.detach())
        torch.sum(step.params["means3D"].grad.abs(), dim=None, out=slot["chk"].view(()))

    def late_prefetch():  # the whole e2e step, but the next step's H2D is queued behind this step's backward
        slot = step.slots[step.pipe_i & 1]
        if slot["d2h"] is not None:
            slot["d2h"].synchronize()
        cur.wait_event(slot["ready"])
        color = step._fwd_bwd(slot["cam"], slot["g"])
        slot["out"].copy_(color.detach())
        torch.sum(step.params["means3D"].grad.abs(), dim=None, out=slot["chk"].view(()))
        slot["free"].record(cur)
        d2h = torch.cuda.Event()
        with torch.cuda.stream(step.copy_stream):
            step.copy_stream.wait_event(slot["free"])
            slot["color_host"].copy_(slot["out"], non_blocking=True)
            slot["check_host"].copy_(slot["chk"], non_blocking=True)
            d2h.record(step.copy_stream)
        slot["d2h"] = d2h
        step._prefetch(step.slots[(step.pipe_i + 1) & 1])
        step.pipe_i += 1

    def prime():
        step._prefetch(step.slots[step.pipe_i & 1])

    for name, fn in (("no_h2d", no_h2d), ("no_d2h", no_d2h), ("slot_tensors_only", slot_tensors_only),
                     ("resident_plus_outputs", resident_plus_outputs)):
        torch.cuda.synchronize()
        out[name] = bench.timed(fn, steps, 5, 1, None)[0] / steps
    prime()
    torch.cuda.synchronize()
    out["late_prefetch"] = bench.timed(late_prefetch, steps, 5, 1, None)[0] / steps
    out["end_to_end_again"] = bench.timed(step.end_to_end, steps, 5, 1, None)[0] / steps
    print(json.dumps({k: round(v, 4) for k, v in out.items()}))


if __name__ == "__main__":
    main()
