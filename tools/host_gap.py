#!/usr/bin/env python
"""Where does a fwd+bwd step spend time outside our kernels?  Host timeline of the bench step.

    python tools/host_gap.py [--steps 40]

Prints, for a few loop variants (stage profiler on/off, sync per step or not), the device time per step
(CUDA events) next to the host time spent inside forward() and backward().
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"))

import torch  # noqa: E402

import bench  # noqa: E402
import scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--workload", default="c3")
    args = ap.parse_args()
    import diff_gaussian_rasterization as mod
    from diff_gaussian_rasterization import _abi

    dev = torch.device("cuda", 0)
    sc = scenes.make_scene(args.workload)
    step = bench.Step(mod, sc, dev)

    def loop(name, sync_each, prof):
        _abi.profile_enable(prof)
        for _ in range(10):
            step.resident()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(args.steps):
            step.resident()
            if sync_each:
                torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        print(f"{name:40s} device {e0.elapsed_time(e1) / args.steps:7.3f} ms/step   host {(t1 - t0) * 1e3 / args.steps:7.3f} ms/step")
        _abi.profile_enable(False)

    def with_sampler(period):
        smp = bench.ClockSampler(0, period_s=period)
        smp.start()
        loop(f"no sync, NVML sampler every {period * 1e3:.0f} ms", False, False)
        print("    ", smp.stop())

    loop("no sync, stage profiler on", False, True)
    loop("no sync, stage profiler off", False, False)
    loop("sync per step, profiler off", True, False)
    loop("no sync, stage profiler off (again)", False, False)
    with_sampler(0.02)
    with_sampler(0.1)
    loop("no sync, stage profiler off (again)", False, False)

    # per-step device times of a no-sync loop (outliers = allocator / driver hiccups)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record()
    for i in range(args.steps):
        step.resident()
        evs[i + 1].record()
    torch.cuda.synchronize()
    per = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    print("per-step device ms: min %.3f median %.3f max %.3f" % (min(per), sorted(per)[len(per) // 2], max(per)),
          " ".join("%.2f" % x for x in per))

    # host time inside the two ABI calls
    import ctypes as C
    lib = _abi.lib
    f_orig, b_orig = lib.ggs_rasterize_forward, lib.ggs_rasterize_backward
    acc = {"f": 0.0, "b": 0.0, "n": 0}

    class Wrap:
        def __init__(self, fn, key):
            self.fn, self.key = fn, key

        def __call__(self, *a):
            t = time.perf_counter()
            r = self.fn(*a)
            acc[self.key] += time.perf_counter() - t
            return r

    mod._lib = type("L", (), {"ggs_rasterize_forward": Wrap(f_orig, "f"), "ggs_rasterize_backward": Wrap(b_orig, "b"),
                              "ggs_integrate_points": lib.ggs_integrate_points, "ggs_mark_visible": lib.ggs_mark_visible})()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step.resident()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    n = args.steps
    print(f"host per step {1e3 * (t1 - t0) / n:.3f} ms: inside ggs_rasterize_forward {1e3 * acc['f'] / n:.3f} ms "
          f"(includes the wait for num_rendered), inside ggs_rasterize_backward {1e3 * acc['b'] / n:.3f} ms, "
          f"python/torch outside {1e3 * ((t1 - t0) - acc['f'] - acc['b']) / n:.3f} ms")

    # the same with the GPU idle at every call: pure host cost of each call
    acc.update(f=0.0, b=0.0)
    tp = 0.0
    for _ in range(args.steps):
        torch.cuda.synchronize()
        t = time.perf_counter()
        step.resident()
        tp += time.perf_counter() - t
    print(f"GPU idle at call: step host time {1e3 * tp / n:.3f} ms, forward call {1e3 * acc['f'] / n:.3f} ms, "
          f"backward call {1e3 * acc['b'] / n:.3f} ms")


if __name__ == "__main__":
    main()
