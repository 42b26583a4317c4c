"""Single-GPU step time of every view of an N-view batch (the views a view-parallel run gives its ranks): how much of
the N-GPU step is the slowest view rather than the exchange.   python tools/view_spread.py c4 8 [steps]"""
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

    workload, n_views = sys.argv[1], int(sys.argv[2])
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    dev = torch.device("cuda", 0)
    out = {}
    for v in [None] + list(range(n_views)):
        sc = bench.make_view(workload, 0, 1) if v is None else bench.make_view(workload, v, n_views)
        step = bench.Step(mod, sc, dev, allreduce=False)
        for _ in range(5):
            step.resident()
        ms, _ = bench.timed(step.resident, steps, 3, 1, None)
        out["solo" if v is None else f"view{v}"] = round(ms / steps, 4)
        del step
    vals = [out[f"view{v}"] for v in range(n_views)]
    out["max_over_solo"] = round(max(vals) / out["solo"], 4)
    out["mean_over_solo"] = round(sum(vals) / len(vals) / out["solo"], 4)
    print(json.dumps({workload: out}))


if __name__ == "__main__":
    main()
