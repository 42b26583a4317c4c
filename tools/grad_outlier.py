#!/usr/bin/env python
"""Which gradient element decides the max-norm error against the reference, and is the difference a bias or
atomic-order noise?  Runs both implementations a few times on one scene and prints the worst elements.
    python tools/grad_outlier.py [--scene c1 --sh-degree 1 --seed 5] [--runs 4]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import helpers as H  # noqa: E402

import scenes  # noqa: E402
from oracle import ref_loader  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="c1")
    ap.add_argument("--sh-degree", type=int, default=1)
    ap.add_argument("--seed", type=int, default=5)
    ap.add_argument("--runs", type=int, default=4)
    a = ap.parse_args()
    import diff_gaussian_rasterization as ours

    ref = ref_loader.load()
    sc = scenes.make_scene(name=a.scene, sh_degree=a.sh_degree, seed=a.seed)
    dev = "cuda:0"
    O = [H.run_module(ours, sc, dev)[1] for _ in range(a.runs)]
    R = [H.run_module(ref, sc, dev)[1] for _ in range(a.runs)]
    cancel = H.grad_cancellation(ours, sc, dev)
    cancel = {k: v.reshape(-1) for k, v in cancel.items()}
    for k in H.GRAD_NAMES:
        o = torch.stack([x[k].double().cpu() for x in O])
        r = torch.stack([x[k].double().cpu() for x in R])
        den = float(r[0].abs().max()) if r[0].numel() else 0.0
        if den == 0:
            continue
        errs = [[float((o[i] - r[j]).abs().max()) / den for j in range(a.runs)] for i in range(a.runs)]
        on = max(float((o[i] - o[0]).abs().max()) / den for i in range(a.runs))
        rn = max(float((r[i] - r[0]).abs().max()) / den for i in range(a.runs))
        flat = [e for row in errs for e in row]
        print(f"{k:12s} ours-vs-ref min {min(flat):.3e} max {max(flat):.3e}   ours noise {on:.3e}  ref noise {rn:.3e}")
        if max(flat) > 5e-5:
            d = (o.mean(0) - r.mean(0)).abs()
            idx = int(d.reshape(-1).argmax())
            row = idx // d[0].numel() if d.dim() > 1 else idx
            print("   worst element", idx, "gaussian", row, "scale of tensor", den, "sum |contributions| >=",
                  float(cancel[k].reshape(-1)[idx]), "-> fp32 bound 16 eps A =",
                  16 * 5.96e-8 * float(cancel[k].reshape(-1)[idx]), "excess", H.grad_excess(O[0][k].reshape(-1), R[0][k].reshape(-1), cancel[k]))
            print("   ours:", [f"{float(x.reshape(-1)[idx]):.9e}" for x in o])
            print("   ref :", [f"{float(x.reshape(-1)[idx]):.9e}" for x in r])
            for kk in ("means3D", "scales", "opacities", "means2D"):
                oo = O[0][kk].double().cpu()[row].tolist()
                rr = R[0][kk].double().cpu()[row].tolist()
                print(f"   {kk:10s} ours {oo}\n              ref  {rr}")
            print("   gaussian scale", sc.scales[row].tolist(), "opacity", float(sc.opacities[row]), "rot",
                  sc.rotations[row].tolist())


if __name__ == "__main__":
    main()
