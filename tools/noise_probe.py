"""Run-to-run gradient noise of the reference (atomic ordering) next to ours-vs-reference, per tensor."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import helpers as H
import scenes
from oracle import ref_loader
import diff_gaussian_rasterization as ours
sys.path.insert(0, os.path.join(H.ROOT, "tests"))
from test_parity_gpu import REF_CASES

ref = ref_loader.load()
names = sys.argv[1:] or list(REF_CASES)
for tag in names:
    sc = scenes.make_scene(**REF_CASES[tag])
    for rd in (True, False):
        runs_r = [H.run_module(ref, sc, "cuda:0", require_depth=rd)[1] for _ in range(4)]
        runs_o = [H.run_module(ours, sc, "cuda:0", require_depth=rd)[1] for _ in range(3)]
        line = []
        for k in ("means3D", "opacities", "scales", "rotations", "shs"):
            rr = max(H.grad_rel_err(runs_r[i][k], runs_r[0][k]) for i in range(1, 4))
            orr = max(H.grad_rel_err(o[k], runs_r[0][k]) for o in runs_o)
            oo = max(H.grad_rel_err(runs_o[i][k], runs_o[0][k]) for i in range(1, 3))
            line.append(f"{k}: ref-ref {rr:.1e} ours-ref {orr:.1e} ours-ours {oo:.1e}")
        print(tag, "depth" if rd else "nodepth", " | ".join(line), flush=True)
