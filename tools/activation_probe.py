"""One GPU: fused activate_gaussians (forward + backward) against the reference's torch op chain, 1 M Gaussians."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch

import diff_gaussian_rasterization as dgr
from test_activations_gpu import _reference

dev = torch.device("cuda:0")
P = 1_000_000
g = torch.Generator().manual_seed(3)
raw = [(torch.randn(P, 3, generator=g) * 0.6 - 4.4), torch.randn(P, 1, generator=g) * 2, torch.randn(P, 4, generator=g)]
f = (torch.rand(P, 1, generator=g) * 0.02).to(dev)
ups = [torch.randn(P, w, generator=g).to(dev) for w in (3, 1, 4)]
for name, fn in (("fused", dgr.activate_gaussians), ("torch", _reference)):
    ps = [t.to(dev).requires_grad_(True) for t in raw]

    def step():
        for p in ps:
            p.grad = None
        out = fn(ps[0], ps[1], ps[2], f)
        torch.autograd.backward(list(out), ups)

    for _ in range(5):
        step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(50):
        step()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 50 * 1e3:.0f} us per forward+backward at P = {P}", flush=True)
