"""One GPU: time ggs_sh_gradient_from_views (the SH half of view_sum_kernel, all reads local) for several view counts."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"))
import torch

from diff_gaussian_rasterization import _abi

dev = torch.device("cuda:0")
P, SHM, SHD = 1_000_000, 16, 3
g = torch.Generator().manual_seed(1)
means = torch.randn(P, 3, generator=g).to(dev)
out = torch.empty(P, SHM, 3, device=dev)
for V in (1, 2, 4, 8):
    rgb = torch.randn(V, P, 3, generator=g).to(dev)
    cam = (torch.randn(V, 3, generator=g) * 4).to(dev)

    def call():
        _abi.check(_abi.lib.ggs_sh_gradient_from_views(P, SHD, SHM, V, _abi.ptr(means), _abi.ptr(cam), _abi.ptr(rgb),
                                                       _abi.ptr(out), _abi.stream_handle(dev)), "sh")

    for _ in range(3):
        call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"V={V}: {ms * 1e3:.0f} us  ({(P * SHM * 12 + V * P * 12) / ms / 1e6:.0f} GB/s)", flush=True)
