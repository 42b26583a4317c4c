"""Golden vectors for SSIM from the reference's own torch code (run in the build container, where /root/reference
exists; the vectors are committed, the reference is never read at test time).

utils/loss_utils.py cannot be imported as a module (it imports CUDA extensions at the top), so the three functions that
state the algorithm — gaussian (:34-37), create_window (:40-45), _ssim (:52-73) — are cut out of its source with `ast`
and executed as they are.  Output: tests/golden/ssim/same.npz (seeded [1,3,48,64] images, the mean SSIM with the
reference's padding = window_size // 2, and its gradient with respect to img1)."""
import ast
import os
from math import exp  # noqa: F401  (used by the reference's gaussian)

import numpy as np
import torch
import torch.nn.functional as F  # noqa: F401
from torch.autograd import Variable  # noqa: F401

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/utils/loss_utils.py"


def reference_functions():
    tree = ast.parse(open(SRC).read())
    wanted = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("gaussian", "create_window", "_ssim")]
    assert len(wanted) == 3
    ns = {"torch": torch, "F": F, "Variable": Variable, "exp": exp}
    exec(compile(ast.Module(body=wanted, type_ignores=[]), SRC, "exec"), ns)
    return ns


def main():
    ns = reference_functions()
    g = torch.Generator().manual_seed(20260)
    img1 = torch.rand(1, 3, 48, 64, generator=g)
    img2 = (img1 + 0.1 * torch.randn(1, 3, 48, 64, generator=g)).clamp(0, 1)
    out = {}
    for dtype, tag in ((torch.float32, "f32"), (torch.float64, "f64")):
        a = img1.to(dtype).clone().requires_grad_(True)
        window = ns["create_window"](11, 3).to(dtype)
        val = ns["_ssim"](a, img2.to(dtype), window, 11, 3, True)
        val.backward()
        out["value_" + tag] = val.detach().numpy()
        out["grad_" + tag] = a.grad.numpy()
    os.makedirs(os.path.join(HERE, "ssim"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "ssim", "same.npz"), img1=img1.numpy(), img2=img2.numpy(), **out)
    print({k: (v.shape, float(np.abs(v).max())) for k, v in out.items()})


if __name__ == "__main__":
    main()
