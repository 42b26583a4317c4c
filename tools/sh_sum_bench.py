"""One GPU: the SH half of the view-parallel exchange (ggs_sh_gradient_from_views = view_sum_kernel on local data) for
V views of P Gaussians: time and fraction of the HBM peak on its algorithmic bytes (12 V + 12 + 12 SHM bytes per Gaussian).
    python tools/sh_sum_bench.py [P] [V]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "geometry-grounded-gaussian-splatting_b200"))
import torch  # noqa: E402


def main():
    from diff_gaussian_rasterization import _abi

    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    V = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    means = torch.randn(P, 3, device=dev, generator=g)
    campos = torch.randn(V, 3, device=dev, generator=g) * 4
    rgb = torch.randn(V, P, 3, device=dev, generator=g)
    rgb[:, ::13] = 0  # culled in a view
    d_sh = torch.empty(P, 16, 3, device=dev)
    flush = torch.empty(64 * 2**20, device=dev)

    def run():
        _abi.check(_abi.lib.ggs_sh_gradient_from_views(P, 3, 16, V, _abi.ptr(means), _abi.ptr(campos), _abi.ptr(rgb),
                                                       _abi.ptr(d_sh), _abi.stream_handle(dev)), "sh")

    for _ in range(5):
        run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(20):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    alg = P * (12 * V + 12 + 192)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6543.4
    print(json.dumps({"P": P, "V": V, "us": ms * 1e3, "alg_bytes": alg, "gbs": alg / ms / 1e6, "frac_hbm": alg / ms / 1e6 / peak,
                      "lib": os.environ.get("GGS_RASTERIZER_LIB", "in-tree")}))


if __name__ == "__main__":
    main()
