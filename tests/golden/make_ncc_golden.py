"""Golden vectors for the patch NCC from the reference's OWN kernel source, executed on the host (run in the build
container, where /root/reference exists; the vectors are committed, the reference is never read at test time).

The reference has no CPU path and no GPU was available, so `forward_mode_differentiation_CUDA<3, true>`
(submodules/warp-patch-ncc/cuda_warp_patch_ncc/warp_patch_ncc_impl.cu:18-262, with its mathUtils.h) is compiled for
the host under tests/native/cuda_emu.h: the kernel text is read from /root/reference at build time into a temporary
directory (nothing of it is copied into the repository), a few type / intrinsic shims are put in front of it, and it is
launched with the reference's own geometry ((P + 255) / 256 blocks of 256).  Host arithmetic is IEEE without
contraction, i.e. the reference's algorithm without --use_fast_math.  Output: tests/golden/ncc/<scene>.npz."""
import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = os.path.dirname(HERE)
sys.path.insert(0, TESTS)
REF = "/root/reference/submodules/warp-patch-ncc/cuda_warp_patch_ncc"

SHIMS = r'''#pragma once
#include <math.h>
#include <string.h>
struct float3 { float x, y, z; };
struct int3 { int x, y, z; };
struct int4 { int x, y, z, w; };
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float3 make_float3(float x, float y, float z) { return float3{x, y, z}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline int2 make_int2(int x, int y) { return int2{x, y}; }
inline int3 make_int3(int x, int y, int z) { return int3{x, y, z}; }
inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
inline float min(float a, float b) { return fminf(a, b); }
inline float max(float a, float b) { return fmaxf(a, b); }
inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
inline unsigned __float_as_uint(float f) { unsigned i; memcpy(&i, &f, 4); return i; }
inline float __uint_as_float(unsigned i) { float f; memcpy(&f, &i, 4); return f; }
inline int atomicMin(int* a, int v) { int o = *a; if (v < o) *a = v; return o; }
inline int atomicMax(int* a, int v) { int o = *a; if (v > o) *a = v; return o; }
inline unsigned atomicMin(unsigned* a, unsigned v) { unsigned o = *a; if (v < o) *a = v; return o; }
inline unsigned atomicMax(unsigned* a, unsigned v) { unsigned o = *a; if (v > o) *a = v; return o; }
'''
COOP = '''#pragma once
namespace cooperative_groups {
struct grid_group { unsigned long long thread_rank() const { return (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; } };
inline grid_group this_grid() { return grid_group(); }
}
'''
HARNESS = r'''#include "cuda_emu.h"
#include "shims.h"
#include "kernel_part.inc"
extern "C" void run_reference_ncc(int P, const float* depths, const float* normals, const int* uvs, const float* R,
                                  const float* T, const float* image_r, const float* image_n, const float* k, int h_r,
                                  int w_r, int h_n, int w_n, float* ncc, float* gd, float* gn, bool* valid) {
    emu_launch(dim3((P + 255) / 256), dim3(256), [=] {
        forward_mode_differentiation_CUDA<3, true>(P, depths, reinterpret_cast<const float3*>(normals),
                                                   reinterpret_cast<const int2*>(uvs), R, T, image_r, image_n, k[0], k[1],
                                                   k[2], k[3], k[4], k[5], k[6], k[7], h_r, w_r, h_n, w_n, ncc, gd,
                                                   reinterpret_cast<float3*>(gn), valid);
    });
}
'''
SCENES = {
    "same_size": dict(n=4000),
    "other_size": dict(height=96, width=128, neighbour_size=(110, 150), seed=11, shift=(12.0, 6.0), n=4000),
    "big_shift": dict(seed=13, shift=(-40.0, 25.0), n=3000),
}
INTR = ("fx_r", "fy_r", "cx_r", "cy_r", "fx_n", "fy_n", "cx_n", "cy_n")


def build(tmp):
    src = open(os.path.join(REF, "warp_patch_ncc_impl.cu")).read()
    with open(os.path.join(tmp, "kernel_part.inc"), "w") as fh:  # the kernel template, without the <<< >>> launcher
        fh.write(src[: src.index("void forward_mode_differentiation(")])
    for name, text in (("shims.h", SHIMS), ("cooperative_groups.h", COOP),
                       ("math_constants.h", "#pragma once\n#define CUDART_INF_F INFINITY\n"), ("harness.cpp", HARNESS)):
        with open(os.path.join(tmp, name), "w") as fh:
            fh.write(text)
    so = os.path.join(tmp, "ref_ncc_host.so")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-ffp-contract=off", "-pthread", "-shared", "-fPIC", "-D__CUDACC__",
                           "-D__inline__=inline", "-I" + tmp, "-I" + os.path.join(TESTS, "native"), "-I" + REF, "-o", so,
                           os.path.join(tmp, "harness.cpp")])
    return C.CDLL(so)


def main():
    import ncc_scenes

    with tempfile.TemporaryDirectory() as tmp:
        lib = build(tmp)
        os.makedirs(os.path.join(HERE, "ncc"), exist_ok=True)
        for tag, kw in SCENES.items():
            sc = ncc_scenes.make_pair(**kw)
            P = int(sc["depths"].shape[0])
            out = [torch.zeros(P), torch.zeros(P), torch.zeros(P, 3), torch.zeros(P, dtype=torch.bool)]
            intr = torch.tensor([sc[k] for k in INTR], dtype=torch.float32)
            p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
            lib.run_reference_ncc(P, p(sc["depths"]), p(sc["normals"]), p(sc["uvs"]), p(sc["R"]), p(sc["T"]),
                                  p(sc["image_r"]), p(sc["image_n"]), p(intr), *[int(v) for v in sc["image_r"].shape],
                                  *[int(v) for v in sc["image_n"].shape], p(out[0]), p(out[1]), p(out[2]), p(out[3]))
            np.savez_compressed(os.path.join(HERE, "ncc", tag + ".npz"), ncc=out[0].numpy(), grad_depths=out[1].numpy(),
                                grad_normals=out[2].numpy(), valid=out[3].numpy(),
                                scene=np.array(repr(sorted(kw.items()))))
            print(tag, P, int(out[3].sum()), float(out[0].max()))


if __name__ == "__main__":
    main()
