// The patch-NCC kernel itself (launched by ggs_patch_ncc.cu).  Kept in a header without CUDA runtime includes so that
// the test suite can also run this source on the host under tests/native/cuda_emu.h (partial blocks, shared-memory
// staging and barrier placement are checked there without a GPU).
#pragma once
#include <stdint.h>

#include "ggs_patch_ncc_point.cuh"

namespace ggs {
namespace ncc {

constexpr int THREADS = 128;

__global__ void __launch_bounds__(THREADS)
warp_patch_kernel(Pair pair, const float* __restrict__ R, const float* __restrict__ T, const long long P,
                  const float* __restrict__ depths,
                  const float* __restrict__ normals, const int2* __restrict__ uvs,
                  const float* __restrict__ image_r, const float* __restrict__ image_n, float* __restrict__ ncc,
                  float* __restrict__ grad_depths, float* __restrict__ grad_normals, uint8_t* __restrict__ valid) {
    __shared__ float rows[THREADS * 3];
    __shared__ float samples[SAMPLES * THREADS];  // [sample][thread]: conflict-free columns
    const long long first = static_cast<long long>(blockIdx.x) * THREADS;
    const int here = static_cast<int>(min(static_cast<long long>(THREADS), P - first));
    const int t = threadIdx.x;
    for (int i = t; i < here * 3; i += THREADS) rows[i] = __ldg(normals + first * 3 + i);
    __syncthreads();
    PointResult r = {0.f, 0.f, 0.f, 0.f, 0.f, false};
    for (int i = 0; i < 9; i++) pair.R[i] = __ldg(R + i);  // uniform addresses: one broadcast transaction each
    for (int i = 0; i < 3; i++) pair.T[i] = __ldg(T + i);
    if (t < here) {
        const int2 uv = __ldg(uvs + first + t);
        const V3 n = v3(rows[3 * t], rows[3 * t + 1], rows[3 * t + 2]);
        r = evaluate_point<THREADS>(pair, uv.x, uv.y, __ldg(depths + first + t), n, image_r, image_n, samples + t);
        ncc[first + t] = r.ncc;
        grad_depths[first + t] = r.d_depth;
        valid[first + t] = r.valid ? 1 : 0;
    }
    __syncthreads();
    if (t < here) {
        rows[3 * t] = r.d_nx;
        rows[3 * t + 1] = r.d_ny;
        rows[3 * t + 2] = r.d_nz;
    }
    __syncthreads();
    for (int i = t; i < here * 3; i += THREADS) grad_normals[first * 3 + i] = rows[i];
}

}  // namespace ncc
}  // namespace ggs
