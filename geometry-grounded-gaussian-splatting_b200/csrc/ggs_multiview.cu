// View-parallel training: the SH-coefficient gradient without an all-reduce of [P, SHM, 3].
//
// For one view the SH gradient of Gaussian p is rank one (render_backward.cu:76-160):
//     dL/dsh[p][k][c] = basis_k(dir(p, campos)) * dL/dRGB[p][c]          (after the colour clamp mask)
// so a rank only needs the 3 floats dL/dRGB[p] of every other view (and that view's camera position) to form the
// SUM over views locally.  At SH degree 3 that replaces an all-reduce of 192 bytes per Gaussian by an all-gather of
// 12 bytes per Gaussian and view -- 81 % of the parameter-gradient bytes leave the wire.
// One thread per Gaussian; the [P][SHM*3] output rows are written as warp slabs through shared memory
// (coalesced 16-byte stores), as in ggs_gaussian_backward.cu.
#include "ggs_kernels.cuh"
#include "ggs_math.cuh"

namespace ggs {

constexpr int MV_THREADS = 128;

__global__ void __launch_bounds__(MV_THREADS)
    sh_gradient_from_views_kernel(int P, int SHD, int SHM, int V, const float* __restrict__ means3D,
                                  const float* __restrict__ campos, const float* __restrict__ dL_drgb,
                                  float* __restrict__ dL_dsh) {
    extern __shared__ float s_slab[];  // [warps][32][RS]
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int F = SHM * 3;
    const int RS = F | 1;
    float* const swarp = s_slab + (size_t)warp * 32 * RS;
    float* const srow = swarp + lane * RS;
    const size_t g0 = (size_t)blockIdx.x * blockDim.x + warp * 32;
    const int ng = g0 < (size_t)P ? (int)min((size_t)32, (size_t)P - g0) : 0;

    if (idx < (size_t)P) {
        for (int k = 0; k < F; k++) srow[k] = 0.f;
        const float3 mean = {__ldg(means3D + 3 * idx), __ldg(means3D + 3 * idx + 1), __ldg(means3D + 3 * idx + 2)};
        for (int v = 0; v < V; v++) {
            const float* gr = dL_drgb + ((size_t)v * P + idx) * 3;
            const float3 d = {__ldg(gr), __ldg(gr + 1), __ldg(gr + 2)};
            if (d.x == 0.f && d.y == 0.f && d.z == 0.f) continue;  // culled (or untouched) in this view
            const float3 dir_orig = {mean.x - __ldg(campos + 3 * v), mean.y - __ldg(campos + 3 * v + 1),
                                     mean.z - __ldg(campos + 3 * v + 2)};
            // same normalisation as the per-view kernels (render_forward.cu:31-32)
            const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
            const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
            auto add = [&](int k, float w) {
                srow[3 * k] += w * d.x;
                srow[3 * k + 1] += w * d.y;
                srow[3 * k + 2] += w * d.z;
            };
            add(0, kSH_C0);
            if (SHD > 0) {
                add(1, -kSH_C1 * y);
                add(2, kSH_C1 * z);
                add(3, -kSH_C1 * x);
                if (SHD > 1) {
                    const float xx = x * x, yy = y * y, zz = z * z;
                    const float xy = x * y, yz = y * z, xz = x * z;
                    add(4, kSH_C2[0] * xy);
                    add(5, kSH_C2[1] * yz);
                    add(6, kSH_C2[2] * (2.f * zz - xx - yy));
                    add(7, kSH_C2[3] * xz);
                    add(8, kSH_C2[4] * (xx - yy));
                    if (SHD > 2) {
                        add(9, kSH_C3[0] * y * (3.f * xx - yy));
                        add(10, kSH_C3[1] * xy * z);
                        add(11, kSH_C3[2] * y * (4.f * zz - xx - yy));
                        add(12, kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                        add(13, kSH_C3[4] * x * (4.f * zz - xx - yy));
                        add(14, kSH_C3[5] * z * (xx - yy));
                        add(15, kSH_C3[6] * x * (xx - 3.f * yy));
                    }
                }
            }
        }
    }
    __syncwarp();
    if ((F & 3) == 0) {
        const int Q = F >> 2;
        float4* dst = reinterpret_cast<float4*>(dL_dsh + g0 * F);
        int gi = 0, rem = lane;
        for (int e = lane; e < ng * Q; e += 32, rem += 32) {
            while (rem >= Q) {
                rem -= Q;
                gi++;
            }
            const float* sr = swarp + gi * RS + rem * 4;
            dst[e] = float4{sr[0], sr[1], sr[2], sr[3]};
        }
    } else {
        float* dst = dL_dsh + g0 * F;
        int gi = 0, rem = lane;
        for (int e = lane; e < ng * F; e += 32, rem += 32) {
            while (rem >= F) {
                rem -= F;
                gi++;
            }
            dst[e] = swarp[gi * RS + rem];
        }
    }
}

void launch_sh_gradient_from_views(int P, int SHD, int SHM, int V, const float* means3D, const float* campos,
                                   const float* dL_drgb, float* dL_dsh, cudaStream_t stream) {
    if (P == 0) return;
    const size_t smem = (size_t)(MV_THREADS / 32) * 32 * ((SHM * 3) | 1) * sizeof(float);
    static size_t configured = 48 * 1024;
    if (smem > configured) {
        cudaFuncSetAttribute(sh_gradient_from_views_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        configured = smem;
    }
    sh_gradient_from_views_kernel<<<(P + MV_THREADS - 1) / MV_THREADS, MV_THREADS, smem, stream>>>(
        P, SHD, SHM, V, means3D, campos, dL_drgb, dL_dsh);
}

}  // namespace ggs
