// View-parallel training, the PUSH half of the gradient exchange over peer memory (NVLink / NVSwitch).
//
// Measured on 2 x B200: a kernel that PULLS its operands out of a peer's memory is bound by NVLink round trips
// (94 us to sum 22 MB, 237 us for the SH part reading 12 MB of colour gradients), while posted peer STORES run at
// link speed.  So every byte crosses the link as a store, issued by the rank that produced it:
//   * push_rgb_kernel      right after the tile backward, on a side stream, concurrent with the per-Gaussian
//                          backward: the masked colour gradient dL/dRGB [P,3] of this view (the factor of the
//                          rank-one SH gradient) goes into slot `rank` of EVERY rank's colour stage -- one
//                          multimem.st per 16 bytes when the buffer has an NVSwitch multicast mapping (the switch
//                          replicates), `world` peer stores otherwise;
//   * push_slices_kernel   after the per-Gaussian backward: slice s of this rank's 11-floats-per-Gaussian region goes
//                          into slot `rank` of rank s's reduce stage (a reduce-scatter by stores).
// After a node barrier, view_sum_kernel (ggs_multiview.cu) works on LOCAL memory only -- it sums the `world` slots of
// its slice and forms the SH gradient from the `world` colour slots -- and stores the summed slice into every rank's
// result region (multimem.st / peer stores).  The reference has no counterpart (no distributed code).
#include "ggs_kernels.cuh"

namespace ggs {

__device__ __forceinline__ void multimem_st4(float4* mc, const float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}

constexpr int PUSH_THREADS = 256;

// dL/dRGB of Gaussian p = accumulator columns 7..9 rounded to float once, zeroed per channel where the forward
// clamped the colour and for culled Gaussians -- exactly what gaussian_backward_kernel calls dL_dRGB
// (render_backward.cu:76-80).  A warp's 32 rows are one 384-byte segment: staged through shared memory and stored
// as 24 x 16 bytes per destination.
__global__ void __launch_bounds__(PUSH_THREADS)
    push_rgb_kernel(int P, const double* __restrict__ grad_acc, const uint8_t* __restrict__ clamped,
                    const int* __restrict__ radii, const float* __restrict__ campos, const PushTargets t) {
    __shared__ __align__(16) float s_seg[PUSH_THREADS / 32][96];
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t g0 = (size_t)blockIdx.x * blockDim.x + warp * 32;
    float3 d = {0.f, 0.f, 0.f};
    if (idx < (size_t)P && radii[idx] > 0) {
        const double* row = grad_acc + idx * GRAD_STRIDE;
        const uint8_t mask = clamped[idx];
        d.x = (mask & 1) ? 0.f : (float)__ldg(row + 7);
        d.y = (mask & 2) ? 0.f : (float)__ldg(row + 8);
        d.z = (mask & 4) ? 0.f : (float)__ldg(row + 9);
    }
    s_seg[warp][3 * lane] = d.x;
    s_seg[warp][3 * lane + 1] = d.y;
    s_seg[warp][3 * lane + 2] = d.z;
    __syncwarp();
    const int ng = g0 < (size_t)P ? (int)min((size_t)32, (size_t)P - g0) : 0;
    if (ng == 32) {
        if (lane < 24) {
            const float4 v = reinterpret_cast<const float4*>(s_seg[warp])[lane];
            const size_t q = g0 * 3 / 4 + lane;  // g0 is a multiple of 32: 16-byte aligned
            if (t.mc_rgb) {
                multimem_st4(reinterpret_cast<float4*>(t.mc_rgb) + q, v);
            } else {
#pragma unroll
                for (int r = 0; r < MAX_VIEWS; r++)
                    if (r < t.world) reinterpret_cast<float4*>(t.rgb[r])[q] = v;
            }
        }
    } else {
        for (int e = lane; e < ng * 3; e += 32)
            for (int r = 0; r < t.world; r++) t.rgb[r][g0 * 3 + e] = s_seg[warp][e];
    }
    if (idx < 3) {
        const float c = campos[idx];
        for (int r = 0; r < t.world; r++) t.campos[r][idx] = c;
    }
}

// blockIdx.y = destination rank s; the blocks of row s copy slice s of `send` (float4 units [lo4[s], lo4[s+1]))
// into dest[s] (slot `rank` of rank s's reduce stage), grid-stride, two columns in flight per thread.
__global__ void __launch_bounds__(PUSH_THREADS)
    push_slices_kernel(const float4* __restrict__ send, const PushSlices t) {
    const int s = blockIdx.y;
    const size_t lo = t.lo4[s], hi = t.lo4[s + 1];
    float4* const dst = t.dest[s];
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + stride < hi; i += 2 * stride) {
        const float4 a = __ldg(send + i);
        const float4 b = __ldg(send + i + stride);
        dst[i - lo] = a;
        dst[i + stride - lo] = b;
    }
    if (i < hi) dst[i - lo] = __ldg(send + i);
}

void launch_push_rgb(int P, const double* grad_acc, const uint8_t* clamped, const int* radii, const float* campos,
                     const PushTargets& t, cudaStream_t stream) {
    if (P == 0) return;
    push_rgb_kernel<<<(P + PUSH_THREADS - 1) / PUSH_THREADS, PUSH_THREADS, 0, stream>>>(P, grad_acc, clamped, radii,
                                                                                       campos, t);
}

void launch_push_slices(const float4* send, const PushSlices& t, cudaStream_t stream) {
    size_t longest = 0;
    for (int s = 0; s < t.world; s++) longest = max(longest, t.lo4[s + 1] - t.lo4[s]);
    if (longest == 0) return;
    // enough threads to keep the link busy, split over the destinations
    size_t bx = (longest + 2 * PUSH_THREADS - 1) / (2 * PUSH_THREADS);
    const size_t cap = (size_t)(148 * 8 + t.world - 1) / t.world;
    if (bx > cap) bx = cap;
    if (bx == 0) bx = 1;
    push_slices_kernel<<<dim3((unsigned)bx, (unsigned)t.world, 1), PUSH_THREADS, 0, stream>>>(send, t);
}

}  // namespace ggs
