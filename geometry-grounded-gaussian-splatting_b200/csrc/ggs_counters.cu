// Measurement only: algorithmic operation counts of the tile blend for the roofline (SURVEY.md 8(d) "ALGORITHMIC
// ops"), taken from a finished forward's state by re-walking every pixel's list with the reference's accept / stop
// rules (render_forward.cu:477-532).  Not on any product path; bench.py calls it once, outside the timed region.
//   walked          A      = sum over pixels of the list positions visited until the pixel stops
//   accepted        Cpairs = sum over pixels of the (pixel, Gaussian) pairs that pass the power / alpha tests
//   accepted_active          the same over the pixels the bisection runs on (final T <= 0.45)
//   active_pixels   Gpx
#include "ggs_kernels.cuh"

namespace ggs {

__global__ void __launch_bounds__(256)
    count_blend_ops_kernel(int W, int H, int gx, const uint2* __restrict__ ranges, const uint32_t* __restrict__ list,
                           const SplatRecord* __restrict__ records, unsigned long long* __restrict__ counters) {
    const uint32_t tile = blockIdx.x;
    const uint32_t px = (tile % gx) * TILE_X + threadIdx.x % TILE_X;
    const uint32_t py = (tile / gx) * TILE_Y + threadIdx.x / TILE_X;
    const bool inside = px < (uint32_t)W && py < (uint32_t)H;
    const uint2 range = ranges[tile];
    unsigned long long walked = 0, accepted = 0;
    float T = 1.f;
    if (inside) {
        const float2 pixf = {(float)px, (float)py};
        for (uint32_t i = range.x; i < range.y; i++) {
            walked++;
            const float4* rec = reinterpret_cast<const float4*>(records + __ldg(list + i));
            const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1);
            const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
            const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, q1.y * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1.f - alpha);
            if (test_T < 0.0001f) break;
            accepted++;
            T = test_T;
        }
    }
    const bool active = inside && T <= MIN_TRANSMITTANCE;
    unsigned long long v[4] = {walked, accepted, active ? accepted : 0ull, active ? 1ull : 0ull};
#pragma unroll
    for (int k = 0; k < 4; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        if ((threadIdx.x & 31) == 0 && v[k]) atomicAdd(counters + k, v[k]);
    }
}

void launch_count_blend_ops(const FrameParams& fp, const uint2* ranges, const uint32_t* list,
                            const SplatRecord* records, unsigned long long* counters, cudaStream_t stream) {
    cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned long long), stream);
    count_blend_ops_kernel<<<fp.gx * fp.gy, TILE_PIXELS, 0, stream>>>(fp.W, fp.H, fp.gx, ranges, list, records, counters);
}

}  // namespace ggs
