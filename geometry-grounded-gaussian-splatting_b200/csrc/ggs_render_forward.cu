// Stage 3 of the forward: per-tile front-to-back blend of colour / alpha / normal, followed by
// the five-iteration 8-way bisection for the geometry-grounded median depth.
// Semantics follow DGR/cuda_rasterizer/render_forward.cu:391-671 (renderCUDA) exactly:
//   * a list entry is rejected when power > 0 or alpha < 1/255; the entry that would push T below
//     1e-4 ends the pixel and is NOT blended; `contributor` counts list positions;
//   * mDepthinit is the ray depth of the last entry blended while T > 0.5 (MEDIAN_DEPTH_INIT);
//   * bisection: bracket [d0-0.4, d0+0.4] clamped at 0, active iff T <= 0.45, 9 samples in the
//     first iteration and 7 new ones afterwards, linear interpolation at the end.
// One CTA per 16x16 tile; list entries are gathered as 64-byte SplatRecords into shared memory.
#include "ggs_kernels.cuh"

namespace ggs {

struct RenderParams {
    int W, H, gx;
    float focal_x, focal_y;
    const float* bg;  // device [3]
};

constexpr int BATCH = TILE_PIXELS;

template <bool GEOMETRY>
__global__ void __launch_bounds__(TILE_PIXELS)
    render_forward_kernel(const RenderParams rp, const uint2* __restrict__ ranges,
                          const uint32_t* __restrict__ point_list, const SplatRecord* __restrict__ records,
                          float* __restrict__ out_color, float* __restrict__ out_alpha,
                          float* __restrict__ out_normal, float* __restrict__ out_mdepth,
                          uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ max_contributors) {
    __shared__ float4 s_q0[BATCH];
    __shared__ float4 s_q1[BATCH];
    __shared__ float4 s_q2[BATCH];
    __shared__ float4 s_q3[BATCH];
    __shared__ uint32_t s_red[TILE_PIXELS / 32];
    __shared__ uint32_t s_block_max;

    const int W = rp.W, H = rp.H;
    const uint32_t tile_id = blockIdx.y * rp.gx + blockIdx.x;
    const uint32_t tid = threadIdx.y * TILE_X + threadIdx.x;
    const uint2 pix = {blockIdx.x * TILE_X + threadIdx.x, blockIdx.y * TILE_Y + threadIdx.y};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {static_cast<float>(pix.x), static_cast<float>(pix.y)};
    const float2 pixnf = {(pixf.x - static_cast<float>(W - 1) / 2.f) / rp.focal_x,
                          (pixf.y - static_cast<float>(H - 1) / 2.f) / rp.focal_y};
    const float rln = rnorm3df(pixnf.x, pixnf.y, 1.f);

    const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
    bool done = !inside;

    const uint2 range = ranges[tile_id];
    const int rounds = ((range.y - range.x + BATCH - 1) / BATCH);
    int toDo = range.y - range.x;

    float T = 1.0f;
    uint32_t contributor = 0;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float Normal[3] = {0.f, 0.f, 0.f};
    float mDepthinit = 0.f;

    for (int i = 0; i < rounds; i++, toDo -= BATCH) {
        const int block_done = __syncthreads_and(done);
        if (block_done) break;

        const int progress = i * BATCH + tid;
        if (range.x + progress < range.y) {
            const uint32_t id = __ldg(point_list + range.x + progress);
            const float4* src = reinterpret_cast<const float4*>(records + id);
            s_q0[tid] = __ldg(src);
            s_q1[tid] = __ldg(src + 1);
            s_q2[tid] = __ldg(src + 2);
            s_q3[tid] = __ldg(src + 3);
        }
        __syncthreads();

        const int lim = min(BATCH, toDo);
        for (int j = 0; !done && j < lim; j++) {
            contributor++;
            const float4 q0 = s_q0[j];
            const float4 q1 = s_q1[j];
            const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
            const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, q1.y * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1.f - alpha);
            if (test_T < 0.0001f) {
                done = true;
                continue;
            }
            const float aT = alpha * T;
            const float4 q2 = s_q2[j];
            const float4 q3 = s_q3[j];
            C[0] += q2.z * aT;
            C[1] += q2.w * aT;
            C[2] += q3.x * aT;
            if constexpr (GEOMETRY) {
                const float t = q1.z * d.x + q1.w * d.y + q2.x;
                Normal[0] += q3.y * aT;
                Normal[1] += q3.z * aT;
                Normal[2] += q3.w * aT;
                mDepthinit = T > 0.5f ? t : mDepthinit;
            }
            T = test_T;
            last_contributor = contributor;
        }
    }

    // block-wide max of last_contributor
    {
        const uint32_t wmax = __reduce_max_sync(0xffffffffu, last_contributor);
        if ((tid & 31) == 0) s_red[tid >> 5] = wmax;
        __syncthreads();
        if (tid == 0) {
            uint32_t m = 0;
#pragma unroll
            for (int w = 0; w < TILE_PIXELS / 32; w++) m = max(m, s_red[w]);
            s_block_max = m;
        }
        __syncthreads();
    }
    const uint32_t max_contributor = s_block_max;

    float mDepth = 0.f;
    if constexpr (GEOMETRY) {
        const int rounds2 = (max_contributor + BATCH - 1) / BATCH;
        float T_p[SPLIT + 1];
        float depth_min = fmaxf(mDepthinit - SAMPLE_RANGE, 0.f);
        float depth_max = fmaxf(mDepthinit + SAMPLE_RANGE, 0.f);
        bool in_range = T <= MIN_TRANSMITTANCE;

#pragma unroll 1
        for (int it = 0; it < SPLIT_ITERATIONS; it++) {
            const bool first = (it == 0);
            constexpr float ONE_OVER_SPLIT = 1.f / static_cast<float>(SPLIT);
            if (first) {
                T_p[0] = 1.f;
                T_p[SPLIT] = 1.f;
            }
#pragma unroll
            for (int s = 1; s < SPLIT; s++) T_p[s] = 1.f;
            const float interval = (depth_max - depth_min) * ONE_OVER_SPLIT;
            bool bdone = !in_range;
            int toDo2 = max_contributor;
            contributor = 0;
            for (int i = 0; i < rounds2; i++, toDo2 -= BATCH) {
                __syncthreads();
                const uint32_t progress = i * BATCH + tid;
                if (progress < max_contributor) {
                    const uint32_t id = __ldg(point_list + range.x + progress);
                    const float4* src = reinterpret_cast<const float4*>(records + id);
                    s_q0[tid] = __ldg(src);
                    s_q1[tid] = __ldg(src + 1);
                    s_q2[tid] = __ldg(src + 2);
                }
                __syncthreads();
                const int lim = min(BATCH, toDo2);
                for (int j = 0; !bdone && j < lim; j++) {
                    contributor++;
                    bdone = contributor >= last_contributor;
                    const float4 q0 = s_q0[j];
                    const float4 q1 = s_q1[j];
                    const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, q1.y * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float2 rp2 = *reinterpret_cast<const float2*>(&s_q2[j]);
                    const float t_peak = q1.z * d.x + q1.w * d.y + rp2.x;
                    const float rsigma = rp2.y;
                    const bool ball = rsigma > 0;
                    if (first) {
#pragma unroll
                        for (int s = 0; s < SPLIT + 1; s++) {
                            const float ts = depth_min + interval * s;
                            const float delta = (ts - t_peak) * rsigma;
                            const float g = ball ? expf(-0.5f * delta * delta) : 0.f;
                            const float one_minus_gaussian = 1.f - alpha * g;
                            const float rvacancy = rsqrtf(one_minus_gaussian);
                            T_p[s] *= (ts > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
                        }
                    } else {
#pragma unroll
                        for (int s = 1; s < SPLIT; s++) {
                            const float ts = depth_min + interval * s;
                            const float delta = (ts - t_peak) * rsigma;
                            const float g = ball ? expf(-0.5f * delta * delta) : 0.f;
                            const float one_minus_gaussian = 1.f - alpha * g;
                            const float rvacancy = rsqrtf(one_minus_gaussian);
                            T_p[s] *= (ts > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
                        }
                    }
                }
            }
            if (first) in_range = (T_p[0] >= 0.5f) && (T_p[SPLIT] <= 0.5f) && in_range;
            int start_id = 0;
#pragma unroll
            for (int p = 1; p < SPLIT; p++) start_id = T_p[p] >= 0.5f ? p : start_id;
            depth_max = depth_min + (start_id + 1) * interval;
            depth_min = depth_min + (start_id + 0) * interval;
            float lo = T_p[0], hi = T_p[1];
#pragma unroll
            for (int p = 1; p < SPLIT; p++) {
                lo = (start_id == p) ? T_p[p] : lo;
                hi = (start_id == p) ? T_p[p + 1] : hi;
            }
            T_p[0] = lo;
            T_p[SPLIT] = hi;
        }
        const float w_max = __saturatef((T_p[0] - 0.5f) / (T_p[0] - T_p[SPLIT]));
        const float w_min = 1.f - w_max;
        mDepth = in_range ? w_max * depth_max + w_min * depth_min : 0.f;
    }

    if (inside) {
        n_contrib[pix_id] = last_contributor;
        const size_t HW = (size_t)H * W;
#pragma unroll
        for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix_id] = C[ch] + T * __ldg(rp.bg + ch);
        out_alpha[pix_id] = 1.f - T;
        if constexpr (GEOMETRY) {
            out_mdepth[pix_id] = mDepth * rln;
            const float len_normal = 1.f - T;
#pragma unroll
            for (int ch = 0; ch < 3; ch++)
                out_normal[ch * HW + pix_id] = last_contributor ? Normal[ch] / len_normal : 0.f;
        }
    }
    if (tid == 0) max_contributors[tile_id] = max_contributor;
}

void launch_render_forward(const FrameParams& fp, const uint2* ranges, const uint32_t* point_list,
                           const SplatRecord* records, const RenderOut& out, bool require_depth,
                           cudaStream_t stream) {
    RenderParams rp;
    rp.W = fp.W;
    rp.H = fp.H;
    rp.gx = fp.gx;
    rp.focal_x = fp.focal_x;
    rp.focal_y = fp.focal_y;
    rp.bg = fp.view.bg;
    const dim3 grid(fp.gx, fp.gy, 1);
    const dim3 block(TILE_X, TILE_Y, 1);
    if (require_depth)
        render_forward_kernel<true><<<grid, block, 0, stream>>>(rp, ranges, point_list, records, out.color, out.alpha,
                                                               out.normal, out.mdepth, out.n_contrib,
                                                               out.max_contributor);
    else
        render_forward_kernel<false><<<grid, block, 0, stream>>>(rp, ranges, point_list, records, out.color,
                                                                out.alpha, out.normal, out.mdepth, out.n_contrib,
                                                                out.max_contributor);
}

}  // namespace ggs
