// Stage 3 of the forward: per-tile front-to-back blend of colour / alpha / normal, followed by
// the five-iteration 8-way bisection for the geometry-grounded median depth.
// Semantics follow DGR/cuda_rasterizer/render_forward.cu:391-671 (renderCUDA) exactly:
//   * a list entry is rejected when power > 0 or alpha < 1/255; the entry that would push T below
//     1e-4 ends the pixel and is NOT blended; `contributor` counts list positions;
//   * mDepthinit is the ray depth of the last entry blended while T > 0.5 (MEDIAN_DEPTH_INIT);
//   * bisection: bracket [d0-0.4, d0+0.4] clamped at 0, active iff T <= 0.45, 9 samples in the
//     first iteration and 7 new ones afterwards, linear interpolation at the end.
//
// Design (one CTA per 16x16 tile, 8 warps, each warp owns an 8x4 pixel block):
//   * list entries are gathered as 64-byte SplatRecords into shared memory, 256 per batch;
//   * pass 1 leaves, per warp, a bitmask of the list positions that at least one of its pixels blended.
//     The five bisection passes walk only those positions (about one third of the list), instead of
//     re-testing every entry against every pixel as the reference does;
//   * in a bisection pass a splat whose depth profile is more than 16 sigma away from the whole current
//     bracket contributes exp() == 0 exactly (ex2.approx.ftz underflows), so its factor is exactly
//     (1 - alpha) in front of the bracket and exactly 1 behind it: no exp / rsqrt is issued for it unless
//     some pixel of the warp is close to it.  Brackets shrink 8x per iteration, so most splats are "far"
//     in iterations 2-5.  Both shortcuts are exact: outputs are bit-identical to the plain evaluation.
#include "ggs_kernels.cuh"

namespace ggs {

struct RenderParams {
    int W, H, gx;
    float focal_x, focal_y;
    const float* bg;  // device [3]
};

constexpr int BATCH = TILE_PIXELS;

constexpr int REC_QUADS = 5;
constexpr int MASK_WORDS = 256;  // per warp: acceptance bits for the first 8192 list positions
constexpr uint32_t FULL = 0xffffffffu;
constexpr float FAR_SIGMAS = 16.f;  // exp(-0.5 * 16^2) = 2^-184: ex2.approx.ftz returns +0

// One depth sample of the vacancy-corrected transmittance product (render_forward.cu:603-621).
__device__ __forceinline__ float sample_factor(float ts, float t_peak, float rsigma, bool ball, float alpha) {
    const float delta = (ts - t_peak) * rsigma;
    const float g = ball ? expf(-0.5f * delta * delta) : 0.f;
    const float one_minus_gaussian = 1.f - alpha * g;
    const float rvacancy = rsqrtf(one_minus_gaussian);
    return (ts > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
}

template <bool GEOMETRY>
__global__ void __launch_bounds__(TILE_PIXELS)
    render_forward_kernel(const RenderParams rp, const uint2* __restrict__ ranges,
                          const uint32_t* __restrict__ point_list, const SplatRecord* __restrict__ records,
                          float* __restrict__ out_color, float* __restrict__ out_alpha,
                          float* __restrict__ out_normal, float* __restrict__ out_mdepth,
                          uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ max_contributors,
                          uint32_t* __restrict__ accept_mask, uint32_t* __restrict__ warp_words) {
    // staged records, 80-byte stride (4 payload quads + 1 pad): conflict-free 16-byte stores by 8 consecutive
    // threads, and one address computation per entry on the (broadcast) read side
    __shared__ float4 s_rec[BATCH * REC_QUADS];
    __shared__ uint32_t s_mask[GEOMETRY ? TILE_PIXELS / 32 : 1][GEOMETRY ? MASK_WORDS : 1];
    __shared__ uint32_t s_red[TILE_PIXELS / 32];
    __shared__ uint32_t s_block_max;

    const int W = rp.W, H = rp.H;
    const uint32_t tile_id = blockIdx.y * rp.gx + blockIdx.x;
    const uint32_t tid = threadIdx.x;
    const uint32_t warp = tid >> 5, lane = tid & 31;
    // warp w owns the 8x4 pixel block (w & 1, w >> 1) of the tile
    const uint2 pix = {blockIdx.x * TILE_X + (warp & 1) * 8 + (lane & 7), blockIdx.y * TILE_Y + (warp >> 1) * 4 + (lane >> 3)};
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
    // first word of this tile in the global acceptance mask (tiles are laid out in list order, one spare word each)
    const uint32_t mask_base = range.x / 32 + tile_id;

    float T = 1.0f;
    uint32_t contributor = 0;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float Normal[3] = {0.f, 0.f, 0.f};
    float mDepthinit = 0.f;
    uint32_t mask_words = 0;  // words of s_mask[warp] that pass 1 produced (warp-uniform)

    for (int i = 0; i < rounds; i++, toDo -= BATCH) {
        const int block_done = __syncthreads_and(done);
        if (block_done) break;

        const int progress = i * BATCH + tid;
        if (range.x + progress < range.y) {
            const uint32_t id = __ldg(point_list + range.x + progress);
            const float4* src = reinterpret_cast<const float4*>(records + id);
            float4* dst = s_rec + tid * REC_QUADS;
            dst[0] = __ldg(src);
            dst[1] = __ldg(src + 1);
            dst[2] = __ldg(src + 2);
            dst[3] = __ldg(src + 3);
        }
        __syncthreads();

        const int lim = min(BATCH, toDo);
        for (int w = 0; w * 32 < lim; w++) {
            if (__all_sync(FULL, done)) break;
            uint32_t bits = 0;
            const int jend = min(32, lim - w * 32);
            for (int jj = 0; jj < jend; jj++) {
                const int j = w * 32 + jj;
                bool blended = false;
                if (!done) {
                    contributor++;
                    const float4* rec = s_rec + j * REC_QUADS;
                    const float4 q0 = rec[0];
                    const float4 q1 = rec[1];
                    const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    if (power <= 0.0f) {
                        const float alpha = fminf(0.99f, q1.y * expf(power));
                        if (alpha >= 1.0f / 255.0f) {
                            const float test_T = T * (1.f - alpha);
                            if (test_T < 0.0001f) {
                                done = true;
                            } else {
                                const float aT = alpha * T;
                                const float4 q2 = rec[2];
                                const float4 q3 = rec[3];
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
                                blended = true;
                            }
                        }
                    }
                }
                if (__any_sync(FULL, blended)) bits |= 1u << jj;
            }
            const uint32_t gw = i * (BATCH / 32) + w;
            if (lane == 0) {
                if (GEOMETRY && gw < MASK_WORDS) s_mask[warp][gw] = bits;
                accept_mask[(size_t)(mask_base + gw) * 8 + warp] = bits;  // kept for the backward
            }
            mask_words = gw + 1;
        }
    }

    // block-wide max of last_contributor
    {
        const uint32_t wmax = __reduce_max_sync(FULL, last_contributor);
        if (lane == 0) s_red[warp] = wmax;
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
        // rsqrt.approx(1.0): the vacancy factor of a splat whose Gaussian term underflowed
        float c1;
        {
            const float one = 1.0f;
            asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(c1) : "f"(one));
        }

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
            const float depth_hi = depth_min + interval * SPLIT;  // the last sample depth, as the loop forms it
            // last list position any pixel of this warp still needs
            const uint32_t warp_last = __reduce_max_sync(FULL, in_range ? last_contributor : 0u);
            for (int i = 0; i < rounds2; i++) {
                __syncthreads();
                const uint32_t progress = i * BATCH + tid;
                if (progress < max_contributor) {
                    const uint32_t id = __ldg(point_list + range.x + progress);
                    const float4* src = reinterpret_cast<const float4*>(records + id);
                    float4* dst = s_rec + tid * REC_QUADS;
                    dst[0] = __ldg(src);
                    dst[1] = __ldg(src + 1);
                    dst[2] = __ldg(src + 2);
                }
                __syncthreads();
                for (int w = 0; w < BATCH / 32; w++) {
                    const uint32_t base = i * BATCH + w * 32;
                    if (base >= warp_last) break;
                    const uint32_t gw = i * (BATCH / 32) + w;
                    uint32_t bits = gw < mask_words ? (gw < MASK_WORDS ? s_mask[warp][gw] : FULL) : 0u;
                    uint32_t keep = 0;  // positions some pixel of the warp can still need in later iterations
                    while (bits) {
                        const int jj = __ffs(bits) - 1;
                        const uint32_t bit = bits & (0u - bits);
                        bits ^= bit;
                        const uint32_t pos = base + jj + 1;  // 1-based list position == `contributor`
                        if (pos > warp_last) break;
                        const float4* rec = s_rec + (w * 32 + jj) * REC_QUADS;
                        const float4 q0 = rec[0];
                        const float4 q1 = rec[1];
                        const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                        const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                        const float alpha = fminf(0.99f, q1.y * expf(power));
                        const bool act = in_range && pos <= last_contributor && !(power > 0.0f) &&
                                         !(alpha < 1.0f / 255.0f);
                        if (!__any_sync(FULL, act)) continue;
                        const float4 q2 = rec[2];
                        const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                        const float rsigma = q2.y;
                        const bool ball = rsigma > 0;
                        // Exact shortcut.  Every sample depth lies in [depth_min, depth_hi] (rounding is monotone),
                        // so if that whole interval is more than 16 sigma in front of / behind the splat's peak
                        // depth then |delta| > 16 at every sample and exp(-delta^2/2) is exactly 0.
                        const float front = depth_min - t_peak;
                        const float back = t_peak - depth_hi;
                        const bool far = ball && fmaxf(front, back) * rsigma > FAR_SIGMAS;
                        const bool far_front = far && front > 0.f;
                        // may be dropped for good only with some slack: the recomputed depth_hi of a later
                        // iteration can sit a few ulps above this one
                        const bool behind_for_good = far && !far_front && back > 1e-5f * depth_hi;
                        if (__any_sync(FULL, act && !behind_for_good)) keep |= bit;
                        if (__any_sync(FULL, act && !far)) {
                            if (act) {
                                if (first) {
#pragma unroll
                                    for (int s = 0; s < SPLIT + 1; s++)
                                        T_p[s] *= sample_factor(depth_min + interval * s, t_peak, rsigma, ball, alpha);
                                } else {
#pragma unroll
                                    for (int s = 1; s < SPLIT; s++)
                                        T_p[s] *= sample_factor(depth_min + interval * s, t_peak, rsigma, ball, alpha);
                                }
                            }
                        } else if (act) {
                            // exp underflows to 0 at every sample: factor = ((1 - alpha) | 1) * rsqrt(1)
                            const float f = (far_front ? (1.f - alpha) : 1.0f) * c1;
                            if (far_front || c1 != 1.0f) {
                                if (first) {
#pragma unroll
                                    for (int s = 0; s < SPLIT + 1; s++) T_p[s] *= f;
                                } else {
#pragma unroll
                                    for (int s = 1; s < SPLIT; s++) T_p[s] *= f;
                                }
                            }
                        }
                    }
                    // positions that were behind every pixel's bracket stay behind it (brackets only shrink):
                    // drop them from the warp's list.  Unvisited positions (beyond warp_last) are never needed.
                    if (c1 == 1.0f && lane == 0 && gw < MASK_WORDS && gw < mask_words) s_mask[warp][gw] = keep;
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
    if (lane == 0) warp_words[tile_id * (TILE_PIXELS / 32) + warp] = mask_words;
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
    const dim3 block(TILE_PIXELS, 1, 1);
    if (require_depth)
        render_forward_kernel<true><<<grid, block, 0, stream>>>(rp, ranges, point_list, records, out.color, out.alpha,
                                                               out.normal, out.mdepth, out.n_contrib,
                                                               out.max_contributor, out.accept_mask, out.warp_words);
    else
        render_forward_kernel<false><<<grid, block, 0, stream>>>(rp, ranges, point_list, records, out.color,
                                                                out.alpha, out.normal, out.mdepth, out.n_contrib,
                                                                out.max_contributor, out.accept_mask, out.warp_words);
}

}  // namespace ggs
