// Median depth sampled at arbitrary sub-pixel positions (`GaussianRasterizer.sample_depth`, used by the
// multi-view geometric / NCC losses, utils/loss_utils.py:160), forward and backward.
// Semantics follow DGR/cuda_rasterizer/sample_forward.cu:430-700 (sampleDepthCUDA) and
// sample_backward.cu:42-358, driven like rasterizer_impl.cu:1042-1395: query points are binned by the tile they
// project into (ggs_points.cu), a block owns (tile, round of 512 points) and every thread carries two points
// through the tile's depth-sorted Gaussian list: one pass for the transmittance / last contributor / initial
// depth, then the same five-iteration 8-way bisection as the image forward (ggs_render_forward.cu).
// The backward writes the same 17-value double-precision accumulator rows as the image backward
// (ggs_common.cuh), so the per-Gaussian chain rule is shared (ggs_gaussian_backward.cu).
//
// Records are staged as the first three quads of the 64-byte SplatRecord (mean, conic, opacity, ray plane).
#include "ggs_kernels.cuh"
#include "ggs_math.cuh"
#include "ggs_tile.cuh"

namespace ggs {

constexpr int PB = POINT_BATCH;

// Tile of a block: smallest tile with tile_offsets[tile] > block_id (inclusive scan of rounds per tile); -1 if none.
__device__ __forceinline__ int tile_of_block(const uint32_t* __restrict__ tile_offsets, int tiles, uint32_t block_id) {
    if (block_id >= tile_offsets[tiles - 1]) return -1;
    int lo = 0, hi = tiles;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (tile_offsets[mid] > block_id)
            hi = mid;
        else
            lo = mid + 1;
    }
    return lo;
}

__device__ __forceinline__ void stage3(float4* s_rec, int tid, const SplatRecord* records, uint32_t id) {
    const float4* src = reinterpret_cast<const float4*>(records + id);
    float4* dst = s_rec + tid * 3;
    dst[0] = __ldg(src);
    dst[1] = __ldg(src + 1);
    dst[2] = __ldg(src + 2);
}

// SDF = false: sample_depth (5 iterations, bracket +-0.4, outputs the camera-space point).
// SDF = true : evaluate_sdf (sample_forward.cu:171-428: 6 iterations, bracket +-0.8; outputs the median ray depth
//              and its difference to the query point's own ray depth).
template <int ITERS, bool SDF>
__global__ void __launch_bounds__(TILE_PIXELS)
    sample_depth_forward_kernel(int tiles, int W, int H, float focal_x, float focal_y,
                                const uint32_t* __restrict__ tile_offsets, const uint2* __restrict__ gaussian_ranges,
                                const uint2* __restrict__ point_ranges, const uint32_t* __restrict__ gaussian_list,
                                const uint32_t* __restrict__ point_list, const SplatRecord* __restrict__ records,
                                const float2* __restrict__ points2D, const float* __restrict__ point_ts,
                                uint32_t* __restrict__ n_contrib, float* __restrict__ median_depth,
                                uint32_t* __restrict__ block_max_contributor, float* __restrict__ output,
                                uint8_t* __restrict__ inside_output) {
    __shared__ float4 s_rec[TILE_PIXELS * 3];
    __shared__ int s_tile;
    __shared__ uint32_t s_red[TILE_WARPS];

    const uint32_t block_id = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) s_tile = tile_of_block(tile_offsets, tiles, block_id);
    __syncthreads();
    const int tile_id = s_tile;
    if (tile_id < 0) return;
    const uint32_t tile_offset = (tile_id == 0) ? 0 : tile_offsets[tile_id - 1];
    const int p_round = block_id - tile_offset;
    const uint2 p_range = point_ranges[tile_id];
    const uint2 range = gaussian_ranges[tile_id];
    const int rounds = ((range.y - range.x + TILE_PIXELS - 1) / TILE_PIXELS);

    bool done[PB] = {false};
    float Depth[PB] = {0.f};
    float T[PB];
    uint32_t point_idx[PB];
    float2 point_xy[PB];
    uint32_t last_contributor[PB] = {0};
    uint32_t contributor = 0;
    int point_done = 0;
    int point_num_round = 0;
#pragma unroll
    for (int p = 0; p < PB; p++) {
        T[p] = 1.f;
        point_idx[p] = 0;
        point_xy[p] = {0.f, 0.f};
        const int progress = (p_round * PB + p) * TILE_PIXELS + tid;
        if (p_range.x + progress < p_range.y) {
            const uint32_t pid = __ldg(point_list + p_range.x + progress);
            point_idx[p] = pid;
            point_xy[p] = points2D[pid];
            point_num_round++;
        }
    }
    bool all_done = point_num_round == 0;
    int toDo = range.y - range.x;

    // ---- pass 1: transmittance, last contributor, initial depth (sample_forward.cu:486-550) ----
    for (int i = 0; i < rounds; i++, toDo -= TILE_PIXELS) {
        if (__syncthreads_and(all_done)) break;
        const int progress = i * TILE_PIXELS + tid;
        if (range.x + progress < range.y) stage3(s_rec, tid, records, __ldg(gaussian_list + range.x + progress));
        __syncthreads();
        const int lim = min(TILE_PIXELS, toDo);
        for (int j = 0; !all_done && j < lim; j++) {
            contributor++;
            const float4 q0 = s_rec[j * 3], q1 = s_rec[j * 3 + 1], q2 = s_rec[j * 3 + 2];
#pragma unroll
            for (int p = 0; p < PB; p++) {
                if (p >= point_num_round || done[p]) continue;
                const float2 d = {q0.x - point_xy[p].x, q0.y - point_xy[p].y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, q1.y * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T[p] * (1.f - alpha);
                if (test_T < 0.0001f) {
                    done[p] = true;
                    point_done++;
                    continue;
                }
                const float t = q1.z * d.x + q1.w * d.y + q2.x;
                Depth[p] = T[p] > 0.5f ? t : Depth[p];
                T[p] = test_T;
                last_contributor[p] = contributor;
            }
            all_done = point_done == point_num_round;
        }
    }

    // block-wide max of last_contributor
    uint32_t max_contributor;
    {
        uint32_t m = 0;
#pragma unroll
        for (int p = 0; p < PB; p++) m = max(m, last_contributor[p]);
        m = __reduce_max_sync(0xffffffffu, m);
        __syncthreads();
        if ((tid & 31) == 0) s_red[tid >> 5] = m;
        __syncthreads();
        m = 0;
#pragma unroll
        for (int w = 0; w < TILE_WARPS; w++) m = max(m, s_red[w]);
        max_contributor = m;
    }

    // ---- five bisection iterations (sample_forward.cu:552-672) ----
    float depth_min[PB], depth_max[PB];
    float T_p[PB][SPLIT + 1];
    bool inside_range[PB];
#pragma unroll
    for (int p = 0; p < PB; p++) {
        constexpr float RANGE = SDF ? SAMPLE_RANGE * 2.f : SAMPLE_RANGE;
        depth_min[p] = fmaxf(Depth[p] - RANGE, 0.f);
        depth_max[p] = fmaxf(Depth[p] + RANGE, 0.f);
        inside_range[p] = p < point_num_round && T[p] <= MIN_TRANSMITTANCE;
#pragma unroll
        for (int s = 0; s < SPLIT + 1; s++) T_p[p][s] = 1.f;
    }
    const int rounds2 = (max_contributor + TILE_PIXELS - 1) / TILE_PIXELS;
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
        const bool first = it == 0;
        constexpr float ONE_OVER_SPLIT = 1.f / static_cast<float>(SPLIT);
        float interval[PB];
        point_done = 0;
        contributor = 0;
#pragma unroll
        for (int p = 0; p < PB; p++) {
            done[p] = !inside_range[p];
            if (p < point_num_round) point_done += done[p] ? 1 : 0;
            interval[p] = (depth_max[p] - depth_min[p]) * ONE_OVER_SPLIT;
#pragma unroll
            for (int s = 1; s < SPLIT; s++) T_p[p][s] = 1.f;
            if (first) {
                T_p[p][0] = 1.f;
                T_p[p][SPLIT] = 1.f;
            }
        }
        all_done = point_done == point_num_round;
        int toDo2 = max_contributor;
        for (int i = 0; i < rounds2; i++, toDo2 -= TILE_PIXELS) {
            __syncthreads();
            const uint32_t progress = i * TILE_PIXELS + tid;
            if (progress < max_contributor) stage3(s_rec, tid, records, __ldg(gaussian_list + range.x + progress));
            __syncthreads();
            const int lim = min(TILE_PIXELS, toDo2);
            for (int j = 0; !all_done && j < lim; j++) {
                contributor++;
                const float4 q0 = s_rec[j * 3], q1 = s_rec[j * 3 + 1], q2 = s_rec[j * 3 + 2];
#pragma unroll
                for (int p = 0; p < PB; p++) {
                    if (p >= point_num_round || done[p]) continue;
                    const float2 d = {q0.x - point_xy[p].x, q0.y - point_xy[p].y};
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, q1.y * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                    const float rsigma = q2.y;
                    const bool ball = rsigma > 0;
#pragma unroll
                    for (int s = 0; s < SPLIT + 1; s++) {
                        if (!first && (s == 0 || s == SPLIT)) continue;
                        const float ts = depth_min[p] + interval[p] * s;
                        const float delta = (ts - t_peak) * rsigma;
                        const float g = ball ? expf(-0.5f * delta * delta) : 0.f;
                        const float one_minus_gaussian = 1.f - alpha * g;
                        const float rvacancy = rsqrtf(one_minus_gaussian);
                        T_p[p][s] *= (ts > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
                    }
                    done[p] = contributor >= last_contributor[p];
                    point_done += (contributor == last_contributor[p]) ? 1 : 0;
                }
                all_done = point_done == point_num_round;
            }
        }
#pragma unroll
        for (int p = 0; p < PB; p++) {
            if (first) inside_range[p] = (T_p[p][0] >= 0.5f) && (T_p[p][SPLIT] <= 0.5f) && inside_range[p];
            int start_id = 0;
#pragma unroll
            for (int s = 1; s < SPLIT; s++) start_id = T_p[p][s] >= 0.5f ? s : start_id;
            depth_max[p] = depth_min[p] + (start_id + 1) * interval[p];
            depth_min[p] = depth_min[p] + (start_id + 0) * interval[p];
            float lo = T_p[p][0], hi = T_p[p][1];
#pragma unroll
            for (int s = 1; s < SPLIT; s++) {
                lo = (start_id == s) ? T_p[p][s] : lo;
                hi = (start_id == s) ? T_p[p][s + 1] : hi;
            }
            T_p[p][0] = lo;
            T_p[p][SPLIT] = hi;
        }
    }

#pragma unroll
    for (int p = 0; p < PB; p++) {
        if (p >= point_num_round) continue;
        const float w_max = __saturatef((T_p[p][0] - 0.5f) / (T_p[p][0] - T_p[p][SPLIT]));
        const float w_min = 1.f - w_max;
        const float mDepth = inside_range[p] ? w_max * depth_max[p] + w_min * depth_min[p] : 0.f;
        if constexpr (SDF) {
            output[point_idx[p]] = mDepth - point_ts[point_idx[p]];
            median_depth[point_idx[p]] = mDepth;
        } else {
            const float2 pixnf = {(point_xy[p].x - static_cast<float>(W - 1) / 2.f) / focal_x,
                                  (point_xy[p].y - static_cast<float>(H - 1) / 2.f) / focal_y};
            const float rln = rnorm3df(pixnf.x, pixnf.y, 1.f);
            const float depth = mDepth * rln;
            float* o = output + 3 * (size_t)point_idx[p];
            o[0] = pixnf.x * depth;
            o[1] = pixnf.y * depth;
            o[2] = depth;
            median_depth[point_idx[p]] = mDepth;
            n_contrib[point_idx[p]] = last_contributor[p];
        }
        inside_output[point_idx[p]] = inside_range[p] ? 1 : 0;
    }
    if (!SDF && tid == 0) block_max_contributor[block_id] = max_contributor;
}

void launch_sample_depth_forward(const FrameParams& fp, int tiles, uint32_t max_blocks, const uint32_t* tile_offsets,
                                 const uint2* gaussian_ranges, const uint2* point_ranges, const uint32_t* gaussian_list,
                                 const uint32_t* point_list, const SplatRecord* records, const float2* points2D,
                                 uint32_t* n_contrib, float* median_depth, uint32_t* block_max_contributor,
                                 float* output, uint8_t* inside, cudaStream_t stream) {
    if (max_blocks == 0) return;
    sample_depth_forward_kernel<SPLIT_ITERATIONS, false><<<max_blocks, TILE_PIXELS, 0, stream>>>(
        tiles, fp.W, fp.H, fp.focal_x, fp.focal_y, tile_offsets, gaussian_ranges, point_ranges, gaussian_list,
        point_list, records, points2D, nullptr, n_contrib, median_depth, block_max_contributor, output, inside);
}

void launch_evaluate_sdf(const FrameParams& fp, int tiles, uint32_t max_blocks, const uint32_t* tile_offsets,
                         const uint2* gaussian_ranges, const uint2* point_ranges, const uint32_t* gaussian_list,
                         const uint32_t* point_list, const SplatRecord* records, const float2* points2D,
                         const float* point_ts, float* out_depth, float* out_sdf, uint8_t* inside,
                         cudaStream_t stream) {
    if (max_blocks == 0) return;
    // the reference instantiates SPLIT_ITERATIONS + 1 iterations here (sample_forward.cu:792)
    sample_depth_forward_kernel<SPLIT_ITERATIONS + 1, true><<<max_blocks, TILE_PIXELS, 0, stream>>>(
        tiles, fp.W, fp.H, fp.focal_x, fp.focal_y, tile_offsets, gaussian_ranges, point_ranges, gaussian_list,
        point_list, records, points2D, point_ts, nullptr, out_depth, nullptr, out_sdf, inside);
}

// ---- backward (sample_backward.cu:77-358) ---------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(TILE_PIXELS)
    sample_depth_backward_kernel(int tiles, int W, int H, float focal_x, float focal_y,
                                 const uint32_t* __restrict__ block_max_contributor,
                                 const uint32_t* __restrict__ tile_offsets, const uint2* __restrict__ gaussian_ranges,
                                 const uint2* __restrict__ point_ranges, const uint32_t* __restrict__ gaussian_list,
                                 const uint32_t* __restrict__ point_list, const SplatRecord* __restrict__ records,
                                 const float2* __restrict__ points2D, const uint32_t* __restrict__ n_contrib,
                                 const float* __restrict__ median_depth, const uint8_t* __restrict__ inside,
                                 const float* __restrict__ dL_doutputs, float2* __restrict__ dL_dpoints2D,
                                 double* __restrict__ grad_acc) {
    __shared__ float4 s_rec[TILE_PIXELS * 3];
    __shared__ uint32_t s_id[TILE_PIXELS];
    __shared__ int s_tile;

    const uint32_t block_id = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    const int lane = tid & 31;
    if (tid == 0) s_tile = tile_of_block(tile_offsets, tiles, block_id);
    __syncthreads();
    const int tile_id = s_tile;
    if (tile_id < 0) return;
    const uint32_t max_contributor = block_max_contributor[block_id];
    const uint32_t tile_offset = (tile_id == 0) ? 0 : tile_offsets[tile_id - 1];
    const int p_round = block_id - tile_offset;
    const uint2 p_range = point_ranges[tile_id];
    const uint2 range = gaussian_ranges[tile_id];
    const float ddelx_dx = 0.5f * W;
    const float ddely_dy = 0.5f * H;

    uint32_t point_idx[PB];
    float2 point_xy[PB];
    float dL_dDepth[PB];
    float2 dL_dpoint_xy[PB];
    uint32_t last_contributor[PB] = {0};
    float mDepth[PB];
    bool inside_range[PB] = {false};
    bool done[PB] = {false};
    int point_done = 0;
    int point_num_round = 0;
#pragma unroll
    for (int p = 0; p < PB; p++) {
        point_idx[p] = 0;
        point_xy[p] = {0.f, 0.f};
        dL_dDepth[p] = 0.f;
        dL_dpoint_xy[p] = {0.f, 0.f};
        mDepth[p] = 0.f;
        const int progress = (p_round * PB + p) * TILE_PIXELS + tid;
        if (p_range.x + progress < p_range.y) {
            const uint32_t pid = __ldg(point_list + p_range.x + progress);
            point_idx[p] = pid;
            last_contributor[p] = n_contrib[pid];
            const float3 dL_doutput = {dL_doutputs[3 * (size_t)pid], dL_doutputs[3 * (size_t)pid + 1],
                                       dL_doutputs[3 * (size_t)pid + 2]};
            point_xy[p] = points2D[pid];
            mDepth[p] = median_depth[pid];
            inside_range[p] = inside[pid] != 0;
            done[p] = (last_contributor[p] == 0) || !inside_range[p];
            point_done += done[p] ? 1 : 0;
            const float2 pixnf = {(point_xy[p].x - static_cast<float>(W - 1) / 2.f) / focal_x,
                                  (point_xy[p].y - static_cast<float>(H - 1) / 2.f) / focal_y};
            const float rln = rnorm3df(pixnf.x, pixnf.y, 1.f);
            const float rln2 = 1.f / (pixnf.x * pixnf.x + pixnf.y * pixnf.y + 1.f);
            const float depth = mDepth[p] * rln;
            const float dL_ddepth = dL_doutput.x * pixnf.x + dL_doutput.y * pixnf.y + dL_doutput.z;
            dL_dDepth[p] = rln * dL_ddepth;
            const float aux = (dL_doutput.x * pixnf.x + dL_doutput.y * pixnf.y + dL_doutput.z) * rln2;
            const float2 dL_dpixnf = {(dL_doutput.x - aux * pixnf.x) * depth, (dL_doutput.y - aux * pixnf.y) * depth};
            dL_dpoint_xy[p] = {dL_dpixnf.x / focal_x, dL_dpixnf.y / focal_y};
            point_num_round++;
        }
    }

    bool all_done = point_done == point_num_round;
    float dT_dtm[PB] = {0.f};
    float dL_dmt_dT_dtm[PB] = {0.f};
    int toDo = max_contributor;
    const int rounds = (max_contributor + TILE_PIXELS - 1) / TILE_PIXELS;
    uint32_t contributor = 0;
    // ---- pass A: dT/dt_m (sample_backward.cu:172-215).  NOTE: the reference walks every entry here (no
    // `!all_done` guard) but updates only points that are not done, which is what this loop does too ----
    for (int i = 0; i < rounds; i++, toDo -= TILE_PIXELS) {
        __syncthreads();
        const uint32_t progress = i * TILE_PIXELS + tid;
        if (progress < max_contributor) stage3(s_rec, tid, records, __ldg(gaussian_list + range.x + progress));
        __syncthreads();
        const int lim = min(TILE_PIXELS, toDo);
        for (int j = 0; j < lim; j++) {
            contributor++;
            if (all_done) continue;
            const float4 q0 = s_rec[j * 3], q1 = s_rec[j * 3 + 1], q2 = s_rec[j * 3 + 2];
#pragma unroll
            for (int p = 0; p < PB; p++) {
                if (p >= point_num_round || done[p]) continue;
                const float2 d = {q0.x - point_xy[p].x, q0.y - point_xy[p].y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, q1.y * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                const float rsigma = q2.y;
                const float t_delta = (mDepth[p] - t_peak) * rsigma;
                const float G_exp = expf(-0.5f * t_delta * t_delta);
                const float Gt = alpha * G_exp;
                dT_dtm[p] += -0.25f * Gt / (1.f - Gt) * fabsf(t_delta) * rsigma;
                done[p] = contributor >= last_contributor[p];
                point_done += (contributor == last_contributor[p]) ? 1 : 0;
            }
            all_done = point_done == point_num_round;
        }
    }
#pragma unroll
    for (int p = 0; p < PB; p++) dL_dmt_dT_dtm[p] = dL_dDepth[p] / fmaxf(-dT_dtm[p], 1e-7f);

    toDo = max_contributor;
    contributor = 0;
    point_done = 0;
#pragma unroll
    for (int p = 0; p < PB; p++) {
        done[p] = (last_contributor[p] == 0) || !inside_range[p];
        if (p < point_num_round) point_done += done[p] ? 1 : 0;
    }

    // ---- pass B: gradients (sample_backward.cu:229-352) ----
    for (int i = 0; i < rounds; i++, toDo -= TILE_PIXELS) {
        __syncthreads();
        const uint32_t progress = i * TILE_PIXELS + tid;
        if (progress < max_contributor) {
            const uint32_t id = __ldg(gaussian_list + range.x + progress);
            s_id[tid] = id;
            stage3(s_rec, tid, records, id);
        }
        __syncthreads();
        const int lim = min(TILE_PIXELS, toDo);
        for (int j = 0; j < lim; j++) {
            contributor++;
            const float4 q0 = s_rec[j * 3], q1 = s_rec[j * 3 + 1];
            const float4 con_o = {q0.z, q0.w, q1.x, q1.y};
            float G[PB];
            bool valid[PB];
            bool any_valid = false;
#pragma unroll
            for (int p = 0; p < PB; p++) {
                const float2 d = {q0.x - point_xy[p].x, q0.y - point_xy[p].y};
                const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                G[p] = expf(power);
                const float alpha = con_o.w * G[p];
                valid[p] = p < point_num_round && !(done[p] || (power > 0.0f) || (alpha < 1.0f / 255.0f));
                any_valid |= valid[p];
            }
            if (!__any_sync(0xffffffffu, any_valid)) continue;

            const float4 q2 = s_rec[j * 3 + 2];
            const float4 ray_plane = {q1.z, q1.w, q2.x, q2.y};
            float dL_dt_local = 0.f, dL_drsigma_local = 0.f;
            float2 dL_dray_planes_local = {0.f, 0.f};
            float2 dL_dmean2D_local = {0.f, 0.f};
            float4 dL_dconic2D_local = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int p = 0; p < PB; p++) {
                if (!valid[p]) continue;
                const float2 d = {q0.x - point_xy[p].x, q0.y - point_xy[p].y};
                const float alpha = fminf(0.99f, con_o.w * G[p]);
                const float t_peak = ray_plane.x * d.x + ray_plane.y * d.y + ray_plane.z;
                const float rsigma = ray_plane.w;
                const float t_delta = (mDepth[p] - t_peak) * rsigma;
                const float G_exp = expf(-0.5f * t_delta * t_delta);
                const float Gt = alpha * G_exp;
                float dL_dGt = dL_dmt_dT_dtm[p] * 0.25f / (1.f - Gt);
                dL_dGt = mDepth[p] > t_peak ? dL_dGt : -dL_dGt;
                dL_dGt = rsigma > 0 ? dL_dGt : 0.f;
                const float dL_dopa = dL_dGt * G_exp - dL_dmt_dT_dtm[p] * (t_delta > 0 ? 0.5f / (1.f - alpha) : 0.f);
                const float dL_ddelta = -dL_dGt * Gt * t_delta;
                dL_drsigma_local += dL_ddelta * (mDepth[p] - t_peak);
                const float dL_dt_peak = -dL_ddelta * rsigma;
                dL_dt_local += dL_dt_peak;
                dL_dray_planes_local.x += dL_dt_peak * d.x;
                dL_dray_planes_local.y += dL_dt_peak * d.y;
                const float dL_dG = con_o.w * dL_dopa;
                const float gdx = G[p] * d.x;
                const float gdy = G[p] * d.y;
                const float dG_ddelx = -gdx * con_o.x - gdy * con_o.y;
                const float dG_ddely = -gdy * con_o.z - gdx * con_o.y;
                const float dL_ddelx = dL_dG * dG_ddelx + dL_dt_peak * ray_plane.x;
                const float dL_ddely = dL_dG * dG_ddely + dL_dt_peak * ray_plane.y;
                dL_dmean2D_local.x += dL_ddelx;
                dL_dmean2D_local.y += dL_ddely;
                dL_dpoint_xy[p].x -= dL_ddelx;
                dL_dpoint_xy[p].y -= dL_ddely;
                dL_dconic2D_local.x += -0.5f * gdx * d.x * dL_dG;
                dL_dconic2D_local.y += -0.5f * gdx * d.y * dL_dG;
                dL_dconic2D_local.z += -0.5f * gdy * d.y * dL_dG;
                dL_dconic2D_local.w += G[p] * dL_dopa;
                done[p] = contributor >= last_contributor[p];
                point_done += (contributor == last_contributor[p]) ? 1 : 0;
            }
            dL_dmean2D_local.x *= ddelx_dx;
            dL_dmean2D_local.y *= ddely_dy;
            // warp sums, then one lane per component adds into the Gaussian's accumulator row (ggs_common.cuh)
            const float s0 = warp_sum(dL_dmean2D_local.x), s1 = warp_sum(dL_dmean2D_local.y);
            const float s3 = warp_sum(dL_dconic2D_local.x), s4 = warp_sum(dL_dconic2D_local.y),
                        s5 = warp_sum(dL_dconic2D_local.z), s6 = warp_sum(dL_dconic2D_local.w);
            const float s13 = warp_sum(dL_dray_planes_local.x), s14 = warp_sum(dL_dray_planes_local.y),
                        s15 = warp_sum(dL_dt_local), s16 = warp_sum(dL_drsigma_local);
            double* row = grad_acc + (size_t)s_id[j] * GRAD_STRIDE;
            float v = 0.f;
            int c = -1;
            switch (lane) {
                case 0: v = s0; c = 0; break;
                case 1: v = s1; c = 1; break;
                case 2: v = s3; c = 3; break;
                case 3: v = s4; c = 4; break;
                case 4: v = s5; c = 5; break;
                case 5: v = s6; c = 6; break;
                case 6: v = s13; c = 13; break;
                case 7: v = s14; c = 14; break;
                case 8: v = s15; c = 15; break;
                case 9: v = s16; c = 16; break;
                default: break;
            }
            if (c >= 0) atomicAdd(row + c, static_cast<double>(v));
        }
    }
#pragma unroll
    for (int p = 0; p < PB; p++)
        if (p < point_num_round)
            dL_dpoints2D[point_idx[p]] = {dL_dpoint_xy[p].x * ddelx_dx, dL_dpoint_xy[p].y * ddely_dy};
}

void launch_sample_depth_backward(const FrameParams& fp, int tiles, uint32_t max_blocks,
                                  const uint32_t* block_max_contributor, const uint32_t* tile_offsets,
                                  const uint2* gaussian_ranges, const uint2* point_ranges,
                                  const uint32_t* gaussian_list, const uint32_t* point_list,
                                  const SplatRecord* records, const float2* points2D, const uint32_t* n_contrib,
                                  const float* median_depth, const uint8_t* inside, const float* dL_doutputs,
                                  float2* dL_dpoints2D, double* grad_acc, cudaStream_t stream) {
    if (max_blocks == 0) return;
    sample_depth_backward_kernel<<<max_blocks, TILE_PIXELS, 0, stream>>>(
        tiles, fp.W, fp.H, fp.focal_x, fp.focal_y, block_max_contributor, tile_offsets, gaussian_ranges, point_ranges,
        gaussian_list, point_list, records, points2D, n_contrib, median_depth, inside, dL_doutputs, dL_dpoints2D,
        grad_acc);
}

// Gradient of the projected query position back to the 3D query point (sample_backward.cu:42-75); rows of points
// that fell outside the image get zeros.
__global__ void points_backward_kernel(int PN, const float* __restrict__ points3D, const float* __restrict__ proj,
                                       const uint32_t* __restrict__ touched, const float2* __restrict__ dL_dpoints2D,
                                       float* __restrict__ dL_dpoints3D) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)PN) return;
    float3 r = {0.f, 0.f, 0.f};
    if (touched[idx] != 0) {
        const float3 m = {points3D[3 * idx], points3D[3 * idx + 1], points3D[3 * idx + 2]};
        const float4 m_hom = xform_point_4x4(m, proj);
        const float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
        const float2 g = dL_dpoints2D[idx];
        r.x = (proj[0] * m_w - proj[3] * mul1) * g.x + (proj[1] * m_w - proj[3] * mul2) * g.y;
        r.y = (proj[4] * m_w - proj[7] * mul1) * g.x + (proj[5] * m_w - proj[7] * mul2) * g.y;
        r.z = (proj[8] * m_w - proj[11] * mul1) * g.x + (proj[9] * m_w - proj[11] * mul2) * g.y;
    }
    dL_dpoints3D[3 * idx] = r.x;
    dL_dpoints3D[3 * idx + 1] = r.y;
    dL_dpoints3D[3 * idx + 2] = r.z;
}

void launch_points_backward(const FrameParams& fp, int PN, const float* points3D, const uint32_t* touched,
                            const float2* dL_dpoints2D, float* dL_dpoints3D, cudaStream_t stream) {
    if (PN == 0) return;
    points_backward_kernel<<<(PN + 255) / 256, 256, 0, stream>>>(PN, points3D, fp.view.proj, touched, dL_dpoints2D,
                                                                dL_dpoints3D);
}

}  // namespace ggs
