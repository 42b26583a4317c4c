// Point queries against one view (mesh-extraction opacity test, `GaussianRasterizer.integrate`).
// Semantics follow DGR/cuda_rasterizer/sample_forward.cu:10-53 (preprocessPointsCUDA) and :55-169
// (evaluateTransmittanceCUDA), driven like rasterizer_impl.cu:594-815: points are binned by the
// tile they project into, a block owns (tile, round of 512 points), each thread carries two points
// through the tile's depth-sorted Gaussian list.
//
// Unlike the reference there is no block->tile expansion table: the grid is an upper bound and a
// block binary-searches its tile in the inclusive scan of per-tile round counts.
#include "ggs_kernels.cuh"
#include "ggs_math.cuh"

namespace ggs {

__global__ void __launch_bounds__(256)
    preprocess_points_kernel(const FrameParams fp, int PN, const float* __restrict__ points3D,
                             float2* __restrict__ points2D, float* __restrict__ ts, uint32_t* __restrict__ touched) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)PN) return;
    touched[idx] = 0;
    const float3 p_orig = {__ldg(points3D + 3 * idx), __ldg(points3D + 3 * idx + 1), __ldg(points3D + 3 * idx + 2)};
    const float3 p_view = xform_point_4x3(p_orig, fp.view.view);
    if (p_view.z <= 0.2f) return;
    const float4 p_hom = xform_point_4x4(p_orig, fp.view.proj);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
    const float2 point_image = {ndc_to_pix(p_proj.x, fp.W), ndc_to_pix(p_proj.y, fp.H)};
    if (point_image.x < 0 || point_image.x > fp.W - 1 || point_image.y < 0 || point_image.y > fp.H - 1) return;
    ts[idx] = norm3df(p_view.x, p_view.y, p_view.z);
    points2D[idx] = point_image;
    touched[idx] = 1;
}

void launch_preprocess_points(const FrameParams& fp, int PN, const float* points3D, float2* points2D, float* depths,
                              uint32_t* touched, uint8_t*, cudaStream_t stream) {
    if (PN == 0) return;
    preprocess_points_kernel<<<(PN + 255) / 256, 256, 0, stream>>>(fp, PN, points3D, points2D, depths, touched);
}

// rounds per tile = ceil(points / 512)   (rasterizer_impl.cu:163-172)
__global__ void point_blocks_kernel(int tiles, const uint2* __restrict__ point_ranges,
                                    uint32_t* __restrict__ tile_rounds) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= tiles) return;
    constexpr int PER_ROUND = TILE_PIXELS * POINT_BATCH;
    const uint2 r = point_ranges[idx];
    tile_rounds[idx] = (r.y - r.x + PER_ROUND - 1) / PER_ROUND;
}

void launch_point_blocks(int tiles, const uint2* point_ranges, uint32_t* tile_rounds, cudaStream_t stream) {
    point_blocks_kernel<<<(tiles + 255) / 256, 256, 0, stream>>>(tiles, point_ranges, tile_rounds);
}

__global__ void __launch_bounds__(TILE_PIXELS)
    integrate_kernel(int tiles, const uint32_t* __restrict__ tile_offsets, const uint2* __restrict__ gaussian_ranges,
                     const uint2* __restrict__ point_ranges, const uint32_t* __restrict__ gaussian_list,
                     const uint32_t* __restrict__ point_list, const SplatRecord* __restrict__ records,
                     const float2* __restrict__ points2D, const float* __restrict__ point_ts,
                     float* __restrict__ out_transmittance, uint8_t* __restrict__ inside) {
    __shared__ float4 s_q0[TILE_PIXELS];
    __shared__ float4 s_q1[TILE_PIXELS];
    __shared__ float2 s_q2[TILE_PIXELS];
    __shared__ int s_tile;

    const uint32_t block_id = blockIdx.x;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) {
        // smallest tile with tile_offsets[tile] > block_id
        int lo = 0, hi = tiles;
        const uint32_t total = tile_offsets[tiles - 1];
        if (block_id >= total) {
            lo = -1;
        } else {
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (tile_offsets[mid] > block_id)
                    hi = mid;
                else
                    lo = mid + 1;
            }
        }
        s_tile = lo;
    }
    __syncthreads();
    const int tile_id = s_tile;
    if (tile_id < 0) return;
    const uint32_t tile_offset = (tile_id == 0) ? 0 : tile_offsets[tile_id - 1];
    const int p_round = block_id - tile_offset;
    const uint2 p_range = point_ranges[tile_id];
    const uint2 range = gaussian_ranges[tile_id];
    const int rounds = ((range.y - range.x + TILE_PIXELS - 1) / TILE_PIXELS);

    bool done[POINT_BATCH] = {false};
    float T_point[POINT_BATCH];
    float T_gaussian[POINT_BATCH];
    uint32_t point_idx[POINT_BATCH];
    float2 point_xy[POINT_BATCH];
    float point_t[POINT_BATCH];
    int point_done = 0;
    int point_num_round = 0;
#pragma unroll
    for (int p = 0; p < POINT_BATCH; p++) {
        const int progress = (p_round * POINT_BATCH + p) * TILE_PIXELS + tid;
        if (p_range.x + progress < p_range.y) {
            T_point[p] = 1.f;
            T_gaussian[p] = 1.f;
            const uint32_t pid = __ldg(point_list + p_range.x + progress);
            point_idx[p] = pid;
            point_xy[p] = points2D[pid];
            point_t[p] = point_ts[pid];
            point_num_round++;
        }
    }
    bool all_done = point_num_round == 0;
    int toDo = range.y - range.x;
    for (int i = 0; i < rounds; i++, toDo -= TILE_PIXELS) {
        const int block_done = __syncthreads_and(all_done);
        if (block_done) break;
        const int progress = i * TILE_PIXELS + tid;
        if (range.x + progress < range.y) {
            const uint32_t id = __ldg(gaussian_list + range.x + progress);
            const float4* src = reinterpret_cast<const float4*>(records + id);
            s_q0[tid] = __ldg(src);
            s_q1[tid] = __ldg(src + 1);
            s_q2[tid] = __ldg(reinterpret_cast<const float2*>(src + 2));
        }
        __syncthreads();
        const int lim = min(TILE_PIXELS, toDo);
        for (int j = 0; !all_done && j < lim; j++) {
            const float4 q0 = s_q0[j];
            const float4 q1 = s_q1[j];
            const float2 q2 = s_q2[j];
#pragma unroll
            for (int p = 0; p < POINT_BATCH; p++) {
                if (p >= point_num_round || done[p]) continue;
                const float2 d = {q0.x - point_xy[p].x, q0.y - point_xy[p].y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                if (power > 0.0f) continue;
                const float alpha = fminf(0.99f, q1.y * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = T_gaussian[p] * (1.f - alpha);
                if (test_T < 0.0001f) {
                    done[p] = true;
                    point_done++;
                    continue;
                }
                const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                const float rsigma = q2.y;
                const float delta = (t_peak - point_t[p]) * rsigma;
                const float g = rsigma > 0 ? expf(-0.5f * delta * delta) : 0.f;
                const float one_minus_gaussian = 1.f - alpha * g;
                const float rvacancy = rsqrtf(one_minus_gaussian);
                T_point[p] *= (point_t[p] > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
                T_gaussian[p] = test_T;
            }
            all_done = point_done == point_num_round;
        }
        __syncthreads();
    }
#pragma unroll
    for (int p = 0; p < POINT_BATCH; p++) {
        if (p < point_num_round) {
            out_transmittance[point_idx[p]] = T_point[p];
            inside[point_idx[p]] = 1;
        }
    }
}

void launch_integrate(int tiles, uint32_t max_blocks, const uint32_t* tile_offsets, const uint2* gaussian_ranges,
                      const uint2* point_ranges, const uint32_t* gaussian_list, const uint32_t* point_list,
                      const SplatRecord* records, const float2* points2D, const float* point_depths,
                      float* out_transmittance, uint8_t* inside, cudaStream_t stream) {
    if (max_blocks == 0) return;
    integrate_kernel<<<max_blocks, TILE_PIXELS, 0, stream>>>(tiles, tile_offsets, gaussian_ranges, point_ranges,
                                                            gaussian_list, point_list, records, points2D,
                                                            point_depths, out_transmittance, inside);
}

}  // namespace ggs
