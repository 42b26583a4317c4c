// Point queries against one prepared view: `integrate` (opacity of mesh-extraction vertices), `sample_depth`
// (median depth at sub-pixel positions, forward + backward) and `evaluate_sdf`.
//
// What is computed per query point is the reference's arithmetic, statement for statement, in the reference's
// list order (DGR/cuda_rasterizer/sample_forward.cu:55-169 evaluateTransmittanceCUDA, :171-428 evaluateSDFCUDA,
// :430-700 sampleDepthCUDA; sample_backward.cu:77-358) -- a query's result depends only on its own position and on
// its tile's depth-sorted Gaussian list.  How the work is laid out is new:
//
//   * queries are GROUPED IN SPACE.  The reference sorts them by tile only, so the 32 queries of a warp are scattered
//     over the tile's 16x16 pixels and every one of them has to look at every list entry.  Here the sort key is
//     (tile, Morton code of the quarter-pixel cell inside the tile): a warp owns 32 neighbours -- a footprint of a few
//     pixels -- and one CTA the 128 queries of four such warps.  No count is read back by the host: queries that
//     miss the image get the largest key and sort behind everything else;
//   * per 32 list entries each lane tests ONE entry against the warp's bounding box (exact minimum of the conic's
//     quadratic form over the box against the alpha >= 1/255 threshold, with rounding slack).  Only the few per
//     cent of entries that can reach the box are evaluated per query;
//   * list entries are gathered as the first three quads of the 64-byte SplatRecord with cp.async into a double
//     buffer shared by the CTA (48-byte stride: conflict-free 16-byte stores, broadcast reads);
//   * sample_depth / evaluate_sdf keep the per-warp acceptance mask of the first pass in shared memory and run the
//     bisection passes on it, with the exact far-splat shortcuts of the image kernel (ggs_render_forward.cu): a splat
//     whose depth profile is more than 6.5 sigma from the whole bracket has 1 - alpha * exp() == 1 exactly, so its
//     factor is (1 - alpha) in front of the bracket and 1 behind it, the leading run of such entries is carried over as one
//     product, and entries behind every bracket are dropped from the mask.  Per-query float products keep their order;
//   * the backward reduces the ten per-(warp, entry) partial sums with one transposing butterfly (12 shuffles) and
//     ten lanes issue one double-precision RED each into the Gaussian's accumulator row (ggs_common.cuh), which the
//     shared per-Gaussian chain rule (ggs_gaussian_backward.cu) consumes.
#include <algorithm>

#include "ggs_kernels.cuh"
#include "ggs_math.cuh"
#include "ggs_tile.cuh"

namespace ggs {

namespace {

constexpr int Q_WARPS = QUERY_CTA / 32;
constexpr int Q_THREADS = QUERY_CTA;
constexpr int Q_BATCH = QUERY_CTA;  // list entries per staged batch (one per thread)
constexpr int Q_QUADS = 3;          // staged quads per record: (mean, conic.xy) (conic.z, opacity, ray.xy) (ray.zw, -, -)
constexpr uint32_t Q_FULL = 0xffffffffu;
constexpr uint32_t Q_NEVER = 0xffffffffu;
constexpr uint32_t NO_TILE = 0xffffffffu;
constexpr int Q_MASK_WORDS = 256;  // words of a warp's acceptance mask kept (and thinned) in shared memory

// ---- projection of the query points -----------------------------------------------------------------------------
// preprocessPointsCUDA (sample_forward.cu:10-53) + the tile choice of createWithKeys (rasterizer_impl.cu:109-137).
// xy = (-1, -1) marks a query that is not evaluated (behind the near plane or outside the image).
// 6 bits -> the even bits of 12 (Morton interleave)
__device__ __forceinline__ uint32_t spread6(uint32_t v) {
    v = (v | (v << 4)) & 0x30fu;
    v = (v | (v << 2)) & 0x333u;
    v = (v | (v << 1)) & 0x555u;
    return v;
}

__global__ void __launch_bounds__(256)
    query_project_kernel(const FrameParams fp, int PN, const float* __restrict__ points3D, float2* __restrict__ xy,
                         float* __restrict__ ts, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)PN) return;
    uint32_t key = 0xffffffffu;
    float2 out = {-1.f, -1.f};
    float t = 0.f;
    const float3 p_orig = {__ldg(points3D + 3 * idx), __ldg(points3D + 3 * idx + 1), __ldg(points3D + 3 * idx + 2)};
    const float3 p_view = xform_point_4x3(p_orig, fp.view.view);
    if (p_view.z > 0.2f) {
        const float4 p_hom = xform_point_4x4(p_orig, fp.view.proj);
        const float p_w = 1.0f / (p_hom.w + 0.0000001f);
        const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
        const float2 pi = {ndc_to_pix(p_proj.x, fp.W), ndc_to_pix(p_proj.y, fp.H)};
        if (!(pi.x < 0 || pi.x > fp.W - 1 || pi.y < 0 || pi.y > fp.H - 1)) {
            t = norm3df(p_view.x, p_view.y, p_view.z);
            out = pi;
            const int tx = min(fp.gx - 1, max((int)0, (int)((pi.x + 0.5f) / TILE_X)));
            const int ty = min(fp.gy - 1, max((int)0, (int)((pi.y + 0.5f) / TILE_Y)));
            // quarter-pixel cell inside the tile, Morton order (x in the even bits)
            const int cx = min(63, max(0, (int)((pi.x + 0.5f - tx * TILE_X) * 4.f)));
            const int cy = min(63, max(0, (int)((pi.y + 0.5f - ty * TILE_Y) * 4.f)));
            key = ((uint32_t)(ty * fp.gx + tx) << QUERY_CELL_BITS) | spread6((uint32_t)cx) | (spread6((uint32_t)cy) << 1);
        }
    }
    xy[idx] = out;
    ts[idx] = t;
    keys[idx] = key;
    vals[idx] = (uint32_t)idx;
}

// Query ranges per tile from the sorted keys; tiles without queries keep (0, 0) (zeroed by the caller).
__global__ void __launch_bounds__(256)
    query_ranges_kernel(const uint32_t* __restrict__ keys, uint32_t n, uint2* __restrict__ ranges) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const uint32_t k = __ldg(keys + idx);
    if (k == 0xffffffffu) return;  // not evaluated: sorted behind every tile
    const uint32_t cur = k >> QUERY_CELL_BITS;
    if (idx == 0) {
        ranges[cur].x = 0;
    } else {
        const uint32_t prev = __ldg(keys + idx - 1) >> QUERY_CELL_BITS;
        if (prev != cur) {
            ranges[prev].y = idx;
            ranges[cur].x = idx;
        }
    }
    const uint32_t nk = idx + 1 < n ? __ldg(keys + idx + 1) : 0xffffffffu;
    if (nk == 0xffffffffu) ranges[cur].y = idx + 1;
}

__global__ void query_rounds_kernel(int tiles, const uint2* __restrict__ ranges, uint32_t* __restrict__ rounds) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= tiles) return;
    const uint2 r = ranges[idx];
    rounds[idx] = (r.y - r.x + QUERY_CTA - 1) / QUERY_CTA;
}

// block -> tile table: smallest tile with offsets[tile] > block (offsets = inclusive scan of rounds per tile)
__global__ void query_block_table_kernel(uint32_t max_blocks, int tiles, const uint32_t* __restrict__ offsets,
                                         uint32_t* __restrict__ table) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= max_blocks) return;
    uint32_t t = NO_TILE;
    if (b < __ldg(offsets + tiles - 1)) {
        int lo = 0, hi = tiles;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(offsets + mid) > b)
                hi = mid;
            else
                lo = mid + 1;
        }
        t = (uint32_t)lo;
    }
    table[b] = t;
}

// ---- shared device pieces ---------------------------------------------------------------------------------------
struct Box {
    float x0, y0, x1, y1;
};

// Can the splat reach alpha >= 1/255 at any position of the box?  Conservative: `false` only when every position is
// provably rejected (the general-rectangle form of may_touch in ggs_render_forward.cu).
__device__ __forceinline__ bool may_touch_box(const float4 q0, const float4 q1, const Box b) {
    const float A = q0.z, B = q0.w, C = q1.x, o = q1.y;
    const float dx1 = q0.x - b.x0, dx0 = q0.x - b.x1;  // d = mean - position ranges over [dx0, dx1] x [dy0, dy1]
    const float dy1 = q0.y - b.y0, dy0 = q0.y - b.y1;
    const float det = A * C - B * B;
    if (!(A > 0.f && C > 0.f && det > 0.f)) return true;  // not positive definite (or NaN): no claim
    const float L = __logf(255.f * o);
    if (L < -0.01f) return false;  // o < 1/255: power <= 0 can never lift alpha over the threshold
    if (dx0 <= 0.f && dx1 >= 0.f && dy0 <= 0.f && dy1 >= 0.f) return true;
    const float rA = __frcp_rn(A), rC = __frcp_rn(C);
    const float ya = fminf(fmaxf(-B * dx0 * rC, dy0), dy1);
    const float yb = fminf(fmaxf(-B * dx1 * rC, dy0), dy1);
    const float xa = fminf(fmaxf(-B * dy0 * rA, dx0), dx1);
    const float xb = fminf(fmaxf(-B * dy1 * rA, dx0), dx1);
    const float e0 = A * dx0 * dx0 + 2.f * B * dx0 * ya + C * ya * ya;
    const float e1 = A * dx1 * dx1 + 2.f * B * dx1 * yb + C * yb * yb;
    const float e2 = A * xa * xa + 2.f * B * xa * dy0 + C * dy0 * dy0;
    const float e3 = A * xb * xb + 2.f * B * xb * dy1 + C * dy1 * dy1;
    const float qmin = fminf(fminf(e0, e1), fminf(e2, e3));
    const float Dx = fmaxf(fabsf(dx0), fabsf(dx1)), Dy = fmaxf(fabsf(dy0), fabsf(dy1));
    const float M = A * Dx * Dx + 2.f * fabsf(B) * Dx * Dy + C * Dy * Dy;
    return !(0.5f * qmin - 1e-5f * M > L + 0.01f);
}

// One depth sample of the vacancy-corrected transmittance product (sample_forward.cu:612-618).
__device__ __forceinline__ float depth_sample_factor(float ts, float t_peak, float rsigma, bool ball, float alpha) {
    const float delta = (ts - t_peak) * rsigma;
    const float g = ball ? expf(-0.5f * delta * delta) : 0.f;
    const float one_minus_gaussian = 1.f - alpha * g;
    const float rvacancy = rsqrtf(one_minus_gaussian);
    return (ts > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
}

__device__ __forceinline__ void stage3q(float4* buf, int slot, const SplatRecord* records, uint32_t id) {
    const float4* src = reinterpret_cast<const float4*>(records + id);
    float4* dst = buf + slot * Q_QUADS;
    cp_async16(dst, src);
    cp_async16(dst + 1, src + 1);
    cp_async16(dst + 2, src + 2);
}

// The queries of this CTA: 128 consecutive entries of its tile's sorted query list, one per thread.
struct Job {
    uint2 grange;  // the tile's Gaussian list
    bool valid;    // this thread holds a query
    uint32_t pid;  // its index in the caller's point array
};
__device__ __forceinline__ bool locate_job(const uint32_t* __restrict__ block_tile,
                                           const uint32_t* __restrict__ tile_offsets,
                                           const uint2* __restrict__ gaussian_ranges,
                                           const uint2* __restrict__ query_ranges,
                                           const uint32_t* __restrict__ query_list, Job& job) {
    const uint32_t tile = __ldg(block_tile + blockIdx.x);
    if (tile == NO_TILE) return false;
    const uint32_t first = tile ? __ldg(tile_offsets + tile - 1) : 0u;
    const uint2 pr = query_ranges[tile];
    const uint32_t sp = pr.x + (blockIdx.x - first) * QUERY_CTA + threadIdx.x;
    job.grange = gaussian_ranges[tile];
    job.valid = sp < pr.y;
    job.pid = job.valid ? __ldg(query_list + sp) : 0u;
    return true;
}

// Bounding box of the warp's queries (positions are >= 0, so their float bits order like unsigned integers).
__device__ __forceinline__ Box warp_box(bool valid, float2 p) {
    const uint32_t lo = 0u, hi = 0x7f800000u;
    Box b;
    b.x0 = __uint_as_float(__reduce_min_sync(Q_FULL, valid ? __float_as_uint(p.x) : hi));
    b.y0 = __uint_as_float(__reduce_min_sync(Q_FULL, valid ? __float_as_uint(p.y) : hi));
    b.x1 = __uint_as_float(__reduce_max_sync(Q_FULL, valid ? __float_as_uint(p.x) : lo));
    b.y1 = __uint_as_float(__reduce_max_sync(Q_FULL, valid ? __float_as_uint(p.y) : lo));
    return b;
}

// Transposing butterfly: every lane brings N values; afterwards the lane with scatter_component<N>(lane) == c holds
// the warp-wide sum of value c.  ceil(N/2) + ceil(N/4) + ... shuffles instead of 5 N.
template <int N, int BIT>
__device__ __forceinline__ float reduce_scatter(const float (&v)[N], const int lane) {
    constexpr int Hn = (N + 1) / 2;
    float a[Hn];
    const bool hi = lane & BIT;
#pragma unroll
    for (int i = 0; i < Hn; i++) {
        const float lo_v = v[i];
        const float hi_v = (i + Hn < N) ? v[(i + Hn < N) ? i + Hn : 0] : 0.f;
        const float keep = hi ? hi_v : lo_v;
        const float send = hi ? lo_v : hi_v;
        a[i] = keep + __shfl_xor_sync(Q_FULL, send, BIT);
    }
    if constexpr (BIT == 1) {
        static_assert(Hn == 1, "at most 32 values");
        return a[0];
    } else {
        return reduce_scatter<Hn, BIT / 2>(a, lane);
    }
}
template <int N>
__device__ __forceinline__ int scatter_component(const int lane) {
    int n[6];
    n[0] = N;
#pragma unroll
    for (int k = 0; k < 5; k++) n[k + 1] = (n[k] + 1) / 2;
    int idx = 0;
    bool ok = true;
#pragma unroll
    for (int k = 4; k >= 0; k--) {
        if (lane & (16 >> k)) idx += n[k + 1];
        ok = ok && idx < n[k];
    }
    return ok ? idx : -1;
}

// ---- integrate --------------------------------------------------------------------------------------------------
// COUNT: measurement only -- counters[0] += (query, entry) pairs evaluated after the warp-level culling,
// counters[1] += pairs that pass the power / alpha tests and update the transmittance (the algorithmic pairs:
// 2 ex2 + 1 rsqrt each on top of the alpha ex2), counters[2] += list entries the CTAs staged.
template <bool COUNT>
__global__ void __launch_bounds__(Q_THREADS, 8)
    integrate_kernel(const uint32_t* __restrict__ block_tile, const uint32_t* __restrict__ tile_offsets,
                     const uint2* __restrict__ gaussian_ranges, const uint2* __restrict__ query_ranges,
                     const uint32_t* __restrict__ gaussian_list, const uint32_t* __restrict__ query_list,
                     const SplatRecord* __restrict__ records, const float2* __restrict__ points2D,
                     const float* __restrict__ point_ts, float* __restrict__ out_transmittance,
                     uint8_t* __restrict__ inside, unsigned long long* __restrict__ counters) {
    __shared__ __align__(16) float4 s_rec[2 * Q_BATCH * Q_QUADS];
    uint32_t n_eval = 0, n_acc = 0, n_staged = 0;
    Job job;
    if (!locate_job(block_tile, tile_offsets, gaussian_ranges, query_ranges, query_list, job)) return;
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    const float2 pxy = job.valid ? points2D[job.pid] : float2{0.f, 0.f};
    const float pt = job.valid ? point_ts[job.pid] : 0.f;
    const Box box = warp_box(job.valid, pxy);

    const uint32_t n = job.grange.y - job.grange.x;
    const int rounds = (n + Q_BATCH - 1) / Q_BATCH;
    const uint32_t* const list = gaussian_list + job.grange.x;
    bool done = !job.valid;
    float T_point = 1.f, T_gaussian = 1.f;

    // rsqrt.approx(1.0): the vacancy factor of a pair whose Gaussian term does not reach 1 - alpha * g (see below)
    float c1;
    {
        const float one = 1.0f;
        asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(c1) : "f"(one));
    }
    const uint32_t s_rec_addr = static_cast<uint32_t>(__cvta_generic_to_shared(s_rec));
    constexpr uint32_t Q_STRIDE = Q_QUADS * sizeof(float4);
    uint32_t id_next = 0;
    if (tid < n) stage3q(s_rec, tid, records, __ldg(list + tid));
    cp_async_commit();
    if (Q_BATCH + tid < n) id_next = __ldg(list + Q_BATCH + tid);
    for (int i = 0; i < rounds; i++) {
        cp_async_wait_all();
        if (__syncthreads_and(done)) break;
        const uint32_t buf_addr = s_rec_addr + (i & 1) * (Q_BATCH * Q_STRIDE);
        if (i + 1 < rounds) {
            const uint32_t e = (i + 1) * Q_BATCH + tid;
            if (e < n) stage3q(s_rec + ((i + 1) & 1) * (Q_BATCH * Q_QUADS), tid, records, id_next);
            cp_async_commit();
            if (e + Q_BATCH < n) id_next = __ldg(list + e + Q_BATCH);
        }
        const int lim = min((uint32_t)Q_BATCH, n - i * Q_BATCH);
        if constexpr (COUNT) n_staged += tid < (uint32_t)lim ? 1u : 0u;
        for (int w = 0; w * 32 < lim; w++) {
            if (__all_sync(Q_FULL, done)) break;
            const int jl = w * 32 + lane;
            const uint32_t word_addr = buf_addr + w * (32 * Q_STRIDE);
            bool c = false;
            if (jl < lim) c = may_touch_box(lds128(word_addr + lane * Q_STRIDE), lds128(word_addr + lane * Q_STRIDE + 16), box);
            uint32_t cand = __ballot_sync(Q_FULL, c);
            while (cand) {
                const int jj = __ffs(cand) - 1;
                cand &= cand - 1;
                // The reference's per-query chain of `continue`s (sample_forward.cu:118-160), predicated: a query takes
                // part in an entry iff it is not done and passes the power / alpha tests; the entry that would push its
                // T below 1e-4 ends it without being applied.
                const uint32_t rec = word_addr + jj * Q_STRIDE;
                const float4 q0 = lds128(rec);
                const float4 q1 = lds128(rec + 16);
                const float2 d = {q0.x - pxy.x, q0.y - pxy.y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                const float alpha = fminf(0.99f, q1.y * expf(power));
                const bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                if constexpr (COUNT) n_eval += done ? 0u : 1u;
                if (!__any_sync(Q_FULL, ok)) continue;
                const float test_T = T_gaussian * (1.f - alpha);
                const bool stop = ok && test_T < 0.0001f;
                done = done || stop;
                const bool upd = ok && !stop;
                const float4 q2 = lds128(rec + 32);
                const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                const float rsigma = q2.y;
                const float delta = (t_peak - pt) * rsigma;
                // From 6.5 sigma on 1 - alpha * exp(-delta^2/2) rounds to exactly 1.0f (ggs_common.cuh), as it does
                // for a splat without depth extent: the factor is then (1 - alpha | 1) * rsqrt.approx(1) and neither
                // exp nor rsqrt is issued unless some query of the warp is closer.  A warp's 32 queries are
                // neighbours in the image, not in depth, but most of a ray's splats are far from any one of them.
                const bool quiet = !(rsigma > 0) || fabsf(delta) > QUIET_SIGMAS;
                float factor;
                if (__any_sync(Q_FULL, upd && !quiet)) {
                    const float g = rsigma > 0 ? expf(-0.5f * delta * delta) : 0.f;
                    const float one_minus_gaussian = 1.f - alpha * g;
                    const float rvacancy = rsqrtf(one_minus_gaussian);
                    factor = (pt > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
                } else {
                    factor = (pt > t_peak ? (1.f - alpha) : 1.f) * c1;
                }
                if (upd) {
                    T_point *= factor;
                    T_gaussian = test_T;
                    if constexpr (COUNT) n_acc++;
                }
            }
        }
    }
    cp_async_wait_all();
    if (job.valid) {
        out_transmittance[job.pid] = T_point;
        inside[job.pid] = 1;
    }
    if constexpr (COUNT) {
        const uint32_t e = __reduce_add_sync(Q_FULL, n_eval), a = __reduce_add_sync(Q_FULL, n_acc),
                       st = __reduce_add_sync(Q_FULL, n_staged);
        if (lane == 0) {
            atomicAdd(counters, (unsigned long long)e);
            atomicAdd(counters + 1, (unsigned long long)a);
            atomicAdd(counters + 2, (unsigned long long)st);
        }
    }
}

// ---- sample_depth / evaluate_sdf forward --------------------------------------------------------------------------
// SDF = false: sample_depth (ITERS = 5, bracket +-0.4; outputs the camera-space point, keeps n_contrib / median depth
//              for the backward).
// SDF = true : evaluate_sdf (ITERS = 6, bracket +-0.8; outputs the median ray depth and its difference to the query's
//              own ray depth).
template <int ITERS, bool SDF>
__global__ void __launch_bounds__(Q_THREADS, 8)
    sample_forward_kernel(int W, int H, float focal_x, float focal_y, const uint32_t* __restrict__ block_tile,
                          const uint32_t* __restrict__ tile_offsets, const uint2* __restrict__ gaussian_ranges,
                          const uint2* __restrict__ query_ranges, const uint32_t* __restrict__ gaussian_list,
                          const uint32_t* __restrict__ query_list, const SplatRecord* __restrict__ records,
                          const float2* __restrict__ points2D, const float* __restrict__ point_ts,
                          uint32_t* __restrict__ n_contrib, float* __restrict__ median_depth,
                          float* __restrict__ output, float* __restrict__ output_sdf,
                          uint8_t* __restrict__ inside_output) {
    __shared__ __align__(16) float4 s_rec[2 * Q_BATCH * Q_QUADS];
    __shared__ uint32_t s_mask[Q_WARPS][Q_MASK_WORDS];
    __shared__ uint32_t s_first[Q_WARPS];
    __shared__ uint32_t s_last[Q_WARPS];
    Job job;
    if (!locate_job(block_tile, tile_offsets, gaussian_ranges, query_ranges, query_list, job)) return;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float2 pxy = job.valid ? points2D[job.pid] : float2{0.f, 0.f};
    const Box box = warp_box(job.valid, pxy);

    const uint32_t n = job.grange.y - job.grange.x;
    const int rounds = (n + Q_BATCH - 1) / Q_BATCH;
    const uint32_t* const list = gaussian_list + job.grange.x;
    bool done = !job.valid;
    float T = 1.f;
    float Depth = 0.f;
    uint32_t last_contributor = 0;
    uint32_t mask_words = 0;  // words of the warp's mask the first pass produced (warp-uniform)

    // ---- pass 1: transmittance, last contributor, initial depth (sample_forward.cu:486-550) ----
    {
        uint32_t id_next = 0;
        if (tid < n) stage3q(s_rec, tid, records, __ldg(list + tid));
        cp_async_commit();
        if (Q_BATCH + tid < n) id_next = __ldg(list + Q_BATCH + tid);
        for (int i = 0; i < rounds; i++) {
            cp_async_wait_all();
            if (__syncthreads_and(done)) break;
            const float4* const buf = s_rec + (i & 1) * (Q_BATCH * Q_QUADS);
            if (i + 1 < rounds) {
                const uint32_t e = (i + 1) * Q_BATCH + tid;
                if (e < n) stage3q(s_rec + ((i + 1) & 1) * (Q_BATCH * Q_QUADS), tid, records, id_next);
                cp_async_commit();
                if (e + Q_BATCH < n) id_next = __ldg(list + e + Q_BATCH);
            }
            const int lim = min((uint32_t)Q_BATCH, n - i * Q_BATCH);
            for (int w = 0; w * 32 < lim; w++) {
                if (__all_sync(Q_FULL, done)) break;
                const int jl = w * 32 + lane;
                bool c = false;
                if (jl < lim) c = may_touch_box(buf[jl * Q_QUADS], buf[jl * Q_QUADS + 1], box);
                uint32_t cand = __ballot_sync(Q_FULL, c);
                uint32_t bits = 0;
                while (cand) {
                    const int jj = __ffs(cand) - 1;
                    cand &= cand - 1;
                    bool blended = false;
                    if (!done) {
                        const float4* rec = buf + (w * 32 + jj) * Q_QUADS;
                        const float4 q0 = rec[0];
                        const float4 q1 = rec[1];
                        const float2 d = {q0.x - pxy.x, q0.y - pxy.y};
                        const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                        if (power <= 0.0f) {
                            const float alpha = fminf(0.99f, q1.y * expf(power));
                            if (alpha >= 1.0f / 255.0f) {
                                const float test_T = T * (1.f - alpha);
                                if (test_T < 0.0001f) {
                                    done = true;
                                } else {
                                    const float t = q1.z * d.x + q1.w * d.y + rec[2].x;
                                    Depth = T > 0.5f ? t : Depth;
                                    T = test_T;
                                    last_contributor = i * Q_BATCH + w * 32 + jj + 1;  // 1-based list position
                                    blended = true;
                                }
                            }
                        }
                    }
                    if (__any_sync(Q_FULL, blended)) bits |= 1u << jj;
                }
                const uint32_t gw = i * (Q_BATCH / 32) + w;
                if (lane == 0 && gw < Q_MASK_WORDS) s_mask[warp][gw] = bits;
                mask_words = gw + 1;
            }
        }
        cp_async_wait_all();
    }

    // ---- bisection (sample_forward.cu:552-676) ----
    constexpr float RANGE = SDF ? SAMPLE_RANGE * 2.f : SAMPLE_RANGE;
    float T_p[SPLIT + 1];
    float depth_min = fmaxf(Depth - RANGE, 0.f);
    float depth_max = fmaxf(Depth + RANGE, 0.f);
    bool in_range = job.valid && T <= MIN_TRANSMITTANCE;
    float c1;  // rsqrt.approx(1.0): the shortcuts rely on it being exactly 1 (checked at run time)
    {
        const float one = 1.0f;
        asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(c1) : "f"(one));
    }
    const bool exact_one = (c1 == 1.0f);
    // every list position <= pre_pos multiplies each sample of this query by the same constants in every remaining
    // iteration; their product, formed in list order, is pre_T
    float pre_T = 1.f;
    uint32_t pre_pos = in_range ? 0u : Q_NEVER;

#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
        const bool first = (it == 0);
        constexpr float ONE_OVER_SPLIT = 1.f / static_cast<float>(SPLIT);
        if (first) {
            T_p[0] = 1.f;
            T_p[SPLIT] = 1.f;
        }
#pragma unroll
        for (int s = 1; s < SPLIT; s++) T_p[s] = pre_T;
        const float interval = (depth_max - depth_min) * ONE_OVER_SPLIT;
        const float depth_hi = depth_min + interval * SPLIT;
        const uint32_t warp_last = __reduce_max_sync(Q_FULL, in_range ? last_contributor : 0u);
        const uint32_t warp_first = __reduce_min_sync(Q_FULL, pre_pos);
        if (lane == 0) {
            s_first[warp] = warp_last ? min(warp_first, warp_last) : Q_NEVER;
            s_last[warp] = warp_last;
        }
        __syncthreads();  // also: every warp is done with the staging buffers of the previous pass
        uint32_t block_first = Q_NEVER, block_last = 0;
#pragma unroll
        for (int w = 0; w < Q_WARPS; w++) {
            block_first = min(block_first, s_first[w]);
            block_last = max(block_last, s_last[w]);
        }
        const int b0 = block_last ? block_first / Q_BATCH : 0;
        const int b1 = (block_last + Q_BATCH - 1) / Q_BATCH;
        bool pfx = exact_one;  // still extending this query's constant prefix

        uint32_t id_next = 0;
        if (b0 < b1) {
            const uint32_t e = b0 * Q_BATCH + tid;
            if (e < block_last) stage3q(s_rec, tid, records, __ldg(list + e));
            cp_async_commit();
            if (e + Q_BATCH < block_last) id_next = __ldg(list + e + Q_BATCH);
        }
        for (int i = b0; i < b1; i++) {
            cp_async_wait_all();
            __syncthreads();
            const int k = i - b0;
            const float4* const buf = s_rec + (k & 1) * (Q_BATCH * Q_QUADS);
            if (i + 1 < b1) {
                const uint32_t e = (i + 1) * Q_BATCH + tid;
                if (e < block_last) stage3q(s_rec + ((k + 1) & 1) * (Q_BATCH * Q_QUADS), tid, records, id_next);
                cp_async_commit();
                if (e + Q_BATCH < block_last) id_next = __ldg(list + e + Q_BATCH);
            }
            for (int w = 0; w < Q_BATCH / 32; w++) {
                const uint32_t base = i * Q_BATCH + w * 32;
                if (base >= warp_last) break;
                if (base + 32 <= warp_first) continue;
                const uint32_t gw = i * (Q_BATCH / 32) + w;
                uint32_t bits = gw < mask_words ? (gw < Q_MASK_WORDS ? s_mask[warp][gw] : Q_FULL) : 0u;
                uint32_t keep = 0;  // positions some query of the warp can still need in later iterations
                while (bits) {
                    const int jj = __ffs(bits) - 1;
                    const uint32_t bit = bits & (0u - bits);
                    bits ^= bit;
                    const uint32_t pos = base + jj + 1;  // 1-based list position == `contributor`
                    if (pos > warp_last) break;
                    if (pos <= warp_first) continue;
                    const float4* rec = buf + (w * 32 + jj) * Q_QUADS;
                    const float4 q0 = rec[0];
                    const float4 q1 = rec[1];
                    const float2 d = {q0.x - pxy.x, q0.y - pxy.y};
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    const float alpha = fminf(0.99f, q1.y * expf(power));
                    const bool open = pos > pre_pos;  // not yet folded into this query's prefix (false if !in_range)
                    const bool act = open && pos <= last_contributor && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    if (!__any_sync(Q_FULL, act)) {
                        if (pfx && open) pre_pos = pos;  // contributes to none of this query's samples, ever
                        continue;
                    }
                    const float2 q2 = *reinterpret_cast<const float2*>(rec + 2);
                    const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                    const float rsigma = q2.y;
                    const bool ball = rsigma > 0;
                    // Every sample depth lies in [depth_min, depth_hi]; if that whole interval is more than 6.5 sigma
                    // in front of / behind the splat's peak depth, 1 - alpha * exp(-delta^2/2) is exactly 1 at every sample.
                    const float front = depth_min - t_peak;
                    const float back = t_peak - depth_hi;
                    const float reach = fmaxf(front, back) * rsigma;  // |delta| of every sample is at least this
                    const bool quiet = ball && reach > QUIET_SIGMAS;  // 1 - alpha * exp() rounds to 1 (ggs_common.cuh)
                    // ... and stays so: brackets only shrink, up to a few ulps of the depth (see ggs_render_forward.cu).
                    // Nothing downstream needs the 16-sigma rule here: the backward of sample_depth walks its own lists.
                    const bool far = quiet && rsigma * depth_hi < 5e5f;
                    const bool far_front = far && front > 0.f;  // stays so: depth_min never decreases
                    const bool behind_for_good = far && !far_front && back > 1e-5f * depth_hi;
                    const bool constant = !act || far_front || behind_for_good;
                    const bool fold = pfx && open && constant;
                    if (__any_sync(Q_FULL, act && !behind_for_good && !fold)) keep |= bit;
                    if (__any_sync(Q_FULL, act && !quiet)) {
                        if (act) {
                            if (first) {
#pragma unroll
                                for (int s = 0; s < SPLIT + 1; s++)
                                    T_p[s] *= depth_sample_factor(depth_min + interval * s, t_peak, rsigma, ball, alpha);
                            } else {
#pragma unroll
                                for (int s = 1; s < SPLIT; s++)
                                    T_p[s] *= depth_sample_factor(depth_min + interval * s, t_peak, rsigma, ball, alpha);
                            }
                        }
                    } else if (act) {
                        // 1 - alpha * exp() is exactly 1 at every sample: factor = ((1 - alpha) | 1) * rsqrt(1)
                        const bool in_front = front > 0.f;  // every active lane is quiet here: in front or behind
                        const float f = (in_front ? (1.f - alpha) : 1.0f) * c1;
                        if (in_front || !exact_one) {
                            if (first) {
#pragma unroll
                                for (int s = 0; s < SPLIT + 1; s++) T_p[s] *= f;
                            } else {
#pragma unroll
                                for (int s = 1; s < SPLIT; s++) T_p[s] *= f;
                            }
                        }
                    }
                    if (fold) {
                        pre_pos = pos;
                        pre_T = T_p[1];  // all samples of this iteration are still equal here
                    } else if (open) {
                        pfx = false;
                    }
                }
                if (exact_one && lane == 0 && gw < Q_MASK_WORDS && gw < mask_words) s_mask[warp][gw] = keep;
            }
        }
        if (first) {
            in_range = (T_p[0] >= 0.5f) && (T_p[SPLIT] <= 0.5f) && in_range;
            if (!in_range) pre_pos = Q_NEVER;
        }
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
    cp_async_wait_all();
    if (!job.valid) return;
    const float w_max = __saturatef((T_p[0] - 0.5f) / (T_p[0] - T_p[SPLIT]));
    const float w_min = 1.f - w_max;
    const float mDepth = in_range ? w_max * depth_max + w_min * depth_min : 0.f;
    if constexpr (SDF) {
        output_sdf[job.pid] = mDepth - point_ts[job.pid];
        output[job.pid] = mDepth;
    } else {
        const float2 pixnf = {(pxy.x - static_cast<float>(W - 1) / 2.f) / focal_x,
                              (pxy.y - static_cast<float>(H - 1) / 2.f) / focal_y};
        const float rln = rnorm3df(pixnf.x, pixnf.y, 1.f);
        const float depth = mDepth * rln;
        output[3 * (size_t)job.pid] = pixnf.x * depth;
        output[3 * (size_t)job.pid + 1] = pixnf.y * depth;
        output[3 * (size_t)job.pid + 2] = depth;
        median_depth[job.pid] = mDepth;
        n_contrib[job.pid] = last_contributor;
    }
    inside_output[job.pid] = in_range ? 1 : 0;
}

// ---- sample_depth backward (sample_backward.cu:77-358) ------------------------------------------------------------
// accumulator slots (ggs_common.cuh) of the ten per-entry sums: mean2D x, y; conic x, y, z; opacity; ray plane x, y, t, rsigma
__device__ __forceinline__ int backward_slot(int c) {
    return c < 2 ? c : (c < 6 ? c + 1 : c + 7);
}

__global__ void __launch_bounds__(Q_THREADS, 6)
    sample_backward_kernel(int W, int H, float focal_x, float focal_y, const uint32_t* __restrict__ block_tile,
                           const uint32_t* __restrict__ tile_offsets, const uint2* __restrict__ gaussian_ranges,
                           const uint2* __restrict__ query_ranges, const uint32_t* __restrict__ gaussian_list,
                           const uint32_t* __restrict__ query_list, const SplatRecord* __restrict__ records,
                           const float2* __restrict__ points2D, const uint32_t* __restrict__ n_contrib,
                           const float* __restrict__ median_depth, const uint8_t* __restrict__ inside,
                           const float* __restrict__ dL_doutputs, float2* __restrict__ dL_dpoints2D,
                           double* __restrict__ grad_acc) {
    __shared__ __align__(16) float4 s_rec[2 * Q_BATCH * Q_QUADS];
    __shared__ uint32_t s_id[2 * Q_BATCH];
    __shared__ uint32_t s_last[Q_WARPS];
    Job job;
    if (!locate_job(block_tile, tile_offsets, gaussian_ranges, query_ranges, query_list, job)) return;
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const int lane = tid & 31;
    const float2 pxy = job.valid ? points2D[job.pid] : float2{0.f, 0.f};
    const Box box = warp_box(job.valid, pxy);
    const uint32_t* const list = gaussian_list + job.grange.x;
    const float ddelx_dx = 0.5f * W;
    const float ddely_dy = 0.5f * H;

    uint32_t last_contributor = 0;
    float mDepth = 0.f, dL_dDepth = 0.f;
    float2 dL_dpoint_xy = {0.f, 0.f};
    bool live = false;
    if (job.valid) {
        last_contributor = n_contrib[job.pid];
        const float3 dL_doutput = {dL_doutputs[3 * (size_t)job.pid], dL_doutputs[3 * (size_t)job.pid + 1],
                                   dL_doutputs[3 * (size_t)job.pid + 2]};
        mDepth = median_depth[job.pid];
        live = !((last_contributor == 0) || inside[job.pid] == 0);
        const float2 pixnf = {(pxy.x - static_cast<float>(W - 1) / 2.f) / focal_x,
                              (pxy.y - static_cast<float>(H - 1) / 2.f) / focal_y};
        const float rln = rnorm3df(pixnf.x, pixnf.y, 1.f);
        const float rln2 = 1.f / (pixnf.x * pixnf.x + pixnf.y * pixnf.y + 1.f);
        const float depth = mDepth * rln;
        const float dL_ddepth = dL_doutput.x * pixnf.x + dL_doutput.y * pixnf.y + dL_doutput.z;
        dL_dDepth = rln * dL_ddepth;
        const float aux = (dL_doutput.x * pixnf.x + dL_doutput.y * pixnf.y + dL_doutput.z) * rln2;
        const float2 dL_dpixnf = {(dL_doutput.x - aux * pixnf.x) * depth, (dL_doutput.y - aux * pixnf.y) * depth};
        dL_dpoint_xy = {dL_dpixnf.x / focal_x, dL_dpixnf.y / focal_y};
    }
    const uint32_t warp_last = __reduce_max_sync(Q_FULL, live ? last_contributor : 0u);
    if (lane == 0) s_last[warp] = warp_last;
    __syncthreads();
    uint32_t block_last = 0;
#pragma unroll
    for (int w = 0; w < Q_WARPS; w++) block_last = max(block_last, s_last[w]);
    const int rounds = (block_last + Q_BATCH - 1) / Q_BATCH;

    // ---- pass A: dT/dt_m at the stored median depth (sample_backward.cu:172-215) ----
    float dT_dtm = 0.f;
    {
        uint32_t id_next = 0;
        if (tid < block_last) stage3q(s_rec, tid, records, __ldg(list + tid));
        cp_async_commit();
        if (Q_BATCH + tid < block_last) id_next = __ldg(list + Q_BATCH + tid);
        for (int i = 0; i < rounds; i++) {
            cp_async_wait_all();
            __syncthreads();
            const float4* const buf = s_rec + (i & 1) * (Q_BATCH * Q_QUADS);
            if (i + 1 < rounds) {
                const uint32_t e = (i + 1) * Q_BATCH + tid;
                if (e < block_last) stage3q(s_rec + ((i + 1) & 1) * (Q_BATCH * Q_QUADS), tid, records, id_next);
                cp_async_commit();
                if (e + Q_BATCH < block_last) id_next = __ldg(list + e + Q_BATCH);
            }
            for (int w = 0; w < Q_BATCH / 32; w++) {
                const uint32_t base = i * Q_BATCH + w * 32;
                if (base >= warp_last) break;
                const uint32_t jl = w * 32 + lane;
                bool c = false;
                if (base + lane < warp_last) c = may_touch_box(buf[jl * Q_QUADS], buf[jl * Q_QUADS + 1], box);
                uint32_t cand = __ballot_sync(Q_FULL, c);
                while (cand) {
                    const int jj = __ffs(cand) - 1;
                    cand &= cand - 1;
                    const uint32_t pos = base + jj + 1;
                    if (!(live && pos <= last_contributor)) continue;
                    const float4* rec = buf + (w * 32 + jj) * Q_QUADS;
                    const float4 q0 = rec[0];
                    const float4 q1 = rec[1];
                    const float2 d = {q0.x - pxy.x, q0.y - pxy.y};
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, q1.y * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float2 q2 = *reinterpret_cast<const float2*>(rec + 2);
                    const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                    const float rsigma = q2.y;
                    const float t_delta = (mDepth - t_peak) * rsigma;
                    const float G_exp = expf(-0.5f * t_delta * t_delta);
                    const float Gt = alpha * G_exp;
                    dT_dtm += -0.25f * Gt / (1.f - Gt) * fabsf(t_delta) * rsigma;
                }
            }
        }
    }
    const float dL_dmt_dT_dtm = dL_dDepth / fmaxf(-dT_dtm, 1e-7f);
    __syncthreads();  // pass B reuses the staging buffers

    // ---- pass B: gradients (sample_backward.cu:229-352) ----
    const int my_comp = scatter_component<10>(lane);
    const int my_slot = my_comp >= 0 ? backward_slot(my_comp) : 0;
    {
        uint32_t id_next = 0;
        if (tid < block_last) {
            const uint32_t id = __ldg(list + tid);
            s_id[tid] = id;
            stage3q(s_rec, tid, records, id);
        }
        cp_async_commit();
        if (Q_BATCH + tid < block_last) id_next = __ldg(list + Q_BATCH + tid);
        for (int i = 0; i < rounds; i++) {
            cp_async_wait_all();
            __syncthreads();
            const float4* const buf = s_rec + (i & 1) * (Q_BATCH * Q_QUADS);
            const uint32_t* const ids = s_id + (i & 1) * Q_BATCH;
            if (i + 1 < rounds) {
                const uint32_t e = (i + 1) * Q_BATCH + tid;
                if (e < block_last) {
                    s_id[((i + 1) & 1) * Q_BATCH + tid] = id_next;
                    stage3q(s_rec + ((i + 1) & 1) * (Q_BATCH * Q_QUADS), tid, records, id_next);
                }
                cp_async_commit();
                if (e + Q_BATCH < block_last) id_next = __ldg(list + e + Q_BATCH);
            }
            for (int w = 0; w < Q_BATCH / 32; w++) {
                const uint32_t base = i * Q_BATCH + w * 32;
                if (base >= warp_last) break;
                const uint32_t jl = w * 32 + lane;
                bool c = false;
                if (base + lane < warp_last) c = may_touch_box(buf[jl * Q_QUADS], buf[jl * Q_QUADS + 1], box);
                uint32_t cand = __ballot_sync(Q_FULL, c);
                while (cand) {
                    const int jj = __ffs(cand) - 1;
                    cand &= cand - 1;
                    const uint32_t pos = base + jj + 1;
                    const int j = w * 32 + jj;
                    const float4* rec = buf + j * Q_QUADS;
                    const float4 q0 = rec[0];
                    const float4 q1 = rec[1];
                    const float4 con_o = {q0.z, q0.w, q1.x, q1.y};
                    const float2 d = {q0.x - pxy.x, q0.y - pxy.y};
                    const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                    const float G = expf(power);
                    const bool valid = live && pos <= last_contributor && !(power > 0.0f) &&
                                       !(con_o.w * G < 1.0f / 255.0f);
                    if (!__any_sync(Q_FULL, valid)) continue;
                    float v[10];
#pragma unroll
                    for (int k = 0; k < 10; k++) v[k] = 0.f;
                    if (valid) {
                        const float2 q2 = *reinterpret_cast<const float2*>(rec + 2);
                        const float4 ray_plane = {q1.z, q1.w, q2.x, q2.y};
                        const float alpha = fminf(0.99f, con_o.w * G);
                        const float t_peak = ray_plane.x * d.x + ray_plane.y * d.y + ray_plane.z;
                        const float rsigma = ray_plane.w;
                        const float t_delta = (mDepth - t_peak) * rsigma;
                        const float G_exp = expf(-0.5f * t_delta * t_delta);
                        const float Gt = alpha * G_exp;
                        float dL_dGt = dL_dmt_dT_dtm * 0.25f / (1.f - Gt);
                        dL_dGt = mDepth > t_peak ? dL_dGt : -dL_dGt;
                        dL_dGt = rsigma > 0 ? dL_dGt : 0.f;
                        const float dL_dopa = dL_dGt * G_exp - dL_dmt_dT_dtm * (t_delta > 0 ? 0.5f / (1.f - alpha) : 0.f);
                        const float dL_ddelta = -dL_dGt * Gt * t_delta;
                        v[9] = dL_ddelta * (mDepth - t_peak);
                        const float dL_dt_peak = -dL_ddelta * rsigma;
                        v[8] = dL_dt_peak;
                        v[6] = dL_dt_peak * d.x;
                        v[7] = dL_dt_peak * d.y;
                        const float dL_dG = con_o.w * dL_dopa;
                        const float gdx = G * d.x;
                        const float gdy = G * d.y;
                        const float dG_ddelx = -gdx * con_o.x - gdy * con_o.y;
                        const float dG_ddely = -gdy * con_o.z - gdx * con_o.y;
                        const float dL_ddelx = dL_dG * dG_ddelx + dL_dt_peak * ray_plane.x;
                        const float dL_ddely = dL_dG * dG_ddely + dL_dt_peak * ray_plane.y;
                        v[0] = dL_ddelx * ddelx_dx;
                        v[1] = dL_ddely * ddely_dy;
                        dL_dpoint_xy.x -= dL_ddelx;
                        dL_dpoint_xy.y -= dL_ddely;
                        v[2] = -0.5f * gdx * d.x * dL_dG;
                        v[3] = -0.5f * gdx * d.y * dL_dG;
                        v[4] = -0.5f * gdy * d.y * dL_dG;
                        v[5] = G * dL_dopa;
                    }
                    const float sum = reduce_scatter<10, 16>(v, lane);
                    if (my_comp >= 0) atomicAdd(grad_acc + (size_t)ids[j] * GRAD_STRIDE + my_slot, static_cast<double>(sum));
                }
            }
        }
    }
    cp_async_wait_all();
    if (job.valid) dL_dpoints2D[job.pid] = {dL_dpoint_xy.x * ddelx_dx, dL_dpoint_xy.y * ddely_dy};
}

// Gradient of the projected query position back to the 3D query point (sample_backward.cu:42-75); rows of queries
// that were not evaluated get zeros.
__global__ void points_backward_kernel(int PN, const float* __restrict__ points3D, const float* __restrict__ proj,
                                       const float2* __restrict__ points2D, const float2* __restrict__ dL_dpoints2D,
                                       float* __restrict__ dL_dpoints3D) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)PN) return;
    float3 r = {0.f, 0.f, 0.f};
    if (points2D[idx].x >= 0.f) {
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

}  // namespace

// ---- host side --------------------------------------------------------------------------------------------------
QueryLayout query_layout(size_t PN, int tiles, bool sample) {
    QueryLayout L;
    size_t o = 0;
    auto take = [&](size_t bytes) {
        const size_t at = o;
        o = align_up(o + bytes, 256);
        return at;
    };
    L.xy = take(PN * 8);
    L.ts = take(PN * 4);
    L.keys_a = take(PN * 4);
    L.keys_b = take(PN * 4);
    L.vals_a = take(PN * 4);
    L.vals_b = take(PN * 4);
    L.hist = take(MAX_PASSES * RADIX * 4);
    int tile_bits = 1;
    while ((1u << tile_bits) <= (uint32_t)tiles) tile_bits++;  // tiles <= 2^tile_bits - 1: the all-ones key stays unique
    L.sort_bits = QUERY_CELL_BITS + tile_bits;
    const int passes = (L.sort_bits + RADIX_BITS - 1) / RADIX_BITS;
    L.status_bytes = (size_t)passes * ((PN + SORT_TILE - 1) / SORT_TILE) * RADIX * 4;
    L.status = take(L.status_bytes);
    L.counters = take(MAX_PASSES * 4);
    L.ranges = take((size_t)tiles * 8);
    L.rounds = take((size_t)tiles * 4);
    L.offsets = take((size_t)tiles * 4);
    L.scan_status = take(((size_t)tiles / SCAN_TILE + 2) * 8);
    L.scan_counter = take(256);
    L.scan_total = take(256);
    // sum over tiles of ceil(n_t / QUERY_CTA) <= (tiles with queries) + PN / QUERY_CTA
    L.max_blocks = (uint32_t)(std::min<size_t>((size_t)tiles, PN) + PN / QUERY_CTA);
    L.block_tile = take((size_t)L.max_blocks * 4);
    L.n_contrib = take(sample ? PN * 4 : 0);
    L.median_depth = take(sample ? PN * 4 : 0);
    L.sorted_in_b = (passes & 1) != 0;
    L.bytes = o;
    return L;
}

void launch_query_binning(const FrameParams& fp, int PN, const float* points3D, void* ws, const QueryLayout& L,
                          cudaStream_t stream) {
    auto at = [&](size_t off) { return static_cast<char*>(ws) + off; };
    const int tiles = fp.gx * fp.gy;
    uint32_t* keys_a = reinterpret_cast<uint32_t*>(at(L.keys_a));
    uint32_t* keys_b = reinterpret_cast<uint32_t*>(at(L.keys_b));
    uint32_t* vals_a = reinterpret_cast<uint32_t*>(at(L.vals_a));
    uint32_t* vals_b = reinterpret_cast<uint32_t*>(at(L.vals_b));
    query_project_kernel<<<(PN + 255) / 256, 256, 0, stream>>>(fp, PN, points3D, reinterpret_cast<float2*>(at(L.xy)),
                                                              reinterpret_cast<float*>(at(L.ts)), keys_a, vals_a);
    launch_radix_sort32(keys_a, keys_b, vals_a, vals_b, (uint32_t)PN, L.sort_bits,
                        reinterpret_cast<uint32_t*>(at(L.hist)), reinterpret_cast<uint32_t*>(at(L.status)),
                        L.status_bytes, reinterpret_cast<uint32_t*>(at(L.counters)), stream);
    const uint32_t* keys = L.sorted_in_b ? keys_b : keys_a;
    uint2* ranges = reinterpret_cast<uint2*>(at(L.ranges));
    cudaMemsetAsync(ranges, 0, (size_t)tiles * 8, stream);
    query_ranges_kernel<<<(PN + 255) / 256, 256, 0, stream>>>(keys, (uint32_t)PN, ranges);
    uint32_t* rounds = reinterpret_cast<uint32_t*>(at(L.rounds));
    uint32_t* offsets = reinterpret_cast<uint32_t*>(at(L.offsets));
    query_rounds_kernel<<<(tiles + 255) / 256, 256, 0, stream>>>(tiles, ranges, rounds);
    launch_scan(rounds, offsets, tiles, reinterpret_cast<uint64_t*>(at(L.scan_status)),
                reinterpret_cast<uint32_t*>(at(L.scan_counter)), reinterpret_cast<uint32_t*>(at(L.scan_total)), stream);
    if (L.max_blocks)
        query_block_table_kernel<<<(L.max_blocks + 255) / 256, 256, 0, stream>>>(
            L.max_blocks, tiles, offsets, reinterpret_cast<uint32_t*>(at(L.block_tile)));
}

namespace {
struct QueryPtrs {
    const uint32_t *block_tile, *offsets, *list;
    const uint2* ranges;
    const float2* xy;
    const float* ts;
    uint32_t* n_contrib;
    float* median_depth;
};
QueryPtrs query_ptrs(void* ws, const QueryLayout& L) {
    char* b = static_cast<char*>(ws);
    QueryPtrs p;
    p.block_tile = reinterpret_cast<const uint32_t*>(b + L.block_tile);
    p.offsets = reinterpret_cast<const uint32_t*>(b + L.offsets);
    p.list = reinterpret_cast<const uint32_t*>(b + (L.sorted_in_b ? L.vals_b : L.vals_a));
    p.ranges = reinterpret_cast<const uint2*>(b + L.ranges);
    p.xy = reinterpret_cast<const float2*>(b + L.xy);
    p.ts = reinterpret_cast<const float*>(b + L.ts);
    p.n_contrib = reinterpret_cast<uint32_t*>(b + L.n_contrib);
    p.median_depth = reinterpret_cast<float*>(b + L.median_depth);
    return p;
}
}  // namespace

void launch_integrate(const QueryLayout& L, void* ws, const uint2* gaussian_ranges, const uint32_t* gaussian_list,
                      const SplatRecord* records, float* out_transmittance, uint8_t* inside,
                      unsigned long long* op_counters, cudaStream_t stream) {
    if (L.max_blocks == 0) return;
    const QueryPtrs q = query_ptrs(ws, L);
    if (op_counters)
        integrate_kernel<true><<<L.max_blocks, Q_THREADS, 0, stream>>>(q.block_tile, q.offsets, gaussian_ranges, q.ranges,
                                                                      gaussian_list, q.list, records, q.xy, q.ts,
                                                                      out_transmittance, inside, op_counters);
    else
        integrate_kernel<false><<<L.max_blocks, Q_THREADS, 0, stream>>>(q.block_tile, q.offsets, gaussian_ranges,
                                                                       q.ranges, gaussian_list, q.list, records, q.xy,
                                                                       q.ts, out_transmittance, inside, nullptr);
}

void launch_sample_depth_forward(const FrameParams& fp, const QueryLayout& L, void* ws, const uint2* gaussian_ranges,
                                 const uint32_t* gaussian_list, const SplatRecord* records, float* output,
                                 uint8_t* inside, cudaStream_t stream) {
    if (L.max_blocks == 0) return;
    const QueryPtrs q = query_ptrs(ws, L);
    sample_forward_kernel<SPLIT_ITERATIONS, false><<<L.max_blocks, Q_THREADS, 0, stream>>>(
        fp.W, fp.H, fp.focal_x, fp.focal_y, q.block_tile, q.offsets, gaussian_ranges, q.ranges, gaussian_list, q.list,
        records, q.xy, q.ts, q.n_contrib, q.median_depth, output, nullptr, inside);
}

void launch_evaluate_sdf(const FrameParams& fp, const QueryLayout& L, void* ws, const uint2* gaussian_ranges,
                         const uint32_t* gaussian_list, const SplatRecord* records, float* out_depth, float* out_sdf,
                         uint8_t* inside, cudaStream_t stream) {
    if (L.max_blocks == 0) return;
    const QueryPtrs q = query_ptrs(ws, L);
    // the reference instantiates SPLIT_ITERATIONS + 1 iterations here (sample_forward.cu:792)
    sample_forward_kernel<SPLIT_ITERATIONS + 1, true><<<L.max_blocks, Q_THREADS, 0, stream>>>(
        fp.W, fp.H, fp.focal_x, fp.focal_y, q.block_tile, q.offsets, gaussian_ranges, q.ranges, gaussian_list, q.list,
        records, q.xy, q.ts, nullptr, nullptr, out_depth, out_sdf, inside);
}

void launch_sample_depth_backward(const FrameParams& fp, const QueryLayout& L, void* ws, const uint2* gaussian_ranges,
                                  const uint32_t* gaussian_list, const SplatRecord* records, const uint8_t* inside,
                                  const float* dL_doutputs, float2* dL_dpoints2D, double* grad_acc,
                                  cudaStream_t stream) {
    if (L.max_blocks == 0) return;
    const QueryPtrs q = query_ptrs(ws, L);
    sample_backward_kernel<<<L.max_blocks, Q_THREADS, 0, stream>>>(
        fp.W, fp.H, fp.focal_x, fp.focal_y, q.block_tile, q.offsets, gaussian_ranges, q.ranges, gaussian_list, q.list,
        records, q.xy, q.n_contrib, q.median_depth, inside, dL_doutputs, dL_dpoints2D, grad_acc);
}

void launch_points_backward(const FrameParams& fp, int PN, const float* points3D, const float2* points2D,
                            const float2* dL_dpoints2D, float* dL_dpoints3D, cudaStream_t stream) {
    if (PN == 0) return;
    points_backward_kernel<<<(PN + 255) / 256, 256, 0, stream>>>(PN, points3D, fp.view.proj, points2D, dL_dpoints2D,
                                                                dL_dpoints3D);
}

}  // namespace ggs
