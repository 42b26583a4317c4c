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
//   * list entries are gathered as 64-byte SplatRecords into shared memory, 256 per batch, with cp.async
//     into a double buffer: the gather of batch i+1 overlaps the arithmetic on batch i, one barrier per batch;
//   * pass 1: each lane first tests ONE entry of a 32-entry chunk against the warp's 8x4 pixel rectangle
//     (exact minimum of the conic's quadratic form over the rectangle against the alpha >= 1/255 threshold,
//     with slack); only entries that can touch the rectangle are evaluated per pixel.  The pass leaves, per
//     warp, a bitmask of the list positions that at least one of its pixels blended;
//   * the five bisection passes walk only those positions.  A splat whose depth profile is more than 6.5 sigma
//     away from the whole current bracket has 1 - alpha * exp() == 1 exactly (ggs_common.cuh), so its factor is
//     exactly (1 - alpha) in front of the bracket and exactly 1 behind it: no exp / rsqrt is issued for it unless some
//     pixel of the warp is close to it.  Brackets only shrink, so such a splat stays quiet:
//       - the leading run of quiet / inactive entries of a pixel multiplies every sample by the same floats in
//         the same order in every iteration: its product is carried over (pre_T, pre_pos) instead of being
//         recomputed, and the warp starts its walk after the shortest such prefix of its pixels;
//       - entries behind every pixel's bracket are dropped from the warp's list for good.
//     All shortcuts keep the order of the per-pixel floating-point products: outputs are bit-identical to
//     the plain evaluation.  A second mask is thinned with the 16-sigma rule only (exp() == 0 exactly) and exported
//     for the backward's dT/dt pass, whose terms still see a quiet splat.
#include "ggs_kernels.cuh"
#include "ggs_tile.cuh"

namespace ggs {

struct RenderParams {
    int W, H, gx;
    float focal_x, focal_y;
    const float* bg;  // device [3]
    const uint32_t* tile_order;  // TileLayout::order or nullptr
};

// A CTA renders HALF a tile (16x8 pixels, 4 warps; blockIdx.z = half): the barrier that hands over a staged batch
// then couples 4 warps instead of 8 (warps of a tile have unequal work per batch: 23 % of the stall samples of the
// full-tile kernel were block barriers) and twice as many CTAs per SM progress independently.  Warp ids stay those of
// the full tile (half * 4 + warp), so the masks handed to the backward keep their layout.
constexpr int WARPS = TILE_WARPS / 2;
constexpr int THREADS = WARPS * 32;
constexpr int BATCH = THREADS;  // list entries per staged batch (one per thread)
constexpr uint32_t FULL = 0xffffffffu;
constexpr uint32_t NEVER = 0xffffffffu;
constexpr size_t REC_BYTES = (size_t)2 * BATCH * REC_QUADS * sizeof(float4);
constexpr size_t MASK_BYTES = (size_t)2 * WARPS * MASK_WORDS * sizeof(uint32_t);  // s_mask + s_near

// Can the splat (q0 = mean.xy, conic.xy; q1 = conic.z, opacity, ..) reach alpha >= 1/255 on any pixel centre of
// the rectangle [x0, x0+7] x [y0, y0+3]?  Conservative: `false` only when every pixel is provably rejected.
//   alpha = min(0.99, o * exp(power)) >= 1/255  needs  -power = q(d)/2 <= log(255 o),  q(d) = A dx^2 + 2 B dx dy + C dy^2.
// The minimum of the convex q over the rectangle of offsets d is 0 when the rectangle contains the centre and
// otherwise lies on one of the four edges, where it is a clamped 1-D parabola minimum.
__device__ __forceinline__ bool may_touch(const float4 q0, const float4 q1, float x0, float y0) {
    const float A = q0.z, B = q0.w, C = q1.x, o = q1.y;
    const float dx1 = q0.x - x0, dx0 = dx1 - 7.f;  // d = mean - pixel ranges over [dx0, dx1] x [dy0, dy1]
    const float dy1 = q0.y - y0, dy0 = dy1 - 3.f;
    const float det = A * C - B * B;
    if (!(A > 0.f && C > 0.f && det > 0.f)) return true;  // not positive definite (or NaN): no claim
    const float L = __logf(255.f * o);
    if (L < -0.01f) return false;  // o < 1/255: power <= 0 can never lift alpha over the threshold
    if (dx0 <= 0.f && dx1 >= 0.f && dy0 <= 0.f && dy1 >= 0.f) return true;
    const float rA = __frcp_rn(A), rC = __frcp_rn(C);
    float qmin;
    {
        const float ya = fminf(fmaxf(-B * dx0 * rC, dy0), dy1);
        const float yb = fminf(fmaxf(-B * dx1 * rC, dy0), dy1);
        const float xa = fminf(fmaxf(-B * dy0 * rA, dx0), dx1);
        const float xb = fminf(fmaxf(-B * dy1 * rA, dx0), dx1);
        const float e0 = A * dx0 * dx0 + 2.f * B * dx0 * ya + C * ya * ya;
        const float e1 = A * dx1 * dx1 + 2.f * B * dx1 * yb + C * yb * yb;
        const float e2 = A * xa * xa + 2.f * B * xa * dy0 + C * dy0 * dy0;
        const float e3 = A * xb * xb + 2.f * B * xb * dy1 + C * dy1 * dy1;
        qmin = fminf(fminf(e0, e1), fminf(e2, e3));
    }
    // rounding slack: both this evaluation and the per-pixel `power` are sums of terms bounded by M
    const float Dx = fmaxf(fabsf(dx0), fabsf(dx1)), Dy = fmaxf(fabsf(dy0), fabsf(dy1));
    const float M = A * Dx * Dx + 2.f * fabsf(B) * Dx * Dy + C * Dy * Dy;
    return !(0.5f * qmin - 1e-5f * M > L + 0.01f);
}

#ifdef GGS_WALK_STATS
// Development build only (make variant DEFS=-DGGS_WALK_STATS): what the bisection walk spends its list positions on.
__device__ unsigned long long g_walk_stats[SPLIT_ITERATIONS][8];
#define WALK_STAT(k, v) ws[k] += (v)
#else
#define WALK_STAT(k, v)
#endif

constexpr uint32_t REC_STRIDE = REC_QUADS * sizeof(float4);
#ifndef GGS_CLASSIFY_ITERATIONS
#define GGS_CLASSIFY_ITERATIONS 2
#endif
constexpr int CLASSIFY_ITERATIONS = GGS_CLASSIFY_ITERATIONS;  // bisection iterations that pre-classify a word's entries

// One depth sample of the vacancy-corrected transmittance product (render_forward.cu:603-621).
__device__ __forceinline__ float sample_factor(float ts, float t_peak, float rsigma, bool ball, float alpha) {
    const float delta = (ts - t_peak) * rsigma;
    const float g = ball ? expf(-0.5f * delta * delta) : 0.f;
    const float one_minus_gaussian = 1.f - alpha * g;
    const float rvacancy = rsqrtf(one_minus_gaussian);
    return (ts > t_peak ? (1.f - alpha) : one_minus_gaussian) * rvacancy;
}

template <bool GEOMETRY>
__global__ void __launch_bounds__(THREADS, 8)
    render_forward_kernel(const RenderParams rp, const uint2* __restrict__ ranges,
                          const uint32_t* __restrict__ point_list, const SplatRecord* __restrict__ records,
                          float* __restrict__ out_color, float* __restrict__ out_alpha,
                          float* __restrict__ out_normal, float* __restrict__ out_mdepth,
                          uint32_t* __restrict__ n_contrib, uint32_t* __restrict__ max_contributors,
                          uint32_t* __restrict__ accept_mask, uint32_t* __restrict__ warp_words,
                          uint32_t* __restrict__ near_mask, uint32_t* __restrict__ near_first) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* const s_rec = reinterpret_cast<float4*>(smem_raw);  // [2][BATCH * REC_QUADS]
    const uint32_t s_rec_addr = static_cast<uint32_t>(__cvta_generic_to_shared(smem_raw));
    // s_mask: list positions this warp still has to visit (thinned with the QUIET rule); s_near: what the backward's
    // dT/dt pass has to visit (thinned with the FAR rule only: a quiet splat still has a non-zero exp() there)
    uint32_t(*const s_mask)[MASK_WORDS] = reinterpret_cast<uint32_t(*)[MASK_WORDS]>(smem_raw + REC_BYTES);
    // (s_near[w] is s_mask[WARPS + w]: one base pointer)
    __shared__ uint32_t s_red[WARPS];

    const int W = rp.W, H = rp.H;
    // which half of which tile: in grid order the CTAs walk TileLayout::order (longest lists first), the halves of a
    // tile next to each other
    uint32_t tile_x = blockIdx.x, tile_y = blockIdx.y, half = blockIdx.z;
    if (rp.tile_order) {
        const uint32_t lin = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        const uint32_t t = __ldg(rp.tile_order + lin / (TILE_WARPS / WARPS));
        half = lin % (TILE_WARPS / WARPS);
        tile_y = t / (uint32_t)rp.gx;
        tile_x = t - tile_y * (uint32_t)rp.gx;
    }
    const uint32_t tile_id = tile_y * rp.gx + tile_x;
    const uint32_t tid = threadIdx.x;
    const uint32_t warp = tid >> 5, lane = tid & 31;  // warp: within this CTA (staging, shared arrays)
    const uint32_t twarp = half * WARPS + warp;  // within the tile (pixels, masks)
    // warp w of the tile owns its 8x4 pixel block (w & 1, w >> 1)
    const uint2 pix0 = {tile_x * TILE_X + (twarp & 1) * 8, tile_y * TILE_Y + (twarp >> 1) * 4};
    const uint2 pix = {pix0.x + (lane & 7), pix0.y + (lane >> 3)};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {static_cast<float>(pix.x), static_cast<float>(pix.y)};

    const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
    bool done = !inside;

    const uint2 range = ranges[tile_id];
    const uint32_t n = range.y - range.x;
    const int rounds = (n + BATCH - 1) / BATCH;
    const uint32_t* const list = point_list + range.x;
    // first word of this tile in the global acceptance masks (tiles are laid out in list order, one spare word each)
    const uint32_t mask_base = range.x / 32 + tile_id;

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float mDepthinit = 0.f;
    uint32_t mask_words = 0;  // words of the warp's mask that pass 1 produced (warp-uniform)

    // ---- pass 1: colour / normal / alpha ---------------------------------------------------------------
    {
        float C[3] = {0.f, 0.f, 0.f};
        float Normal[3] = {0.f, 0.f, 0.f};
        const float rx0 = static_cast<float>(pix0.x), ry0 = static_cast<float>(pix0.y);
        uint32_t id_next = 0;
        if (tid < n) stage_record(s_rec, tid, records, __ldg(list + tid), 4);
        cp_async_commit();
        if (BATCH + tid < n) id_next = __ldg(list + BATCH + tid);

        for (int i = 0; i < rounds; i++) {
            cp_async_wait_all();
            if (__syncthreads_and(done)) break;
            const uint32_t buf_addr = s_rec_addr + (i & 1) * (BATCH * REC_STRIDE);
            if (i + 1 < rounds) {
                const uint32_t e = (i + 1) * BATCH + tid;
                if (e < n) stage_record(s_rec + ((i + 1) & 1) * (BATCH * REC_QUADS), tid, records, id_next, 4);
                cp_async_commit();
                if (e + BATCH < n) id_next = __ldg(list + e + BATCH);
            }
            const int lim = min((uint32_t)BATCH, n - i * BATCH);
            for (int w = 0; w * 32 < lim; w++) {
                if (__all_sync(FULL, done)) break;
                // lane-parallel candidate test: one entry per lane against the warp's pixel rectangle
                const int jl = w * 32 + lane;
                bool c = false;
                const uint32_t word_addr = buf_addr + w * (32 * REC_STRIDE);
                if (jl < lim) c = may_touch(lds128(word_addr + lane * REC_STRIDE), lds128(word_addr + lane * REC_STRIDE + 16), rx0, ry0);
                uint32_t cand = __ballot_sync(FULL, c);
                uint32_t bits = 0;
                while (cand) {
                    const int jj = __ffs(cand) - 1;
                    cand &= cand - 1;
                    bool blended = false;
                    if (!done) {
                        const uint32_t rec = word_addr + jj * REC_STRIDE;
                        const float4 q0 = lds128(rec);
                        const float4 q1 = lds128(rec + 16);
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
                                    const float4 q2 = lds128(rec + 32);
                                    const float4 q3 = lds128(rec + 48);
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
                                    last_contributor = i * BATCH + w * 32 + jj + 1;  // 1-based list position
                                    blended = true;
                                }
                            }
                        }
                    }
                    if (__any_sync(FULL, blended)) bits |= 1u << jj;
                }
                const uint32_t gw = i * (BATCH / 32) + w;
                if (lane == 0) {
                    if (GEOMETRY && gw < MASK_WORDS) {
                        s_mask[warp][gw] = bits;
                        s_mask[WARPS + warp][gw] = bits;
                    }
                    accept_mask[(size_t)(mask_base + gw) * 8 + twarp] = bits;  // kept for the backward
                }
                mask_words = gw + 1;
            }
        }
        cp_async_wait_all();

        // colour, alpha and normal are final here (render_forward.cu:648-670)
        if (inside) {
            n_contrib[pix_id] = last_contributor;
            const size_t HW = (size_t)H * W;
#pragma unroll
            for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix_id] = C[ch] + T * __ldg(rp.bg + ch);
            out_alpha[pix_id] = 1.f - T;
            if constexpr (GEOMETRY) {
                const float len_normal = 1.f - T;
#pragma unroll
                for (int ch = 0; ch < 3; ch++)
                    out_normal[ch * HW + pix_id] = last_contributor ? Normal[ch] / len_normal : 0.f;
            }
        }
    }

    // block-wide max of last_contributor
    uint32_t max_contributor;
    {
        const uint32_t wmax = __reduce_max_sync(FULL, last_contributor);
        if (lane == 0) s_red[warp] = wmax;
        __syncthreads();
        uint32_t m = 0;
#pragma unroll
        for (int w = 0; w < WARPS; w++) m = max(m, s_red[w]);
        max_contributor = m;
    }
    if (tid == 0) atomicMax(max_contributors + tile_id, max_contributor);  // over the two halves (zeroed at launch)
    if (lane == 0) warp_words[tile_id * TILE_WARPS + twarp] = mask_words;
    if constexpr (!GEOMETRY) return;

    // ---- passes 2-6: median depth by 8-way bisection ---------------------------------------------------
    const float rx0 = static_cast<float>(pix0.x), ry0 = static_cast<float>(pix0.y);
    const float2 pixnf = {(pixf.x - static_cast<float>(W - 1) / 2.f) / rp.focal_x,
                          (pixf.y - static_cast<float>(H - 1) / 2.f) / rp.focal_y};
    const float rln = rnorm3df(pixnf.x, pixnf.y, 1.f);
    float T_p[SPLIT + 1];
    float depth_min = fmaxf(mDepthinit - SAMPLE_RANGE, 0.f);
    float depth_max = fmaxf(mDepthinit + SAMPLE_RANGE, 0.f);
    bool in_range = T <= MIN_TRANSMITTANCE;
    // rsqrt.approx(1.0): the vacancy factor of a splat whose Gaussian term underflowed.  The carry-over of
    // constant prefixes and the thinning rely on it being exactly 1 (it is on sm_100; checked at run time).
    float c1;
    {
        const float one = 1.0f;
        asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(c1) : "f"(one));
    }
    const bool exact_one = (c1 == 1.0f);
    // Per pixel: every list position <= pre_pos multiplies each sample by the same constants in every remaining
    // iteration; their product, formed in list order, is pre_T.
    float pre_T = 1.f;
    uint32_t pre_pos = in_range ? 0u : NEVER;

#pragma unroll 1
    for (int it = 0; it < SPLIT_ITERATIONS; it++) {
        const bool first = (it == 0);
        constexpr float ONE_OVER_SPLIT = 1.f / static_cast<float>(SPLIT);
        if (first) {
            T_p[0] = 1.f;
            T_p[SPLIT] = 1.f;
        }
#pragma unroll
        for (int s = 1; s < SPLIT; s++) T_p[s] = pre_T;
        const float interval = (depth_max - depth_min) * ONE_OVER_SPLIT;
        const float depth_hi = depth_min + interval * SPLIT;  // the last sample depth, as the loop forms it
        // list positions (1-based) any pixel of this warp still needs: (warp_first, warp_last]
        const uint32_t warp_last = __reduce_max_sync(FULL, in_range ? last_contributor : 0u);
        const uint32_t warp_first = __reduce_min_sync(FULL, pre_pos);
        // far end of the brackets of the warp's live pixels (depths are >= 0: their bits order like the floats)
        const float warp_hi = __uint_as_float(__reduce_max_sync(FULL, in_range ? __float_as_uint(depth_hi) : 0u));
        bool pfx = exact_one;  // still extending this pixel's constant prefix
#ifdef GGS_WALK_STATS
        unsigned int ws[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
        // The warp walks ITS list positions on its own: after pass 1 and the thinning of earlier iterations they are a
        // few dozen out of the tile's hundreds, so only their records are gathered -- lane j copies entry j of a mask
        // word with cp.async into the warp's private double buffer (a quarter of the staging area pass 1 used; the
        // block barrier behind pass 1 separates the two uses), the next non-empty word is in flight while this one is
        // processed, and no block barrier couples the warps of the CTA any more.
        const uint32_t word_end = (warp_last + 31) / 32;  // words [warp_first / 32, word_end) hold (warp_first, warp_last]
        auto word_bits = [&](uint32_t gw) -> uint32_t {
            uint32_t b = gw < mask_words ? (gw < MASK_WORDS ? s_mask[warp][gw] : FULL) : 0u;
            const uint32_t base = gw * 32;
            if (warp_last - base < 32u) b &= (1u << (warp_last - base)) - 1u;  // nothing lives beyond warp_last
            if (warp_first > base) b &= FULL << (warp_first - base);           // bits below are folded everywhere
            return b;
        };
        auto next_word = [&](uint32_t gw, uint32_t& b) -> uint32_t {  // first word >= gw with a position to visit
            b = 0;
            while (gw < word_end && !(b = word_bits(gw))) gw++;
            return gw;
        };
        auto gather = [&](uint32_t gw, uint32_t b, int slot) {
            if ((b >> lane) & 1u)
                stage_record(s_rec + (warp * 2 + slot) * (32 * REC_QUADS), lane, records, __ldg(list + gw * 32 + lane), 3);
            cp_async_commit();
        };
        uint32_t bits = 0;
        uint32_t gw = warp_last ? next_word(warp_first / 32, bits) : word_end;
        if (gw < word_end) gather(gw, bits, 0);
        for (int slot = 0; gw < word_end; slot ^= 1) {
            uint32_t bits_next = 0;
            const uint32_t gw_next = next_word(gw + 1, bits_next);
            if (gw_next < word_end) {
                gather(gw_next, bits_next, slot ^ 1);
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                cp_async_wait_all();
            }
            __syncwarp();
            {
                const uint32_t base = gw * 32;
                const uint32_t upto = warp_last - base >= 32u ? FULL : (1u << (warp_last - base)) - 1u;
                const uint32_t word_addr = s_rec_addr + (warp * 2 + slot) * (32 * REC_STRIDE);
                uint32_t keep = 0;    // positions some pixel of the warp can still need in later iterations
                uint32_t drop16 = 0;  // positions that leave the backward's list
                if (it < CLASSIFY_ITERATIONS && bits) {
                    // Most of what pass 1 accepted lies BEHIND the median depth (a pixel blends on until T < 1e-4), far
                    // behind every bracket of the warp once those are 0.8 or 0.1 wide.  Lane j examines entry j of the
                    // word: the peak depth is linear in the pixel, so its minimum over the warp's 8x4 rectangle is at
                    // a corner; if that is more than 6.5 sigma behind the far end of every live bracket the entry is
                    // quiet behind the bracket for every pixel -- factor exactly 1, now and later -- and is never
                    // visited (and leaves the backward's list too if even 16 sigma away).  The margins (2 % + 0.1
                    // sigma here, 6.5 against 6.004 in the rule itself) cover the rounding differences between this
                    // bound and the per-pixel evaluation for splats at least 4e-6 of the depth thick.
                    bool behind = false, behind16 = false;
                    if ((bits >> lane) & 1u) {
                        const uint32_t rec = word_addr + lane * REC_STRIDE;
                        const float4 q0 = lds128(rec);
                        const float4 q1 = lds128(rec + 16);
                        const float4 q2 = lds128(rec + 32);
                        const float dx0 = q0.x - rx0, dy0 = q0.y - ry0;  // d = mean - pixel, pixel in [x0, x0+7] x [y0, y0+3]
                        const float t_min = fminf(q1.z * dx0, q1.z * (dx0 - 7.f)) + fminf(q1.w * dy0, q1.w * (dy0 - 3.f)) + q2.x;
                        const float rs = q2.y;
                        const float gap = (t_min - warp_hi) * rs;
                        const bool thick = rs > 0.f && rs * warp_hi < 2.5e5f;
                        behind = thick && gap > QUIET_SIGMAS * 1.02f + 0.1f;
                        behind16 = thick && gap > FAR_SIGMAS * 1.02f + 0.1f && t_min - warp_hi > 2e-5f * warp_hi;
                    }
                    bits &= ~__ballot_sync(FULL, behind);
                    drop16 |= __ballot_sync(FULL, behind16);
                }
                WALK_STAT(6, 1);
                while (bits) {
                    const uint32_t bit = bits & (0u - bits);
                    bits ^= bit;
                    const int jj = __ffs(bit) - 1;
                    const uint32_t pos = base + jj + 1;  // 1-based list position == `contributor`
                    const uint32_t rec = word_addr + jj * REC_STRIDE;
                    const float4 q0 = lds128(rec);
                    const float4 q1 = lds128(rec + 16);
                    const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    const float alpha = fminf(0.99f, q1.y * expf(power));
                    const bool open = pos > pre_pos;  // not yet folded into this pixel's prefix (false if !in_range)
                    const bool live = in_range && pos <= last_contributor && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    const bool act = open && live;
                    WALK_STAT(0, 1);
                    if (!__any_sync(FULL, act)) {
                        WALK_STAT(1, 1);
                        if (pfx && open) pre_pos = pos;  // contributes to none of this pixel's samples, ever
                        continue;  // (the backward's list keeps what it had for this position)
                    }
                    const float4 q2 = lds128(rec + 32);
                    const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                    const float rsigma = q2.y;
                    const bool ball = rsigma > 0;
                    // Exact shortcuts (ggs_common.cuh).  Every sample depth lies in [depth_min, depth_hi] (rounding is
                    // monotone), so a splat whose peak depth is r sigmas outside that interval is at least that far
                    // from every sample.  QUIET (6.5): its factor is exactly (1 - alpha) in front of the bracket and
                    // exactly 1 behind it.  Brackets only shrink, up to a few ulps of the depth; 6.5 against 6.004
                    // leaves 0.49 sigma, twice what 4 ulps are worth for a splat whose sigma is 2e-6 of the depth;
                    // thinner ones are never quiet.  FAR (16): exp() == 0 exactly, also in the backward's dT/dt sum
                    // (which re-derives the median depth through a multiply that may move it by an ulp).
                    const float front = depth_min - t_peak;
                    const float back = t_peak - depth_hi;
                    const bool in_front = front > 0.f;
                    const float reach = (in_front ? front : back) * rsigma;  // |delta| of every sample is at least this
                    const bool quiet = ball && reach > QUIET_SIGMAS && rsigma * depth_hi < 5e5f;
                    // constant factor in every remaining iteration: (1 - alpha) in front, 1 behind, none if inactive
                    const bool fold = pfx && open && (!act || quiet);
                    if (__any_sync(FULL, act && !quiet)) {
                        // needed again by this kernel, and by the backward
                        keep |= bit;
                        WALK_STAT(2, 1);
                        WALK_STAT(4, __popc(__ballot_sync(FULL, act)));
                        WALK_STAT(5, __popc(__ballot_sync(FULL, act && !quiet)));
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
                    } else {
                        // 1 - alpha * exp() is exactly 1 at every sample of every active pixel: factor
                        // ((1 - alpha) | 1) * rsqrt(1).  This kernel's list: a quiet splat behind the bracket is never
                        // needed again, one in front is needed until every pixel has folded it into its prefix.
                        if (__any_sync(FULL, act && in_front && !fold)) keep |= bit;
                        // The backward's list: its dT/dt pass adds one term per (pixel, entry), in any order, so an
                        // entry leaves that list once the term is exactly 0 for every pixel it is live on (pixels
                        // that folded it into their prefix still vote) -- far in front, or far behind with some slack
                        // (the recomputed depth_hi of a later iteration can sit a few ulps above this one).
                        const bool gone = quiet && reach > FAR_SIGMAS && (in_front || back > 1e-5f * depth_hi);
                        if (!__any_sync(FULL, live && !gone)) drop16 |= bit;
                        if (act) {
                            WALK_STAT(3, 1);
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
                    }
                    if (fold) {
                        pre_pos = pos;
                        pre_T = T_p[1];  // all samples of this iteration are still equal here
                    } else if (open) {
                        pfx = false;
                    }
                }
                if (exact_one && lane == 0 && gw < MASK_WORDS && gw < mask_words) {
                    // (bits below warp_first in this word were skipped, not examined: they are folded everywhere)
                    s_mask[warp][gw] = keep;
                    // the backward's list keeps what was not examined, except positions beyond warp_last
                    s_mask[WARPS + warp][gw] &= ~drop16 & upto;
                }
            }
            __syncwarp();  // every lane is done with this slot before the gather after next refills it
            gw = gw_next;
            bits = bits_next;
        }
#ifdef GGS_WALK_STATS
        ws[3] = __reduce_max_sync(FULL, ws[3]);  // constant-path entries (counted by the active lanes only)
        if (lane == 0)
            for (int k = 0; k < 8; k++)
                if (ws[k]) atomicAdd(&g_walk_stats[it][k], (unsigned long long)ws[k]);
#endif
        if (first) {
            in_range = (T_p[0] >= 0.5f) && (T_p[SPLIT] <= 0.5f) && in_range;
            if (!in_range) pre_pos = NEVER;
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
    const float w_max = __saturatef((T_p[0] - 0.5f) / (T_p[0] - T_p[SPLIT]));
    const float w_min = 1.f - w_max;
    const float mDepth = in_range ? w_max * depth_max + w_min * depth_min : 0.f;
    if (inside) out_mdepth[pix_id] = mDepth * rln;

    // what the backward's dT/dt pass has to visit: (near_first, ...] restricted to the thinned mask
    // what the backward's dT/dt pass has to visit: the 16-sigma mask, and the number of leading list positions it
    // has cleared (the backward starts staging at the first batch any of its warps needs)
    __syncwarp();
    const uint32_t kept_words = min(mask_words, (uint32_t)MASK_WORDS);
    uint32_t first_set = kept_words * 32u;  // nothing set among the kept words: positions beyond use the accept mask
    for (uint32_t gw = lane; gw < kept_words; gw += 32) {
        const uint32_t word = s_mask[WARPS + warp][gw];
        near_mask[(size_t)(mask_base + gw) * 8 + twarp] = word;
        if (word) first_set = min(first_set, gw * 32u + (uint32_t)__ffs(word) - 1u);
    }
    first_set = __reduce_min_sync(FULL, first_set);
    if (lane == 0) near_first[tile_id * TILE_WARPS + twarp] = first_set;
}

void launch_render_forward(const FrameParams& fp, const uint2* ranges, const uint32_t* point_list,
                           const SplatRecord* records, const RenderOut& out, bool require_depth,
                           cudaStream_t stream, bool cleared) {
    RenderParams rp;
    rp.W = fp.W;
    rp.H = fp.H;
    rp.gx = fp.gx;
    rp.focal_x = fp.focal_x;
    rp.focal_y = fp.focal_y;
    rp.bg = fp.view.bg;
    rp.tile_order = out.tile_order;
    const dim3 grid(fp.gx, fp.gy, TILE_WARPS / WARPS);
    const dim3 block(THREADS, 1, 1);
    if (!cleared) cudaMemsetAsync(out.max_contributor, 0, (size_t)fp.gx * fp.gy * sizeof(uint32_t), stream);
    static SmemOptIn opted_geometry, opted_colour;
    opt_in_shared_memory(render_forward_kernel<true>, REC_BYTES + MASK_BYTES, opted_geometry);
    opt_in_shared_memory(render_forward_kernel<false>, REC_BYTES, opted_colour);
    if (require_depth)
        render_forward_kernel<true><<<grid, block, REC_BYTES + MASK_BYTES, stream>>>(
            rp, ranges, point_list, records, out.color, out.alpha, out.normal, out.mdepth, out.n_contrib,
            out.max_contributor, out.accept_mask, out.warp_words, out.near_mask, out.near_first);
    else
        render_forward_kernel<false><<<grid, block, REC_BYTES, stream>>>(
            rp, ranges, point_list, records, out.color, out.alpha, out.normal, out.mdepth, out.n_contrib,
            out.max_contributor, out.accept_mask, out.warp_words, out.near_mask, out.near_first);
}

#ifdef GGS_WALK_STATS
void read_walk_stats(unsigned long long* out, bool reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_walk_stats, sizeof(g_walk_stats));
    if (reset) {
        static unsigned long long zero[SPLIT_ITERATIONS][8] = {};
        cudaMemcpyToSymbol(g_walk_stats, zero, sizeof(zero));
    }
}
#endif

}  // namespace ggs
