// Backward of the tile blend.  Semantics follow DGR/cuda_rasterizer/render_backward.cu:716-1069:
//   pass A (geometry only): front-to-back, dT/dt_m at the stored median depth, then
//           dL_dmt_dT_dtm = dL/dmdepth * rln / max(-dT/dt_m, 1e-7);
//   pass B: back-to-front over [0, max_contributor), recomputing alpha, with the colour / normal
//           "accumulated behind" recurrences, depth-term gradients and dL/dalpha.
// One CTA per 16x16 tile; warp w owns an 8x4 pixel block, as in the forward.  Pass B walks only the list
// positions that the forward recorded as blended by at least one pixel of the warp (accept_mask): about a
// third of the list.  Pass A walks fewer still: every term of dT/dt_m carries exp(-delta^2/2) of the splat's
// depth profile at the median depth, which is exactly 0 (ex2.approx.ftz underflow) beyond 16 sigma.  The forward's
// bisection exports the list positions it saw within 16 sigma of some pixel's bracket the last time it examined
// them (near_mask); what is left of those is tested against the median depth itself before any alpha arithmetic.
// Adding an exact zero changes nothing, so skipping those entries is exact.  Both passes are driven by the masks and
// warp-private: lane l fetches mask word l of a group of 32, the non-empty words are visited in order, lane j copies
// the record of entry j with cp.async into the warp's own double buffer while the previous word is processed -- only
// the records a warp needs are gathered and no block barrier couples the eight warps of a tile inside a pass.
// Each (warp, list entry) produces 17 partial sums.  Instead of 17 five-step shuffle trees and 17
// scalar atomics from one lane (the reference), the warp runs ONE transposing butterfly (20
// shuffles) that leaves component c on its own lane, and those 17 lanes issue a single predicated
// RED (double precision, see ggs_common.cuh) into the entry's contiguous 144-byte accumulator row.
#include "ggs_kernels.cuh"
#include "ggs_tile.cuh"

namespace ggs {

struct RenderBwdParams {
    int W, H, gx;
    float focal_x, focal_y;
    const float* bg;  // device [3]
    const uint32_t* tile_order;  // TileLayout::order or nullptr
};


constexpr int REC_QUADS_B = REC_QUADS;
constexpr size_t BWD_REC_BYTES = (size_t)2 * STAGE_QUADS * sizeof(float4);
constexpr uint32_t FULL_B = 0xffffffffu;

// Butterfly reduce-scatter of 17 values (v[17] must be 0).  Returns the warp-wide sum of component
// warp_component(lane) on lanes where that is >= 0.
__device__ __forceinline__ float warp_reduce_scatter17(const float (&v)[18], const int lane) {
    float a[10], b[6], c[4], d[2];
    {
        const bool hi = lane & 16;
#pragma unroll
        for (int i = 0; i < 9; i++) {
            const float keep = hi ? v[i + 9] : v[i];
            const float send = hi ? v[i] : v[i + 9];
            a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
        a[9] = 0.f;
    }
    {
        const bool hi = lane & 8;
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const float keep = hi ? a[i + 5] : a[i];
            const float send = hi ? a[i] : a[i + 5];
            b[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
        b[5] = 0.f;
    }
    {
        const bool hi = lane & 4;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const float keep = hi ? b[i + 3] : b[i];
            const float send = hi ? b[i] : b[i + 3];
            c[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
        c[3] = 0.f;
    }
    {
        const bool hi = lane & 2;
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const float keep = hi ? c[i + 2] : c[i];
            const float send = hi ? c[i] : c[i + 2];
            d[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
        }
    }
    const bool hi = lane & 1;
    const float keep = hi ? d[1] : d[0];
    const float send = hi ? d[0] : d[1];
    return keep + __shfl_xor_sync(0xffffffffu, send, 1);
}

__device__ __forceinline__ int warp_component(const int lane) {
    const int b1 = lane & 1, b2 = (lane >> 1) & 1, b4 = (lane >> 2) & 1, b8 = (lane >> 3) & 1, b16 = (lane >> 4) & 1;
    const int ci = b1 + 2 * b2;
    const int bi = ci + 3 * b4;
    const int ai = bi + 5 * b8;
    const int vi = ai + 9 * b16;
    const bool ok = (ci < 3) && (bi < 5) && (ai < 9) && (vi < GRAD_LIVE);
    return ok ? vi : -1;
}

template <bool GEOMETRY>
#ifndef GGS_BWD_MIN_BLOCKS
#define GGS_BWD_MIN_BLOCKS 3
#endif
__global__ void __launch_bounds__(TILE_PIXELS, GGS_BWD_MIN_BLOCKS)
    render_backward_kernel(const RenderBwdParams rp, const uint2* __restrict__ ranges,
                           const uint32_t* __restrict__ point_list, const SplatRecord* __restrict__ records,
                           const float* __restrict__ alphas, const float* __restrict__ normalmap,
                           const float* __restrict__ mdepth, const uint32_t* __restrict__ n_contrib,
                           const uint32_t* __restrict__ max_contributors, const uint32_t* __restrict__ accept_mask,
                           const uint32_t* __restrict__ warp_words, const uint32_t* __restrict__ near_mask,
                           const uint32_t* __restrict__ near_first, const float* __restrict__ dL_dpixels,
                           const float* __restrict__ dL_dpixel_mdepths, const float* __restrict__ dL_dalphas,
                           const float* __restrict__ dL_dpixel_normals, double* __restrict__ grad_acc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* const s_rec = reinterpret_cast<float4*>(smem_raw);                     // [2][STAGE_QUADS]

    const int W = rp.W, H = rp.H;
    const size_t HW = (size_t)H * W;
    uint32_t tile_x = blockIdx.x, tile_y = blockIdx.y;
    if (rp.tile_order) {  // in grid order the CTAs walk TileLayout::order: longest lists first
        const uint32_t t = __ldg(rp.tile_order + blockIdx.y * gridDim.x + blockIdx.x);
        tile_y = t / (uint32_t)rp.gx;
        tile_x = t - tile_y * (uint32_t)rp.gx;
    }
    const uint32_t tile_id = tile_y * rp.gx + tile_x;
    const uint32_t tid = threadIdx.x;
    const uint32_t warp = tid >> 5;
    const int lane = tid & 31;
    const uint2 pix = {tile_x * TILE_X + (warp & 1) * 8 + (lane & 7), tile_y * TILE_Y + (warp >> 1) * 4 + (lane >> 3)};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {static_cast<float>(pix.x), static_cast<float>(pix.y)};
    const float2 pixnf = {(pixf.x - static_cast<float>(W - 1) / 2.f) / rp.focal_x,
                          (pixf.y - static_cast<float>(H - 1) / 2.f) / rp.focal_y};

    const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
    const uint2 range = ranges[tile_id];
    const uint32_t mask_base = range.x / 32 + tile_id;
    const uint32_t* const list = point_list + range.x;
    const uint32_t mask_words = warp_words[tile_id * (TILE_PIXELS / 32) + warp];
    const uint32_t* my_mask = accept_mask + (size_t)mask_base * 8 + warp;  // word gw at my_mask[gw * 8]

    const float w_final = inside ? alphas[pix_id] : 0.f;
    const float T_final = 1.f - w_final;
    float mDepth = 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0u;
    const uint32_t warp_last = __reduce_max_sync(FULL_B, last_contributor);

    float accum_rec[3] = {0.f, 0.f, 0.f};
    float dL_dpixel[3] = {0.f, 0.f, 0.f};
    float dL_dfinalT = 0.f;
    float dL_dpixel_mt = 0.f;
    float accum_normal_rec[3] = {0.f, 0.f, 0.f};
    float dL_dpixel_normal[3] = {0.f, 0.f, 0.f};

    if (inside) {
#pragma unroll
        for (int ch = 0; ch < 3; ch++) dL_dpixel[ch] = dL_dpixels[ch * HW + pix_id];
        dL_dfinalT = -dL_dalphas[pix_id];
#pragma unroll
        for (int ch = 0; ch < 3; ch++) dL_dfinalT += __ldg(rp.bg + ch) * dL_dpixel[ch];
        if constexpr (GEOMETRY) {
            const float inv_w = 1 / w_final;
            const float rln = rnorm3df(pixnf.x, pixnf.y, 1.f);
            dL_dpixel_mt = dL_dpixel_mdepths[pix_id] * rln;
            float normaln[3];
#pragma unroll
            for (int ch = 0; ch < 3; ch++) {
                dL_dpixel_normal[ch] = dL_dpixel_normals[ch * HW + pix_id] * inv_w;
                normaln[ch] = normalmap[ch * HW + pix_id];
            }
            dL_dfinalT += dL_dpixel_normal[0] * normaln[0] + dL_dpixel_normal[1] * normaln[1] +
                          dL_dpixel_normal[2] * normaln[2];
            mDepth = mdepth[pix_id] * norm3df(pixnf.x, pixnf.y, 1.f);
        }
    }

    float dL_dmt_dT_dtm = 0.f;

    if constexpr (GEOMETRY) {
        // ---- pass A: dT/dt_m at the stored median depth, front to back --------------------------------
        float dT_dtm = 0.f;
        const bool live = !((mDepth == 0) || (last_contributor == 0) || !inside);
        const uint32_t warp_last_a = __reduce_max_sync(FULL_B, live ? last_contributor : 0u);
        // list positions <= nf need no visit: the forward cleared them all in near_mask
        const uint32_t nf = near_first[tile_id * TILE_WARPS + warp];
        const uint32_t* my_near = near_mask + (size_t)mask_base * 8 + warp;
        // The warp walks its own list positions (near_mask beyond nf, a few dozen of the tile's hundreds): lane l fetches
        // mask word wb + l, the non-empty words are visited in order, and for each of them lane j copies the record of
        // entry j with cp.async into the warp's private double buffer (an eighth of the staging area pass B uses),
        // the next non-empty word in flight while this one is processed.  No block barrier inside the pass.
        const uint32_t word_end = min((warp_last_a + 31) / 32, mask_words);
        float4* const wbuf = s_rec + warp * (2 * 32 * REC_QUADS_B);
        auto gather = [&](uint32_t gw, uint32_t b, int slot) {
            if ((b >> lane) & 1u)
                stage_record(wbuf + slot * (32 * REC_QUADS_B), lane, records, __ldg(list + gw * 32 + lane), 3);
            cp_async_commit();
        };
        for (uint32_t wb = warp_last_a ? nf / 32 : word_end; wb < word_end; wb += 32) {
            const uint32_t gwl = wb + lane;
            uint32_t mine = 0;
            if (gwl < word_end) {
                mine = __ldg((gwl < MASK_WORDS ? my_near : my_mask) + (size_t)gwl * 8);
                const uint32_t base = gwl * 32;
                if (warp_last_a - base < 32u) mine &= (1u << (warp_last_a - base)) - 1u;  // pos <= warp_last_a
                if (nf > base) mine &= 0xffffffffu << (nf - base);                        // pos > nf
            }
            uint32_t words = __ballot_sync(FULL_B, mine != 0);
            if (!words) continue;
            int k = __ffs(words) - 1;
            words &= words - 1;
            uint32_t bits = __shfl_sync(FULL_B, mine, k);
            gather(wb + k, bits, 0);
            for (int slot = 0; k >= 0; slot ^= 1) {
                const int k_next = words ? __ffs(words) - 1 : -1;
                words &= words - 1;
                const uint32_t bits_next = __shfl_sync(FULL_B, mine, k_next < 0 ? 0 : k_next);
                if (k_next >= 0) {
                    gather(wb + k_next, bits_next, slot ^ 1);
                    asm volatile("cp.async.wait_group 1;" ::: "memory");
                } else {
                    cp_async_wait_all();
                }
                __syncwarp();
                const uint32_t base = (wb + k) * 32;
                const float4* const buf = wbuf + slot * (32 * REC_QUADS_B);
                while (bits) {
                    const int jj = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const uint32_t pos = base + jj + 1;
                    const float4* rec = buf + jj * REC_QUADS_B;
                    const float4 q0 = rec[0];
                    const float4 q1 = rec[1];
                    const float4 q2 = rec[2];
                    const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                    const float t_peak = q1.z * d.x + q1.w * d.y + q2.x;
                    const float rsigma = q2.y;
                    const float t_delta = (mDepth - t_peak) * rsigma;
                    // The forward's list is exact but not minimal (it stops examining an entry once every pixel has
                    // folded it into its constant prefix): beyond 16 sigma exp(-delta^2/2) underflows to +0 (it does
                    // from 13.3 on) and the term below is exactly 0 -- decided before the alpha arithmetic.
                    const bool near = live && pos <= last_contributor &&
                                      !(fabsf(t_delta) > FAR_SIGMAS && fabsf(t_delta) < 3e38f);  // (0 * inf is not 0)
                    if (!__any_sync(FULL_B, near)) continue;
                    const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                    const float alpha = fminf(0.99f, q1.y * expf(power));
                    const bool act = near && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                    if (act) {
                        const float G_exp = expf(-0.5f * t_delta * t_delta);
                        const float Gt = alpha * G_exp;
                        dT_dtm += -0.25f * Gt / (1.f - Gt) * fabsf(t_delta) * rsigma;
                    }
                }
                __syncwarp();  // every lane is done with this slot before the gather after next refills it
                k = k_next;
                bits = bits_next;
            }
        }
        __syncthreads();  // pass B reuses the staging buffers
        dL_dmt_dT_dtm = dL_dpixel_mt / fmaxf(-dT_dtm, 1e-7f);
    }

    // ---- pass B: back to front ----------------------------------------------------------------------
    float last_alpha = 0.f;
    float last_color[3] = {0.f, 0.f, 0.f};
    float last_normal[3] = {0.f, 0.f, 0.f};

    const float ddelx_dx = 0.5f * W;
    const float ddely_dy = 0.5f * H;
    const int my_comp = warp_component(lane);

    // The warp walks its own accepted list positions back to front (accept_mask: ~80 of the tile's hundreds), with the
    // same private gather as pass A: lane l fetches mask word wb + l, the non-empty words are visited from the back,
    // lane j copies the four quads of entry j's record with cp.async into the warp's double buffer and keeps the
    // entry's Gaussian index for the atomics; the next word is in flight while this one is processed.
    const uint32_t word_end_b = min((warp_last + 31) / 32, mask_words);
    float4* const wbuf_b = s_rec + warp * (2 * 32 * REC_QUADS_B);
    uint32_t id_slot0 = 0u, id_slot1 = 0u;  // Gaussian index of entry `lane` of the word in slot 0 / 1
    auto gather_b = [&](uint32_t gw, uint32_t b, int slot) {
        if ((b >> lane) & 1u) {
            const uint32_t id = __ldg(list + gw * 32 + lane);
            if (slot)
                id_slot1 = id;
            else
                id_slot0 = id;
            stage_record(wbuf_b + slot * (32 * REC_QUADS_B), lane, records, id, 4);
        }
        cp_async_commit();
    };
    for (int wb = word_end_b ? (int)((word_end_b - 1) / 32 * 32) : -1; wb >= 0; wb -= 32) {
        const uint32_t gwl = wb + lane;
        uint32_t mine = 0;
        if (gwl < word_end_b) {
            mine = __ldg(my_mask + (size_t)gwl * 8);
            const uint32_t base = gwl * 32;
            if (warp_last - base < 32u) mine &= (1u << (warp_last - base)) - 1u;  // idx = base + jj < warp_last
        }
        uint32_t words = __ballot_sync(FULL_B, mine != 0);
        if (!words) continue;
        int k = 31 - __clz(words);
        words ^= 1u << k;
        uint32_t bits = __shfl_sync(FULL_B, mine, k);
        gather_b(wb + k, bits, 0);
        for (int slot = 0; k >= 0; slot ^= 1) {
            const int k_next = words ? 31 - __clz(words) : -1;
            if (k_next >= 0) words ^= 1u << k_next;
            const uint32_t bits_next = __shfl_sync(FULL_B, mine, k_next < 0 ? 0 : k_next);
            if (k_next >= 0) {
                gather_b(wb + k_next, bits_next, slot ^ 1);
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                cp_async_wait_all();
            }
            __syncwarp();
            const uint32_t base = (wb + k) * 32;
            const float4* const buf = wbuf_b + slot * (32 * REC_QUADS_B);
            const uint32_t ids_mine = slot ? id_slot1 : id_slot0;
            while (bits) {
                const int jj = 31 - __clz(bits);
                bits ^= 1u << jj;
                const uint32_t idx = base + jj;  // 0-based list position == the reference's `contributor`
                const int j = jj;
                const float4* rec = buf + j * REC_QUADS_B;
                const float4 q0 = rec[0];
                const float4 q1 = rec[1];
                const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                const float4 con_o = {q0.z, q0.w, q1.x, q1.y};
                const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
                const float G = expf(power);
                const float alpha = fminf(0.99f, con_o.w * G);
                const bool valid = !((idx >= last_contributor) || (power > 0.0f) || (alpha < 1.0f / 255.0f));
                if (!__any_sync(FULL_B, valid)) continue;

                float v[18];
#pragma unroll
                for (int k = 0; k < 18; k++) v[k] = 0.f;

                if (valid) {
                    T = T / (1.f - alpha);
                    const float blending_weight = alpha * T;
                    const float4 q2 = rec[2];
                    const float4 q3 = rec[3];
                    const float col[3] = {q2.z, q2.w, q3.x};
                    float dL_dopa = 0.0f;
#pragma unroll
                    for (int ch = 0; ch < 3; ch++) {
                        const float c = col[ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        const float dL_dchannel = dL_dpixel[ch];
                        dL_dopa += (c - accum_rec[ch]) * dL_dchannel;
                        v[7 + ch] = blending_weight * dL_dchannel;
                    }
                    float dL_dt = 0.f;
                    float dL_dopa_sigma = 0.f;
                    const float2 rpl = {q1.z, q1.w};
                    if constexpr (GEOMETRY) {
                        const float nrm[3] = {q3.y, q3.z, q3.w};
#pragma unroll
                        for (int ch = 0; ch < 3; ch++) {
                            const float n = nrm[ch];
                            accum_normal_rec[ch] =
                                last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
                            last_normal[ch] = n;
                            const float dL_dchannel = dL_dpixel_normal[ch];
                            dL_dopa += (n - accum_normal_rec[ch]) * dL_dchannel;
                            v[10 + ch] = blending_weight * dL_dchannel;
                        }
                        const float t_peak = rpl.x * d.x + rpl.y * d.y + q2.x;
                        const float rsigma = q2.y;
                        const float t_delta = (mDepth - t_peak) * rsigma;
                        const float G_exp = expf(-0.5f * t_delta * t_delta);
                        const float Gt = alpha * G_exp;
                        float dL_dGt = dL_dmt_dT_dtm * 0.25f / (1.f - Gt);
                        dL_dGt = mDepth > t_peak ? dL_dGt : -dL_dGt;
                        dL_dGt = rsigma > 0 ? dL_dGt : 0.f;
                        dL_dopa_sigma = dL_dGt * G_exp - dL_dmt_dT_dtm * (t_delta > 0 ? 0.5f / (1.f - alpha) : 0.f);
                        const float dL_ddelta = -dL_dGt * Gt * t_delta;
                        v[16] = dL_ddelta * (mDepth - t_peak);
                        const float dL_dt_peak = -dL_ddelta * rsigma;
                        dL_dt = dL_dt_peak;
                        v[15] = dL_dt;
                        v[13] = dL_dt * d.x;
                        v[14] = dL_dt * d.y;
                    }
                    dL_dopa *= T;
                    if constexpr (GEOMETRY) dL_dopa += dL_dopa_sigma;
                    dL_dopa += -T_final / (1.f - alpha) * dL_dfinalT;
                    last_alpha = alpha;

                    const float dL_dG = con_o.w * dL_dopa;
                    const float gdx = G * d.x;
                    const float gdy = G * d.y;
                    const float dG_ddelx = -gdx * con_o.x - gdy * con_o.y;
                    const float dG_ddely = -gdy * con_o.z - gdx * con_o.y;
                    float dL_ddelx = dL_dG * dG_ddelx;
                    float dL_ddely = dL_dG * dG_ddely;
                    if constexpr (GEOMETRY) {
                        dL_ddelx += dL_dt * rpl.x;
                        dL_ddely += dL_dt * rpl.y;
                    }
                    v[0] = dL_ddelx * ddelx_dx;
                    v[1] = dL_ddely * ddely_dy;
                    v[2] = fabsf(v[0]) + fabsf(v[1]);
                    v[3] = -0.5f * gdx * d.x * dL_dG;
                    v[4] = -0.5f * gdx * d.y * dL_dG;
                    v[5] = -0.5f * gdy * d.y * dL_dG;
                    v[6] = G * dL_dopa;
                }
                const float sum = warp_reduce_scatter17(v, lane);
                const uint32_t id = __shfl_sync(FULL_B, ids_mine, jj);
                if (my_comp >= 0) {
                    if (GEOMETRY || my_comp < 10)
                        atomicAdd(grad_acc + (size_t)id * GRAD_STRIDE + my_comp, static_cast<double>(sum));
                }
            }
            __syncwarp();  // every lane is done with this slot before the gather after next refills it
            k = k_next;
            bits = bits_next;
        }
    }
}

void launch_render_backward(const FrameParams& fp, const uint2* ranges, const uint32_t* point_list,
                            const SplatRecord* records, const RenderBwdIn& in, double* grad_acc,
                            bool require_depth, cudaStream_t stream) {
    RenderBwdParams rp;
    rp.W = fp.W;
    rp.H = fp.H;
    rp.gx = fp.gx;
    rp.focal_x = fp.focal_x;
    rp.focal_y = fp.focal_y;
    rp.bg = fp.view.bg;
    rp.tile_order = in.tile_order;
    const dim3 grid(fp.gx, fp.gy, 1);
    const dim3 block(TILE_PIXELS, 1, 1);
    constexpr size_t SMEM = BWD_REC_BYTES;  // eight warps x two slots x 32 records
    static SmemOptIn opted_geometry, opted_colour;
    opt_in_shared_memory(render_backward_kernel<true>, SMEM, opted_geometry);
    opt_in_shared_memory(render_backward_kernel<false>, SMEM, opted_colour);
    if (require_depth)
        render_backward_kernel<true><<<grid, block, SMEM, stream>>>(
            rp, ranges, point_list, records, in.alpha, in.normal, in.mdepth, in.n_contrib, in.max_contributor,
            in.accept_mask, in.warp_words, in.near_mask, in.near_first, in.dL_dcolor, in.dL_dmdepth, in.dL_dalpha,
            in.dL_dnormal, grad_acc);
    else
        render_backward_kernel<false><<<grid, block, SMEM, stream>>>(
            rp, ranges, point_list, records, in.alpha, in.normal, in.mdepth, in.n_contrib, in.max_contributor,
            in.accept_mask, in.warp_words, in.near_mask, in.near_first, in.dL_dcolor, in.dL_dmdepth, in.dL_dalpha,
            in.dL_dnormal, grad_acc);
}

}  // namespace ggs
