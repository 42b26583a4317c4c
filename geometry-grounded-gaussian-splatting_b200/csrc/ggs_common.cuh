// Shared definitions of the B200-native Gaussian rasterizer kernels (sm_100a).
//
// Compile-time configuration mirrors the reference's shipped configuration
// (DGR/cuda_rasterizer/config.h:15-39): MEDIAN_DEPTH_INIT, TRAINING=1, 16x16 tiles.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ggs {

constexpr int TILE_X = 16;
constexpr int TILE_Y = 16;
constexpr int TILE_PIXELS = TILE_X * TILE_Y;
constexpr float MIN_TRANSMITTANCE = 0.45f;  // config.h:31
constexpr int SPLIT = 8;                    // config.h:32
constexpr float SAMPLE_RANGE = 0.4f;        // config.h:34 (TRAINING)
constexpr int SPLIT_ITERATIONS = 5;         // config.h:35 (TRAINING)
constexpr int POINT_BATCH = 2;              // config.h:27 SAMPLE_BATCH_SIZE
// Depth-profile shortcuts of the bisection passes (image and point-query kernels), in units of the splat's depth sigma:
//   FAR:   beyond 16 sigma exp(-delta^2/2) = 2^-184 and ex2.approx.ftz returns +0 (it does from 13.3 sigma on): the
//          splat's term is exactly absent from the forward AND from the backward's dT/dt sum -> it leaves the lists.
//   QUIET: from 6.5 sigma on the Gaussian term no longer reaches the factor: g <= 2^-30.4 (ex2.approx is good to
//          2^-22), so alpha * g < 2^-25 and 1 - alpha * g rounds to exactly 1.0f, with or without multiply-add
//          contraction (the smallest |delta| with that property is 6.004).  No exp / rsqrt is issued for it, the
//          forward is bit-identical; its tiny backward term is kept, so it stays in the exported masks.
constexpr float FAR_SIGMAS = 16.f;
#ifndef GGS_QUIET_SIGMAS
#define GGS_QUIET_SIGMAS 6.5f
#endif
constexpr float QUIET_SIGMAS = GGS_QUIET_SIGMAS;
static_assert(QUIET_SIGMAS >= 6.25f && QUIET_SIGMAS <= FAR_SIGMAS, "1 - alpha * g must round to 1");
constexpr int MASK_WORDS = 128;  // per warp, words of the acceptance mask the forward keeps (and thins) in shared memory

// Per-Gaussian record consumed by the tile kernels: 16 floats = one 64-byte, 64-byte aligned
// gather unit (two full 32-byte sectors, no wasted bytes).
//   q0 = (mean2D.x, mean2D.y, conic.x, conic.y)
//   q1 = (conic.z, opacity*coef, ray_plane.x, ray_plane.y)
//   q2 = (ray_plane.z, ray_plane.w, r, g)
//   q3 = (b, normal.x, normal.y, normal.z)
struct __align__(16) SplatRecord {
    float4 q0, q1, q2, q3;
};
static_assert(sizeof(SplatRecord) == 64, "record must be 64 bytes");

// Per-Gaussian gradient accumulator written by the tile backward with warp-aggregated
// reductions: 17 live values, padded to 18, held in DOUBLE precision (144 bytes, 16-byte aligned).
// The per-warp partial sums are float (as in the reference); only the cross-warp / cross-tile accumulation
// through global atomics is done in double.  Several warps add partial sums of opposite sign into the same
// row in a timing-dependent order; in float that order noise reaches 1e-4 of a gradient on splats whose
// contributions cancel (measured: tools/grad_outlier.py), in double it is below 1e-12.
//   [0..2]  dL/dmean2D (x, y, |x|+|y|)
//   [3..6]  dL/dconic (x, y, z), dL/d(opacity*coef)
//   [7..9]  dL/dcolor
//   [10..12] dL/dnormal
//   [13..16] dL/dray_plane (x, y, tc, rsigma)
constexpr int GRAD_STRIDE = 18;
constexpr int GRAD_LIVE = 17;

// Camera tensors stay in device memory (as in the reference, which takes raw device pointers);
// nothing here forces a host round trip.
struct View {
    const float* view;    // [16] transposed world->view (reference layout, scene/cameras.py:70)
    const float* proj;    // [16] transposed full projection
    const float* campos;  // [3]
    const float* bg;      // [3]
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per DEVICE (and per function); a process may drive several
// devices (and several host threads may launch at once), so the largest size opted into so far is kept per device,
// atomically.  `state`: one zero-initialised array per kernel.
constexpr int MAX_DEVICES = 64;
struct SmemOptIn {
    unsigned long long bytes[MAX_DEVICES];
};
template <typename Kernel>
inline void opt_in_shared_memory(Kernel kernel, size_t smem, SmemOptIn& state) {
    if (smem <= 48 * 1024) return;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        return;
    }
    unsigned long long seen = __atomic_load_n(&state.bytes[dev], __ATOMIC_ACQUIRE);
    if (smem <= seen) return;
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    while (seen < smem && !__atomic_compare_exchange_n(&state.bytes[dev], &seen, (unsigned long long)smem, false,
                                                       __ATOMIC_RELEASE, __ATOMIC_ACQUIRE)) {
    }
}

// Reference host helper (rasterizer_impl.cu:37-50): NOT ceil(log2(n)).
inline uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb)
            msb += step;
        else
            msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// ---- workspace layouts (private to this library; offsets exported for tests) -------------

struct GeomLayout {
    size_t records;        // SplatRecord[P]
    size_t depths;         // f32[P]
    size_t tiles_touched;  // u32[P]
    size_t point_offsets;  // u32[P]: inclusive scan of tiles_touched taken in DEPTH ORDER (see order_vals_a)
    size_t clamped;        // u8[P]  (bit c set when colour channel c was clamped)
    size_t rects;          // uint2[P]: (min.x | min.y<<16, max.x | max.y<<16) tile rectangle
    size_t scan_status;    // u64[scan tiles + 1]
    size_t scan_counter;   // u32
    size_t total;          // u32 D mirror
    // Depth pre-sort of the Gaussians (stable, by the float bits of the depth; culled Gaussians last): the key
    // duplication walks the Gaussians in this order, so that the pair sort only has to order the tile bits.
    size_t order_keys_a, order_keys_b;  // u32[P]  (depth bits; culled Gaussians 0xffffffff)
    size_t order_vals_a, order_vals_b;  // u32[P]  (the sorted order lands in order_vals_a)
    size_t order_hist;                  // u32[MAX_PASSES][RADIX]
    size_t order_status;                // u32[4][sort tiles][RADIX]
    size_t order_counters;              // u32[MAX_PASSES]
    size_t order_status_bytes;
    size_t bytes;
};

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 12;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int MAX_PASSES = 8;
constexpr int ORDER_PASSES = 4;  // 32 depth bits

constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_TILE = SCAN_ITEMS * SCAN_THREADS;

inline GeomLayout geom_layout(size_t P) {
    GeomLayout L;
    size_t o = 0;
    auto take = [&](size_t bytes) {
        size_t at = o;
        o = align_up(o + bytes, 256);
        return at;
    };
    L.records = take(P * sizeof(SplatRecord));
    L.depths = take(P * 4);
    L.tiles_touched = take(P * 4);
    L.point_offsets = take(P * 4);
    L.clamped = take(P);
    L.rects = take(P * 8);
    size_t tiles = (P + SCAN_TILE - 1) / SCAN_TILE + 1;
    L.scan_status = take(tiles * 8);
    L.scan_counter = take(256);
    L.total = take(256);
    L.order_keys_a = take(P * 4);
    L.order_keys_b = take(P * 4);
    L.order_vals_a = take(P * 4);
    L.order_vals_b = take(P * 4);
    L.order_hist = take(MAX_PASSES * RADIX * 4);
    L.order_status_bytes = (size_t)ORDER_PASSES * ((P + SORT_TILE - 1) / SORT_TILE) * RADIX * 4;
    L.order_status = take(L.order_status_bytes);
    L.order_counters = take(MAX_PASSES * 4);
    L.bytes = o;
    return L;
}

struct BinLayout {
    size_t keys_a, keys_b;  // u64[D]
    size_t vals_a, vals_b;  // u32[D]
    size_t hist;            // u32[MAX_PASSES][RADIX]  (global digit histograms -> exclusive offsets)
    size_t status;          // u32[passes][tiles][RADIX] decoupled look-back descriptors
    size_t counters;        // u32[MAX_PASSES] dynamic tile ids
    size_t accept_mask;     // u32[(D/32 + tiles) * 8]: per (tile, warp) bitmask of blended list positions
    size_t near_mask;       // same shape: accept_mask thinned to the positions whose depth profile can still be
                            // non-zero at the pixel's median depth (what the backward's dT/dt pass must visit)
    size_t bytes;
    size_t status_bytes;
};

inline BinLayout bin_layout(size_t D, int passes, size_t tiles_for_mask = 0) {
    BinLayout L;
    size_t o = 0;
    auto take = [&](size_t bytes) {
        size_t at = o;
        o = align_up(o + bytes, 256);
        return at;
    };
    L.keys_a = take(D * 8);
    L.keys_b = take(D * 8);
    L.vals_a = take(D * 4);
    L.vals_b = take(D * 4);
    L.hist = take(MAX_PASSES * RADIX * 4);
    size_t tiles = (D + SORT_TILE - 1) / SORT_TILE;
    L.status_bytes = (size_t)passes * tiles * RADIX * 4;
    L.status = take(L.status_bytes);
    L.counters = take(MAX_PASSES * 4);
    L.accept_mask = take(tiles_for_mask ? (D / 32 + tiles_for_mask + 1) * 8 * 4 : 0);
    L.near_mask = take(tiles_for_mask ? (D / 32 + tiles_for_mask + 1) * 8 * 4 : 0);
    L.bytes = o;
    return L;
}

struct TileLayout {
    size_t ranges;           // uint2[tiles]
    size_t max_contributor;  // u32[tiles]
    size_t warp_words;       // u32[tiles][8]: words of accept_mask each warp of the tile produced
    size_t near_first;       // u32[tiles][8]: list positions <= this are exactly zero in the warp's dT/dt pass
    size_t order;            // u32[tiles]: tile ids by descending list length -- the order the render kernels' CTAs take
                             // them in (longest first, so that the grid's tail is made of short tiles)
    size_t bytes;
};
inline TileLayout tile_layout(size_t tiles) {
    TileLayout L;
    L.ranges = 0;
    L.max_contributor = align_up(tiles * 8, 256);
    L.warp_words = L.max_contributor + align_up(tiles * 4, 256);
    L.near_first = L.warp_words + align_up(tiles * 32, 256);
    L.order = L.near_first + align_up(tiles * 32, 256);
    L.bytes = L.order + align_up(tiles * 4, 256);
    return L;
}

struct ImgLayout {
    size_t n_contrib;  // u32[W*H]
    size_t bytes;
};
inline ImgLayout img_layout(size_t npix) {
    ImgLayout L;
    L.n_contrib = 0;
    L.bytes = align_up(npix * 4, 256);
    return L;
}

#ifdef __CUDACC__
// A warp's 32 rows of F = 4 Q floats ([32][F], contiguous in global memory) -> shared-memory rows of stride RS
// (odd: conflict-free when every lane later walks its own row).  Rows whose bit in `rows` is clear are skipped.
// The loads of a batch are issued together before any is stored: with the dependent one-load-per-iteration form
// half of the stall samples of the per-Gaussian kernels sat on this copy (one DRAM round trip per 512 bytes per warp).
// QC > 0: Q known at compile time (12 at SH degree 3): row / column of an element by constant division instead of
// the running subtraction (which was 15 % of the per-Gaussian backward's instructions).
template <int QC>
__device__ __forceinline__ void load_rows_to_shared_impl(const float4* __restrict__ src, int Qr, int ng, uint32_t rows,
                                                         float* __restrict__ swarp, int RS, int lane) {
    constexpr int BATCH = 6;
    const int Q = QC > 0 ? QC : Qr;
    const int total = ng * Q;
    for (int e0 = lane; e0 < total; e0 += 32 * BATCH) {
        float4 v[BATCH];
#pragma unroll
        for (int j = 0; j < BATCH; j++) {
            const int e = e0 + 32 * j;
            const int gi = e / Q;
            if (e < total && ((rows >> gi) & 1u)) v[j] = __ldg(src + e);
        }
#pragma unroll
        for (int j = 0; j < BATCH; j++) {
            const int e = e0 + 32 * j;
            const int gi = e / Q, rem = e - gi * Q;
            if (e < total && ((rows >> gi) & 1u)) {
                float* d = swarp + gi * RS + rem * 4;
                d[0] = v[j].x;
                d[1] = v[j].y;
                d[2] = v[j].z;
                d[3] = v[j].w;
            }
        }
    }
}
__device__ __forceinline__ void load_rows_to_shared(const float4* __restrict__ src, int Q, int ng, uint32_t rows,
                                                    float* __restrict__ swarp, int RS, int lane) {
    if (Q == 12)
        load_rows_to_shared_impl<12>(src, Q, ng, rows, swarp, RS, lane);
    else
        load_rows_to_shared_impl<0>(src, Q, ng, rows, swarp, RS, lane);
}

// The same copy without registers in between and without waiting: 4-byte cp.async (the odd row stride leaves the rows
// 4-byte aligned only), every warp instruction one 128-byte line of the slab.  The caller overlaps whatever does not
// need the rows and then waits -- cp.async.wait_all + __syncwarp at a convergent point, or, where the consumers have
// diverged from the lanes that issued part of their row, an mbarrier every lane's copies arrive on (below).
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}
template <int FC>
__device__ __forceinline__ void copy_rows_to_shared_async_impl(const float* __restrict__ src, int Fr, int ng,
                                                               uint32_t rows, float* __restrict__ swarp, int RS,
                                                               int lane) {
    const int F = FC > 0 ? FC : Fr;
    const int total = ng * F;
#pragma unroll 4
    for (int e = lane; e < total; e += 32) {
        const int gi = e / F, rem = e - gi * F;
        if ((rows >> gi) & 1u) cp_async4(swarp + gi * RS + rem, src + e);
    }
}
// src: the warp's [ng][F] floats; rows: bit gi set = row gi wanted
__device__ __forceinline__ void copy_rows_to_shared_async(const float* __restrict__ src, int F, int ng, uint32_t rows,
                                                          float* __restrict__ swarp, int RS, int lane) {
    if (F == 48)
        copy_rows_to_shared_async_impl<48>(src, F, ng, rows, swarp, RS, lane);
    else
        copy_rows_to_shared_async_impl<0>(src, F, ng, rows, swarp, RS, lane);
}
__device__ __forceinline__ void cp_async_wait_everything() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// A warp's mbarrier for such copies: init for 32 arrivals (one lane, then __syncwarp), every lane lets its copies
// arrive on it (also a lane that issued none), consumers wait for phase 0 -- they may have diverged by then.
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(bar));
    asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(a), "r"(arrivals) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_on(uint64_t* bar) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(bar));
    asm volatile("cp.async.mbarrier.arrive.noinc.shared.b64 [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(bar));
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(a),
        "r"(parity)
        : "memory");
}

// The way back: shared-memory rows of stride RS -> [ng][4 Q] floats contiguous in global memory, 16-byte stores.
template <int QC>
__device__ __forceinline__ void store_rows_from_shared_impl(float4* __restrict__ dst, int Qr, int ng,
                                                            const float* __restrict__ swarp, int RS, int lane) {
    const int Q = QC > 0 ? QC : Qr;
    const int total = ng * Q;
    for (int e = lane; e < total; e += 32) {
        const int gi = e / Q, rem = e - gi * Q;
        const float* sr = swarp + gi * RS + rem * 4;
        dst[e] = float4{sr[0], sr[1], sr[2], sr[3]};
    }
}
__device__ __forceinline__ void store_rows_from_shared(float4* __restrict__ dst, int Q, int ng,
                                                       const float* __restrict__ swarp, int RS, int lane) {
    if (Q == 12)
        store_rows_from_shared_impl<12>(dst, Q, ng, swarp, RS, lane);
    else
        store_rows_from_shared_impl<0>(dst, Q, ng, swarp, RS, lane);
}
#endif

}  // namespace ggs
