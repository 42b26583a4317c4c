// Onesweep-style least-significant-digit radix sort of (u64 key, u32 value) pairs, written for
// this library (no CUB).  Replaces cub::DeviceRadixSort::SortPairs at
// DGR/cuda_rasterizer/rasterizer_impl.cu:406-412 (stable, ascending, bits [begin_bit, end_bit)).
//
//   1. histogram_kernel : one read of the keys builds the 256-bin digit histogram of every pass.
//   2. offsets_kernel   : exclusive scan of each histogram -> global bin offsets.
//   3. onesweep_kernel  : one launch per 8-bit digit.  A block ranks a tile of 3072 pairs with
//      warp match-any multisplit (stable), resolves the tile's global bin bases through a chained
//      scan with decoupled look-back (one thread per digit), reorders the tile through shared
//      memory and writes each digit run coalesced.
// Every pass moves 12 B in + 12 B out per pair: HBM-bound.
#include "ggs_kernels.cuh"

namespace ggs {

constexpr uint32_t FLAG_AGG = 1u << 30;
constexpr uint32_t FLAG_PREFIX = 2u << 30;
constexpr uint32_t VALUE_MASK = (1u << 30) - 1;
// blocks of the digit-histogram kernel: every block flushes up to passes x 256 bins with global atomics
#ifndef GGS_HIST_BLOCKS
#define GGS_HIST_BLOCKS (148 * 2)
#endif
#ifndef GGS_LOOK_WINDOW
#define GGS_LOOK_WINDOW 8
#endif
constexpr int LOOK_WINDOW = GGS_LOOK_WINDOW;

__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Digit extraction of one pass (see SortPlan).  vshift < 0: plain digit (key >> shift) & mask.
struct DigitSpec {
    int shift;
    uint32_t mask;
    int vshift;  // >= 0: digit of the virtual key ((tile & tile_mask) << nb) | (depth_top_byte - emin)
    uint32_t emin, tile_mask;
    int nb;
};
__device__ __forceinline__ uint32_t digit_of(const uint64_t key, const DigitSpec& s) {
    if (s.vshift < 0) return (uint32_t)(key >> s.shift) & s.mask;
    const uint32_t v = ((((uint32_t)(key >> 32)) & s.tile_mask) << s.nb) | ((((uint32_t)key) >> 24) - s.emin);
    return (v >> s.vshift) & s.mask;
}
__device__ __forceinline__ uint32_t digit_of(const uint32_t key, const DigitSpec& s) {
    return (key >> s.shift) & s.mask;  // 32-bit tile keys: plain digits only
}
struct PassSpecs {
    DigitSpec d[MAX_PASSES];
};

static DigitSpec pass_spec(const SortPlan& p, int pass) {
    DigitSpec s;
    s.emin = p.emin;
    s.nb = p.nb;
    s.tile_mask = p.tile_bits >= 32 ? 0xffffffffu : ((1u << p.tile_bits) - 1u);
    if (!p.compact || pass < 3) {
        s.shift = p.begin_bit + RADIX_BITS * pass;
        const int width = p.end_bit - s.shift < RADIX_BITS ? p.end_bit - s.shift : RADIX_BITS;
        s.mask = (1u << width) - 1u;
        s.vshift = -1;
    } else {
        s.shift = 0;
        s.vshift = RADIX_BITS * (pass - 3);
        const int vbits = p.tile_bits + p.nb;
        const int width = vbits - s.vshift < RADIX_BITS ? vbits - s.vshift : RADIX_BITS;
        s.mask = (1u << width) - 1u;
    }
    return s;
}

SortPlan plain_sort_plan(int begin_bit, int end_bit) {
    SortPlan p;
    p.begin_bit = begin_bit;
    p.end_bit = end_bit;
    p.passes = (end_bit - begin_bit + RADIX_BITS - 1) / RADIX_BITS;
    p.compact = false;
    p.emin = 0;
    p.nb = 0;
    p.tile_bits = end_bit > 32 ? end_bit - 32 : 0;
    return p;
}

SortPlan rasterizer_sort_plan(int end_bit, uint32_t depth_min_bits, uint32_t depth_max_bits) {
    SortPlan p = plain_sort_plan(0, end_bit);
    if (end_bit <= 32 || depth_max_bits < depth_min_bits) return p;
    const uint32_t emin = depth_min_bits >> 24, emax = depth_max_bits >> 24;
    int nb = 0;
    while ((1u << nb) < emax - emin + 1) nb++;
    const int compact_passes = 3 + (p.tile_bits + nb + RADIX_BITS - 1) / RADIX_BITS;
    if (nb <= 4 && compact_passes < p.passes) {
        p.compact = true;
        p.emin = emin;
        p.nb = nb;
        p.passes = compact_passes;
    }
    return p;
}

template <typename KeyT>
__global__ void __launch_bounds__(256) histogram_kernel(const KeyT* __restrict__ keys, uint32_t n,
                                                        const PassSpecs specs, int passes, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[MAX_PASSES][RADIX];
    for (int i = threadIdx.x; i < MAX_PASSES * RADIX; i += blockDim.x) (&s_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        // only the planned digits take part (the last one may be narrower than 8 bits)
        const KeyT k = __ldg(keys + i);
#pragma unroll
        for (int p = 0; p < MAX_PASSES; p++) {
            if (p < passes) atomicAdd(&s_hist[p][digit_of(k, specs.d[p])], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x) {
        const uint32_t c = (&s_hist[0][0])[i];
        if (c) atomicAdd(hist + i, c);
    }
}

// one block, 256 threads: exclusive scan of each pass' histogram in place
__global__ void __launch_bounds__(RADIX) offsets_kernel(uint32_t* __restrict__ hist, int passes) {
    __shared__ uint32_t s_warp[RADIX / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int p = 0; p < passes; p++) {
        const uint32_t c = hist[p * RADIX + threadIdx.x];
        uint32_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        uint32_t off = 0;
#pragma unroll
        for (int w = 0; w < RADIX / 32; w++)
            if (w < warp) off += s_warp[w];
        hist[p * RADIX + threadIdx.x] = off + incl - c;
        __syncthreads();
    }
}

// MATCH: rank with match.any (one instruction, ~10 cycles when most lanes hold different digits, as for depth
// mantissa bits -- but measured ~65 cycles when lanes share digits in runs, as for tile bits of depth-sorted
// pairs: the ADU pipe became the bottleneck of those passes).  !MATCH: eight ballots, constant cost.
// KeyT = uint64_t: the general sort.  KeyT = uint32_t: tile keys of pairs that were emitted in depth order; with EMIT64
// the pass is the last one and writes the full reference key (tile << 32 | depth bits of the pair's Gaussian) to
// keys64_out instead of the 32-bit key -- the 64-bit keys never move through the passes.
template <bool MATCH, typename KeyT, bool EMIT64>
__global__ void __launch_bounds__(SORT_THREADS) onesweep_kernel(const KeyT* __restrict__ keys_in,
                                                                KeyT* __restrict__ keys_out,
                                                                const uint32_t* __restrict__ vals_in,
                                                                uint32_t* __restrict__ vals_out, uint32_t n,
                                                                const DigitSpec spec,
                                                                const uint32_t* __restrict__ bin_offsets,
                                                                uint32_t* __restrict__ status,
                                                                uint32_t* __restrict__ counter,
                                                                uint64_t* __restrict__ keys64_out,
                                                                const uint32_t* __restrict__ depth_bits,
                                                                const uint32_t* __restrict__ n_dev) {
    constexpr int WARPS = SORT_THREADS / 32;
    if (n_dev) n = min(n, __ldg(n_dev));  // launched for the capacity, the count lives on the device
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp_hist[WARPS][RADIX];
    __shared__ uint32_t s_bin_start[RADIX];
    __shared__ uint32_t s_global_base[RADIX];
    __shared__ uint32_t s_scan[WARPS];
    __shared__ KeyT s_keys[SORT_TILE];
    __shared__ uint32_t s_vals[SORT_TILE];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(counter, 1u);
    for (int i = tid; i < WARPS * RADIX; i += SORT_THREADS) (&s_warp_hist[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t base = tile * SORT_TILE;
    if (base >= n) return;  // (whole block: tiles beyond the count publish nothing and nobody looks back at them)
    const uint32_t count = min((uint32_t)SORT_TILE, n - base);

    // ---- load (warp-striped) and rank --------------------------------------------------------
    KeyT key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
    const uint32_t warp_base = warp * 32 * SORT_ITEMS;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const uint32_t li = warp_base + i * 32 + lane;
        if (li < count) {
            key[i] = __ldg(keys_in + base + li);
            val[i] = __ldg(vals_in + base + li);
        } else {
            key[i] = ~KeyT(0);
            val[i] = 0;
        }
    }
    const uint32_t lanemask_lt = (1u << lane) - 1;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        // padding (beyond the end of the input) must rank after every real pair: it takes the largest digit
        const uint32_t d = (warp_base + i * 32 + lane < count) ? digit_of(key[i], spec) : spec.mask;
        uint32_t peers;
        if constexpr (MATCH) {
            peers = __match_any_sync(0xffffffffu, d);
        } else {
            peers = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < RADIX_BITS; b++) {
                const bool one = (d >> b) & 1u;
                const uint32_t bal = __ballot_sync(0xffffffffu, one);
                peers &= one ? bal : ~bal;
            }
        }
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader) {
            old = s_warp_hist[warp][d];
            s_warp_hist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        rank[i] = old + __popc(peers & lanemask_lt);
        __syncwarp();
    }
    __syncthreads();

    // ---- per-digit: warp offsets, tile count, publish aggregate, local bin start ---------------
    const uint32_t d = tid;  // SORT_THREADS == RADIX
    uint32_t cnt = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
        const uint32_t c = s_warp_hist[w][d];
        s_warp_hist[w][d] = cnt;
        cnt += c;
    }
    uint32_t* my_status = status + (size_t)tile * RADIX + d;
    if (tile == 0)
        st_relaxed_u32(my_status, FLAG_PREFIX | cnt);
    else
        st_relaxed_u32(my_status, FLAG_AGG | cnt);

    uint32_t incl = cnt;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, s);
        if (lane >= s) incl += o;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < WARPS; w++)
        if (w < warp) off += s_scan[w];
    const uint32_t bin_start = off + incl - cnt;
    s_bin_start[d] = bin_start;

    // ---- decoupled look-back: one thread per digit, LOOK_WINDOW predecessors in flight per round ------------
    // When the tiles of a wave publish their aggregates at about the same time, a walk that reads one predecessor
    // per L2 round trip needs ~sqrt(2 x tiles) dependent loads (measured: 23 us of a pass over 326 tiles).  The
    // window issues its loads together and consumes them in order; it stops at the first word not yet published.
    uint32_t excl = 0;
    if (tile > 0) {
        int look = (int)tile - 1;
        bool found = false;
        while (!found) {
            uint32_t st[LOOK_WINDOW];
#pragma unroll
            for (int j = 0; j < LOOK_WINDOW; j++)
                st[j] = (look - j >= 0) ? ld_relaxed_u32(status + (size_t)(look - j) * RADIX + d) : FLAG_PREFIX;
            bool stop = false;
            int adv = 0;
#pragma unroll
            for (int j = 0; j < LOOK_WINDOW; j++) {
                if (!stop) {
                    if (st[j] & FLAG_PREFIX) {
                        excl += st[j] & VALUE_MASK;
                        found = true;
                        stop = true;
                    } else if (st[j] & FLAG_AGG) {
                        excl += st[j] & VALUE_MASK;
                        adv++;
                    } else {
                        stop = true;
                    }
                }
            }
            look -= adv;
        }
        st_relaxed_u32(my_status, FLAG_PREFIX | (excl + cnt));
    }
    s_global_base[d] = __ldg(bin_offsets + d) + excl - bin_start;
    __syncthreads();

    // ---- reorder through shared memory, then coalesced digit runs ------------------------------
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const uint32_t dg = (warp_base + i * 32 + lane < count) ? digit_of(key[i], spec) : spec.mask;
        const uint32_t pos = s_bin_start[dg] + s_warp_hist[warp][dg] + rank[i];
        s_keys[pos] = key[i];
        s_vals[pos] = val[i];
    }
    __syncthreads();
    for (uint32_t p = tid; p < count; p += SORT_THREADS) {
        const KeyT k = s_keys[p];
        const uint32_t dg = digit_of(k, spec);
        const uint32_t out = s_global_base[dg] + p;
        const uint32_t v = s_vals[p];
        if constexpr (EMIT64)
            keys64_out[out] = ((uint64_t)k << 32) | (uint64_t)__ldg(depth_bits + v);
        else
            keys_out[out] = k;
        vals_out[out] = v;
    }
}

int launch_radix_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, uint32_t n,
                      const SortPlan& plan, uint32_t* hist, uint32_t* status, size_t status_bytes,
                      uint32_t* counters, cudaStream_t stream) {
    if (n == 0) return 0;
    const int passes = plan.passes;
    PassSpecs specs;
    for (int p = 0; p < MAX_PASSES; p++) specs.d[p] = pass_spec(plan, p < passes ? p : 0);
    const uint32_t tiles = (n + SORT_TILE - 1) / SORT_TILE;
    cudaMemsetAsync(hist, 0, MAX_PASSES * RADIX * 4, stream);
    cudaMemsetAsync(status, 0, status_bytes, stream);
    cudaMemsetAsync(counters, 0, MAX_PASSES * 4, stream);
    const int hist_blocks = (int)min((uint32_t)GGS_HIST_BLOCKS, (n + 255) / 256);
    histogram_kernel<uint64_t><<<hist_blocks, 256, 0, stream>>>(keys_a, n, specs, passes, hist);
    offsets_kernel<<<1, RADIX, 0, stream>>>(hist, passes);
    uint64_t* kin = keys_a;
    uint64_t* kout = keys_b;
    uint32_t* vin = vals_a;
    uint32_t* vout = vals_b;
    for (int p = 0; p < passes; p++) {
        // tile-bit digits (virtual key, or plain digits above bit 32) come in runs: rank them with ballots
        const bool tile_digit = specs.d[p].vshift >= 0 || specs.d[p].shift >= 32;
        if (tile_digit)
            onesweep_kernel<false, uint64_t, false><<<tiles, SORT_THREADS, 0, stream>>>(
                kin, kout, vin, vout, n, specs.d[p], hist + p * RADIX, status + (size_t)p * tiles * RADIX, counters + p,
                nullptr, nullptr, nullptr);
        else
            onesweep_kernel<true, uint64_t, false><<<tiles, SORT_THREADS, 0, stream>>>(
                kin, kout, vin, vout, n, specs.d[p], hist + p * RADIX, status + (size_t)p * tiles * RADIX, counters + p,
                nullptr, nullptr, nullptr);
        uint64_t* tk = kin;
        kin = kout;
        kout = tk;
        uint32_t* tv = vin;
        vin = vout;
        vout = tv;
    }
    return passes & 1;
}

// Stable sort of (u32 key, u32 value) pairs on key bits [0, bits): the point-query binning (tile | sub-tile cell) and
// the depth pre-sort of the Gaussians.  8-byte pairs instead of 12: a pass moves a third less.
int launch_radix_sort32(uint32_t* keys_a, uint32_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, uint32_t n, int bits,
                        uint32_t* hist, uint32_t* status, size_t status_bytes, uint32_t* counters,
                        cudaStream_t stream, int match_passes, bool cleared) {
    if (n == 0 || bits <= 0) return 0;
    const int passes = (bits + RADIX_BITS - 1) / RADIX_BITS;
    PassSpecs specs;
    for (int p = 0; p < MAX_PASSES; p++) {
        DigitSpec d = {};
        const int q = p < passes ? p : 0;
        d.shift = RADIX_BITS * q;
        const int width = bits - d.shift < RADIX_BITS ? bits - d.shift : RADIX_BITS;
        d.mask = (1u << width) - 1u;
        d.vshift = -1;
        specs.d[p] = d;
    }
    const uint32_t tiles = (n + SORT_TILE - 1) / SORT_TILE;
    if (!cleared) {
        cudaMemsetAsync(hist, 0, MAX_PASSES * RADIX * 4, stream);
        cudaMemsetAsync(status, 0, status_bytes, stream);
        cudaMemsetAsync(counters, 0, MAX_PASSES * 4, stream);
    }
    const int hist_blocks = (int)min((uint32_t)GGS_HIST_BLOCKS, (n + 255) / 256);
    histogram_kernel<uint32_t><<<hist_blocks, 256, 0, stream>>>(keys_a, n, specs, passes, hist);
    offsets_kernel<<<1, RADIX, 0, stream>>>(hist, passes);
    uint32_t *kin = keys_a, *kout = keys_b, *vin = vals_a, *vout = vals_b;
    for (int p = 0; p < passes; p++) {
        // digits whose values are mixed across a warp rank with match.any (cheap there): the low digit of spatially
        // coherent keys, the three mantissa digits of depth keys; digits that come in runs rank with ballots (see
        // onesweep_kernel)
        if (p < match_passes)
            onesweep_kernel<true, uint32_t, false><<<tiles, SORT_THREADS, 0, stream>>>(
                kin, kout, vin, vout, n, specs.d[p], hist + p * RADIX, status + (size_t)p * tiles * RADIX, counters + p,
                nullptr, nullptr, nullptr);
        else
            onesweep_kernel<false, uint32_t, false><<<tiles, SORT_THREADS, 0, stream>>>(
                kin, kout, vin, vout, n, specs.d[p], hist + p * RADIX, status + (size_t)p * tiles * RADIX, counters + p,
                nullptr, nullptr, nullptr);
        uint32_t* tk = kin;
        kin = kout;
        kout = tk;
        uint32_t* tv = vin;
        vin = vout;
        vout = tv;
    }
    return passes & 1;
}

int tile_sort_passes(int tile_bits) { return (tile_bits + RADIX_BITS - 1) / RADIX_BITS; }

void launch_tile_sort(uint32_t* tk_a, uint32_t* tk_b, uint32_t* vals_a, uint32_t* vals_b, uint32_t n,
                      const uint32_t* n_dev, int tile_bits, const uint32_t* depth_bits, uint64_t* keys64_out,
                      uint32_t* hist, uint32_t* status, uint32_t* counters, cudaStream_t stream, bool cleared) {
    if (n == 0) return;
    const int passes = tile_sort_passes(tile_bits);
    const uint32_t tiles = (n + SORT_TILE - 1) / SORT_TILE;
    if (!cleared) {
        cudaMemsetAsync(status, 0, (size_t)passes * tiles * RADIX * 4, stream);
        cudaMemsetAsync(counters, 0, MAX_PASSES * 4, stream);
    }
    offsets_kernel<<<1, RADIX, 0, stream>>>(hist, passes);  // the histograms come from the key duplication
    // final values must land in vals_a: an even number of passes starts in *_a, an odd one in *_b
    uint32_t* kin = (passes & 1) ? tk_b : tk_a;
    uint32_t* kout = (passes & 1) ? tk_a : tk_b;
    uint32_t* vin = (passes & 1) ? vals_b : vals_a;
    uint32_t* vout = (passes & 1) ? vals_a : vals_b;
    for (int p = 0; p < passes; p++) {
        DigitSpec spec = {};
        spec.shift = RADIX_BITS * p;
        const int width = tile_bits - spec.shift < RADIX_BITS ? tile_bits - spec.shift : RADIX_BITS;
        spec.mask = (1u << width) - 1u;
        spec.vshift = -1;
        uint32_t* st = status + (size_t)p * tiles * RADIX;
        // tile digits of depth-ordered pairs come in runs: rank with ballots (see onesweep_kernel)
        if (p + 1 == passes && keys64_out)
            onesweep_kernel<false, uint32_t, true><<<tiles, SORT_THREADS, 0, stream>>>(
                kin, kout, vin, vout, n, spec, hist + p * RADIX, st, counters + p, keys64_out, depth_bits, n_dev);
        else
            onesweep_kernel<false, uint32_t, false><<<tiles, SORT_THREADS, 0, stream>>>(
                kin, kout, vin, vout, n, spec, hist + p * RADIX, st, counters + p, nullptr, nullptr, n_dev);
        uint32_t* tk = kin;
        kin = kout;
        kout = tk;
        uint32_t* tv = vin;
        vin = vout;
        vout = tv;
    }
}

}  // namespace ggs
