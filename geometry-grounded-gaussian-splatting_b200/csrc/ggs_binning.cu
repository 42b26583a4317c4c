// Stage 2 of the forward: prefix sum of tile counts, key duplication, tile ranges.
// Hand-written replacements for the reference's cub::DeviceScan::InclusiveSum
// (rasterizer_impl.cu:380), duplicateWithKeys (:70-107), createWithKeys (:109-137) and
// identifyTileRanges (:142-161).  The radix sort lives in ggs_sort.cu.
#include "ggs_kernels.cuh"

namespace ggs {

// ------------------------------------------------------------------------------------------------
// Single-pass inclusive scan with decoupled look-back.  Descriptor: bits 63..62 = state
// (0 invalid, 1 aggregate, 2 inclusive prefix), low 32 bits = value.
// ------------------------------------------------------------------------------------------------
constexpr uint64_t ST_AGG = 1ull << 62;
constexpr uint64_t ST_PREFIX = 2ull << 62;

__device__ __forceinline__ uint64_t ld_relaxed(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <bool GATHER>
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const uint32_t* __restrict__ in,
                                                            const uint32_t* __restrict__ order,
                                                            uint32_t* __restrict__ out, int n,
                                                            uint64_t* __restrict__ status,
                                                            uint32_t* __restrict__ counter,
                                                            uint32_t* __restrict__ total) {
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_warp[SCAN_THREADS / 32];
    __shared__ uint32_t s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(counter, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;

    uint32_t v[SCAN_ITEMS];
    const uint32_t* src = GATHER ? order : in;
    if (base + SCAN_ITEMS <= n) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(src + base));
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(src + base) + 1);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
        v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < n) ? __ldg(src + base + i) : 0u;
    }
    if constexpr (GATHER) {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++) v[i] = (base + i < n) ? __ldg(in + v[i]) : 0u;
    }
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        sum += v[i];
        v[i] = sum;
    }
    // warp inclusive scan of the per-thread sums
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++)
        if (w < warp) warp_off += s_warp[w];
    const uint32_t thread_excl = warp_off + incl - sum;

    // look-back by the last warp: 32 predecessors per round (one per lane), consumed in order up to the first
    // inclusive prefix; a serial walk costs ~sqrt(2 x tiles) dependent L2 round trips when a wave starts together
    if (warp == SCAN_THREADS / 32 - 1) {
        const uint32_t aggregate = __shfl_sync(0xffffffffu, warp_off + incl, 31);
        uint32_t prefix = 0;
        if (tile == 0) {
            if (lane == 31) st_relaxed(status, ST_PREFIX | aggregate);
        } else {
            if (lane == 31) st_relaxed(status + tile, ST_AGG | aggregate);
            int look = (int)tile - 1;
            while (true) {
                const int at = look - lane;
                const uint64_t d = at >= 0 ? ld_relaxed(status + at) : ST_PREFIX;  // before tile 0: prefix 0
                const uint32_t is_prefix = __ballot_sync(0xffffffffu, (d & ST_PREFIX) != 0);
                const uint32_t is_invalid = __ballot_sync(0xffffffffu, (d & (ST_PREFIX | ST_AGG)) == 0);
                const int first_prefix = is_prefix ? __ffs(is_prefix) - 1 : 32;
                const int first_invalid = is_invalid ? __ffs(is_invalid) - 1 : 32;
                // usable lanes: [0, take) plus the prefix lane itself when it comes before any invalid word
                const bool done = first_prefix < first_invalid;
                const int take = done ? first_prefix + 1 : first_invalid;
                prefix += __reduce_add_sync(0xffffffffu, lane < take ? (uint32_t)d : 0u);
                if (done) break;
                look -= take;
            }
            if (lane == 31) st_relaxed(status + tile, ST_PREFIX | (uint64_t)(prefix + aggregate));
        }
        if (lane == 31) {
            s_prefix = prefix;
            if ((int)(tile + 1) * SCAN_TILE >= n) *total = prefix + aggregate;
        }
    }
    __syncthreads();
    const uint32_t off = s_prefix + thread_excl;
    if (base + SCAN_ITEMS <= n) {
        uint4 a = {v[0] + off, v[1] + off, v[2] + off, v[3] + off};
        uint4 b = {v[4] + off, v[5] + off, v[6] + off, v[7] + off};
        reinterpret_cast<uint4*>(out + base)[0] = a;
        reinterpret_cast<uint4*>(out + base)[1] = b;
    } else {
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++)
            if (base + i < n) out[base + i] = v[i] + off;
    }
}

void launch_scan(const uint32_t* in, uint32_t* out, int n, uint64_t* status, uint32_t* counter, uint32_t* total,
                 cudaStream_t stream, const uint32_t* order, bool cleared) {
    const int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (!cleared) {
        cudaMemsetAsync(status, 0, (size_t)(tiles + 1) * 8, stream);
        cudaMemsetAsync(counter, 0, 4, stream);
        cudaMemsetAsync(total, 0, 4, stream);
    }
    if (n == 0) return;
    if (order)
        scan_kernel<true><<<tiles, SCAN_THREADS, 0, stream>>>(in, order, out, n, status, counter, total);
    else
        scan_kernel<false><<<tiles, SCAN_THREADS, 0, stream>>>(in, order, out, n, status, counter, total);
}

// ------------------------------------------------------------------------------------------------
// Key duplication, output-parallel: each block owns 256 consecutive Gaussians OF THE DEPTH ORDER and walks the
// contiguous slice of the key array they expand into; every output slot finds its Gaussian by
// a binary search over the block's 257 offsets held in shared memory.  Writes are fully
// coalesced and perfectly balanced (the reference loops over a splat's tiles in one thread).
// Reference key = (tile << 32) | f32 bits of depth, value = Gaussian index; slot order inside one
// Gaussian is y-major then x (rasterizer_impl.cu:96-104).  Only the 32-bit tile half is materialised here: the
// last pass of the tile sort writes the full 64-bit keys (launch_tile_sort).
// The pairs leave this kernel ordered by (depth, Gaussian index, tile); the stable pair sort on the tile bits alone
// then yields the reference's order (tile, depth, Gaussian index) -- SortPairs over bits [0, 32 + tile bits) of keys
// emitted in Gaussian order (rasterizer_impl.cu:406-412) -- without ever moving the pairs for the depth bits.
// ------------------------------------------------------------------------------------------------
constexpr int DUP_THREADS = 256;

__global__ void __launch_bounds__(DUP_THREADS) duplicate_keys_kernel(int P, const uint32_t* __restrict__ order,
                                                                     const uint32_t* __restrict__ offsets,
                                                                     const uint2* __restrict__ rects, int gx,
                                                                     uint32_t* __restrict__ tile_keys,
                                                                     uint32_t* __restrict__ vals, uint32_t capacity,
                                                                     int passes, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_off[DUP_THREADS + 1];
    __shared__ uint32_t s_g[DUP_THREADS];
    __shared__ uint2 s_rect[DUP_THREADS];
    __shared__ uint32_t s_hist[4][RADIX];  // digit histograms of this block's keys (tile ids < 2^32: <= 4 digits)
    const int g0 = blockIdx.x * DUP_THREADS;
    const int tid = threadIdx.x;
    const int gi = g0 + tid;
    for (int i = tid; i < passes * RADIX; i += DUP_THREADS) (&s_hist[0][0])[i] = 0;
    if (tid == 0) s_off[0] = (g0 == 0) ? 0u : __ldg(offsets + g0 - 1);
    s_off[tid + 1] = (gi < P) ? __ldg(offsets + gi) : 0xffffffffu;
    if (gi < P) {
        const uint32_t g = __ldg(order + gi);
        s_g[tid] = g;
        s_rect[tid] = __ldg(rects + g);  // culled Gaussians own no slots: their (unwritten) rects are never used
    }
    __syncthreads();
    const uint32_t begin = s_off[0];
    const int last = min(DUP_THREADS, P - g0);
    // `capacity`: slots of the key arrays.  It is the exact pair count unless the workspace was sized from a hint
    // before the count was known; a count beyond it is detected by the host, which bins again (ggs_api.cu)
    const uint32_t end = min(s_off[last], capacity);
    // Each thread owns FOUR consecutive slots, aligned to 16 bytes in the output (a warp covers 128 slots, one 512-byte
    // run per array): one binary search for the first of them, the others are found by walking on -- the slots of
    // a Gaussian are consecutive, y-major then x, so the tile coordinates advance without a division.  Quads that
    // straddle the block's slice are written slot by slot (the rest of them belongs to the neighbouring blocks).
    const uint32_t q_begin = begin / 4, q_end = (end + 3) / 4;  // quads that touch [begin, end)
    // (the second tile-key array of the ping-pong starts `capacity` words behind the first: not always 16-byte aligned)
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(tile_keys) | reinterpret_cast<uintptr_t>(vals)) & 15) == 0;
    for (uint32_t qb = (tid & ~31u); q_begin + qb < q_end; qb += DUP_THREADS) {  // whole warps: the votes need every lane
        const uint32_t o0 = (q_begin + qb + (tid & 31u)) * 4;
        uint32_t tile[4], val[4];
        bool live[4];
        // largest j in [0,last) with s_off[j] <= o, for the first live slot of the quad
        const uint32_t of = max(o0, begin);
        int lo = 0;
        uint32_t tx = 0, ty = 0, minx = 0, maxx = 1, next_off = 0;
        const bool any_live = of < end && of < o0 + 4;
        if (any_live) {
            int hi = last;
#pragma unroll
            for (int st = 0; st < 9; st++) {
                const int mid = (lo + hi) >> 1;
                if (lo + 1 < hi) {
                    if (s_off[mid] <= of)
                        lo = mid;
                    else
                        hi = mid;
                }
            }
            // Gaussians without slots share their offset with the next one: s_off[j] <= of < s_off[j + 1] holds for
            // the LAST j with s_off[j] <= of, which the search returns
            const uint2 rc = s_rect[lo];
            minx = rc.x & 0xffffu;
            maxx = rc.y & 0xffffu;
            const uint32_t w = maxx - minx;
            const uint32_t local = of - s_off[lo];
            const uint32_t row = local / w;
            ty = (rc.x >> 16) + row;
            tx = minx + (local - row * w);
            next_off = s_off[lo + 1];
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t o = o0 + k;
            live[k] = o >= begin && o < end;
            tile[k] = 0;
            val[k] = 0;
            if (live[k]) {
                if (o >= next_off) {
                    // on to the next Gaussian that owns slots (those in between own none)
                    do {
                        lo++;
                    } while (s_off[lo + 1] <= o);
                    const uint2 rc = s_rect[lo];
                    minx = rc.x & 0xffffu;
                    maxx = rc.y & 0xffffu;
                    ty = rc.x >> 16;
                    tx = minx;
                    next_off = s_off[lo + 1];
                }
                tile[k] = ty * (uint32_t)gx + tx;
                val[k] = s_g[lo];
                if (++tx == maxx) {
                    tx = minx;
                    ty++;
                }
            }
        }
        if (live[0] && live[3] && vec_ok) {
            *reinterpret_cast<uint4*>(tile_keys + o0) = uint4{tile[0], tile[1], tile[2], tile[3]};
            *reinterpret_cast<uint4*>(vals + o0) = uint4{val[0], val[1], val[2], val[3]};
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (live[k]) {
                    tile_keys[o0 + k] = tile[k];
                    vals[o0 + k] = val[k];
                }
            }
        }
        // digit histograms: the low digit of neighbouring slots differs (consecutive tiles), the higher ones are
        // mostly warp-uniform -- one add for the warp then, per-lane atomics otherwise
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (live[k]) atomicAdd(&s_hist[0][tile[k] & (RADIX - 1)], 1u);
            for (int p = 1; p < passes; p++) {
                const uint32_t d = (tile[k] >> (RADIX_BITS * p)) & (RADIX - 1);
                const uint32_t lives = __ballot_sync(0xffffffffu, live[k]);
                if (lives == 0) break;
                const int leader = __ffs(lives) - 1;
                const uint32_t d0 = __shfl_sync(0xffffffffu, d, leader);
                if (__all_sync(0xffffffffu, !live[k] || d == d0)) {
                    if ((tid & 31) == leader) atomicAdd(&s_hist[p][d0], (uint32_t)__popc(lives));
                } else if (live[k]) {
                    atomicAdd(&s_hist[p][d], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < passes * RADIX; i += DUP_THREADS) {
        const uint32_t c = (&s_hist[0][0])[i];
        if (c) atomicAdd(hist + i, c);
    }
}

void launch_duplicate_keys(int P, const uint32_t* order, const uint32_t* offsets, const uint2* rects, int gx,
                           uint32_t* tile_keys, uint32_t* vals, uint32_t capacity, int passes, uint32_t* hist,
                           cudaStream_t stream, bool cleared) {
    if (!cleared) cudaMemsetAsync(hist, 0, MAX_PASSES * RADIX * 4, stream);
    if (P == 0 || capacity == 0) return;
    duplicate_keys_kernel<<<(P + DUP_THREADS - 1) / DUP_THREADS, DUP_THREADS, 0, stream>>>(
        P, order, offsets, rects, gx, tile_keys, vals, capacity, passes > 4 ? 4 : passes, hist);
}

// ------------------------------------------------------------------------------------------------
// Tile ranges from the sorted tile keys (rasterizer_impl.cu:142-161); untouched tiles stay (0,0).  The keys are the
// 32-bit tile halves the sort moved: the reference's 64-bit keys (tile << 32 | depth bits) are not materialised on the
// hot path -- nothing downstream reads their depth half -- and are formed on demand (emit_keys64_kernel below).
// ------------------------------------------------------------------------------------------------
// Eight keys per thread (two 16-byte loads + the predecessor).
constexpr int RANGE_ITEMS = 8;
__global__ void __launch_bounds__(256) tile_ranges_kernel(const uint32_t* __restrict__ keys, uint32_t n,
                                                          const uint32_t* __restrict__ n_dev,
                                                          uint2* __restrict__ ranges) {
    if (n_dev) n = min(n, __ldg(n_dev));  // launched for the capacity, the count lives on the device
    const uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) * RANGE_ITEMS;
    if (base >= n) return;
    uint32_t t[RANGE_ITEMS + 1];
    if (base + RANGE_ITEMS <= n) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(keys + base));
        const uint4 b = __ldg(reinterpret_cast<const uint4*>(keys + base) + 1);
        t[1] = a.x; t[2] = a.y; t[3] = a.z; t[4] = a.w;
        t[5] = b.x; t[6] = b.y; t[7] = b.z; t[8] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < RANGE_ITEMS; i++) t[i + 1] = base + i < n ? __ldg(keys + base + i) : 0u;
    }
    t[0] = base ? __ldg(keys + base - 1) : 0u;
#pragma unroll
    for (int i = 0; i < RANGE_ITEMS; i++) {
        const uint32_t idx = base + i;
        if (idx >= n) break;
        const uint32_t cur = t[i + 1];
        if (idx == 0) {
            ranges[cur].x = 0;
        } else if (cur != t[i]) {
            ranges[t[i]].y = idx;
            ranges[cur].x = idx;
        }
        if (idx == n - 1) ranges[cur].y = n;
    }
}

void launch_tile_ranges(const uint32_t* sorted_tile_keys, uint32_t n, const uint32_t* n_dev, uint2* ranges, int tiles,
                        cudaStream_t stream, bool cleared) {
    if (!cleared) cudaMemsetAsync(ranges, 0, (size_t)tiles * sizeof(uint2), stream);
    if (n == 0) return;
    const uint32_t per_block = 256 * RANGE_ITEMS;
    tile_ranges_kernel<<<(n + per_block - 1) / per_block, 256, 0, stream>>>(sorted_tile_keys, n, n_dev, ranges);
}

// Longest-processing-time-first order of the tiles for the render kernels.  A tile's cost follows its list length; the
// hardware hands CTAs out in grid order, so with row-major tiles whichever long lists happen to sit in the last rows
// run alone at the end.  One block, one read of the lengths (kept in registers): 1024 cost classes on a log scale
// (exponent and five bits below the leading one, i.e. 3 % steps), descending, counting sort.  The order inside a
// class is whatever the shared-memory atomics give -- no result depends on it.
constexpr int ORDER_THREADS = 1024;
constexpr int ORDER_KEEP = 8;  // costs per thread kept in registers: tiles beyond 8192 are read again
__device__ __forceinline__ uint32_t cost_class(uint32_t c) {
    const int e = 31 - __clz(c | 1u);
    const uint32_t m = e >= 5 ? (c >> (e - 5)) & 31u : (c << (5 - e)) & 31u;
    return (uint32_t)(ORDER_THREADS - 1) - (((uint32_t)e << 5) | m);
}
__global__ void __launch_bounds__(ORDER_THREADS) tile_order_kernel(const uint2* __restrict__ ranges,
                                                                   const uint32_t* __restrict__ lengths, int tiles,
                                                                   uint32_t* __restrict__ order) {
    __shared__ uint32_t s_count[ORDER_THREADS];
    __shared__ uint32_t s_warp[ORDER_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // cost of tile t: its list length, or lengths[t] where the caller knows better
    auto cost = [&](int t) {
        if (lengths) return __ldg(lengths + t);
        const uint2 r = __ldg(ranges + t);
        return r.y - r.x;
    };
    s_count[tid] = 0;
    uint32_t klass[ORDER_KEEP];
#pragma unroll
    for (int k = 0; k < ORDER_KEEP; k++) {
        const int t = tid + k * ORDER_THREADS;
        klass[k] = t < tiles ? cost_class(cost(t)) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ORDER_KEEP; k++)
        if (tid + k * ORDER_THREADS < tiles) atomicAdd(&s_count[klass[k]], 1u);
    for (int t = tid + ORDER_KEEP * ORDER_THREADS; t < tiles; t += ORDER_THREADS) atomicAdd(&s_count[cost_class(cost(t))], 1u);
    __syncthreads();
    // exclusive scan of the class counts: thread t owns class t
    const uint32_t mine = s_count[tid];
    uint32_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = s_warp[lane];
        uint32_t wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= d) wi += up;
        }
        s_warp[lane] = wi - w;
    }
    __syncthreads();
    s_count[tid] = s_warp[warp] + incl - mine;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ORDER_KEEP; k++) {
        const int t = tid + k * ORDER_THREADS;
        if (t < tiles) order[atomicAdd(&s_count[klass[k]], 1u)] = (uint32_t)t;
    }
    for (int t = tid + ORDER_KEEP * ORDER_THREADS; t < tiles; t += ORDER_THREADS)
        order[atomicAdd(&s_count[cost_class(cost(t))], 1u)] = (uint32_t)t;
}

// One launch instead of a handful of small memsets (ggs_kernels.cuh: ClearList), for callers without a grid of
// their own to do it on the way.
__global__ void __launch_bounds__(256) clear_kernel(const ClearList list) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (int s = 0; s < list.count; s++) {
        uint4* const p = static_cast<uint4*>(list.p[s]);
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < list.n16[s]; i += stride) p[i] = uint4{0, 0, 0, 0};
    }
}

void launch_clear(const ClearList& list, cudaStream_t stream) {
    uint32_t most = 0;
    for (int s = 0; s < list.count; s++) most = max(most, list.n16[s]);
    if (most == 0) return;
    const int blocks = (int)min((most + 255u) / 256u, 148u * 8u);
    clear_kernel<<<blocks, 256, 0, stream>>>(list);
}

void launch_tile_order(const uint2* ranges, const uint32_t* lengths, int tiles, uint32_t* order, cudaStream_t stream) {
    if (tiles <= 0) return;
    tile_order_kernel<<<1, ORDER_THREADS, 0, stream>>>(ranges, lengths, tiles, order);
}

// The reference's sorted 64-bit keys, for inspection and the parity tests: (tile << 32) | depth bits of the pair's Gaussian.
__global__ void __launch_bounds__(256) emit_keys64_kernel(const uint32_t* __restrict__ tile_keys,
                                                          const uint32_t* __restrict__ vals,
                                                          const uint32_t* __restrict__ depth_bits, uint32_t n,
                                                          uint64_t* __restrict__ keys64) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keys64[i] = ((uint64_t)__ldg(tile_keys + i) << 32) | (uint64_t)__ldg(depth_bits + __ldg(vals + i));
}

void launch_emit_keys64(const uint32_t* tile_keys, const uint32_t* vals, const uint32_t* depth_bits, uint32_t n,
                        uint64_t* keys64, cudaStream_t stream) {
    if (n == 0) return;
    emit_keys64_kernel<<<(n + 255) / 256, 256, 0, stream>>>(tile_keys, vals, depth_bits, n, keys64);
}

}  // namespace ggs
