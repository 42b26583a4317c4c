// C ABI of the rasterizer (include/ggs_rasterizer.h): argument checking, workspace carving and the
// stage sequence.  Host orchestration follows CudaRasterizer::Rasterizer::forward / backward /
// evaluateTransmittance (DGR/cuda_rasterizer/rasterizer_impl.cu:285-448, :452-592, :594-815).
#include <algorithm>
#include <cstdio>
#include <string>

#include <nvtx3/nvToolsExt.h>  // header-only: the calls are no-ops unless a profiler injects its library

#include "../../include/ggs_rasterizer.h"
#include "ggs_kernels.cuh"

using namespace ggs;

namespace {

thread_local std::string g_error;
thread_local uint32_t* g_pinned = nullptr;
// the event behind the asynchronous read-back of the pair count, one per host thread and device
thread_local cudaEvent_t g_count_event[MAX_DEVICES] = {};

int fail(int code, const std::string& msg) {
    g_error = msg;
    return code;
}

#define GGS_CUDA(expr)                                                                           \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess)                                                                  \
            return fail(GGS_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));      \
    } while (0)

#define GGS_STAGE(name, debug, stream)                                                           \
    do {                                                                                         \
        cudaError_t e__ = cudaGetLastError();                                                    \
        if (e__ == cudaSuccess && (debug)) e__ = cudaStreamSynchronize(stream);                  \
        if (e__ != cudaSuccess)                                                                  \
            return fail(GGS_ERR_CUDA, std::string("stage ") + name + ": " + cudaGetErrorString(e__)); \
    } while (0)

// ---- optional per-stage device timing (ggs_profile_enable / ggs_profile_read) -------------------
enum { EV_F0, EV_PRE, EV_ORDER, EV_SCAN, EV_DUP0, EV_DUP, EV_SORT, EV_RANGES, EV_RENDER, EV_B0, EV_RBWD, EV_GBWD,
       EV_Q0, EV_QBIN, EV_QRUN, EV_COUNT };
struct Profiler {
    bool on = false;
    bool have_fwd = false, have_bwd = false, have_query = false;
    cudaEvent_t ev[EV_COUNT] = {};
    bool created = false;
    int sort_passes = 0;
};
Profiler g_prof;  // process-wide: autograd runs the backward on its own thread

// NVTX ranges around the enqueue of every stage, driven by the same marks as the event timer:
// a mark closes the range of the stage that ends there and opens the next one's.
const char* const kStageAfter[EV_COUNT] = {
    "ggs:preprocess", "ggs:depth_order", "ggs:scan", "ggs:pair_count+workspace", "ggs:duplicate_keys", "ggs:tile_sort",
    "ggs:tile_ranges", "ggs:render_forward", nullptr, "ggs:render_backward", "ggs:gaussian_backward", nullptr,
    "ggs:query_binning", "ggs:query_kernel", nullptr};
thread_local bool g_nvtx_open = false;

void prof_mark(cudaStream_t s, int which) {
    if (g_nvtx_open) nvtxRangePop();
    g_nvtx_open = kStageAfter[which] != nullptr;
    if (g_nvtx_open) nvtxRangePushA(kStageAfter[which]);
    if (!g_prof.on) return;
    if (!g_prof.created) {
        for (int i = 0; i < EV_COUNT; i++) cudaEventCreate(&g_prof.ev[i]);
        g_prof.created = true;
    }
    cudaEventRecord(g_prof.ev[which], s);
}

cudaEvent_t count_event() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return nullptr;
    if (!g_count_event[dev] &&
        cudaEventCreateWithFlags(&g_count_event[dev], cudaEventDisableTiming) != cudaSuccess)
        g_count_event[dev] = nullptr;
    return g_count_event[dev];
}

uint32_t* pinned_word() {
    if (!g_pinned) {
        if (cudaMallocHost(&g_pinned, 64) != cudaSuccess) g_pinned = nullptr;
    }
    return g_pinned;
}

// render CTAs take the tiles longest list first (TileLayout::order); -DGGS_TILE_ORDER=0 keeps row-major order
#ifndef GGS_TILE_ORDER
#define GGS_TILE_ORDER 1
#endif
constexpr bool kTileOrder = GGS_TILE_ORDER != 0;

int sort_end_bit(int tiles) { return 32 + (int)higher_msb((uint32_t)tiles); }
int sort_passes(int begin_bit, int end_bit) { return (end_bit - begin_bit + RADIX_BITS - 1) / RADIX_BITS; }

// geometry_only: the point-query paths (integrate / sample_depth / evaluate_sdf) ignore colour altogether; the
// reference's callers hand them the model's active SH / SG degrees without any colour tensors
// (gaussian_renderer/__init__.py:120-121), so the degrees are not validated there.
int make_frame(const GgsFrame* f, const GgsGaussians* g, FrameParams& fp, GaussianInputs& in,
               bool geometry_only = false) {
    if (!f || !g) return fail(GGS_ERR_INVALID_ARGUMENT, "null frame or gaussians");
    if (f->image_width <= 0 || f->image_height <= 0)
        return fail(GGS_ERR_INVALID_ARGUMENT, "image size must be positive");
    if (g->count < 0 || g->count > 0x7fffffffLL) return fail(GGS_ERR_INVALID_ARGUMENT, "bad Gaussian count");
    if (!f->viewmatrix || !f->projmatrix || !f->campos || (!f->background && !geometry_only))
        return fail(GGS_ERR_INVALID_ARGUMENT, "camera tensors must be device pointers");
    if (g->count > 0 && (!g->means3D || !g->opacities))
        return fail(GGS_ERR_INVALID_ARGUMENT, "means3D and opacities are required");
    if (g->count > 0 && (!g->scales || !g->rotations))
        return fail(GGS_ERR_UNSUPPORTED,
                    "scale/rotation pair required: precomputed 3D covariance is not supported (the reference "
                    "kernel cannot consume it either, render_forward.cu:344-345)");
    fp.P = (int)g->count;
    fp.W = f->image_width;
    fp.H = f->image_height;
    fp.gx = (fp.W + TILE_X - 1) / TILE_X;
    fp.gy = (fp.H + TILE_Y - 1) / TILE_Y;
    if (fp.gx >= 65536 || fp.gy >= 65536) return fail(GGS_ERR_INVALID_ARGUMENT, "image too large");
    fp.SHD = geometry_only ? 0 : f->sh_degree;
    fp.SHM = (g->shs && !geometry_only) ? g->sh_coeffs : 0;
    fp.SGD = geometry_only ? 0 : f->sg_degree;
    fp.SGM = (g->sg_color && !geometry_only) ? g->sg_lobes : 0;
    if (fp.SHD < 0 || fp.SHD > 3) return fail(GGS_ERR_INVALID_ARGUMENT, "sh_degree must be in 0..3");
    if (g->shs && (fp.SHD + 1) * (fp.SHD + 1) > fp.SHM)
        return fail(GGS_ERR_INVALID_ARGUMENT, "sh_coeffs smaller than (sh_degree+1)^2");
    if (fp.SGD < 0 || fp.SGD > fp.SGM) return fail(GGS_ERR_INVALID_ARGUMENT, "sg_degree exceeds allocated lobes");
    fp.tanfovx = f->tanfovx;
    fp.tanfovy = f->tanfovy;
    // rasterizer_impl.cu:318-319
    fp.focal_y = fp.H / (2.0f * fp.tanfovy);
    fp.focal_x = fp.W / (2.0f * fp.tanfovx);
    fp.kernel_size = f->kernel_size;
    fp.scale_modifier = f->scale_modifier;
    fp.view.view = f->viewmatrix;
    fp.view.proj = f->projmatrix;
    fp.view.campos = f->campos;
    fp.view.bg = f->background;
    in.means3D = g->means3D;
    in.opacities = g->opacities;
    in.scales = g->scales;
    in.rotations = g->rotations;
    in.shs = geometry_only ? nullptr : g->shs;
    in.sg_axis = fp.SGM ? g->sg_axis : nullptr;
    in.sg_sharpness = fp.SGM ? g->sg_sharpness : nullptr;
    in.sg_color = fp.SGM ? g->sg_color : nullptr;
    in.colors_precomp = geometry_only ? nullptr : g->colors_precomp;
    return GGS_OK;
}

template <typename T>
T* at(void* base, size_t off) {
    return reinterpret_cast<T*>(static_cast<char*>(base) + off);
}

// Bins the visible Gaussians of a frame: preprocess, depth order, scan, duplicate, sort, ranges.
// On return *D holds the number of (tile, Gaussian) pairs, *bin the binning workspace and *capacity the number of pairs
// that workspace is laid out for (every later use of the workspace derives its offsets from it).
//
// The pair count D is produced on the device by the scan.  The reference blocks the host on it before it can size the
// buffers and launch anything else (rasterizer_impl.cu:384).  capacity_hint == 0 does the same: capacity = D.  With a
// hint (what a caller saw on earlier frames, with head room) the workspace is sized from the hint, the D-dependent
// kernels are launched for the capacity and take the count from device memory, and the host only waits for the
// count -- copied to pinned memory right behind the scan -- AFTER it has queued duplicate / sort / ranges: the device
// never runs dry while the host launches.  If the count turns out to exceed the capacity the three stages are queued
// again on an exact-size workspace (nothing downstream has consumed the short one).
int bin_gaussians(const FrameParams& fp, const GaussianInputs& in, bool want_color, int* radii, void* geom,
                  uint2* ranges, GgsAllocFn alloc, void* user, void** bin, int64_t* D_out, int64_t capacity_hint,
                  int64_t* capacity_out, bool debug, cudaStream_t stream, void* also_clear = nullptr,
                  size_t also_clear_bytes = 0) {
    const GeomLayout GL = geom_layout(fp.P);
    const int tiles = fp.gx * fp.gy;
    prof_mark(stream, EV_F0);
    ClearList cl;
    {
        // everything the stages up to the tile ranges expect zeroed, cleared by the preprocess grid on its way (they
        // used to issue ten small memsets between their kernels): the scan's descriptors, ticket and total, the depth
        // pre-sort's histograms / look-back descriptors / tickets, the tile ranges, and whatever the caller adds
        // (the forward: max_contributor)
        cl.add(at<char>(geom, GL.scan_status), GL.order_keys_a - GL.scan_status);
        cl.add(at<char>(geom, GL.order_hist), GL.bytes - GL.order_hist);
        cl.add(ranges, (size_t)tiles * sizeof(uint2));
        if (also_clear) cl.add(also_clear, also_clear_bytes);
    }
    launch_preprocess(fp, in, want_color, at<SplatRecord>(geom, GL.records), at<float>(geom, GL.depths),
                      at<uint32_t>(geom, GL.tiles_touched), at<uint2>(geom, GL.rects),
                      at<uint8_t>(geom, GL.clamped), radii, at<uint32_t>(geom, GL.order_keys_a),
                      at<uint32_t>(geom, GL.order_vals_a), stream, &cl);
    GGS_STAGE("preprocess", debug, stream);
    prof_mark(stream, EV_PRE);
    // Depth pre-sort of the P Gaussians (4 passes over 8 B x P: 32-bit keys) so that the D pairs only need the tile
    // passes.  ORDER_PASSES is even: the order lands back in order_vals_a.
    launch_radix_sort32(at<uint32_t>(geom, GL.order_keys_a), at<uint32_t>(geom, GL.order_keys_b),
                        at<uint32_t>(geom, GL.order_vals_a), at<uint32_t>(geom, GL.order_vals_b), (uint32_t)fp.P, 32,
                        at<uint32_t>(geom, GL.order_hist), at<uint32_t>(geom, GL.order_status), GL.order_status_bytes,
                        at<uint32_t>(geom, GL.order_counters), stream, 3, true);
    GGS_STAGE("depth_order", debug, stream);
    prof_mark(stream, EV_ORDER);
    const uint32_t* order = at<uint32_t>(geom, GL.order_vals_a);
    uint32_t* const total = at<uint32_t>(geom, GL.total);
    launch_scan(at<uint32_t>(geom, GL.tiles_touched), at<uint32_t>(geom, GL.point_offsets), fp.P,
                at<uint64_t>(geom, GL.scan_status), at<uint32_t>(geom, GL.scan_counter), total, stream, order, true);
    GGS_STAGE("scan", debug, stream);
    prof_mark(stream, EV_SCAN);
    uint32_t* host = pinned_word();
    cudaEvent_t counted = count_event();
    if (!host || !counted) return fail(GGS_ERR_CUDA, "cudaMallocHost / cudaEventCreate failed");
    GGS_CUDA(cudaMemcpyAsync(host, total, 4, cudaMemcpyDeviceToHost, stream));
    GGS_CUDA(cudaEventRecord(counted, stream));

    const int end_bit = sort_end_bit(tiles);
    const int tile_bits = end_bit - 32;
    const int passes = tile_sort_passes(tile_bits);
    g_prof.sort_passes = passes;
    const bool odd = passes & 1;
    uint64_t cap = capacity_hint > 0 ? (uint64_t)std::min<int64_t>(capacity_hint, (1 << 30) - 1) : 0;
    bool counted_yet = false;
    uint32_t D = 0;
    for (bool first_round = true;; first_round = false) {
        if (cap == 0) {
            // no hint (or the hint was too small): the host waits where the reference does
            GGS_CUDA(cudaEventSynchronize(counted));
            D = host[0];
            counted_yet = true;
            if (D >= (1u << 30)) return fail(GGS_ERR_UNSUPPORTED, "more than 2^30 (tile, Gaussian) pairs");
            cap = D;
        }
        const BinLayout BL = bin_layout((size_t)cap, sort_passes(32, end_bit), tiles);
        void* b = alloc(GGS_SLOT_BINNING, BL.bytes, user);
        if (!b) return fail(GGS_ERR_ALLOC, "binning workspace allocation failed");
        *bin = b;
        // The pairs are emitted in depth order, so only the tile bits [32, end_bit) are left to sort (stable), and only
        // the 32-bit tile halves of the keys move through the passes: two u32[cap] ping-pong arrays inside keys_b.  The
        // sorted tile keys always land in the first of them and the sorted values in vals_a; the reference's 64-bit keys
        // (keys_a) are only formed on demand (ggs_materialize_sorted_keys).
        uint32_t* tk_a = at<uint32_t>(b, BL.keys_b);
        uint32_t* tk_b = tk_a + cap;
        uint32_t* v_a = at<uint32_t>(b, BL.vals_a);
        uint32_t* v_b = at<uint32_t>(b, BL.vals_b);
        const uint32_t* n_dev = counted_yet ? nullptr : total;
        {
            ClearList cl;  // digit histograms, look-back descriptors and tickets of the tile sort: adjacent
            cl.add(at<char>(b, BL.hist), BL.accept_mask - BL.hist);
            launch_clear(cl, stream);
        }
        prof_mark(stream, EV_DUP0);
        launch_duplicate_keys(fp.P, order, at<uint32_t>(geom, GL.point_offsets), at<uint2>(geom, GL.rects), fp.gx,
                              odd ? tk_b : tk_a, odd ? v_b : v_a, (uint32_t)cap, passes, at<uint32_t>(b, BL.hist),
                              stream, true);
        GGS_STAGE("duplicate_keys", debug, stream);
        prof_mark(stream, EV_DUP);
        launch_tile_sort(tk_a, tk_b, v_a, v_b, (uint32_t)cap, n_dev, tile_bits, nullptr, nullptr,
                         at<uint32_t>(b, BL.hist), at<uint32_t>(b, BL.status), at<uint32_t>(b, BL.counters), stream,
                         true);
        GGS_STAGE("radix_sort", debug, stream);
        prof_mark(stream, EV_SORT);
        // (the sorted tile keys land in tk_a; the ranges of a first, too short round are cleared again)
        launch_tile_ranges(tk_a, (uint32_t)cap, n_dev, ranges, tiles, stream, first_round);
        GGS_STAGE("tile_ranges", debug, stream);
        prof_mark(stream, EV_RANGES);
        if (counted_yet) break;
        GGS_CUDA(cudaEventSynchronize(counted));  // long since signalled: the device is busy with the three stages
        D = host[0];
        counted_yet = true;
        if (D >= (1u << 30)) return fail(GGS_ERR_UNSUPPORTED, "more than 2^30 (tile, Gaussian) pairs");
        if (D <= cap) break;
        cap = 0;  // the hint was too small: once more, exact
    }
    *D_out = D;
    *capacity_out = (int64_t)cap;
    return GGS_OK;
}

// pairs the binning workspace of a finished forward is laid out for (a state filled in by hand, capacity 0: exact)
int64_t state_capacity(const GgsForwardState* s) { return s->capacity > 0 ? s->capacity : s->num_rendered; }

const uint32_t* sorted_list(void* bin, int64_t D, int tiles) {
    const int passes = sort_passes(32, sort_end_bit(tiles));
    const BinLayout BL = bin_layout((size_t)D, passes, tiles);
    return at<uint32_t>(bin, BL.vals_a);  // bin_gaussians arranges the ping-pong so that the result lands here
}

uint32_t* accept_mask(void* bin, int64_t D, int tiles) {
    const BinLayout BL = bin_layout((size_t)D, sort_passes(32, sort_end_bit(tiles)), tiles);
    return at<uint32_t>(bin, BL.accept_mask);
}

uint32_t* near_mask(void* bin, int64_t D, int tiles) {
    const BinLayout BL = bin_layout((size_t)D, sort_passes(32, sort_end_bit(tiles)), tiles);
    return at<uint32_t>(bin, BL.near_mask);
}

}  // namespace

extern "C" {

int ggs_abi_version(void) { return GGS_ABI_VERSION; }

#ifdef GGS_WALK_STATS
}
namespace ggs { void read_walk_stats(unsigned long long* out, bool reset); }
extern "C" {
// development builds only (not in the header): [5 iterations][8 counters] of the forward's bisection walk
__attribute__((visibility("default"))) void ggs_debug_walk_stats(unsigned long long* out, int reset) {
    ggs::read_walk_stats(out, reset != 0);
}
#endif

const char* ggs_last_error(void) { return g_error.c_str(); }

int64_t ggs_rasterize_forward(const GgsFrame* frame, const GgsGaussians* gaussians, GgsAllocFn alloc, void* user,
                              const GgsForwardOutputs* out, GgsForwardState* state, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FrameParams fp;
    GaussianInputs in;
    int rc = make_frame(frame, gaussians, fp, in);
    if (rc) return rc;
    if (!alloc || !out || !state) return fail(GGS_ERR_INVALID_ARGUMENT, "null alloc/out/state");
    if (!out->color || !out->alpha || !out->normal || !out->mdepth || (fp.P > 0 && !out->radii))
        return fail(GGS_ERR_INVALID_ARGUMENT, "null output tensor");
    if (fp.P > 0 && ((in.shs == nullptr) == (in.colors_precomp == nullptr)))
        return fail(GGS_ERR_INVALID_ARGUMENT, "provide exactly one of shs / colors_precomp");
    const bool debug = frame->debug != 0;
    const size_t npix = (size_t)fp.W * fp.H;
    const int tiles = fp.gx * fp.gy;
    const int64_t capacity_hint = state->capacity;  // in: 0 = none (see include/ggs_rasterizer.h)
    state->geometry = state->binning = state->image = state->tile = nullptr;
    state->num_rendered = 0;
    state->capacity = 0;

    if (fp.P == 0) {
        // rasterize_points.cu:95: nothing is launched, the zero-filled outputs are returned
        GGS_CUDA(cudaMemsetAsync(out->color, 0, 3 * npix * 4, stream));
        GGS_CUDA(cudaMemsetAsync(out->alpha, 0, npix * 4, stream));
        GGS_CUDA(cudaMemsetAsync(out->normal, 0, 3 * npix * 4, stream));
        GGS_CUDA(cudaMemsetAsync(out->mdepth, 0, npix * 4, stream));
        return 0;
    }

    const GeomLayout GL = geom_layout(fp.P);
    const TileLayout TL = tile_layout(tiles);
    const ImgLayout IL = img_layout(npix);
    void* geom = alloc(GGS_SLOT_GEOMETRY, GL.bytes, user);
    void* tile = alloc(GGS_SLOT_TILE, TL.bytes, user);
    void* img = alloc(GGS_SLOT_IMAGE, IL.bytes, user);
    if (!geom || !tile || !img) return fail(GGS_ERR_ALLOC, "workspace allocation failed");
    state->geometry = geom;
    state->tile = tile;
    state->image = img;

    int64_t D = 0, cap = 0;
    void* bin = nullptr;
    rc = bin_gaussians(fp, in, true, out->radii, geom, at<uint2>(tile, TL.ranges), alloc, user, &bin, &D,
                       capacity_hint, &cap, debug, stream, at<char>(tile, TL.max_contributor),
                       (size_t)tiles * sizeof(uint32_t));
    if (rc) return rc;
    state->binning = bin;
    state->num_rendered = D;
    state->capacity = cap;

    if (!frame->require_depth) {
        GGS_CUDA(cudaMemsetAsync(out->normal, 0, 3 * npix * 4, stream));
        GGS_CUDA(cudaMemsetAsync(out->mdepth, 0, npix * 4, stream));
    }
    RenderOut ro;
    ro.color = out->color;
    ro.alpha = out->alpha;
    ro.normal = out->normal;
    ro.mdepth = out->mdepth;
    ro.n_contrib = at<uint32_t>(img, IL.n_contrib);
    ro.max_contributor = at<uint32_t>(tile, TL.max_contributor);
    ro.accept_mask = accept_mask(bin, cap, tiles);
    ro.warp_words = at<uint32_t>(tile, TL.warp_words);
    ro.near_mask = near_mask(bin, cap, tiles);
    ro.near_first = at<uint32_t>(tile, TL.near_first);
    ro.tile_order = nullptr;
    if (kTileOrder) {
        launch_tile_order(at<uint2>(tile, TL.ranges), nullptr, tiles, at<uint32_t>(tile, TL.order), stream);
        ro.tile_order = at<uint32_t>(tile, TL.order);
    }
    launch_render_forward(fp, at<uint2>(tile, TL.ranges), sorted_list(bin, cap, tiles),
                          at<SplatRecord>(geom, GL.records), ro, frame->require_depth != 0, stream,
                          /*cleared=*/true);
    GGS_STAGE("render_forward", debug, stream);
    prof_mark(stream, EV_RENDER);
    g_prof.have_fwd = g_prof.on;
    return D;
}

}  // extern "C"

namespace {

// side stream + events of the early colour-gradient push (one set per device, created on first use)
struct PushLane {
    cudaStream_t stream = nullptr;
    cudaEvent_t ready = nullptr, done = nullptr;
};
PushLane* push_lane() {
    static PushLane lanes[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    PushLane& L = lanes[dev];
    if (!L.stream) {
        if (cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        cudaEventCreateWithFlags(&L.ready, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming);
    }
    return &L;
}

int check_exchange(const GgsPeerExchange* px) {
    if (!px || !px->peer_buffers) return fail(GGS_ERR_INVALID_ARGUMENT, "null peer exchange");
    if (px->world <= 0 || px->world > MAX_VIEWS || px->rank < 0 || px->rank >= px->world)
        return fail(GGS_ERR_INVALID_ARGUMENT, "bad world / rank (at most 16 ranks)");
    if ((px->send_offset | px->recv_offset | px->reduce_floats | px->rgb_stage_offset | px->rgb_stage_stride |
         px->reduce_stage_offset | px->reduce_stage_stride | px->campos_stage_offset) & 3)
        return fail(GGS_ERR_INVALID_ARGUMENT, "exchange offsets must be multiples of 4 floats");
    for (int r = 0; r < px->world; r++)
        if (!px->peer_buffers[r] || (px->peer_buffers[r] & 15))
            return fail(GGS_ERR_INVALID_ARGUMENT, "bad peer buffer");
    if (px->chunks > 1) {
        if (px->chunks > GGS_MAX_CHUNKS || !px->pull_reduce || !px->multicast_base)
            return fail(GGS_ERR_INVALID_ARGUMENT, "chunked sums need pull_reduce, a multicast mapping and <= 8 chunks");
        int64_t prev = 0;
        for (int k = 0; k < px->chunks; k++) {
            if (!px->chunk_events[k]) return fail(GGS_ERR_INVALID_ARGUMENT, "null chunk event");
            if (k + 1 < px->chunks && (px->chunk_end[k] <= prev || (px->chunk_end[k] & 127)))
                return fail(GGS_ERR_INVALID_ARGUMENT, "chunk boundaries must increase in multiples of 128");
            prev = px->chunk_end[k];
        }
    }
    const size_t n4 = (size_t)px->reduce_floats / 4;
    const size_t longest = (n4 + px->world - 1) / px->world;
    if ((size_t)px->reduce_stage_stride / 4 < longest)
        return fail(GGS_ERR_INVALID_ARGUMENT, "reduce stage slots are smaller than a slice");
    return GGS_OK;
}
size_t slice_lo4(const GgsPeerExchange* px, int s) {
    return (size_t)px->reduce_floats / 4 * (size_t)s / (size_t)px->world;
}
float* peer_ptr(const GgsPeerExchange* px, int r, int64_t off) {
    return reinterpret_cast<float*>(px->peer_buffers[r]) + off;
}

int backward_impl(const GgsFrame* frame, const GgsGaussians* gaussians, const GgsForwardState* state,
                  const GgsBackwardInputs* bi, GgsAllocFn alloc, void* user, const GgsGradients* grads,
                  const GgsPeerExchange* px, void* stream_, void* side_stream_ = nullptr) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FrameParams fp;
    GaussianInputs in;
    int rc = make_frame(frame, gaussians, fp, in);
    if (rc) return rc;
    if (!state || !bi || !alloc || !grads) return fail(GGS_ERR_INVALID_ARGUMENT, "null state/inputs/alloc/grads");
    if (fp.P == 0) return GGS_OK;
    if (!state->geometry || !state->binning || !state->image || !state->tile)
        return fail(GGS_ERR_INVALID_ARGUMENT, "forward state is incomplete");
    if (!grads->dL_dmeans3D || !grads->dL_dmeans2D || !grads->dL_dcolors || !grads->dL_dopacity ||
        !grads->dL_dscales || !grads->dL_drotations)
        return fail(GGS_ERR_INVALID_ARGUMENT, "null gradient tensor");
    if (in.shs && !grads->dL_dsh && !grads->dL_drgb && !px)
        return fail(GGS_ERR_INVALID_ARGUMENT, "dL_dsh (or dL_drgb) required with shs");
    if (in.shs && fp.SGM && (!grads->dL_dsg_axis || !grads->dL_dsg_sharpness || !grads->dL_dsg_color))
        return fail(GGS_ERR_INVALID_ARGUMENT, "spherical-Gaussian gradient tensors required");
    const bool debug = frame->debug != 0;
    const size_t npix = (size_t)fp.W * fp.H;
    const int tiles = fp.gx * fp.gy;
    const GeomLayout GL = geom_layout(fp.P);
    const TileLayout TL = tile_layout(tiles);
    const ImgLayout IL = img_layout(npix);
    (void)IL;

    const size_t acc_bytes = (size_t)fp.P * GRAD_STRIDE * sizeof(double);
    double* acc = static_cast<double*>(alloc(GGS_SLOT_BACKWARD, acc_bytes, user));
    if (!acc) return fail(GGS_ERR_ALLOC, "backward workspace allocation failed");
    prof_mark(stream, EV_B0);
    GGS_CUDA(cudaMemsetAsync(acc, 0, acc_bytes, stream));

    RenderBwdIn ri;
    ri.alpha = bi->alpha;
    ri.normal = bi->normal;
    ri.mdepth = bi->mdepth;
    ri.n_contrib = at<uint32_t>(state->image, 0);
    ri.max_contributor = at<uint32_t>(state->tile, TL.max_contributor);
    ri.accept_mask = accept_mask(state->binning, state_capacity(state), tiles);
    ri.warp_words = at<uint32_t>(state->tile, TL.warp_words);
    ri.near_mask = near_mask(state->binning, state_capacity(state), tiles);
    ri.near_first = at<uint32_t>(state->tile, TL.near_first);
    ri.tile_order = kTileOrder ? at<uint32_t>(state->tile, TL.order) : nullptr;
    ri.dL_dcolor = bi->dL_dcolor;
    ri.dL_dmdepth = bi->dL_dmdepth;
    ri.dL_dalpha = bi->dL_dalpha;
    ri.dL_dnormal = bi->dL_dnormal;
    launch_render_backward(fp, at<uint2>(state->tile, TL.ranges), sorted_list(state->binning, state_capacity(state), tiles),
                           at<SplatRecord>(state->geometry, GL.records), ri, acc, frame->require_depth != 0, stream);
    GGS_STAGE("render_backward", debug, stream);
    prof_mark(stream, EV_RBWD);

    // fuse_rgb_push: the per-Gaussian backward stores dL/dRGB into every rank's colour stage itself (multicast); no
    // separate pass over the accumulators on a side stream
    const bool fused_push = px && px->fuse_rgb_push && px->multicast_base && in.shs;
    PushLane* lane = nullptr;
    if (px && !fused_push) {
        // the colour gradient of this view is final here: push it to every rank while the per-Gaussian backward runs
        lane = push_lane();
        if (!lane) return fail(GGS_ERR_CUDA, "could not create the push stream");
        PushTargets t = {};
        t.world = px->world;
        for (int r = 0; r < px->world; r++) {
            t.rgb[r] = peer_ptr(px, r, px->rgb_stage_offset + (int64_t)px->rank * px->rgb_stage_stride);
            t.campos[r] = peer_ptr(px, r, px->campos_stage_offset + (int64_t)px->rank * 4);
        }
        if (px->multicast_base)
            t.mc_rgb = reinterpret_cast<float*>(px->multicast_base) + px->rgb_stage_offset +
                       (int64_t)px->rank * px->rgb_stage_stride;
        // the caller's side stream if it gave one (it then chains the view-sum of the colour gradients behind the
        // push on that stream, concurrent with the per-Gaussian backward), else the library's own
        cudaStream_t side = side_stream_ ? static_cast<cudaStream_t>(side_stream_) : lane->stream;
        GGS_CUDA(cudaEventRecord(lane->ready, stream));
        GGS_CUDA(cudaStreamWaitEvent(side, lane->ready, 0));
        launch_push_rgb(fp.P, acc, at<uint8_t>(state->geometry, GL.clamped), bi->radii, fp.view.campos, t, side);
        GGS_CUDA(cudaEventRecord(lane->done, side));
    }

    GaussianGrads gg;
    gg.dL_dmeans3D = grads->dL_dmeans3D;
    gg.dL_dmeans2D = grads->dL_dmeans2D;
    gg.dL_dcolors = grads->dL_dcolors;
    gg.dL_dopacity = grads->dL_dopacity;
    gg.dL_dscales = grads->dL_dscales;
    gg.dL_drotations = grads->dL_drotations;
    gg.dL_dsh = in.shs ? grads->dL_dsh : nullptr;
    gg.dL_drgb = in.shs ? grads->dL_drgb : nullptr;
    gg.mc_rgb = gg.mc_campos = nullptr;
    const bool sg = in.shs && fp.SGM;
    gg.dL_dsg_axis = sg ? grads->dL_dsg_axis : nullptr;
    gg.dL_dsg_sharpness = sg ? grads->dL_dsg_sharpness : nullptr;
    gg.dL_dsg_color = sg ? grads->dL_dsg_color : nullptr;
    if (fused_push) {
        float* const mc = reinterpret_cast<float*>(px->multicast_base);
        gg.mc_rgb = mc + px->rgb_stage_offset + (int64_t)px->rank * px->rgb_stage_stride;
        gg.mc_campos = mc + px->campos_stage_offset + (int64_t)px->rank * 4;
    }
    if (px && px->chunks > 1) {
        // pieces of the model, an event of the caller behind each: the caller sums piece k over the ranks (node barrier +
        // ggs_peer_sum_ranges on another stream) while piece k + 1 is computed here
        int64_t first = 0;
        for (int k = 0; k < px->chunks; k++) {
            const int64_t last = k + 1 == px->chunks ? (int64_t)fp.P : std::min<int64_t>(px->chunk_end[k], fp.P);
            launch_gaussian_backward(fp, in, bi->radii, at<uint8_t>(state->geometry, GL.clamped), acc, gg, stream, first,
                                     last);
            GGS_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(px->chunk_events[k]), stream));
            first = last;
        }
    } else {
        launch_gaussian_backward(fp, in, bi->radii, at<uint8_t>(state->geometry, GL.clamped), acc, gg, stream);
    }
    GGS_STAGE("gaussian_backward", debug, stream);
    prof_mark(stream, EV_GBWD);
    g_prof.have_bwd = g_prof.on;
    if (px && px->pull_reduce) {
        if (lane) GGS_CUDA(cudaStreamWaitEvent(stream, lane->done, 0));
    } else if (px) {
        // reduce-scatter by stores: slice s of this rank's gradient region -> slot `rank` of rank s's reduce stage
        PushSlices ps = {};
        ps.world = px->world;
        for (int s = 0; s <= px->world; s++) ps.lo4[s] = slice_lo4(px, s);
        for (int s = 0; s < px->world; s++)
            ps.dest[s] = reinterpret_cast<float4*>(
                peer_ptr(px, s, px->reduce_stage_offset + (int64_t)px->rank * px->reduce_stage_stride));
        launch_push_slices(reinterpret_cast<const float4*>(peer_ptr(px, px->rank, px->send_offset)), ps, stream);
        if (lane) GGS_CUDA(cudaStreamWaitEvent(stream, lane->done, 0));
        GGS_STAGE("peer_push", debug, stream);
    }
    return GGS_OK;
}

}  // namespace

extern "C" {

int ggs_rasterize_backward(const GgsFrame* frame, const GgsGaussians* gaussians, const GgsForwardState* state,
                           const GgsBackwardInputs* bi, GgsAllocFn alloc, void* user, const GgsGradients* grads,
                           void* stream) {
    return backward_impl(frame, gaussians, state, bi, alloc, user, grads, nullptr, stream);
}

int ggs_rasterize_backward_peer(const GgsFrame* frame, const GgsGaussians* gaussians, const GgsForwardState* state,
                                const GgsBackwardInputs* bi, GgsAllocFn alloc, void* user, const GgsGradients* grads,
                                const GgsPeerExchange* px, void* stream, void* side_stream) {
    const int rc = check_exchange(px);
    if (rc) return rc;
    if (!grads || reinterpret_cast<const float*>(grads->dL_drotations) == nullptr)
        return fail(GGS_ERR_INVALID_ARGUMENT, "null gradient tensor");
    return backward_impl(frame, gaussians, state, bi, alloc, user, grads, px, stream, side_stream);
}

int ggs_backward_from_accumulators(const GgsFrame* frame, const GgsGaussians* gaussians, const GgsForwardState* state,
                                   const int32_t* radii, const double* accumulators, const GgsGradients* grads,
                                   void* stream_) {
    static_assert(GGS_ACCUMULATOR_STRIDE == GRAD_STRIDE && GGS_ACCUMULATOR_LIVE == GRAD_LIVE, "header out of sync");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FrameParams fp;
    GaussianInputs in;
    int rc = make_frame(frame, gaussians, fp, in);
    if (rc) return rc;
    if (fp.P == 0) return GGS_OK;
    if (!state || !state->geometry || !radii || !accumulators || !grads)
        return fail(GGS_ERR_INVALID_ARGUMENT, "null state/radii/accumulators/grads");
    if (!grads->dL_dmeans3D || !grads->dL_dmeans2D || !grads->dL_dcolors || !grads->dL_dopacity ||
        !grads->dL_dscales || !grads->dL_drotations || (in.shs && !grads->dL_dsh))
        return fail(GGS_ERR_INVALID_ARGUMENT, "null gradient tensor");
    const GeomLayout GL = geom_layout(fp.P);
    GaussianGrads gg;
    gg.dL_dmeans3D = grads->dL_dmeans3D;
    gg.dL_dmeans2D = grads->dL_dmeans2D;
    gg.dL_dcolors = grads->dL_dcolors;
    gg.dL_dopacity = grads->dL_dopacity;
    gg.dL_dscales = grads->dL_dscales;
    gg.dL_drotations = grads->dL_drotations;
    gg.dL_dsh = in.shs ? grads->dL_dsh : nullptr;
    gg.dL_drgb = in.shs ? grads->dL_drgb : nullptr;
    gg.mc_rgb = gg.mc_campos = nullptr;
    const bool sg = in.shs && fp.SGM;
    if (sg && (!grads->dL_dsg_axis || !grads->dL_dsg_sharpness || !grads->dL_dsg_color))
        return fail(GGS_ERR_INVALID_ARGUMENT, "spherical-Gaussian gradient tensors required");
    gg.dL_dsg_axis = sg ? grads->dL_dsg_axis : nullptr;
    gg.dL_dsg_sharpness = sg ? grads->dL_dsg_sharpness : nullptr;
    gg.dL_dsg_color = sg ? grads->dL_dsg_color : nullptr;
    launch_gaussian_backward(fp, in, radii, at<uint8_t>(state->geometry, GL.clamped), accumulators, gg, stream);
    GGS_STAGE("gaussian_backward", frame->debug != 0, stream);
    return GGS_OK;
}

int ggs_sh_gradient_from_views(int64_t count, int32_t sh_degree, int32_t sh_coeffs, int32_t views, const float* means3D,
                               const float* campos, const float* dL_drgb, float* dL_dsh, void* stream_) {
    if (count < 0 || count > 0x7fffffffLL || views <= 0 || views > MAX_VIEWS)
        return fail(GGS_ERR_INVALID_ARGUMENT, "bad count / views (at most 16 views)");
    if (sh_degree < 0 || sh_degree > 3 || (sh_degree + 1) * (sh_degree + 1) > sh_coeffs || sh_coeffs > 16)
        return fail(GGS_ERR_INVALID_ARGUMENT, "bad sh_degree / sh_coeffs");
    if (count == 0) return GGS_OK;
    if (!means3D || !campos || !dL_drgb || !dL_dsh) return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    launch_sh_gradient_from_views((int)count, sh_degree, sh_coeffs, views, means3D, campos, dL_drgb, dL_dsh, stream);
    GGS_STAGE("sh_gradient_from_views", false, stream);
    return GGS_OK;
}

int ggs_view_sum_peer(int64_t count, int32_t sh_degree, int32_t sh_coeffs, const float* means3D,
                      const GgsPeerExchange* px, float* dL_dsh, int32_t parts, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int rc = check_exchange(px);
    if (rc) return rc;
    if (count < 0 || count > 0x7fffffffLL) return fail(GGS_ERR_INVALID_ARGUMENT, "bad count");
    if (sh_degree < 0 || sh_degree > 3 || (sh_degree + 1) * (sh_degree + 1) > sh_coeffs || sh_coeffs > 16)
        return fail(GGS_ERR_INVALID_ARGUMENT, "bad sh_degree / sh_coeffs");
    if (count > 0 && (!means3D || !dL_dsh)) return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    const int world = px->world, rank = px->rank;
    ViewSources src = {};
    PeerJob job = {};
    job.world = world;
    job.lo4 = slice_lo4(px, rank);
    job.hi4 = slice_lo4(px, rank + 1);
    for (int r = 0; r < world; r++) {
        // everything this launch READS is local: slot r of this rank's colour / camera / reduce stages
        src.rgb[r] = peer_ptr(px, rank, px->rgb_stage_offset + (int64_t)r * px->rgb_stage_stride);
        src.campos[r] = peer_ptr(px, rank, px->campos_stage_offset + (int64_t)r * 4);
        // indexed by the flat position i of the slice: local slot base - lo4, or (pull) rank r's send region itself
        job.send[r] = px->pull_reduce
                          ? reinterpret_cast<const float4*>(peer_ptr(px, r, px->send_offset))
                          : reinterpret_cast<const float4*>(peer_ptr(
                                px, rank, px->reduce_stage_offset + (int64_t)r * px->reduce_stage_stride)) -
                                job.lo4;
        job.recv[r] = reinterpret_cast<float4*>(peer_ptr(px, r, px->recv_offset));
    }
    if (px->multicast_base) {
        job.mc_recv = reinterpret_cast<float4*>(reinterpret_cast<float*>(px->multicast_base) + px->recv_offset);
        if (px->pull_reduce)
            job.mc_send = reinterpret_cast<const float4*>(reinterpret_cast<float*>(px->multicast_base) + px->send_offset);
    }
    const size_t per_block = 256;
    const size_t want = (job.hi4 - job.lo4 + per_block - 1) / per_block;
    job.blocks = (int)(want < (size_t)(148 * 4) ? want : (size_t)(148 * 4));
    if (!(parts & GGS_VIEW_SUM_SLICE)) job.blocks = 0;
    launch_view_sum_peer((parts & GGS_VIEW_SUM_SH) ? (int)count : 0, sh_degree, sh_coeffs, means3D, src, dL_dsh, job,
                         stream);
    GGS_STAGE("view_sum_peer", false, stream);
    return GGS_OK;
}

int ggs_peer_sum_ranges(const GgsPeerExchange* px, int32_t count, const int64_t* lo4, const int64_t* hi4,
                        void* stream_) {
    const int rc = check_exchange(px);
    if (rc) return rc;
    if (!px->multicast_base) return fail(GGS_ERR_UNSUPPORTED, "ggs_peer_sum_ranges needs a multicast mapping");
    if (count < 0 || count > 8 || (count > 0 && (!lo4 || !hi4))) return fail(GGS_ERR_INVALID_ARGUMENT, "bad ranges");
    SumRanges r = {};
    r.count = count;
    const int64_t n4 = px->reduce_floats / 4;
    for (int k = 0; k < count; k++) {
        if (lo4[k] < 0 || hi4[k] < lo4[k] || hi4[k] > n4) return fail(GGS_ERR_INVALID_ARGUMENT, "range outside the region");
        r.lo4[k] = (size_t)lo4[k];
        r.hi4[k] = (size_t)hi4[k];
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    launch_sum_ranges_multimem(
        reinterpret_cast<const float4*>(reinterpret_cast<float*>(px->multicast_base) + px->send_offset),
        reinterpret_cast<float4*>(reinterpret_cast<float*>(px->multicast_base) + px->recv_offset), r, stream);
    GGS_STAGE("peer_sum_ranges", false, stream);
    return GGS_OK;
}

int ggs_activate_forward(int64_t count, const float* raw_scaling, const float* raw_opacity, const float* raw_rotation,
                         const float* filter_3D, float* scales, float* opacities, float* rotations, void* stream_) {
    if (count < 0 || count > 0x7fffffffLL) return fail(GGS_ERR_INVALID_ARGUMENT, "bad Gaussian count");
    if (count == 0) return GGS_OK;
    if (!raw_scaling || !raw_opacity || !raw_rotation || !filter_3D || !scales || !opacities || !rotations)
        return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    launch_activate_forward((int)count, raw_scaling, raw_opacity, raw_rotation, filter_3D, scales, opacities,
                            rotations, stream);
    GGS_STAGE("activate_forward", false, stream);
    return GGS_OK;
}

int ggs_activate_backward(int64_t count, const float* raw_scaling, const float* raw_opacity, const float* raw_rotation,
                          const float* filter_3D, const float* dL_dscales, const float* dL_dopacities,
                          const float* dL_drotations, float* dL_draw_scaling, float* dL_draw_opacity,
                          float* dL_draw_rotation, void* stream_) {
    if (count < 0 || count > 0x7fffffffLL) return fail(GGS_ERR_INVALID_ARGUMENT, "bad Gaussian count");
    if (count == 0) return GGS_OK;
    if (!raw_scaling || !raw_opacity || !raw_rotation || !filter_3D || !dL_dscales || !dL_dopacities ||
        !dL_drotations || !dL_draw_scaling || !dL_draw_opacity || !dL_draw_rotation)
        return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    launch_activate_backward((int)count, raw_scaling, raw_opacity, raw_rotation, filter_3D, dL_dscales, dL_dopacities,
                             dL_drotations, dL_draw_scaling, dL_draw_opacity, dL_draw_rotation, stream);
    GGS_STAGE("activate_backward", false, stream);
    return GGS_OK;
}

int ggs_profile_enable(int on) {
    g_prof.on = on != 0;
    g_prof.have_fwd = g_prof.have_bwd = g_prof.have_query = false;
    return GGS_OK;
}

int ggs_profile_read(GgsStageTimes* t) {
    if (!t) return fail(GGS_ERR_INVALID_ARGUMENT, "null stage-time struct");
    *t = GgsStageTimes{};
    auto span = [&](int a, int b, float* out) -> int {
        GGS_CUDA(cudaEventSynchronize(g_prof.ev[b]));
        GGS_CUDA(cudaEventElapsedTime(out, g_prof.ev[a], g_prof.ev[b]));
        return GGS_OK;
    };
    int rc = GGS_OK;
    if (g_prof.have_fwd) {
        if ((rc = span(EV_F0, EV_PRE, &t->preprocess_ms))) return rc;
        if ((rc = span(EV_PRE, EV_ORDER, &t->depth_order_ms))) return rc;
        if ((rc = span(EV_ORDER, EV_SCAN, &t->scan_ms))) return rc;
        if ((rc = span(EV_DUP0, EV_DUP, &t->duplicate_ms))) return rc;
        if ((rc = span(EV_DUP, EV_SORT, &t->sort_ms))) return rc;
        if ((rc = span(EV_SORT, EV_RANGES, &t->ranges_ms))) return rc;
        if ((rc = span(EV_RANGES, EV_RENDER, &t->render_forward_ms))) return rc;
        t->have_forward = 1;
    }
    t->sort_passes = g_prof.sort_passes;
    if (g_prof.have_bwd) {
        if ((rc = span(EV_B0, EV_RBWD, &t->render_backward_ms))) return rc;
        if ((rc = span(EV_RBWD, EV_GBWD, &t->gaussian_backward_ms))) return rc;
        t->have_backward = 1;
    }
    if (g_prof.have_query) {
        if ((rc = span(EV_Q0, EV_QBIN, &t->query_binning_ms))) return rc;
        if ((rc = span(EV_QBIN, EV_QRUN, &t->query_kernel_ms))) return rc;
        t->have_query = 1;
    }
    return GGS_OK;
}

int ggs_mark_visible(const float* means3D, int64_t count, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream_) {
    (void)projmatrix;
    if (count < 0 || count > 0x7fffffffLL) return fail(GGS_ERR_INVALID_ARGUMENT, "bad count");
    if (count == 0) return GGS_OK;
    if (!means3D || !viewmatrix || !present) return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    launch_mark_visible((int)count, means3D, viewmatrix, present, stream);
    GGS_STAGE("mark_visible", false, stream);
    return GGS_OK;
}

int ggs_count_blend_ops(const GgsFrame* frame, const GgsGaussians* gaussians, const GgsForwardState* state,
                        uint64_t* counters, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FrameParams fp;
    GaussianInputs in;
    int rc = make_frame(frame, gaussians, fp, in);
    if (rc) return rc;
    if (!state || !counters) return fail(GGS_ERR_INVALID_ARGUMENT, "null state / counters");
    if (fp.P == 0) {
        GGS_CUDA(cudaMemsetAsync(counters, 0, 32, stream));
        return GGS_OK;
    }
    if (!state->geometry || !state->binning || !state->tile) return fail(GGS_ERR_INVALID_ARGUMENT, "forward state is incomplete");
    const int tiles = fp.gx * fp.gy;
    const GeomLayout GL = geom_layout(fp.P);
    const TileLayout TL = tile_layout(tiles);
    launch_count_blend_ops(fp, at<uint2>(state->tile, TL.ranges), sorted_list(state->binning, state_capacity(state), tiles),
                           at<SplatRecord>(state->geometry, GL.records),
                           reinterpret_cast<unsigned long long*>(counters), stream);
    GGS_STAGE("count_blend_ops", false, stream);
    return GGS_OK;
}

int ggs_materialize_sorted_keys(int64_t P, int32_t width, int32_t height, const GgsForwardState* state, void* stream_) {
    if (!state || P < 0 || width <= 0 || height <= 0) return fail(GGS_ERR_INVALID_ARGUMENT, "bad arguments");
    if (state->num_rendered == 0) return GGS_OK;
    if (!state->geometry || !state->binning) return fail(GGS_ERR_INVALID_ARGUMENT, "forward state is incomplete");
    const int gx = (width + TILE_X - 1) / TILE_X, gy = (height + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;
    const GeomLayout GL = geom_layout((size_t)P);
    const BinLayout BL = bin_layout((size_t)state_capacity(state), sort_passes(32, sort_end_bit(tiles)), tiles);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    launch_emit_keys64(at<uint32_t>(state->binning, BL.keys_b), at<uint32_t>(state->binning, BL.vals_a),
                       at<uint32_t>(state->geometry, GL.depths), (uint32_t)state->num_rendered,
                       at<uint64_t>(state->binning, BL.keys_a), stream);
    GGS_STAGE("materialize_sorted_keys", false, stream);
    return GGS_OK;
}

int ggs_describe_layout(int64_t P, int32_t width, int32_t height, int64_t num_rendered, GgsLayout* L) {
    if (!L || P < 0 || width <= 0 || height <= 0 || num_rendered < 0)
        return fail(GGS_ERR_INVALID_ARGUMENT, "bad layout query");
    const int gx = (width + TILE_X - 1) / TILE_X, gy = (height + TILE_Y - 1) / TILE_Y;
    const int tiles = gx * gy;
    const GeomLayout GL = geom_layout((size_t)P);
    L->geometry_bytes = GL.bytes;
    L->records = GL.records;
    L->depths = GL.depths;
    L->tiles_touched = GL.tiles_touched;
    L->point_offsets = GL.point_offsets;
    L->clamped = GL.clamped;
    L->rects = GL.rects;
    L->depth_order = GL.order_vals_a;
    const int end_bit = sort_end_bit(tiles);
    const int passes = sort_passes(32, end_bit);
    const BinLayout BL = bin_layout((size_t)num_rendered, passes, tiles);
    L->binning_bytes = BL.bytes;
    const bool in_b = false;  // the ping-pong is arranged so that the sorted pairs land in keys_a / vals_a
    L->sorted_keys = in_b ? BL.keys_b : BL.keys_a;
    L->point_list = in_b ? BL.vals_b : BL.vals_a;
    L->unsorted_keys = in_b ? BL.keys_a : BL.keys_b;
    L->unsorted_list = in_b ? BL.vals_a : BL.vals_b;
    const TileLayout TL = tile_layout(tiles);
    L->tile_bytes = TL.bytes;
    L->ranges = TL.ranges;
    L->max_contributor = TL.max_contributor;
    const ImgLayout IL = img_layout((size_t)width * height);
    L->image_bytes = IL.bytes;
    L->n_contrib = IL.n_contrib;
    L->sort_bits = (uint32_t)end_bit;
    L->sort_passes = (uint32_t)passes;
    return GGS_OK;
}

size_t ggs_sort_scratch_bytes(int64_t n, int32_t begin_bit, int32_t end_bit) {
    if (n < 0 || end_bit <= begin_bit) return 0;
    const BinLayout BL = bin_layout((size_t)n, sort_passes(begin_bit, end_bit));
    return BL.bytes;
}

int ggs_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                   int64_t n, int32_t begin_bit, int32_t end_bit, void* scratch, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n < 0 || n >= (1LL << 30)) return fail(GGS_ERR_INVALID_ARGUMENT, "n out of range");
    if (begin_bit < 0 || end_bit > 64 || end_bit <= begin_bit || sort_passes(begin_bit, end_bit) > MAX_PASSES)
        return fail(GGS_ERR_INVALID_ARGUMENT, "bad bit range");
    if (n == 0) return GGS_OK;
    if (!keys_in || !vals_in || !keys_out || !vals_out || !scratch)
        return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    const int passes = sort_passes(begin_bit, end_bit);
    const BinLayout BL = bin_layout((size_t)n, passes);
    GGS_CUDA(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, stream));
    GGS_CUDA(cudaMemcpyAsync(vals_out, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
    const int where = launch_radix_sort(keys_out, at<uint64_t>(scratch, BL.keys_b), vals_out,
                                        at<uint32_t>(scratch, BL.vals_b), (uint32_t)n, plain_sort_plan(begin_bit, end_bit),
                                        at<uint32_t>(scratch, BL.hist), at<uint32_t>(scratch, BL.status),
                                        BL.status_bytes, at<uint32_t>(scratch, BL.counters), stream);
    if (where) {
        GGS_CUDA(cudaMemcpyAsync(keys_out, at<uint64_t>(scratch, BL.keys_b), (size_t)n * 8, cudaMemcpyDeviceToDevice,
                                 stream));
        GGS_CUDA(cudaMemcpyAsync(vals_out, at<uint32_t>(scratch, BL.vals_b), (size_t)n * 4, cudaMemcpyDeviceToDevice,
                                 stream));
    }
    GGS_STAGE("sort_pairs", false, stream);
    return GGS_OK;
}

}  // extern "C"

// ---- ggs_sort_pairs_static: the stand-alone sort without any allocation ----------------------------------------
// No counterpart in the reference: for callers that cannot call
// cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it.  The scratch that
// ggs_sort_pairs takes from its caller lives here in a file-scope device array sized for GGS_SORT_STATIC_MAX_PAIRS
// pairs over the full 64 key bits; every GPU that loads the library holds it (about 3.7 MiB) until the process exits.
namespace {

constexpr size_t static_up(size_t x) { return (x + 255) / 256 * 256; }
constexpr size_t STATIC_SORT_PAIRS = GGS_SORT_STATIC_MAX_PAIRS;
constexpr size_t STATIC_SORT_TILES = (STATIC_SORT_PAIRS + SORT_TILE - 1) / SORT_TILE;
// the second buffer of the key / value ping-pong, the digit histograms, the look-back descriptors of every pass, and
// the tile tickets -- bin_layout's fields without its first buffers (the caller's outputs play that part)
constexpr size_t STATIC_SORT_KEYS = 0;
constexpr size_t STATIC_SORT_VALS = STATIC_SORT_KEYS + static_up(STATIC_SORT_PAIRS * 8);
constexpr size_t STATIC_SORT_HIST = STATIC_SORT_VALS + static_up(STATIC_SORT_PAIRS * 4);
constexpr size_t STATIC_SORT_STATUS = STATIC_SORT_HIST + static_up((size_t)MAX_PASSES * RADIX * 4);
constexpr size_t STATIC_SORT_COUNTERS = STATIC_SORT_STATUS + static_up((size_t)MAX_PASSES * STATIC_SORT_TILES * RADIX * 4);
constexpr size_t STATIC_SORT_BYTES = STATIC_SORT_COUNTERS + static_up((size_t)MAX_PASSES * 4);

static __device__ __align__(256) unsigned char g_static_sort_scratch[STATIC_SORT_BYTES];

// In host code the array's name is a host-side stand-in, not a device address: the device address is asked for once
// per device, with that device current, and kept.  Only written under the lock; a device's slot never changes again.
void* g_static_sort_ptr[MAX_DEVICES] = {};
volatile int g_static_sort_lock = 0;

int static_sort_scratch(unsigned char** out) {
    int dev = -1;
    GGS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return fail(GGS_ERR_INVALID_ARGUMENT, "device index out of range");
    while (__sync_lock_test_and_set(&g_static_sort_lock, 1)) {
    }
    void* p = g_static_sort_ptr[dev];
    cudaError_t e = cudaSuccess;
    if (!p) {
        e = cudaGetSymbolAddress(&p, g_static_sort_scratch);
        if (e == cudaSuccess) g_static_sort_ptr[dev] = p;
    }
    __sync_lock_release(&g_static_sort_lock);
    if (e != cudaSuccess || !p)
        return fail(GGS_ERR_CUDA, std::string("cudaGetSymbolAddress(static sort scratch): ") + cudaGetErrorString(e));
    *out = static_cast<unsigned char*>(p);
    return GGS_OK;
}

}  // namespace

extern "C" int ggs_sort_pairs_static(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                                     uint32_t* vals_out, int64_t n, int32_t begin_bit, int32_t end_bit,
                                     void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (n < 0) return fail(GGS_ERR_INVALID_ARGUMENT, "n out of range");
    if (n > (int64_t)STATIC_SORT_PAIRS)
        return fail(GGS_ERR_INVALID_ARGUMENT,
                    "ggs_sort_pairs_static: n = " + std::to_string(n) + " exceeds GGS_SORT_STATIC_MAX_PAIRS = " +
                        std::to_string(STATIC_SORT_PAIRS) + " (use ggs_sort_pairs with caller scratch)");
    if (begin_bit < 0 || end_bit > 64 || end_bit <= begin_bit || sort_passes(begin_bit, end_bit) > MAX_PASSES)
        return fail(GGS_ERR_INVALID_ARGUMENT, "bad bit range");
    if (n == 0) return GGS_OK;
    if (!keys_in || !vals_in || !keys_out || !vals_out) return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    unsigned char* scratch = nullptr;
    const int rc = static_sort_scratch(&scratch);
    if (rc) return rc;
    const int passes = sort_passes(begin_bit, end_bit);
    const size_t tiles = ((size_t)n + SORT_TILE - 1) / SORT_TILE;
    const size_t status_bytes = (size_t)passes * tiles * RADIX * 4;  // <= the array's share: n and passes are bounded
    uint64_t* keys_b = reinterpret_cast<uint64_t*>(scratch + STATIC_SORT_KEYS);
    uint32_t* vals_b = reinterpret_cast<uint32_t*>(scratch + STATIC_SORT_VALS);
    GGS_CUDA(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * 8, cudaMemcpyDeviceToDevice, stream));
    GGS_CUDA(cudaMemcpyAsync(vals_out, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
    // the array keeps the last call's contents: launch_radix_sort clears the histograms, the descriptors and the
    // tickets on `stream` before its first kernel; the ping-pong buffers are fully written before they are read
    const int where = launch_radix_sort(keys_out, keys_b, vals_out, vals_b, (uint32_t)n,
                                        plain_sort_plan(begin_bit, end_bit),
                                        reinterpret_cast<uint32_t*>(scratch + STATIC_SORT_HIST),
                                        reinterpret_cast<uint32_t*>(scratch + STATIC_SORT_STATUS), status_bytes,
                                        reinterpret_cast<uint32_t*>(scratch + STATIC_SORT_COUNTERS), stream);
    if (where) {
        GGS_CUDA(cudaMemcpyAsync(keys_out, keys_b, (size_t)n * 8, cudaMemcpyDeviceToDevice, stream));
        GGS_CUDA(cudaMemcpyAsync(vals_out, vals_b, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
    }
    GGS_STAGE("sort_pairs_static", false, stream);
    return GGS_OK;
}

// ---- point queries: integrate / sample_depth / evaluate_sdf ----------------------------------------------------
// The reference bins the Gaussians of the view inside every call (rasterizer_impl.cu:637-714, :861-936, :1084-1160),
// i.e. once per (chunk of points, view) in mesh_extract_tetrahedra.py:70-82.  Here that half is its own entry point
// (ggs_view_prepare) whose result any number of query calls can share; the one-shot entry points run both halves.
namespace {

struct ViewTileLayout {  // GGS_SLOT_TILE of a prepared view
    size_t granges, radii, bytes;
};
ViewTileLayout view_tile_layout(int tiles, int P) {
    ViewTileLayout L;
    L.granges = 0;
    L.radii = align_up((size_t)tiles * 8, 256);
    L.bytes = L.radii + align_up((size_t)P * 4, 256);
    return L;
}

int prepare_view(const GgsFrame* frame, const GgsGaussians* gaussians, GgsAllocFn alloc, void* user, GgsViewState* view,
                 FrameParams& fp, cudaStream_t stream) {
    GaussianInputs in;
    int rc = make_frame(frame, gaussians, fp, in, /*geometry_only=*/true);
    if (rc) return rc;
    if (!view || !alloc) return fail(GGS_ERR_INVALID_ARGUMENT, "null view state / alloc");
    *view = GgsViewState{};
    view->count = fp.P;
    if (fp.P == 0) return GGS_OK;
    const int tiles = fp.gx * fp.gy;
    const GeomLayout GL = geom_layout(fp.P);
    const ViewTileLayout TL = view_tile_layout(tiles, fp.P);
    void* geom = alloc(GGS_SLOT_GEOMETRY, GL.bytes, user);
    void* tile = alloc(GGS_SLOT_TILE, TL.bytes, user);
    if (!geom || !tile) return fail(GGS_ERR_ALLOC, "workspace allocation failed");
    int64_t D = 0, cap = 0;
    void* bin = nullptr;
    // no capacity hint: a prepared view is binned once and queried many times, its workspace is exact (capacity == D)
    rc = bin_gaussians(fp, in, false, at<int>(tile, TL.radii), geom, at<uint2>(tile, TL.granges), alloc, user, &bin, &D,
                       0, &cap, frame->debug != 0, stream);
    if (rc) return rc;
    view->geometry = geom;
    view->binning = bin;
    view->tile = tile;
    view->num_rendered = D;
    return GGS_OK;
}

// Camera half of a frame for a query against a prepared view (the Gaussians are not needed again).
int query_frame(const GgsFrame* f, const GgsViewState* view, FrameParams& fp) {
    if (!f || !view) return fail(GGS_ERR_INVALID_ARGUMENT, "null frame or view state");
    GgsGaussians g = {};
    g.count = 0;
    GaussianInputs in;
    const int rc = make_frame(f, &g, fp, in, true);
    if (rc) return rc;
    fp.P = (int)view->count;
    if (view->count > 0 && (!view->geometry || !view->binning || !view->tile))
        return fail(GGS_ERR_INVALID_ARGUMENT, "view state is incomplete");
    return GGS_OK;
}

enum QueryKind { Q_INTEGRATE, Q_SAMPLE, Q_SDF };

// Bins the query points and runs one forward kernel.  `ws_out`: the query workspace (GGS_SLOT_POINT).
int run_query(QueryKind kind, const GgsFrame* frame, const GgsViewState* view, const float* points3D,
              int64_t num_points, GgsAllocFn alloc, void* user, float* out_a, float* out_b, uint8_t* inside,
              void** ws_out, cudaStream_t stream, uint64_t* op_counters = nullptr) {
    FrameParams fp;
    int rc = query_frame(frame, view, fp);
    if (rc) return rc;
    if (num_points < 0 || num_points > 0x7fffffffLL) return fail(GGS_ERR_INVALID_ARGUMENT, "bad point count");
    if (ws_out) *ws_out = nullptr;
    if (fp.P == 0 || num_points == 0) return GGS_OK;  // rasterize_points.cu:335, :528: nothing is launched
    if (!points3D || !out_a || !inside || !alloc || (kind == Q_SDF && !out_b))
        return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    const bool debug = frame->debug != 0;
    const int PN = (int)num_points;
    const int tiles = fp.gx * fp.gy;
    if (tiles >= (1 << (32 - QUERY_CELL_BITS)) - 1) return fail(GGS_ERR_UNSUPPORTED, "image too large for point queries");
    const QueryLayout QL = query_layout((size_t)PN, tiles, kind == Q_SAMPLE);
    void* ws = alloc(GGS_SLOT_POINT, QL.bytes, user);
    if (!ws) return fail(GGS_ERR_ALLOC, "query workspace allocation failed");
    if (ws_out) *ws_out = ws;
    prof_mark(stream, EV_Q0);
    launch_query_binning(fp, PN, points3D, ws, QL, stream);
    GGS_STAGE("query_binning", debug, stream);
    prof_mark(stream, EV_QBIN);
    const GeomLayout GL = geom_layout(fp.P);
    const ViewTileLayout TL = view_tile_layout(tiles, fp.P);
    const uint2* granges = at<uint2>(view->tile, TL.granges);
    const uint32_t* glist = sorted_list(view->binning, view->num_rendered, tiles);
    const SplatRecord* records = at<SplatRecord>(view->geometry, GL.records);
    if (kind == Q_INTEGRATE)
        launch_integrate(QL, ws, granges, glist, records, out_a, inside,
                         reinterpret_cast<unsigned long long*>(op_counters), stream);
    else if (kind == Q_SAMPLE)
        launch_sample_depth_forward(fp, QL, ws, granges, glist, records, out_a, inside, stream);
    else
        launch_evaluate_sdf(fp, QL, ws, granges, glist, records, out_a, out_b, inside, stream);
    GGS_STAGE("query", debug, stream);
    prof_mark(stream, EV_QRUN);
    g_prof.have_query = g_prof.on;
    return GGS_OK;
}

}  // namespace

extern "C" {

int64_t ggs_view_prepare(const GgsFrame* frame, const GgsGaussians* gaussians, GgsAllocFn alloc, void* user,
                         GgsViewState* view, void* stream) {
    FrameParams fp;
    const int rc = prepare_view(frame, gaussians, alloc, user, view, fp, static_cast<cudaStream_t>(stream));
    return rc ? rc : view->num_rendered;
}

int ggs_integrate_view(const GgsFrame* frame, const GgsViewState* view, const float* points3D, int64_t num_points,
                       GgsAllocFn alloc, void* user, float* out_transmittance, uint8_t* inside, void* stream) {
    return run_query(Q_INTEGRATE, frame, view, points3D, num_points, alloc, user, out_transmittance, nullptr, inside,
                     nullptr, static_cast<cudaStream_t>(stream));
}

int ggs_integrate_view_counted(const GgsFrame* frame, const GgsViewState* view, const float* points3D,
                               int64_t num_points, GgsAllocFn alloc, void* user, float* out_transmittance,
                               uint8_t* inside, uint64_t* op_counters, void* stream) {
    if (!op_counters) return fail(GGS_ERR_INVALID_ARGUMENT, "null counters");
    return run_query(Q_INTEGRATE, frame, view, points3D, num_points, alloc, user, out_transmittance, nullptr, inside,
                     nullptr, static_cast<cudaStream_t>(stream), op_counters);
}

int ggs_evaluate_sdf_view(const GgsFrame* frame, const GgsViewState* view, const float* points3D, int64_t num_points,
                          GgsAllocFn alloc, void* user, float* out_depth, float* out_sdf, uint8_t* inside,
                          void* stream) {
    return run_query(Q_SDF, frame, view, points3D, num_points, alloc, user, out_depth, out_sdf, inside, nullptr,
                     static_cast<cudaStream_t>(stream));
}

int64_t ggs_integrate_points(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                             int64_t num_points, GgsAllocFn alloc, void* user, float* out_transmittance,
                             uint8_t* inside, void* stream) {
    if (num_points == 0 || (gaussians && gaussians->count == 0)) return 0;  // rasterize_points.cu:335
    GgsViewState view;
    FrameParams fp;
    int rc = prepare_view(frame, gaussians, alloc, user, &view, fp, static_cast<cudaStream_t>(stream));
    if (rc) return rc;
    rc = ggs_integrate_view(frame, &view, points3D, num_points, alloc, user, out_transmittance, inside, stream);
    return rc ? rc : view.num_rendered;
}

int64_t ggs_evaluate_sdf(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                         int64_t num_points, GgsAllocFn alloc, void* user, float* out_depth, float* out_sdf,
                         uint8_t* inside, void* stream) {
    if (num_points == 0 || (gaussians && gaussians->count == 0)) return 0;
    GgsViewState view;
    FrameParams fp;
    int rc = prepare_view(frame, gaussians, alloc, user, &view, fp, static_cast<cudaStream_t>(stream));
    if (rc) return rc;
    rc = ggs_evaluate_sdf_view(frame, &view, points3D, num_points, alloc, user, out_depth, out_sdf, inside, stream);
    return rc ? rc : view.num_rendered;
}

int64_t ggs_sample_depth_forward(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                                 int64_t num_points, GgsAllocFn alloc, void* user, float* out_points,
                                 uint8_t* inside, GgsSampleState* state, void* stream) {
    if (!state) return fail(GGS_ERR_INVALID_ARGUMENT, "null state");
    *state = GgsSampleState{};
    if (num_points == 0 || (gaussians && gaussians->count == 0)) return 0;  // rasterize_points.cu:528
    GgsViewState view;
    FrameParams fp;
    int rc = prepare_view(frame, gaussians, alloc, user, &view, fp, static_cast<cudaStream_t>(stream));
    if (rc) return rc;
    void* ws = nullptr;
    rc = run_query(Q_SAMPLE, frame, &view, points3D, num_points, alloc, user, out_points, nullptr, inside, &ws,
                   static_cast<cudaStream_t>(stream));
    if (rc) return rc;
    state->geometry = view.geometry;
    state->binning = view.binning;
    state->tile = view.tile;
    state->point = ws;
    state->num_rendered = view.num_rendered;
    state->num_evaluated = num_points;
    return view.num_rendered;
}

int ggs_sample_depth_backward(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                              int64_t num_points, const GgsSampleState* state, const uint8_t* inside,
                              const float* dL_dout_points, GgsAllocFn alloc, void* user, const GgsGradients* grads,
                              float* dL_dpoints3D, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    FrameParams fp;
    GaussianInputs in;
    int rc = make_frame(frame, gaussians, fp, in, /*geometry_only=*/true);
    if (rc) return rc;
    if (num_points < 0 || num_points > 0x7fffffffLL) return fail(GGS_ERR_INVALID_ARGUMENT, "bad point count");
    if (!grads || !state || !alloc) return fail(GGS_ERR_INVALID_ARGUMENT, "null state/grads/alloc");
    const int PN = (int)num_points;
    const size_t P = (size_t)fp.P;
    if (fp.P > 0) {
        if (!grads->dL_dmeans3D || !grads->dL_dmeans2D || !grads->dL_dcolors || !grads->dL_dopacity ||
            !grads->dL_dscales || !grads->dL_drotations)
            return fail(GGS_ERR_INVALID_ARGUMENT, "null gradient tensor");
    }
    if (PN > 0 && !dL_dpoints3D) return fail(GGS_ERR_INVALID_ARGUMENT, "null dL_dpoints3D");
    if (fp.P == 0 || PN == 0) {
        // rasterize_points.cu:600: nothing is launched, the zero-initialised gradients are returned
        if (fp.P > 0) {
            GGS_CUDA(cudaMemsetAsync(grads->dL_dmeans3D, 0, P * 12, stream));
            GGS_CUDA(cudaMemsetAsync(grads->dL_dmeans2D, 0, P * 12, stream));
            GGS_CUDA(cudaMemsetAsync(grads->dL_dcolors, 0, P * 12, stream));
            GGS_CUDA(cudaMemsetAsync(grads->dL_dopacity, 0, P * 4, stream));
            GGS_CUDA(cudaMemsetAsync(grads->dL_dscales, 0, P * 12, stream));
            GGS_CUDA(cudaMemsetAsync(grads->dL_drotations, 0, P * 16, stream));
        }
        if (PN > 0) GGS_CUDA(cudaMemsetAsync(dL_dpoints3D, 0, (size_t)PN * 12, stream));
        return GGS_OK;
    }
    if (!points3D || !inside || !dL_dout_points) return fail(GGS_ERR_INVALID_ARGUMENT, "null argument");
    if (!state->geometry || !state->binning || !state->point || !state->tile)
        return fail(GGS_ERR_INVALID_ARGUMENT, "forward state is incomplete");
    const bool debug = frame->debug != 0;
    const int tiles = fp.gx * fp.gy;
    const GeomLayout GL = geom_layout(fp.P);
    const ViewTileLayout TL = view_tile_layout(tiles, fp.P);
    const QueryLayout QL = query_layout((size_t)PN, tiles, true);

    // scratch: accumulator rows (double) + dL_dpoints2D
    const size_t acc_bytes = align_up(P * GRAD_STRIDE * sizeof(double), 256);
    const size_t p2d_bytes = align_up((size_t)PN * 8, 256);
    char* scratch = static_cast<char*>(alloc(GGS_SLOT_BACKWARD, acc_bytes + p2d_bytes, user));
    if (!scratch) return fail(GGS_ERR_ALLOC, "backward workspace allocation failed");
    double* acc = reinterpret_cast<double*>(scratch);
    float2* dp2d = reinterpret_cast<float2*>(scratch + acc_bytes);
    GGS_CUDA(cudaMemsetAsync(scratch, 0, acc_bytes + p2d_bytes, stream));

    launch_sample_depth_backward(fp, QL, state->point, at<uint2>(state->tile, TL.granges),
                                 sorted_list(state->binning, state->num_rendered, tiles),
                                 at<SplatRecord>(state->geometry, GL.records), inside, dL_dout_points, dp2d, acc, stream);
    GGS_STAGE("sample_depth_backward", debug, stream);

    GaussianGrads gg{};
    gg.dL_dmeans3D = grads->dL_dmeans3D;
    gg.dL_dmeans2D = grads->dL_dmeans2D;
    gg.dL_dcolors = grads->dL_dcolors;
    gg.dL_dopacity = grads->dL_dopacity;
    gg.dL_dscales = grads->dL_dscales;
    gg.dL_drotations = grads->dL_drotations;
    launch_gaussian_backward(fp, in, at<int>(state->tile, TL.radii), at<uint8_t>(state->geometry, GL.clamped), acc, gg,
                             stream);
    GGS_STAGE("gaussian_backward", debug, stream);
    launch_points_backward(fp, PN, points3D, at<float2>(state->point, QL.xy), dp2d, dL_dpoints3D, stream);
    GGS_STAGE("points_backward", debug, stream);
    return GGS_OK;
}

}  // extern "C"
