// Launch entry points of the individual stages (internal C++ interface; the public boundary is
// the C ABI in include/ggs_rasterizer.h).
#pragma once
#include "ggs_common.cuh"

namespace ggs {

struct FrameParams {
    int P, W, H;
    int gx, gy;  // tile grid
    int SHD, SHM, SGD, SGM;
    float tanfovx, tanfovy, focal_x, focal_y;
    float kernel_size, scale_modifier;
    View view;  // device pointers
};

struct GaussianInputs {
    const float* means3D;         // [P,3]
    const float* opacities;       // [P]
    const float* scales;          // [P,3]
    const float* rotations;       // [P,4] (r,x,y,z)
    const float* shs;             // [P,SHM,3] or null
    const float* sg_axis;         // [P,SGM,3] or null
    const float* sg_sharpness;    // [P,SGM] or null
    const float* sg_color;        // [P,SGM,3] or null
    const float* colors_precomp;  // [P,3] or null
};

// ---- stage 1: per-Gaussian preprocess + scan -------------------------------------------------
// The stages of the binning clear their own counters / histograms / look-back descriptors with small memsets, each of
// which costs the stream a drain and a refill.  A caller that knows the whole sequence lists the spans (16-byte
// aligned, up to six) and has them cleared in one go -- by the preprocess grid on its way (nothing in that kernel
// touches them), or by launch_clear -- and passes cleared = true to the stages.
struct ClearList {
    void* p[6];
    uint32_t n16[6];  // 16-byte units
    int count = 0;
    void add(void* ptr, size_t bytes) {
        p[count] = ptr;
        n16[count] = (uint32_t)((bytes + 15) / 16);
        count++;
    }
};
void launch_clear(const ClearList& list, cudaStream_t stream);

void launch_preprocess(const FrameParams& fp, const GaussianInputs& in, bool want_color, SplatRecord* records,
                       float* depths, uint32_t* tiles_touched, uint2* rects, uint8_t* clamped, int* radii,
                       uint32_t* order_keys, uint32_t* order_vals, cudaStream_t stream,
                       const ClearList* clear = nullptr);

// order != null: scans in[order[i]] (the tile counts taken in depth order)
void launch_scan(const uint32_t* in, uint32_t* out, int n, uint64_t* status, uint32_t* counter, uint32_t* total,
                 cudaStream_t stream, const uint32_t* order = nullptr, bool cleared = false);

// ---- stage 2: key duplication, radix sort, tile ranges ----------------------------------------
// order[i] = Gaussian at depth rank i; offsets = inclusive scan of the tile counts in that order.  Emits 32-bit TILE
// keys (the depth half of the reference's 64-bit key is implied by the emission order) and accumulates the 8-bit
// digit histograms of the `passes` tile-sort passes into hist[passes][256] (zeroed here).
// `capacity`: slots of the key arrays -- the pair count, or more when the workspace was sized from a hint before the
// count was known (slots beyond the count stay unwritten; pairs beyond the capacity are dropped, see bin_gaussians).
void launch_duplicate_keys(int P, const uint32_t* order, const uint32_t* offsets, const uint2* rects, int gx,
                           uint32_t* tile_keys, uint32_t* vals, uint32_t capacity, int passes, uint32_t* hist,
                           cudaStream_t stream, bool cleared = false);

// Digit plan of the radix sort.  Plain: 8-bit digits of bits [begin_bit, end_bit).  Compact (rasterizer keys
// tile<<32 | depth bits only): the top byte of a positive float depth takes very few distinct values (sign + 7
// exponent bits: 0.2 .. 100 spans five), so after three plain passes over depth bits 0..23 the remaining order is
// that of the narrow virtual key  (tile << nb) | (depth_top_byte - emin)  with nb <= 4 bits: two more passes
// instead of three.  Same stable order as sorting bits [0, 32 + tile_bits) of the real key; keys are not modified.
struct SortPlan {
    int begin_bit, end_bit;
    int passes;
    bool compact;
    uint32_t emin;  // smallest depth top byte present
    int nb;         // bits that hold (top byte - emin)
    int tile_bits;  // end_bit - 32
};
SortPlan plain_sort_plan(int begin_bit, int end_bit);
// depth_min_bits / depth_max_bits: extremes of the 32 low key bits over all keys (any superset range is fine)
SortPlan rasterizer_sort_plan(int end_bit, uint32_t depth_min_bits, uint32_t depth_max_bits);

// Sorts (keys_in, vals_in); both buffer pairs are clobbered.  Returns 0 if the result is in *_a, 1 if in *_b, when
// the input was in *_a (callers that want the result in a fixed buffer put the input in *_b when plan.passes is odd).
int launch_radix_sort(uint64_t* keys_a, uint64_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, uint32_t n,
                      const SortPlan& plan, uint32_t* hist, uint32_t* status, size_t status_bytes, uint32_t* counters,
                      cudaStream_t stream);

// Stable sort of depth-ordered pairs on their 32-bit tile keys (bits [0, tile_bits)); the histograms are those of
// launch_duplicate_keys.  Input in (tk_a, vals_a) when tile_sort_passes() is even, in (tk_b, vals_b) when odd; the
// sorted keys and values land in (tk_a, vals_a).  With keys64_out the last pass writes the reference's 64-bit keys
// (tile << 32 | depth bits of the pair's Gaussian) there instead of the 32-bit ones (not used on the hot path).
// Stable sort of (u32 key, u32 value) pairs on key bits [0, bits).  Returns 0 if the result is in *_a, 1 if in *_b.
int launch_radix_sort32(uint32_t* keys_a, uint32_t* keys_b, uint32_t* vals_a, uint32_t* vals_b, uint32_t n, int bits,
                        uint32_t* hist, uint32_t* status, size_t status_bytes, uint32_t* counters,
                        cudaStream_t stream, int match_passes = 1, bool cleared = false);
int tile_sort_passes(int tile_bits);
// n_dev (may be null): the pair count in device memory; the launch is then sized for n = the capacity and the kernels
// clamp to min(n, *n_dev).
void launch_tile_sort(uint32_t* tk_a, uint32_t* tk_b, uint32_t* vals_a, uint32_t* vals_b, uint32_t n,
                      const uint32_t* n_dev, int tile_bits, const uint32_t* depth_bits, uint64_t* keys64_out,
                      uint32_t* hist, uint32_t* status, uint32_t* counters, cudaStream_t stream,
                      bool cleared = false);

void launch_tile_ranges(const uint32_t* sorted_tile_keys, uint32_t n, const uint32_t* n_dev, uint2* ranges, int tiles,
                        cudaStream_t stream, bool cleared = false);
// keys64[i] = (tile_keys[i] << 32) | depth_bits[vals[i]]: the reference's sorted keys, on demand (tests, inspection)
void launch_emit_keys64(const uint32_t* tile_keys, const uint32_t* vals, const uint32_t* depth_bits, uint32_t n,
                        uint64_t* keys64, cudaStream_t stream);

// ---- stage 3: tile blend --------------------------------------------------------------------
struct RenderOut {
    float* color;   // [3,H,W]
    float* alpha;   // [1,H,W]
    float* normal;  // [3,H,W]
    float* mdepth;  // [1,H,W]
    uint32_t* n_contrib;
    uint32_t* max_contributor;
    uint32_t* accept_mask;  // see BinLayout::accept_mask
    uint32_t* warp_words;   // see TileLayout::warp_words
    uint32_t* near_mask;    // see BinLayout::near_mask
    uint32_t* near_first;   // see TileLayout::near_first
    const uint32_t* tile_order;  // see TileLayout::order; nullptr = CTAs take the tiles in row-major order
};
// Tile ids by descending list length (counting sort on 1024 length classes, one block): the render kernels hand their
// CTAs out in this order, longest lists first, so that the last wave of the grid is made of the shortest tiles.
// Cost of a tile = its list length, or lengths[t] when `lengths` is given.
void launch_tile_order(const uint2* ranges, const uint32_t* lengths, int tiles, uint32_t* order, cudaStream_t stream);

void launch_render_forward(const FrameParams& fp, const uint2* ranges, const uint32_t* point_list,
                           const SplatRecord* records, const RenderOut& out, bool require_depth,
                           cudaStream_t stream, bool cleared = false);

// ---- backward --------------------------------------------------------------------------------
struct RenderBwdIn {
    const float* alpha;
    const float* normal;
    const float* mdepth;
    const uint32_t* n_contrib;
    const uint32_t* max_contributor;
    const uint32_t* accept_mask;
    const uint32_t* warp_words;
    const uint32_t* near_mask;
    const uint32_t* near_first;
    const uint32_t* tile_order;  // see TileLayout::order; nullptr = row-major
    const float* dL_dcolor;
    const float* dL_dmdepth;
    const float* dL_dalpha;
    const float* dL_dnormal;
};
void launch_render_backward(const FrameParams& fp, const uint2* ranges, const uint32_t* point_list,
                            const SplatRecord* records, const RenderBwdIn& in, double* grad_acc,
                            bool require_depth, cudaStream_t stream);

struct GaussianGrads {
    float* dL_dmeans3D;       // [P,3]
    float* dL_dmeans2D;       // [P,3]
    float* dL_dcolors;        // [P,3]
    float* dL_dopacity;       // [P]
    float* dL_dscales;        // [P,3]
    float* dL_drotations;     // [P,4]
    float* dL_dsh;            // [P,SHM,3] or null
    float* dL_dsg_axis;       // [P,SGM,3] or null
    float* dL_dsg_sharpness;  // [P,SGM] or null
    float* dL_dsg_color;      // [P,SGM,3] or null
    float* dL_drgb;           // [P,3] or null: dL/d(colour) after the clamp mask (the factor of the SH gradient)
    float* mc_rgb;            // or null: NVSwitch multicast address of this view's [P,3] slot in every rank's colour
    float* mc_campos;         //          stage, and of its camera-position slot (view-parallel exchange, fused push)
};
// [first, last): the Gaussians of this launch (first a multiple of 128; the whole model by default)
void launch_gaussian_backward(const FrameParams& fp, const GaussianInputs& in, const int* radii,
                              const uint8_t* clamped, const double* grad_acc, const GaussianGrads& g,
                              cudaStream_t stream, int64_t first = 0, int64_t last = -1);

// ---- point queries (integrate / sample_depth / evaluate_sdf), ggs_query.cu -------------------------------------
// Queries are sorted by (tile, Morton code of the quarter-pixel cell inside the tile); a CTA owns QUERY_CTA consecutive
// queries of one tile.  Everything a query call needs lives in ONE workspace whose layout depends on the number of
// points and tiles only -- no count is read back by the host.
constexpr int QUERY_CTA = 128;
constexpr int QUERY_CELL_BITS = 12;
struct QueryLayout {
    size_t xy, ts;                          // float2[PN] pixel position ((-1,-1): not evaluated), f32[PN] ray depth
    size_t keys_a, keys_b, vals_a, vals_b;  // u32[PN] sort ping-pong: key = tile << 12 | cell, value = point index
    size_t hist, status, counters, status_bytes;
    size_t ranges;                          // uint2[tiles] slice of the sorted queries per tile
    size_t rounds, offsets;                 // u32[tiles] CTAs per tile and their inclusive scan
    size_t scan_status, scan_counter, scan_total;
    size_t block_tile;                      // u32[max_blocks] tile of each CTA (0xffffffff: none)
    size_t n_contrib, median_depth;         // u32[PN], f32[PN]: sample_depth keeps them for its backward
    size_t bytes;
    uint32_t max_blocks;
    int sort_bits;
    bool sorted_in_b;
};
QueryLayout query_layout(size_t PN, int tiles, bool sample);
void launch_query_binning(const FrameParams& fp, int PN, const float* points3D, void* ws, const QueryLayout& L,
                          cudaStream_t stream);
// op_counters (measurement, may be null): u64[3] += pairs evaluated after culling, accepted pairs, list entries staged
void launch_integrate(const QueryLayout& L, void* ws, const uint2* gaussian_ranges, const uint32_t* gaussian_list,
                      const SplatRecord* records, float* out_transmittance, uint8_t* inside,
                      unsigned long long* op_counters, cudaStream_t stream);
void launch_sample_depth_forward(const FrameParams& fp, const QueryLayout& L, void* ws, const uint2* gaussian_ranges,
                                 const uint32_t* gaussian_list, const SplatRecord* records, float* output,
                                 uint8_t* inside, cudaStream_t stream);
void launch_evaluate_sdf(const FrameParams& fp, const QueryLayout& L, void* ws, const uint2* gaussian_ranges,
                         const uint32_t* gaussian_list, const SplatRecord* records, float* out_depth, float* out_sdf,
                         uint8_t* inside, cudaStream_t stream);
void launch_sample_depth_backward(const FrameParams& fp, const QueryLayout& L, void* ws, const uint2* gaussian_ranges,
                                  const uint32_t* gaussian_list, const SplatRecord* records, const uint8_t* inside,
                                  const float* dL_doutputs, float2* dL_dpoints2D, double* grad_acc,
                                  cudaStream_t stream);
void launch_points_backward(const FrameParams& fp, int PN, const float* points3D, const float2* points2D,
                            const float2* dL_dpoints2D, float* dL_dpoints3D, cudaStream_t stream);

// View-parallel training: dL_dsh[p][k][c] = sum_v basis_k(dir(p, campos[v])) * dL_drgb[v][p][c]  (ggs_multiview.cu)
constexpr int MAX_VIEWS = 16;
struct ViewSources {  // where each view's masked colour gradient [P,3] and camera position [3] live
    const float* rgb[MAX_VIEWS];
    const float* campos[MAX_VIEWS];
};
void launch_sh_gradient_from_views(int P, int SHD, int SHM, int V, const float* means3D, const float* campos,
                                   const float* dL_drgb, float* dL_dsh, cudaStream_t stream);

// Fused form over peer memory (one launch): the first `blocks` blocks sum this rank's slice [lo4, hi4) (float4
// units) of the flat gradient region over all ranks and store it into every rank's result region; the others form
// the SH gradient from the peers' colour gradients.  mc_*: NVSwitch multicast addresses of the same regions, or null.
struct PeerJob {
    int blocks, world;
    size_t lo4, hi4;
    const float4* send[MAX_VIEWS];
    float4* recv[MAX_VIEWS];
    const float4* mc_send;
    float4* mc_recv;
};
void launch_view_sum_peer(int P, int SHD, int SHM, const float* means3D, const ViewSources& src, float* dL_dsh,
                          const PeerJob& job, cudaStream_t stream);
// recv[i] = multimem.ld_reduce(send[i]) for the 16-byte units of up to 8 ranges, stored with multimem.st
struct SumRanges {
    int count;
    size_t lo4[8], hi4[8];
};
void launch_sum_ranges_multimem(const float4* mc_send, float4* mc_recv, const SumRanges& r, cudaStream_t stream);

// Push half of the exchange (ggs_peer.cu): every byte crosses the link as a posted store.
struct PushTargets {  // slot `rank` of every rank's colour stage / camera stage, as this process maps them
    int world;
    float* rgb[MAX_VIEWS];
    float* campos[MAX_VIEWS];
    float* mc_rgb;  // multicast address of slot `rank` of the colour stage, or null
};
struct PushSlices {  // dest[s] = slot `rank` of rank s's reduce stage; slice s = float4 units [lo4[s], lo4[s+1])
    int world;
    size_t lo4[MAX_VIEWS + 1];
    float4* dest[MAX_VIEWS];
};
void launch_push_rgb(int P, const double* grad_acc, const uint8_t* clamped, const int* radii, const float* campos,
                     const PushTargets& t, cudaStream_t stream);
void launch_push_slices(const float4* send, const PushSlices& t, cudaStream_t stream);

// Fused caller-side activations (ggs_activations.cu; scene/gaussian_model.py:158-160,203-212)
void launch_activate_forward(int P, const float* raw_scaling, const float* raw_opacity, const float* raw_rotation,
                             const float* filter_3D, float* scales, float* opacities, float* rotations,
                             cudaStream_t stream);
void launch_activate_backward(int P, const float* raw_scaling, const float* raw_opacity, const float* raw_rotation,
                              const float* filter_3D, const float* dL_dscales, const float* dL_dopacities,
                              const float* dL_drotations, float* dL_draw_scaling, float* dL_draw_opacity,
                              float* dL_draw_rotation, cudaStream_t stream);

// measurement only (ggs_counters.cu): counters[4] = walked, accepted, accepted on bisection-active pixels, active pixels
void launch_count_blend_ops(const FrameParams& fp, const uint2* ranges, const uint32_t* list,
                            const SplatRecord* records, unsigned long long* counters, cudaStream_t stream);

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream);

}  // namespace ggs
