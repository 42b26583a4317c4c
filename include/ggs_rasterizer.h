/*
 * ggs_rasterizer.h -- C ABI of the B200-native differentiable Gaussian rasterizer.
 *
 * This is the drop-in boundary for the hot path of XiShuFan/Geometry-Grounded-Gaussian-Splatting's
 * submodules/diff-gaussian-rasterization ("DGR/").  Each entry point replaces one static method
 * of CudaRasterizer::Rasterizer (DGR/cuda_rasterizer/rasterizer.h) that the reference's torch
 * binding (DGR/rasterize_points.cu, exported by DGR/ext.cpp:15-23) calls:
 *
 *   ggs_rasterize_forward   <- Rasterizer::forward                (rasterizer.h:29,  rasterize_points.cu:39-139)
 *   ggs_rasterize_backward  <- Rasterizer::backward               (rasterizer.h:63,  rasterize_points.cu:141-258)
 *   ggs_integrate_points    <- Rasterizer::evaluateTransmittance  (rasterizer.h:111, rasterize_points.cu:279-366)
 *   ggs_mark_visible        <- Rasterizer::markVisible            (rasterizer.h:22,  rasterize_points.cu:260-277)
 *
 * Conventions
 *   - plain C: pointers, sizes and PODs only; no torch, STL or CUDA types in any signature;
 *   - every data pointer is a DEVICE pointer on the current CUDA device unless marked "host";
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it.  Binning the Gaussians of a
 *     view (ggs_rasterize_forward, ggs_view_prepare and the one-shot point queries built on it)
 *     synchronises that stream once to learn how many (tile, Gaussian) pairs exist, where the reference
 *     issues its blocking cudaMemcpy (rasterizer_impl.cu:384, :676).  Queries against a prepared view
 *     never synchronise (the reference blocks a second time per call to count the points, :743);
 *   - the caller owns all memory.  Where the reference takes std::function<char*(size_t)> resize
 *     callbacks (rasterizer.h:30-33) this ABI takes ONE C callback, GgsAllocFn, tagged with the
 *     buffer slot it is asked to provide.  Buffers handed out in the forward must be passed back
 *     unchanged (same address, same contents) to the backward;
 *   - functions return 0 (or a non-negative count) on success and a negative GgsStatus on error;
 *     ggs_last_error() gives the message for the calling thread;
 *   - tensors use the reference's layouts: means3D [P,3], scales [P,3], rotations [P,4] (r,x,y,z),
 *     opacities [P], shs [P,sh_coeffs,3], sg_axis [P,sg_lobes,3], sg_sharpness [P,sg_lobes],
 *     sg_color [P,sg_lobes,3], images channel-major [C,H,W]; viewmatrix / projmatrix are the
 *     transposed 4x4 matrices the reference's cameras hold (scene/cameras.py:70-72).
 *   - threads and devices: a call works on the CURRENT device and enqueues on the stream it is given; several host
 *     threads may call at once, each on its own stream, and one process may drive several devices.  What the library
 *     keeps is keyed accordingly: the last error, the pinned word and the event behind the pair-count read-back per
 *     host thread (and device); shared-memory opt-ins and the side stream of the peer exchange per device; the
 *     optional stage timer (ggs_profile_enable) is process-wide and meant for one stream at a time.  The one
 *     exception is ggs_sort_pairs_static, whose scratch is one array per device (see its comment);
 *   - precomputed 3D covariances are not supported: the reference's own kernel cannot consume them
 *     (render_forward.cu:344-345 dereferences scales+idx unconditionally).
 */
#ifndef GGS_RASTERIZER_H_
#define GGS_RASTERIZER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGS_ABI_VERSION 2

#if defined(__GNUC__)
#define GGS_API __attribute__((visibility("default")))
#else
#define GGS_API
#endif

typedef enum GgsStatus {
    GGS_OK = 0,
    GGS_ERR_INVALID_ARGUMENT = -1,
    GGS_ERR_CUDA = -2,
    GGS_ERR_ALLOC = -3,
    GGS_ERR_UNSUPPORTED = -4
} GgsStatus;

/* Workspace slots requested through GgsAllocFn (the reference's geometryBuffer, binningBuffer,
 * imageBuffer, tileBuffer, pointBuffer, point_binningBuffer resize lambdas). */
typedef enum GgsSlot {
    GGS_SLOT_GEOMETRY = 0,
    GGS_SLOT_BINNING = 1,
    GGS_SLOT_IMAGE = 2,
    GGS_SLOT_TILE = 3,
    GGS_SLOT_POINT = 4,
    GGS_SLOT_POINT_BINNING = 5,
    GGS_SLOT_BACKWARD = 6,
    GGS_SLOT_BLOCKS = 7 /* per (tile, round of points) block state: the reference's duplicatedTileBuffer */
} GgsSlot;

/* Must return a device pointer to at least `bytes` bytes aligned to 256, or NULL on failure. */
typedef void* (*GgsAllocFn)(int slot, size_t bytes, void* user);

/* One camera / image (GaussianRasterizationSettings, DGR/diff_gaussian_rasterization/__init__.py:254-269). */
typedef struct GgsFrame {
    int32_t image_width;
    int32_t image_height;
    float tanfovx;
    float tanfovy;
    float kernel_size;
    float scale_modifier;
    int32_t sh_degree;     /* active SH degree, 0..3 */
    int32_t sg_degree;     /* active spherical-Gaussian lobes */
    int32_t prefiltered;   /* accepted for signature parity; culled Gaussians are simply skipped */
    int32_t require_depth; /* 0: colour + alpha only, depth and normal outputs are left untouched */
    int32_t debug;         /* 1: synchronise and report CUDA errors after every stage */
    const float* viewmatrix; /* [16] */
    const float* projmatrix; /* [16] */
    const float* campos;     /* [3]  */
    const float* background; /* [3]  */
} GgsFrame;

/* The Gaussian model as handed to GaussianRasterizer.forward. */
typedef struct GgsGaussians {
    int64_t count;               /* P */
    const float* means3D;        /* [P,3] */
    const float* opacities;      /* [P]   */
    const float* scales;         /* [P,3] */
    const float* rotations;      /* [P,4] */
    const float* shs;            /* [P,sh_coeffs,3] or NULL */
    int32_t sh_coeffs;           /* allocated coefficients per Gaussian (SHM) */
    const float* sg_axis;        /* [P,sg_lobes,3] or NULL */
    const float* sg_sharpness;   /* [P,sg_lobes]   or NULL */
    const float* sg_color;       /* [P,sg_lobes,3] or NULL */
    int32_t sg_lobes;            /* allocated lobes per Gaussian (SGM) */
    const float* colors_precomp; /* [P,3] or NULL; exactly one of shs / colors_precomp */
} GgsGaussians;

typedef struct GgsForwardOutputs {
    float* color;   /* [3,H,W] */
    float* alpha;   /* [1,H,W] */
    float* normal;  /* [3,H,W] written only when require_depth */
    float* mdepth;  /* [1,H,W] written only when require_depth */
    int32_t* radii; /* [P] */
} GgsForwardOutputs;

/* What the forward handed out and the backward needs again.
 * `capacity` is read by ggs_rasterize_forward as a HINT and written back:
 *   in   0: none.  The host waits for the pair count before it sizes the binning workspace and launches anything
 *           else -- where the reference blocks (rasterizer_impl.cu:384).
 *       >0: pairs to lay the binning workspace out for (e.g. 1.25 x the count of the caller's last frames).  The
 *           count-dependent kernels are launched for that capacity and take the count from device memory; the host
 *           waits for the count only after it has queued them, and queues them again on an exact-size workspace if
 *           the count exceeds the hint.  Results do not depend on the hint.
 *   out  the number of pairs the binning workspace IS laid out for (>= num_rendered); the backward and
 *        ggs_describe_layout take their offsets from it. */
typedef struct GgsForwardState {
    void* geometry;
    void* binning;
    void* image;
    void* tile;
    int64_t num_rendered;
    int64_t capacity;
} GgsForwardState;

typedef struct GgsBackwardInputs {
    const float* alpha;      /* forward outputs */
    const float* normal;
    const float* mdepth;
    const int32_t* radii;
    const float* dL_dcolor;  /* [3,H,W] */
    const float* dL_dmdepth; /* [1,H,W] */
    const float* dL_dalpha;  /* [1,H,W] */
    const float* dL_dnormal; /* [3,H,W] */
} GgsBackwardInputs;

/* Every row of every tensor is written (zeros for culled Gaussians); no pre-zeroing needed. */
typedef struct GgsGradients {
    float* dL_dmeans3D;      /* [P,3] */
    float* dL_dmeans2D;      /* [P,3] (x, y, |x|+|y|) */
    float* dL_dcolors;       /* [P,3] */
    float* dL_dopacity;      /* [P]   */
    float* dL_dscales;       /* [P,3] */
    float* dL_drotations;    /* [P,4] */
    float* dL_dsh;           /* [P,sh_coeffs,3] or NULL */
    float* dL_dsg_axis;      /* [P,sg_lobes,3]  or NULL */
    float* dL_dsg_sharpness; /* [P,sg_lobes]    or NULL */
    float* dL_dsg_color;     /* [P,sg_lobes,3]  or NULL */
    /* Optional (NULL in the reference's contract): dL/d(colour) after the colour clamp mask, [P,3].  It is the
     * factor of the rank-one SH gradient dL_dsh[p][k][c] = basis_k(dir_p) * dL_drgb[p][c]; when it is given and
     * dL_dsh is NULL the SH gradient is not written (view-parallel training forms the sum over views with
     * ggs_sh_gradient_from_views instead of all-reducing [P,sh_coeffs,3]). */
    float* dL_drgb;
} GgsGradients;

GGS_API int ggs_abi_version(void);
GGS_API const char* ggs_last_error(void);

/* Forward.  Returns num_rendered (>= 0) or a negative GgsStatus.  `state` receives the buffers. */
GGS_API int64_t ggs_rasterize_forward(const GgsFrame* frame, const GgsGaussians* gaussians, GgsAllocFn alloc, void* user,
                              const GgsForwardOutputs* out, GgsForwardState* state, void* stream);

/* Backward of ggs_rasterize_forward; asks `alloc` for GGS_SLOT_BACKWARD scratch. */
GGS_API int ggs_rasterize_backward(const GgsFrame* frame, const GgsGaussians* gaussians, const GgsForwardState* state,
                           const GgsBackwardInputs* in, GgsAllocFn alloc, void* user, const GgsGradients* grads,
                           void* stream);

/* Second half of the backward on its own: the per-Gaussian chain rule (the reference's computeCov2DCUDA +
 * preprocessCUDA backward, render_backward.cu:249-713) applied to a caller-provided accumulator array
 * double[P][GGS_ACCUMULATOR_STRIDE] in the layout ggs_rasterize_backward leaves in its GGS_SLOT_BACKWARD scratch:
 * dL/d(mean2D xy, |x|+|y|), conic xyz, opacity*coef, colour rgb, normal xyz, ray-plane xyzw.  The chain is linear in
 * the accumulators, which lets tests measure its conditioning component by component. */
#define GGS_ACCUMULATOR_STRIDE 18
#define GGS_ACCUMULATOR_LIVE 17
GGS_API int ggs_backward_from_accumulators(const GgsFrame* frame, const GgsGaussians* gaussians,
                                   const GgsForwardState* state, const int32_t* radii, const double* accumulators,
                                   const GgsGradients* grads, void* stream);

/* Sum over V views of the SH-coefficient gradient, from each view's masked colour gradient and camera position:
 * dL_dsh[p][k][c] = sum_v basis_k(normalize(means3D[p] - campos[v])) * dL_drgb[v][p][c], rows k >= (sh_degree+1)^2
 * zero.  campos is [V,3], dL_drgb is [V,P,3] (e.g. the all-gather of every rank's GgsGradients.dL_drgb). */
GGS_API int ggs_sh_gradient_from_views(int64_t count, int32_t sh_degree, int32_t sh_coeffs, int32_t views,
                               const float* means3D, const float* campos, const float* dL_drgb, float* dL_dsh,
                               void* stream);

/* ---- point queries against one view ---------------------------------------------------------------
 * The reference bins the view's Gaussians inside every query call (evaluateTransmittance, sampleDepth, evaluateSDF:
 * rasterizer_impl.cu:637-714, :861-936, :1084-1160), i.e. once per (chunk of points, view) of
 * mesh_extract_tetrahedra.py:65-87.  Here that half is its own entry point: ggs_view_prepare runs preprocess, sort and
 * tile ranges once (slots GEOMETRY, BINNING, TILE; colour is not evaluated, sh_degree / sg_degree of the frame are
 * ignored) and the *_view entry points answer any number of query batches against it (slot POINT per call; no host
 * synchronisation).  The one-shot entry points below them (= the reference's signatures) run both halves. */
typedef struct GgsViewState {
    void* geometry;
    void* binning;
    void* tile;
    int64_t num_rendered; /* (tile, Gaussian) pairs */
    int64_t count;        /* Gaussians the view was prepared with */
} GgsViewState;
/* Returns num_rendered (>= 0) or a negative GgsStatus.  The frame's kernel_size applies (the reference's wrappers pass
 * 0 on these paths, __init__.py:382,443,513). */
GGS_API int64_t ggs_view_prepare(const GgsFrame* frame, const GgsGaussians* gaussians, GgsAllocFn alloc, void* user,
                         GgsViewState* view, void* stream);
/* `frame` must describe the same camera and image as at preparation.  Outputs as for the one-shot entry points. */
GGS_API int ggs_integrate_view(const GgsFrame* frame, const GgsViewState* view, const float* points3D, int64_t num_points,
                       GgsAllocFn alloc, void* user, float* out_transmittance, uint8_t* inside, void* stream);
/* Measurement variant: as ggs_integrate_view, and adds to three device counters (caller-zeroed) [0] the (query, entry)
 * pairs evaluated after the warp-level culling, [1] the pairs that pass the power / alpha tests (the algorithmic
 * pairs of SURVEY.md 8(a) A9: 2 ex2 + 1 rsqrt each besides the alpha ex2), [2] the list entries staged. */
GGS_API int ggs_integrate_view_counted(const GgsFrame* frame, const GgsViewState* view, const float* points3D,
                               int64_t num_points, GgsAllocFn alloc, void* user, float* out_transmittance,
                               uint8_t* inside, uint64_t* op_counters, void* stream);
GGS_API int ggs_evaluate_sdf_view(const GgsFrame* frame, const GgsViewState* view, const float* points3D,
                          int64_t num_points, GgsAllocFn alloc, void* user, float* out_depth, float* out_sdf,
                          uint8_t* inside, void* stream);

/* Transmittance of query points seen from `frame` (integrate): out_transmittance [PN] and
 * inside [PN] (uint8 0/1) must be zero-initialised by the caller.  Returns num_rendered. */
GGS_API int64_t ggs_integrate_points(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                             int64_t num_points, GgsAllocFn alloc, void* user, float* out_transmittance,
                             uint8_t* inside, void* stream);

/* Median depth sampled at query points (sample_depth): replaces Rasterizer::sampleDepth / sampleDepthBackward
 * (rasterizer.h:170,197; rasterize_points.cu:459-634; pybind sample_rasterized_depth[_backward], ext.cpp:21-22).
 * out_points [PN,3] (camera-space point on the ray through the query's pixel position at the median depth) and
 * inside [PN] (uint8) must be zero-initialised by the caller: queries that project outside the image are not
 * touched.  `state` receives the buffers the backward needs (slots GEOMETRY, BINNING, POINT, TILE; point_binning and
 * blocks are reserved and stay NULL).  Returns num_rendered.  The reference's Python wrapper passes kernel_size = 0 to the forward and the
 * settings' kernel_size to the backward (__init__.py:513,606); callers wanting parity do the same. */
typedef struct GgsSampleState {
    void* geometry;
    void* binning;
    void* point;
    void* point_binning;
    void* tile;
    void* blocks;
    int64_t num_rendered;  /* (tile, Gaussian) pairs */
    int64_t num_evaluated; /* query points handed in (how many fall inside the image stays on the device) */
    int64_t num_blocks;    /* reserved */
} GgsSampleState;

GGS_API int64_t ggs_sample_depth_forward(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                                 int64_t num_points, GgsAllocFn alloc, void* user, float* out_points,
                                 uint8_t* inside, GgsSampleState* state, void* stream);

/* Signed distance of query points to the rendered median-depth surface along their rays (evaluate_sdf): replaces
 * Rasterizer::evaluateSDF (rasterizer.h:140; rasterize_points.cu:368-457; pybind evaluate_sdf_from_signle_view,
 * ext.cpp:20).  out_depth [PN] = median ray depth at the query's pixel position, out_sdf [PN] = that depth minus the
 * query's own ray depth, inside [PN] (uint8); all three zero-initialised by the caller. */
GGS_API int64_t ggs_evaluate_sdf(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                         int64_t num_points, GgsAllocFn alloc, void* user, float* out_depth, float* out_sdf,
                         uint8_t* inside, void* stream);

/* Gradients of ggs_sample_depth_forward: dL_dout_points [PN,3] in; dL_dpoints3D [PN,3] and the Gaussian
 * gradients (means3D, opacity, scales, rotations of `grads`; its other members may be NULL except dL_dmeans2D and
 * dL_dcolors, which are scratch outputs of the shared chain rule) out.  Every row is written. */
GGS_API int ggs_sample_depth_backward(const GgsFrame* frame, const GgsGaussians* gaussians, const float* points3D,
                              int64_t num_points, const GgsSampleState* state, const uint8_t* inside,
                              const float* dL_dout_points, GgsAllocFn alloc, void* user, const GgsGradients* grads,
                              float* dL_dpoints3D, void* stream);

/* present[i] = 1 iff Gaussian i passes the near-plane test. */
GGS_API int ggs_mark_visible(const float* means3D, int64_t count, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* ---- view-parallel training over peer memory (NVLink / NVSwitch): no NCCL call on the data path --------------
 * Every rank of the node owns an exchange buffer of identical layout that all ranks have mapped
 * (peer_buffers[r] = rank r's buffer as THIS process sees it: torch symmetric memory, or cudaIpcOpenMemHandle).
 * All offsets are in floats from the buffer's base and multiples of 4:
 *   [send_offset, +reduce_floats)      this rank's per-view gradients of the non-SH parameters: point
 *                                      GgsGradients.{dL_drotations, dL_dmeans3D, dL_dscales, dL_dopacity} into it
 *   rgb_stage_offset + r*rgb_stage_stride       slot r: rank r's masked colour gradient dL/dRGB [count,3]
 *   campos_stage_offset + r*4                   slot r: rank r's camera position
 *   reduce_stage_offset + r*reduce_stage_stride slot r: rank r's copy of THIS rank's 1/world slice of the region
 *   [recv_offset, +reduce_floats)      receives the sum over ranks of the region
 * multicast_base: address of an NVSwitch multicast mapping of the same buffers (0 if none); broadcasts then take one
 * multimem.st instead of `world` peer stores.
 *
 * ggs_rasterize_backward_peer = ggs_rasterize_backward + the PUSH half: this rank's dL/dRGB goes to slot `rank` of
 * every rank's colour stage as soon as the tile backward is done (side stream, concurrent with the per-Gaussian
 * backward), and slice s of the send region goes to slot `rank` of rank s's reduce stage (a reduce-scatter by posted
 * stores; pulling operands from a peer is bound by NVLink round trips).  GgsGradients.dL_dsh and dL_drgb may be NULL.
 * Then, after a node barrier on the stream, ggs_view_sum_peer sums the slots of this rank's slice, stores the sum into
 * every rank's recv region and writes dL_dsh = sum over ranks of the rank-one SH gradients (as
 * ggs_sh_gradient_from_views) -- all reads local; after a second barrier the recv region holds the gradients summed
 * over all views.  The reference has no counterpart (no distributed code). */
#define GGS_MAX_CHUNKS 8
typedef struct GgsPeerExchange {
    int32_t world, rank;
    const uint64_t* peer_buffers; /* [world], host array */
    uint64_t multicast_base;
    int64_t send_offset, recv_offset, reduce_floats;
    int64_t rgb_stage_offset, rgb_stage_stride;
    int64_t campos_stage_offset;
    int64_t reduce_stage_offset, reduce_stage_stride;
    /* 0: slices are pushed into the owners' reduce stages and summed from local memory (default).
     * 1: no slice push; the owner of a slice sums it straight out of every rank's send region -- one
     *    multimem.ld_reduce per 16 bytes when multicast_base is set (the switch adds), peer loads otherwise. */
    int32_t pull_reduce;
    /* Pipelining of the sum with the per-Gaussian backward (pull_reduce with a multicast mapping only; 0 or 1: off).
     * ggs_rasterize_backward_peer launches the per-Gaussian backward in `chunks` pieces -- chunk k = Gaussians
     * [chunk_end[k-1], chunk_end[k]), every boundary but the last a multiple of 128 -- and records
     * chunk_events[k] (a cudaEvent_t of the caller) on `stream` behind piece k.  The caller, on another stream, waits
     * for event k, holds a node barrier and calls ggs_peer_sum_ranges on the part of the gradient region chunk k wrote:
     * the links sum chunk k while the SMs compute chunk k + 1. */
    int32_t chunks;
    int64_t chunk_end[GGS_MAX_CHUNKS];
    uint64_t chunk_events[GGS_MAX_CHUNKS];
    /* 1 (needs multicast_base): the per-Gaussian backward itself stores this view's dL/dRGB and camera position into
     * every rank's stage (multimem.st from the kernel that computes them) instead of a separate pass over the gradient
     * accumulators on the side stream.  The colour gradients are then complete when `stream` has finished the
     * backward: hold the node barrier and launch ggs_view_sum_peer(GGS_VIEW_SUM_SH) behind it, not behind the side
     * stream. */
    int32_t fuse_rgb_push;
} GgsPeerExchange;
/* side_stream (may be NULL: the library then uses a stream of its own): the stream the colour push is enqueued on.
 * A caller that passes one can chain, on that stream, a barrier and ggs_view_sum_peer(GGS_VIEW_SUM_SH) behind the
 * push: the SH half of the sum then runs concurrently with the per-Gaussian backward on `stream`, and only
 * ggs_view_sum_peer(GGS_VIEW_SUM_SLICE) is left for after it.  `parts` selects the halves of the launch. */
#define GGS_VIEW_SUM_SLICE 1 /* sum this rank's slice of the gradient region and store it to every rank */
#define GGS_VIEW_SUM_SH 2    /* SH gradient summed over views from the staged colour gradients */
GGS_API int ggs_rasterize_backward_peer(const GgsFrame* frame, const GgsGaussians* gaussians,
                                const GgsForwardState* state, const GgsBackwardInputs* inputs, GgsAllocFn alloc,
                                void* user, const GgsGradients* grads, const GgsPeerExchange* exchange, void* stream,
                                void* side_stream);
GGS_API int ggs_view_sum_peer(int64_t count, int32_t sh_degree, int32_t sh_coeffs, const float* means3D,
                      const GgsPeerExchange* exchange, float* dL_dsh, int32_t parts, void* stream);
/* recv[i] = sum over ranks of send[i] for every 16-byte unit i of the `count` (<= 8) ranges [lo4[j], hi4[j]) of the
 * gradient region (units of 4 floats from send_offset / recv_offset), stored into every rank's recv region: one
 * multimem.ld_reduce + one multimem.st per unit (needs multicast_base).  The ranges a rank passes are ITS share of the
 * units to sum -- the ranks of the node together cover what they want summed, each unit exactly once. */
GGS_API int ggs_peer_sum_ranges(const GgsPeerExchange* exchange, int32_t count, const int64_t* lo4, const int64_t* hi4,
                        void* stream);

/* ---- fused caller-side activations (opt-in, the reference does this in torch) -------------
 * One pass each way over what gaussian_renderer/__init__.py:62-63 obtains from
 * GaussianModel.get_scaling_n_opacity_with_3D_filter (scene/gaussian_model.py:203-212) and GaussianModel.get_rotation
 * (:158-160):  scales = sqrt(exp(raw_scaling)^2 + filter_3D^2) [P,3],
 * opacities = sigmoid(raw_opacity) * sqrt(prod exp(raw_scaling)^2 / prod (exp(raw_scaling)^2 + filter_3D^2)) [P,1],
 * rotations = raw_rotation / max(|raw_rotation|, 1e-12) [P,4].  filter_3D [P,1] carries no gradient (it is a buffer
 * in the reference).  IEEE arithmetic (no fast-math): results within an ulp or two of the torch expressions. */
GGS_API int ggs_activate_forward(int64_t count, const float* raw_scaling, const float* raw_opacity,
                         const float* raw_rotation, const float* filter_3D, float* scales, float* opacities,
                         float* rotations, void* stream);
GGS_API int ggs_activate_backward(int64_t count, const float* raw_scaling, const float* raw_opacity,
                          const float* raw_rotation, const float* filter_3D, const float* dL_dscales,
                          const float* dL_dopacities, const float* dL_drotations, float* dL_draw_scaling,
                          float* dL_draw_opacity, float* dL_draw_rotation, void* stream);

/* ---- optional per-stage device timing ------------------------------------------------------------
 * When enabled (process-wide) the forward / backward record CUDA events on the caller's stream
 * between stages; ggs_profile_read waits for the events of the most recent forward and backward and
 * returns the stage durations in milliseconds.  The reference has no counterpart (its only timing is
 * the torch.cuda.Event pair in train.py:63-64); bench.py uses this for the per-kernel roofline. */
typedef struct GgsStageTimes {
    int32_t have_forward, have_backward;
    float preprocess_ms, scan_ms, duplicate_ms, sort_ms, ranges_ms, render_forward_ms;
    float render_backward_ms, gaussian_backward_ms;
    int32_t sort_passes; /* radix passes over the (tile, Gaussian) pairs the most recent forward ran */
    float depth_order_ms; /* depth pre-sort of the Gaussians (between preprocess and scan) */
    /* most recent point query (integrate / sample_depth forward / evaluate_sdf): projecting + sorting + slicing the
     * query points, and the query kernel itself */
    int32_t have_query;
    float query_binning_ms, query_kernel_ms;
} GgsStageTimes;
GGS_API int ggs_profile_enable(int on);
GGS_API int ggs_profile_read(GgsStageTimes* times);

/* ---- measurement: algorithmic operation counts of the tile blend ----------------------------------------
 * Re-walks every pixel's list of a finished forward with the reference's accept / stop rules
 * (render_forward.cu:477-532) and writes four device counters: [0] list positions visited, [1] accepted
 * (pixel, Gaussian) pairs, [2] accepted pairs of the pixels the median-depth bisection runs on (final T <= 0.45),
 * [3] the number of those pixels.  bench.py turns them into the FP32 / XU pipe roofline; nothing on
 * the product path calls it. */
GGS_API int ggs_count_blend_ops(const GgsFrame* frame, const GgsGaussians* gaussians, const GgsForwardState* state,
                        uint64_t* counters, void* stream);

/* ---- introspection for tests: byte offsets of the arrays inside the opaque workspaces ---------- */
typedef struct GgsLayout {
    uint64_t geometry_bytes, records, depths, tiles_touched, point_offsets, clamped, rects;
    uint64_t binning_bytes, sorted_keys, point_list, unsorted_keys, unsorted_list;
    uint64_t tile_bytes, ranges, max_contributor;
    uint64_t image_bytes, n_contrib;
    uint32_t sort_bits, sort_passes;
    uint64_t depth_order; /* u32[P] in the geometry workspace: Gaussian index at each depth rank (culled last);
                             point_offsets is the inclusive scan of tiles_touched taken in this order */
} GgsLayout;
GGS_API int ggs_describe_layout(int64_t P, int32_t width, int32_t height, int64_t num_rendered, GgsLayout* layout);
/* The forward sorts the (tile, Gaussian) pairs on the 32-bit tile half of the reference's key -- the pairs are emitted
 * in depth order, so the depth half never has to move -- and nothing downstream reads the 64-bit keys
 * (rasterizer_impl.cu:406-414: SortPairs, then identifyTileRanges looks at `key >> 32` only).  This writes them,
 * (tile << 32) | depth bits of the pair's Gaussian in sorted order, into the binning workspace at
 * GgsLayout.sorted_keys for inspection and for the bit-exactness tests against the reference's point_list_keys. */
GGS_API int ggs_materialize_sorted_keys(int64_t P, int32_t width, int32_t height, const GgsForwardState* state,
                                        void* stream);

/* Stand-alone stable radix sort of (u64 key, u32 value) pairs on bits [begin_bit, end_bit) -- the
 * replacement for cub::DeviceRadixSort::SortPairs at rasterizer_impl.cu:406-412.  `scratch` must
 * hold ggs_sort_scratch_bytes(n, begin_bit, end_bit) bytes.  Results land in keys_out / vals_out. */
GGS_API size_t ggs_sort_scratch_bytes(int64_t n, int32_t begin_bit, int32_t end_bit);
GGS_API int ggs_sort_pairs(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                   int64_t n, int32_t begin_bit, int32_t end_bit, void* scratch, void* stream);

/* ggs_sort_pairs without an allocation anywhere -- NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments and results as ggs_sort_pairs minus `scratch`: the scratch is a file-scope
 * `static __device__` array inside the library, sized for GGS_SORT_STATIC_MAX_PAIRS pairs over all 64 key bits
 * (about 3.7 MiB, held by every GPU that loads the library until the process exits).  n beyond the maximum is refused
 * with GGS_ERR_INVALID_ARGUMENT and a message naming the limit.  The call makes no cudaMalloc / cudaFree / allocating
 * library call; it enqueues copies, memsets and kernels on `stream` and returns.  Every call on a device shares the
 * one array, which keeps the previous call's contents (the histograms, look-back descriptors and tile tickets are
 * cleared on `stream` at the start of each call): calls that OVERLAP on the GPU -- two streams, two host threads, or
 * this call beside another still running -- race on it; serialise them on one stream or with events. */
#define GGS_SORT_STATIC_MAX_PAIRS (1 << 18)
GGS_API int ggs_sort_pairs_static(const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                                  uint32_t* vals_out, int64_t n, int32_t begin_bit, int32_t end_bit, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GGS_RASTERIZER_H_ */
