// Stage 1 of the forward: per-Gaussian projection, 2D covariance + Mip 2D filter, ray-plane /
// normal, SH + spherical-Gaussian colour.  One thread per Gaussian; outputs written as one 64-byte
// SplatRecord plus SoA side arrays.  The SH coefficients ([P][SHM*3] rows, 192 bytes at SH3) are moved per
// warp as one contiguous slab through shared memory with coalesced 16-byte loads (rows of culled Gaussians
// are skipped), instead of 48 strided scalar loads per thread.
//
// Behaviour follows DGR/cuda_rasterizer/render_forward.cu:283-386 (preprocessCUDA) with helpers
// :22-78 (colour), :81-243 (computeCov2D) and auxiliary.h:38-49,133-153.  The float pipelines that
// feed radii / tile rects / depth keys keep the reference's expression order (see ggs_math.cuh).
#include "ggs_kernels.cuh"
#include "ggs_math.cuh"

namespace ggs {

__device__ __forceinline__ float3 ld3(const float* p, size_t i) {
    return {__ldg(p + 3 * i), __ldg(p + 3 * i + 1), __ldg(p + 3 * i + 2)};
}

// Colour from SH (degree <= 3) + spherical Gaussians (render_forward.cu:22-78).
__device__ __forceinline__ float3 eval_color(size_t idx, const FrameParams& fp, const float3 pos, const float* sh,
                                             const float* sg_axis, const float* sg_sharpness, const float* sg_color,
                                             uint8_t& clamp_mask) {
    float3 dir = {pos.x - fp.view.campos[0], pos.y - fp.view.campos[1], pos.z - fp.view.campos[2]};
    float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir = {dir.x / len, dir.y / len, dir.z / len};

    // `sh` is this Gaussian's coefficient row (shared memory)
    auto C = [&](int k) -> float3 { return float3{sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]}; };
    float3 s0 = C(0);
    float3 res = {kSH_C0 * s0.x, kSH_C0 * s0.y, kSH_C0 * s0.z};
    auto acc = [&](float w, const float3& c) {
        res.x += w * c.x;
        res.y += w * c.y;
        res.z += w * c.z;
    };
    if (fp.SHD > 0) {
        float x = dir.x, y = dir.y, z = dir.z;
        acc(-kSH_C1 * y, C(1));
        acc(kSH_C1 * z, C(2));
        acc(-kSH_C1 * x, C(3));
        if (fp.SHD > 1) {
            float xx = x * x, yy = y * y, zz = z * z;
            float xy = x * y, yz = y * z, xz = x * z;
            acc(kSH_C2[0] * xy, C(4));
            acc(kSH_C2[1] * yz, C(5));
            acc(kSH_C2[2] * (2.0f * zz - xx - yy), C(6));
            acc(kSH_C2[3] * xz, C(7));
            acc(kSH_C2[4] * (xx - yy), C(8));
            if (fp.SHD > 2) {
                acc(kSH_C3[0] * y * (3.0f * xx - yy), C(9));
                acc(kSH_C3[1] * xy * z, C(10));
                acc(kSH_C3[2] * y * (4.0f * zz - xx - yy), C(11));
                acc(kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy), C(12));
                acc(kSH_C3[4] * x * (4.0f * zz - xx - yy), C(13));
                acc(kSH_C3[5] * z * (xx - yy), C(14));
                acc(kSH_C3[6] * x * (xx - 3.0f * yy), C(15));
            }
        }
    }
    for (int g = 0; g < fp.SGD; g++) {
        size_t o = idx * (size_t)fp.SGM + g;
        float3 ax = ld3(sg_axis, o);
        float3 col = ld3(sg_color, o);
        float sharp = __ldg(sg_sharpness + o);
        float gauss = expf(sharp * ((ax.x * dir.x + ax.y * dir.y + ax.z * dir.z) - 1.0f));
        acc(gauss, col);
    }
    res.x += 0.5f;
    res.y += 0.5f;
    res.z += 0.5f;
    clamp_mask = (res.x < 0 ? 1 : 0) | (res.y < 0 ? 2 : 0) | (res.z < 0 ? 4 : 0);
    return {fmaxf(res.x, 0.f), fmaxf(res.y, 0.f), fmaxf(res.z, 0.f)};
}

struct Projected {
    float3 p_orig;
    float tc, my_radius, coef;
    float2 point_image;
    float3 conic, normal;
    float4 ray_plane;
    uint2 rect_min, rect_max;
};

// Geometry part of the preprocess; false when the Gaussian is culled (near plane, singular covariance, empty
// tile rectangle), in the order the reference tests them.
__device__ __forceinline__ bool project_gaussian(const size_t idx, const FrameParams& fp, const GaussianInputs& in,
                                                 Projected& o) {
    const float* viewmatrix = fp.view.view;
    const float3 p_orig = ld3(in.means3D, idx);
    // near cull (auxiliary.h:133-153)
    const float3 p_view = xform_point_4x3(p_orig, viewmatrix);
    if (p_view.z <= 0.2f) return false;

    const float4 p_hom = xform_point_4x4(p_orig, fp.view.proj);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

    // ---- 2D covariance, ray plane, normal (render_forward.cu:81-243) --------------------------
    // Every float operation below is spelled out with an explicit rounding step (ggs_math.cuh: MUL / ADD /
    // FMA / rcp_approx ...) in the evaluation order the reference binary uses, so conic, opacity, ray plane
    // and normal come out bit-identical to the reference's geometry buffers (DESIGN.md "bit-exact records").
    const float fx = fp.focal_x, fy = fp.focal_y;
    const float v0 = viewmatrix[0], v1 = viewmatrix[1], v2 = viewmatrix[2], v4 = viewmatrix[4], v5 = viewmatrix[5],
                v6 = viewmatrix[6], v8 = viewmatrix[8], v9 = viewmatrix[9], v10 = viewmatrix[10];
    const float tx0 = p_view.x, ty0 = p_view.y, tz = p_view.z;
    const float tc = norm3df(tx0, ty0, tz);  // also the depth key

    const float rz = rcp_approx(tz);
    const float limx = MUL(fp.tanfovx, 1.3f);
    const float limy = MUL(fp.tanfovy, 1.3f);
    const float tx = MUL(tz, fminf(limx, fmaxf(-limx, MUL(tx0, rz))));
    const float ty = MUL(tz, fminf(limy, fmaxf(-limy, MUL(ty0, rz))));
    const float u = MUL(tx, rz);
    const float v = MUL(ty, rz);
    const float r_tz2 = rcp_approx(MUL(tz, tz));

    // J (two non-trivial columns) and T = W * J
    const float J00 = MUL(rz, fx), J02 = MUL(MUL(tx, -fx), r_tz2);
    const float J11 = MUL(rz, fy), J12 = MUL(MUL(ty, -fy), r_tz2);
    const float T00 = FMA(v2, J02, FMA(v0, J00, MUL(0.f, v1)));
    const float T01 = FMA(v6, J02, FMA(v4, J00, MUL(0.f, v5)));
    const float T02 = FMA(v10, J02, FMA(v8, J00, MUL(0.f, v9)));
    const float T10 = FMA(v2, J12, FMA(0.f, v0, MUL(v1, J11)));
    const float T11 = FMA(v6, J12, FMA(0.f, v4, MUL(v5, J11)));
    const float T12 = FMA(v10, J12, FMA(0.f, v8, MUL(v9, J11)));

    const float mod = fp.scale_modifier;
    const float3 sc = ld3(in.scales, idx);
    const float s0 = MUL(sc.x, mod), s1 = MUL(sc.y, mod), s2 = MUL(sc.z, mod);
    const float i0 = __frcp_rn(s0), i1 = __frcp_rn(s1), i2 = __frcp_rn(s2);

    const float4 rot = __ldg(reinterpret_cast<const float4*>(in.rotations) + idx);
    const float qr = rot.x, qx = rot.y, qy = rot.z, qz = rot.w;
    // rotation matrix, R<col><row>
    const float R00 = one_minus_twice(ADD(MUL(qy, qy), MUL(qz, qz)));
    const float R11 = one_minus_twice(FMA(qx, qx, MUL(qz, qz)));
    const float R22 = one_minus_twice(FMA(qx, qx, MUL(qy, qy)));
    const float R01 = twice(FMA(qx, qy, -MUL(qr, qz)));
    const float R02 = twice(FMA(qr, qy, MUL(qx, qz)));
    const float R10 = twice(FMA(qx, qy, MUL(qr, qz)));
    const float R12 = twice(FMA(qy, qz, -MUL(qr, qx)));
    const float R20 = twice(FMA(-qr, qy, MUL(qx, qz)));
    const float R21 = twice(FMA(qy, qz, MUL(qr, qx)));

    // M = (S R) T, only the two columns that reach the 2D covariance; cov = M^T M
    const float B00 = MUL(s0, R00), B01 = MUL(s1, R01), B02 = MUL(s2, R02);
    const float B10 = MUL(s0, R10), B11 = MUL(s1, R11), B12 = MUL(s2, R12);
    const float B20 = MUL(s0, R20), B21 = MUL(s1, R21), B22 = MUL(s2, R22);
    const float M00 = sop3(T00, B00, T01, B10, T02, B20);
    const float M01 = sop3(T00, B01, T01, B11, T02, B21);
    const float M02 = sop3(T00, B02, T01, B12, T02, B22);
    const float M10 = sop3(T10, B00, T11, B10, T12, B20);
    const float M11 = sop3(T10, B01, T11, B11, T12, B21);
    const float M12 = sop3(T10, B02, T11, B12, T12, B22);
    const float cov00 = sop3(M00, M00, M01, M01, M02, M02);
    const float cov01 = sop3(M00, M10, M01, M11, M02, M12);
    const float cov11 = sop3(M10, M10, M11, M11, M12, M12);

    // camera-space inverse covariance  C = Minv^T Minv,  Minv = (S^-1 R) W
    const float A00 = MUL(i0, R00), A01 = MUL(i1, R01), A02 = MUL(i2, R02);
    const float A10 = MUL(i0, R10), A11 = MUL(i1, R11), A12 = MUL(i2, R12);
    const float A20 = MUL(i0, R20), A21 = MUL(i1, R21), A22 = MUL(i2, R22);
    const float N00 = sop3(v0, A00, v4, A10, v8, A20), N01 = sop3(v0, A01, v4, A11, v8, A21),
                N02 = sop3(v0, A02, v4, A12, v8, A22);
    const float N10 = sop3(v1, A00, v5, A10, v9, A20), N11 = sop3(v1, A01, v5, A11, v9, A21),
                N12 = sop3(v1, A02, v5, A12, v9, A22);
    const float N20 = sop3(v2, A00, v6, A10, v10, A20), N21 = sop3(v2, A01, v6, A11, v10, A21),
                N22 = sop3(v2, A02, v6, A12, v10, A22);
    const float C00 = sop3(N00, N00, N01, N01, N02, N02);
    const float C10 = sop3(N00, N10, N01, N11, N02, N12);
    const float C11 = sop3(N10, N10, N11, N11, N12, N12);
    const float C20 = sop3(N00, N20, N01, N21, N02, N22);
    const float C21 = sop3(N10, N20, N11, N21, N12, N22);
    const float C22 = sop3(N20, N20, N21, N21, N22, N22);

    // Mip 2D filter
    const float ks = fp.kernel_size;
    const float ca = ADD(cov00, ks), cc = ADD(cov11, ks), cb = cov01;
    const float cb2 = MUL(cb, cb);
    const float det = FMA(ca, cc, -cb2);
    const float det_0 = fmaxf(FMA(cov11, cov00, -cb2), 1e-6f);
    const float det_1 = fmaxf(det, 1e-6f);
    const float coef = sqrt_approx(MUL(det_0, rcp_approx(det_1)));

    float4 ray_plane;
    float3 normal;
    {
        const float mx = ADD(FMA(u, C00, MUL(v, C10)), C20);
        const float my = ADD(FMA(u, C10, MUL(v, C11)), C21);
        const float mz = ADD(FMA(u, C20, MUL(v, C21)), C22);
        const float vb = ADD(mz, FMA(u, mx, MUL(v, my)));
        const float r_vb = rcp_approx(vb);
        const float ax = MUL(mx, r_vb), ay = MUL(my, r_vb), az = MUL(mz, r_vb);
        const float vv = MUL(v, v), uv = MUL(u, v);
        const float plane_x = FMA(-u, az, FMA(ADD(vv, 1.f), ax, -MUL(uv, ay)));
        const float plane_y = FMA(-v, az, FMA(FMA(u, u, 1.f), ay, -MUL(uv, ax)));
        const float r_len2 = rcp_approx(ADD(FMA(u, u, vv), 1.f));
        const float l = norm3df(tx, ty, tz);
        const float factor = MUL(l, r_len2);
        ray_plane = {MUL(MUL(factor, plane_x), rcp_approx(fx)), MUL(MUL(factor, plane_y), rcp_approx(fy)), tc,
                     sqrt_approx(MUL(vb, r_len2))};
        const float nx = MUL(factor, -plane_x), ny = MUL(factor, -plane_y);
        const float rl = rcp_approx(l);
        const float cx = FMA(-tx, rl, FMA(nx, rz, MUL(0.f, ny)));
        const float cy = FMA(-ty, rl, FMA(0.f, nx, MUL(ny, rz)));
        const float cz = FMA(-tz, rl, FMA(nx, MUL(tx, -r_tz2), MUL(ny, MUL(ty, -r_tz2))));
        const float rs = rsqrt_approx(FMA(cz, cz, FMA(cx, cx, MUL(cy, cy))));
        normal = {MUL(cx, rs), MUL(cy, rs), MUL(cz, rs)};
    }

    // ---- conic, radius, tile rectangle (render_forward.cu:347-368) ----------------------------
    if (det == 0.0f) return false;
    const float det_inv = rcp_approx(det);
    const float3 conic = {MUL(cc, det_inv), MUL(det_inv, -cb), MUL(ca, det_inv)};

    const float mid = MUL(ADD(ca, cc), 0.5f);
    const float sq = sqrt_approx(fmaxf(FMA(mid, mid, -det), 0.1f));
    const float my_radius = ceilf(MUL(sqrt_approx(fmaxf(ADD(mid, sq), ADD(mid, -sq))), 3.f));
    const float2 point_image = {ndc_to_pix(p_proj.x, fp.W), ndc_to_pix(p_proj.y, fp.H)};
    uint2 rect_min, rect_max;
    tile_rect(point_image, my_radius, rect_min, rect_max, fp.gx, fp.gy);
    if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) return false;
    o.p_orig = p_orig;
    o.tc = tc;
    o.my_radius = my_radius;
    o.coef = coef;
    o.point_image = point_image;
    o.conic = conic;
    o.normal = normal;
    o.ray_plane = ray_plane;
    o.rect_min = rect_min;
    o.rect_max = rect_max;
    return true;
}

constexpr int PRE_THREADS = 128;

#ifndef GGS_PRE_MIN_BLOCKS
#define GGS_PRE_MIN_BLOCKS 7  // 72 registers: 89 us instead of 94 at 80 registers / 6 blocks (8 blocks: spills, no gain)
#endif
__global__ void __launch_bounds__(PRE_THREADS, GGS_PRE_MIN_BLOCKS) preprocess_kernel(const FrameParams fp, const GaussianInputs in,
                                                                 const bool want_color, SplatRecord* __restrict__ records,
                                                                 float* __restrict__ depths,
                                                                 uint32_t* __restrict__ tiles_touched,
                                                                 uint2* __restrict__ rects, uint8_t* __restrict__ clamped,
                                                                 int* __restrict__ radii,
                                                                 uint32_t* __restrict__ order_keys,
                                                                 uint32_t* __restrict__ order_vals,
                                                                 const ClearList clear) {
    extern __shared__ float s_slab[];  // [warps][32][RS] SH rows (only when colours come from SH)
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool in_range = idx < (size_t)fp.P;

    // what the later stages of the binning expect zeroed (ClearList in ggs_kernels.cuh): nothing in this kernel
    // touches those spans, so the grid clears them on its way instead of a launch of its own
    for (int s = 0; s < clear.count; s++) {
        uint4* const p = static_cast<uint4*>(clear.p[s]);
        for (uint32_t i = (uint32_t)idx; i < clear.n16[s]; i += gridDim.x * blockDim.x) p[i] = uint4{0, 0, 0, 0};
    }

    Projected pr;
    const bool vis = in_range && project_gaussian(idx, fp, in, pr);
    if (in_range && !vis) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
    }
    if (in_range) {
        // input of the depth pre-sort: culled Gaussians take the largest key and land behind every visible one
        order_keys[idx] = vis ? __float_as_uint(pr.tc) : 0xffffffffu;
        order_vals[idx] = (uint32_t)idx;
    }

    // ---- SH slab of the warp's visible Gaussians -> shared memory -------------------------------------
    const bool use_sh = want_color && in.colors_precomp == nullptr;
    const int F = fp.SHM * 3;
    const int RS = F | 1;  // odd row stride: conflict-free when every lane walks its own row
    float* const swarp = s_slab + (size_t)warp * 32 * RS;
    if (use_sh) {
        const size_t g0 = (size_t)blockIdx.x * blockDim.x + warp * 32;
        const int ng = g0 < (size_t)fp.P ? (int)min((size_t)32, (size_t)fp.P - g0) : 0;
        const uint32_t vm = __ballot_sync(0xffffffffu, vis);
        if (vm) {
            if ((F & 3) == 0) {
                const int Q = F >> 2;
                const float4* src = reinterpret_cast<const float4*>(in.shs + g0 * F);
                load_rows_to_shared(src, Q, ng, vm, swarp, RS, lane);
            } else {
                const float* src = in.shs + g0 * F;
                int gi = 0, rem = lane;
                for (int e = lane; e < ng * F; e += 32, rem += 32) {
                    while (rem >= F) {
                        rem -= F;
                        gi++;
                    }
                    if ((vm >> gi) & 1u) swarp[gi * RS + rem] = __ldg(src + e);
                }
            }
        }
        __syncwarp();
    }
    if (!vis) return;

    // ---- colour -------------------------------------------------------------------------------
    float3 rgb = {0.f, 0.f, 0.f};
    uint8_t mask = 0;
    if (want_color) {
        if (in.colors_precomp != nullptr)
            rgb = ld3(in.colors_precomp, idx);
        else
            rgb = eval_color(idx, fp, pr.p_orig, swarp + lane * RS, in.sg_axis, in.sg_sharpness, in.sg_color, mask);
    }

    depths[idx] = pr.tc;
    radii[idx] = pr.my_radius;
    clamped[idx] = mask;
    rects[idx] = {pr.rect_min.x | (pr.rect_min.y << 16), pr.rect_max.x | (pr.rect_max.y << 16)};
    tiles_touched[idx] = (pr.rect_max.y - pr.rect_min.y) * (pr.rect_max.x - pr.rect_min.x);

    const float opac = MUL(pr.coef, __ldg(in.opacities + idx));
    float4* dst = reinterpret_cast<float4*>(records + idx);
    dst[0] = {pr.point_image.x, pr.point_image.y, pr.conic.x, pr.conic.y};
    dst[1] = {pr.conic.z, opac, pr.ray_plane.x, pr.ray_plane.y};
    dst[2] = {pr.ray_plane.z, pr.ray_plane.w, rgb.x, rgb.y};
    dst[3] = {rgb.z, pr.normal.x, pr.normal.y, pr.normal.z};
}

void launch_preprocess(const FrameParams& fp, const GaussianInputs& in, bool want_color, SplatRecord* records,
                       float* depths, uint32_t* tiles_touched, uint2* rects, uint8_t* clamped, int* radii,
                       uint32_t* order_keys, uint32_t* order_vals, cudaStream_t stream, const ClearList* clear) {
    if (fp.P == 0) {
        if (clear) launch_clear(*clear, stream);
        return;
    }
    const int threads = PRE_THREADS;
    const int blocks = (fp.P + threads - 1) / threads;
    const bool use_sh = want_color && in.colors_precomp == nullptr;
    const size_t smem = use_sh ? (size_t)(PRE_THREADS / 32) * 32 * ((fp.SHM * 3) | 1) * sizeof(float) : 0;
    static SmemOptIn opted;
    opt_in_shared_memory(preprocess_kernel, smem, opted);
    preprocess_kernel<<<blocks, threads, smem, stream>>>(fp, in, want_color, records, depths, tiles_touched, rects,
                                                         clamped, radii, order_keys, order_vals,
                                                         clear ? *clear : ClearList{});
}

// Frustum visibility only (reference checkFrustum, rasterizer_impl.cu:54-66).
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)P) return;
    const float3 p = ld3(means3D, idx);
    const float m2 = __ldg(view + 2), m6 = __ldg(view + 6), m10 = __ldg(view + 10), m14 = __ldg(view + 14);
    const float zv = m2 * p.x + m6 * p.y + m10 * p.z + m14;
    present[idx] = zv > 0.2f;
}

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t stream) {
    if (P == 0) return;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, view, present);
}

}  // namespace ggs
