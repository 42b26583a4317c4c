// Stage 1 of the forward: per-Gaussian projection, 2D covariance + Mip 2D filter, ray-plane /
// normal, SH + spherical-Gaussian colour.  One thread per Gaussian, inputs read through the
// read-only path, outputs written as one 64-byte SplatRecord plus SoA side arrays.
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
__device__ __forceinline__ float3 eval_color(size_t idx, const FrameParams& fp, const float3 pos, const float* shs,
                                             const float* sg_axis, const float* sg_sharpness, const float* sg_color,
                                             uint8_t& clamp_mask) {
    float3 dir = {pos.x - fp.view.campos[0], pos.y - fp.view.campos[1], pos.z - fp.view.campos[2]};
    float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir = {dir.x / len, dir.y / len, dir.z / len};

    const float* sh = shs + idx * (size_t)fp.SHM * 3;
    auto C = [&](int k) -> float3 { return ld3(sh, k); };
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

__global__ void __launch_bounds__(256) preprocess_kernel(const FrameParams fp, const GaussianInputs in,
                                                         const bool want_color, SplatRecord* __restrict__ records,
                                                         float* __restrict__ depths,
                                                         uint32_t* __restrict__ tiles_touched,
                                                         uint2* __restrict__ rects, uint8_t* __restrict__ clamped,
                                                         int* __restrict__ radii) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)fp.P) return;

    radii[idx] = 0;
    tiles_touched[idx] = 0;

    const float* viewmatrix = fp.view.view;
    const float3 p_orig = ld3(in.means3D, idx);
    // near cull (auxiliary.h:133-153)
    const float3 p_view = xform_point_4x3(p_orig, viewmatrix);
    if (p_view.z <= 0.2f) return;

    const float4 p_hom = xform_point_4x4(p_orig, fp.view.proj);
    const float p_w = 1.0f / (p_hom.w + 0.0000001f);
    const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

    // ---- 2D covariance, ray plane, normal (render_forward.cu:81-243) --------------------------
    const float focal_x = fp.focal_x, focal_y = fp.focal_y;
    float3 t = xform_point_4x3(p_orig, viewmatrix);
    const float tc = norm3df(t.x, t.y, t.z);

    const float limx = 1.3f * fp.tanfovx;
    const float limy = 1.3f * fp.tanfovy;
    float u = t.x / t.z;
    float v = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, u)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, v)) * t.z;
    u = t.x / t.z;
    v = t.y / t.z;

    const M3 J = m3_cols(focal_x / t.z, 0.f, -(focal_x * t.x) / (t.z * t.z), 0.f, focal_y / t.z,
                         -(focal_y * t.y) / (t.z * t.z), 0.f, 0.f, 0.f);
    const M3 Wm = m3_cols(viewmatrix[0], viewmatrix[4], viewmatrix[8], viewmatrix[1], viewmatrix[5], viewmatrix[9],
                          viewmatrix[2], viewmatrix[6], viewmatrix[10]);
    const M3 T = Wm * J;

    const float mod = fp.scale_modifier;
    const float3 sc = ld3(in.scales, idx);
    const float scale_local[3] = {mod * sc.x, mod * sc.y, mod * sc.z};
    const M3 S = m3_diag(scale_local[0], scale_local[1], scale_local[2]);
    const M3 S_inv = m3_diag(__frcp_rn(scale_local[0]), __frcp_rn(scale_local[1]), __frcp_rn(scale_local[2]));

    const float4 rot = __ldg(reinterpret_cast<const float4*>(in.rotations) + idx);
    const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    const M3 R = m3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                         2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                         2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));

    const M3 M = S * R * T;
    const M3 cov = transpose(M) * M;
    const M3 M_inv = S_inv * R * Wm;
    const M3 cov_cam_inv = transpose(M_inv) * M_inv;

    const float kernel_size = fp.kernel_size;
    const float cov2D0 = float(cov.c[0][0] + kernel_size);
    const float cov2D1 = float(cov.c[0][1]);
    const float cov2D2 = float(cov.c[1][1] + kernel_size);
    const float det_0 = fmaxf(1e-6f, cov.c[0][0] * cov.c[1][1] - cov.c[0][1] * cov.c[0][1]);
    const float det_1 =
        fmaxf(1e-6f, (cov.c[0][0] + kernel_size) * (cov.c[1][1] + kernel_size) - cov.c[0][1] * cov.c[0][1]);
    const float coef = sqrtf(det_0 / det_1);

    const V3 uvh = {u, v, 1.f};
    const V3 uvh_m = cov_cam_inv * uvh;

    float4 ray_plane;
    float3 normal;
    {
        const float u2 = u * u, v2 = v * v, uv = u * v;
        const float l = norm3df(t.x, t.y, t.z);
        const M3 nJ_inv = m3_cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -u, -v, 0);
        const float vb = dot(uvh_m, uvh);
        const float ray_len2 = u2 + v2 + 1;
        const float factor_normal = l / ray_len2;
        const V3 plane = nJ_inv * (uvh_m / vb);
        const float rsigmat = sqrtf(vb / ray_len2);  // scale/rotation path is always "well conditioned"
        ray_plane = {plane.x * factor_normal / focal_x, plane.y * factor_normal / focal_y, tc, rsigmat};
        const V3 ray_normal_vector = {-plane.x * factor_normal, -plane.y * factor_normal, -1};
        const M3 nJ = m3_cols(1 / t.z, 0.0f, -(t.x) / (t.z * t.z), 0.0f, 1 / t.z, -(t.y) / (t.z * t.z), t.x / l,
                              t.y / l, t.z / l);
        const V3 cam_normal_vector = nJ * ray_normal_vector;
        const V3 nv = normalize(cam_normal_vector);
        normal = {nv.x, nv.y, nv.z};
    }

    // ---- conic, radius, tile rectangle (render_forward.cu:347-368) ----------------------------
    const float3 covv = {cov2D0, cov2D1, cov2D2};
    const float det = (covv.x * covv.z - covv.y * covv.y);
    if (det == 0.0f) return;
    const float det_inv = 1.f / det;
    const float3 conic = {covv.z * det_inv, -covv.y * det_inv, covv.x * det_inv};

    const float mid = 0.5f * (covv.x + covv.z);
    const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
    const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
    const float2 point_image = {ndc_to_pix(p_proj.x, fp.W), ndc_to_pix(p_proj.y, fp.H)};
    uint2 rect_min, rect_max;
    tile_rect(point_image, my_radius, rect_min, rect_max, fp.gx, fp.gy);
    if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) return;

    // ---- colour -------------------------------------------------------------------------------
    float3 rgb = {0.f, 0.f, 0.f};
    uint8_t mask = 0;
    if (want_color) {
        if (in.colors_precomp != nullptr)
            rgb = ld3(in.colors_precomp, idx);
        else
            rgb = eval_color(idx, fp, p_orig, in.shs, in.sg_axis, in.sg_sharpness, in.sg_color, mask);
    }

    depths[idx] = norm3df(p_view.x, p_view.y, p_view.z);
    radii[idx] = my_radius;
    clamped[idx] = mask;
    rects[idx] = {rect_min.x | (rect_min.y << 16), rect_max.x | (rect_max.y << 16)};
    tiles_touched[idx] = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);

    const float opac = __ldg(in.opacities + idx) * coef;
    SplatRecord rec;
    rec.q0 = {point_image.x, point_image.y, conic.x, conic.y};
    rec.q1 = {conic.z, opac, ray_plane.x, ray_plane.y};
    rec.q2 = {ray_plane.z, ray_plane.w, rgb.x, rgb.y};
    rec.q3 = {rgb.z, normal.x, normal.y, normal.z};
    float4* dst = reinterpret_cast<float4*>(records + idx);
    dst[0] = rec.q0;
    dst[1] = rec.q1;
    dst[2] = rec.q2;
    dst[3] = rec.q3;
}

void launch_preprocess(const FrameParams& fp, const GaussianInputs& in, bool want_color, SplatRecord* records,
                       float* depths, uint32_t* tiles_touched, uint2* rects, uint8_t* clamped, int* radii,
                       cudaStream_t stream) {
    if (fp.P == 0) return;
    const int threads = 256;
    const int blocks = (fp.P + threads - 1) / threads;
    preprocess_kernel<<<blocks, threads, 0, stream>>>(fp, in, want_color, records, depths, tiles_touched, rects,
                                                      clamped, radii);
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
