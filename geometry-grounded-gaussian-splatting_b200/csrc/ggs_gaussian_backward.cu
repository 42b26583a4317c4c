// Per-Gaussian backward: one fused kernel for what the reference runs as computeCov2DCUDA
// (DGR/cuda_rasterizer/render_backward.cu:249-656, with the scale/rotation chain :193-244) followed
// by preprocessCUDA (:661-713, colour backward :56-191).  It consumes the 17-float accumulator row
// written by the tile backward and writes EVERY output gradient row exactly once (zeros for culled
// Gaussians), so the host never zero-fills the [P,SHM,3] gradient tensor separately.
//
// Memory access: the SH coefficients and their gradients are [P][SHM*3] rows of 192 bytes (SH3).  One thread
// per Gaussian reading / writing its own row would touch 32 different 128-byte lines per warp instruction, so
// each warp moves the contiguous slab of its 32 Gaussians through shared memory instead: coalesced 16-byte
// loads into rows with an odd stride (conflict-free per-thread access), the per-Gaussian arithmetic on the
// thread's own row, the gradient written back into the same row, and a coalesced store of the whole slab
// (zeros for culled Gaussians; their coefficients are never read).
#include "ggs_kernels.cuh"
#include "ggs_math.cuh"

namespace ggs {

__device__ __forceinline__ float3 ld3g(const float* p, size_t i) {
    return {__ldg(p + 3 * i), __ldg(p + 3 * i + 1), __ldg(p + 3 * i + 2)};
}
__device__ __forceinline__ void st3(float* p, size_t i, float a, float b, float c) {
    p[3 * i] = a;
    p[3 * i + 1] = b;
    p[3 * i + 2] = c;
}

// d normalize(v) / d v applied to dv (auxiliary.h:104-113)
__device__ __forceinline__ float3 dnormvdv(float3 v, float3 dv) {
    const float inv_norm = rnorm3df(v.x, v.y, v.z);
    const float3 vn = {v.x * inv_norm, v.y * inv_norm, v.z * inv_norm};
    float3 r;
    r.x = ((+1.f - vn.x * vn.x) * dv.x - vn.y * vn.x * dv.y - vn.z * vn.x * dv.z) * inv_norm;
    r.y = (-vn.x * vn.y * dv.x + (1.f - vn.y * vn.y) * dv.y - vn.z * vn.y * dv.z) * inv_norm;
    r.z = (-vn.x * vn.z * dv.x - vn.y * vn.z * dv.y + (1.f - vn.z * vn.z) * dv.z) * inv_norm;
    return r;
}

constexpr int GB_THREADS = 128;

#ifndef GGS_GB_MIN_BLOCKS
#define GGS_GB_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(GB_THREADS, GGS_GB_MIN_BLOCKS)
    gaussian_backward_kernel(const FrameParams fp, const GaussianInputs in, const int* __restrict__ radii,
                             const uint8_t* __restrict__ clamped, const double* __restrict__ grad_acc,
                             const GaussianGrads g, const uint32_t first_block) {
    extern __shared__ float s_slab[];  // [warps][32][RS]
    const size_t block = (size_t)blockIdx.x + first_block;  // a launch may cover a piece of the model (ggs_api.cu)
    const size_t idx = block * blockDim.x + threadIdx.x;
    const int SHM = fp.SHM, SGM = fp.SGM;
    const bool has_sh = in.shs != nullptr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool vis = idx < (size_t)fp.P && radii[idx] > 0;

    // ---- SH slab of this warp's 32 Gaussians -> shared memory ---------------------------------------
    const int F = SHM * 3;   // floats per Gaussian
    const int RS = F | 1;    // odd row stride: conflict-free when every lane walks its own row
    float* const swarp = s_slab + (size_t)warp * 32 * RS;
    float* const srow = swarp + lane * RS;
    const size_t g0 = block * blockDim.x + warp * 32;  // first Gaussian of the warp
    const int ng = g0 < (size_t)fp.P ? (int)min((size_t)32, (size_t)fp.P - g0) : 0;
    const uint32_t vm = __ballot_sync(0xffffffffu, vis);
    const bool vec4 = (F & 3) == 0;
    // Asynchronous: the rows are not needed before the colour section, behind the accumulator gather and the whole
    // geometry chain, and the lanes have diverged by then (culled ones are done) -- hence the mbarrier.  With the
    // copy in front as a blocking load the kernel ran at 57 % of HBM with 15 warps per SM and nothing in flight
    // during the arithmetic.
    __shared__ uint64_t s_rows_landed[GB_THREADS / 32];
    const bool rows_in_flight = has_sh && vm;
    if (rows_in_flight) {
        if (lane == 0) mbar_init(&s_rows_landed[warp], 32);
        __syncwarp();
        copy_rows_to_shared_async(in.shs + g0 * F, F, ng, vm, swarp, RS, lane);
        cp_async_arrive_on(&s_rows_landed[warp]);
    }

    float3 rgb_push = {0.f, 0.f, 0.f};  // dL/dRGB of this Gaussian after the clamp mask (0 when culled)
    if (!vis) {
        if (idx < (size_t)fp.P) {
            st3(g.dL_dmeans3D, idx, 0.f, 0.f, 0.f);
            st3(g.dL_dmeans2D, idx, 0.f, 0.f, 0.f);
            st3(g.dL_dcolors, idx, 0.f, 0.f, 0.f);
            if (g.dL_drgb) st3(g.dL_drgb, idx, 0.f, 0.f, 0.f);
            g.dL_dopacity[idx] = 0.f;
            st3(g.dL_dscales, idx, 0.f, 0.f, 0.f);
            reinterpret_cast<float4*>(g.dL_drotations)[idx] = {0.f, 0.f, 0.f, 0.f};
            if (g.dL_dsh)
                for (int k = 0; k < F; k++) srow[k] = 0.f;
            if (g.dL_dsg_axis) {
                for (int k = 0; k < SGM; k++) {
                    st3(g.dL_dsg_axis, idx * (size_t)SGM + k, 0.f, 0.f, 0.f);
                    st3(g.dL_dsg_color, idx * (size_t)SGM + k, 0.f, 0.f, 0.f);
                    g.dL_dsg_sharpness[idx * (size_t)SGM + k] = 0.f;
                }
            }
        }
    } else {
    // ---- gather the accumulator row -----------------------------------------------------------
    // 17 doubles (see ggs_common.cuh), rounded to float once
    float4 r0, r1, r2, r3, r4;
    {
        const double2* row = reinterpret_cast<const double2*>(grad_acc + idx * GRAD_STRIDE);
        double2 d[9];
#pragma unroll
        for (int k = 0; k < 9; k++) d[k] = __ldg(row + k);
        r0 = {(float)d[0].x, (float)d[0].y, (float)d[1].x, (float)d[1].y};
        r1 = {(float)d[2].x, (float)d[2].y, (float)d[3].x, (float)d[3].y};
        r2 = {(float)d[4].x, (float)d[4].y, (float)d[5].x, (float)d[5].y};
        r3 = {(float)d[6].x, (float)d[6].y, (float)d[7].x, (float)d[7].y};
        r4 = {(float)d[8].x, 0.f, 0.f, 0.f};
    }
    const float3 dL_dmean2D = {r0.x, r0.y, r0.z};
    const float4 dL_dconic = {r0.w, r1.x, r1.y, r1.z};
    const float3 dL_dcolor = {r1.w, r2.x, r2.y};
    const V3 dL_dnormal = {r2.z, r2.w, r3.x};
    const float h_x = fp.focal_x, h_y = fp.focal_y;
    float4 dL_dray_plane = {r3.y, r3.z, r3.w, r4.x};
    dL_dray_plane.x /= h_x;
    dL_dray_plane.y /= h_y;
    const float dL_dtc = dL_dray_plane.z;
    const float dL_drsigma = dL_dray_plane.w;

    const float* view_matrix = fp.view.view;
    const float3 mean = ld3g(in.means3D, idx);

    // ================= computeCov2D backward (render_backward.cu:290-655) =======================
    float3 t = xform_point_4x3(mean, view_matrix);
    const float rtc = rnorm3df(t.x, t.y, t.z);
    const float3 dL_dt = {t.x * rtc * dL_dtc, t.y * rtc * dL_dtc, t.z * rtc * dL_dtc};

    const float limx = 1.3f * fp.tanfovx;
    const float limy = 1.3f * fp.tanfovy;
    float u = t.x / t.z;
    float v = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, u)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, v)) * t.z;
    const float x_grad_mul = u < -limx || u > limx ? 0 : 1;
    const float y_grad_mul = v < -limy || v > limy ? 0 : 1;
    u = t.x / t.z;
    v = t.y / t.z;

    const M3 J = m3_cols(h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z), 0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z), 0,
                         0, 0);
    const M3 Wm = m3_cols(view_matrix[0], view_matrix[4], view_matrix[8], view_matrix[1], view_matrix[5],
                          view_matrix[9], view_matrix[2], view_matrix[6], view_matrix[10]);
    const M3 T = Wm * J;

    const float mod = fp.scale_modifier;
    const float3 sc = ld3g(in.scales, idx);
    const V3 scale_local = {mod * sc.x, mod * sc.y, mod * sc.z};
    const M3 S = m3_diag(scale_local.x, scale_local.y, scale_local.z);
    const M3 S_inv = m3_diag(__frcp_rn(scale_local.x), __frcp_rn(scale_local.y), __frcp_rn(scale_local.z));
    const float4 rot = __ldg(reinterpret_cast<const float4*>(in.rotations) + idx);
    const float qr = rot.x, qx = rot.y, qy = rot.z, qz = rot.w;
    const M3 R = m3_cols(1.f - 2.f * (qy * qy + qz * qz), 2.f * (qx * qy - qr * qz), 2.f * (qx * qz + qr * qy),
                         2.f * (qx * qy + qr * qz), 1.f - 2.f * (qx * qx + qz * qz), 2.f * (qy * qz - qr * qx),
                         2.f * (qx * qz - qr * qy), 2.f * (qy * qz + qr * qx), 1.f - 2.f * (qx * qx + qy * qy));

    M3 M = S * R * T;
    const M3 cov2D = transpose(M) * M;
    M = S * R;
    const M3 Vrk = transpose(M) * M;
    M3 M_inv = S_inv * R * Wm;
    const M3 cov_cam_inv = transpose(M_inv) * M_inv;
    M_inv = S_inv * R;
    const M3 Vrk_inv = transpose(M_inv) * M_inv;

    const float kernel_size = fp.kernel_size;
    const float det_0 = fmaxf(1e-6f, cov2D.c[0][0] * cov2D.c[1][1] - cov2D.c[0][1] * cov2D.c[0][1]);
    const float det_1 = fmaxf(
        1e-6f, (cov2D.c[0][0] + kernel_size) * (cov2D.c[1][1] + kernel_size) - cov2D.c[0][1] * cov2D.c[0][1]);
    const float coef = sqrtf(det_0 / det_1);

    const V3 uvh = {u, v, 1.f};
    const V3 uvh_m = cov_cam_inv * uvh;
    const float u2 = u * u, v2 = v * v, uv = u * v;

    M3 dL_dVrk;
    float dL_du, dL_dv, dL_dz;
    {
        const float vb = dot(uvh_m, uvh);
        const float l = norm3df(t.x, t.y, t.z);
        const M3 nJ = m3_cols(1 / t.z, 0.0f, -(t.x) / (t.z * t.z), 0.0f, 1 / t.z, -(t.y) / (t.z * t.z), t.x / l,
                              t.y / l, t.z / l);
        const M3 nJ_inv = m3_cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -u, -v, 0);
        const float clamp_vb = fmaxf(vb, 1e-7f);
        const float ray_len2 = u2 + v2 + 1;
        const float ray_len_inv = rsqrtf(ray_len2);
        const float factor_normal = l / ray_len2;
        const V3 uvh_m_vb = uvh_m / clamp_vb;
        const V3 plane = nJ_inv * uvh_m_vb;
        const V3 ray_normal_vector = {-plane.x * factor_normal, -plane.y * factor_normal, -1};
        const V3 cam_normal_vector = nJ * ray_normal_vector;
        const V3 normal_vector = normalize(cam_normal_vector);
        const float rlv = rnorm3df(normal_vector.x, normal_vector.y, normal_vector.z);
        const V3 dL_dcam_normal_vector = (dL_dnormal - normal_vector * dot(normal_vector, dL_dnormal)) * rlv;
        const V3 dL_dray_normal_vector = transpose(nJ) * dL_dcam_normal_vector;
        const M3 dL_dnJ = outer(dL_dcam_normal_vector, ray_normal_vector);
        const float dL_dfactor_normal = plane.x * (-dL_dray_normal_vector.x + dL_dray_plane.x) +
                                        plane.y * (-dL_dray_normal_vector.y + dL_dray_plane.y);
        const float dL_dplane_x = (-dL_dray_normal_vector.x + dL_dray_plane.x) * factor_normal;
        const float dL_dplane_y = (-dL_dray_normal_vector.y + dL_dray_plane.y) * factor_normal;
        const V3 dL_dplane_append = {dL_dplane_x, dL_dplane_y, 0.f};
        const float aux = dL_dplane_x * plane.x + dL_dplane_y * plane.y;
        const V3 W_uvh = Wm * uvh;
        const V3 dL_duvh_plane =
            (2 * (-aux)) * uvh_m_vb + ((cov_cam_inv / clamp_vb) * transpose(nJ_inv)) * dL_dplane_append;
        const float rsigmat = sqrtf(vb / ray_len2);
        const float dL_dray_len2_sigma_x2 = -dL_drsigma * rsigmat / ray_len2;
        const float dL_du_sigma = dL_dray_len2_sigma_x2 * u;
        const float dL_dv_sigma = dL_dray_len2_sigma_x2 * v;

        const float aux_nJ = (-dL_dnJ.c[2][0] * u - dL_dnJ.c[2][1] * v - dL_dnJ.c[2][2]) / ray_len2 * ray_len_inv;
        const float dL_du_nJ = -dL_dnJ.c[0][2] / t.z + dL_dnJ.c[2][0] * ray_len_inv + aux_nJ * u;
        const float dL_dv_nJ = -dL_dnJ.c[1][2] / t.z + dL_dnJ.c[2][1] * ray_len_inv + aux_nJ * v;
        const float dL_dz_nJ =
            (dL_dnJ.c[0][0] + dL_dnJ.c[1][1] - dL_dnJ.c[0][2] * u - dL_dnJ.c[1][2] * v) / (-t.z * t.z);

        const M3 dL_dnJ_inv = outer(dL_dplane_append, uvh_m_vb);
        const float dL_du_plane = dL_duvh_plane.x + (dL_dnJ_inv.c[0][1] + dL_dnJ_inv.c[1][0]) * (-v) +
                                  2 * dL_dnJ_inv.c[1][1] * u - dL_dnJ_inv.c[2][0];
        const float dL_dv_plane = dL_duvh_plane.y + (dL_dnJ_inv.c[0][1] + dL_dnJ_inv.c[1][0]) * (-u) +
                                  2 * dL_dnJ_inv.c[0][0] * v - dL_dnJ_inv.c[2][1];

        const float aux_factor = dL_dfactor_normal * (-t.z / ray_len2 * ray_len_inv);
        const float dL_du_factor = aux_factor * u;
        const float dL_dv_factor = aux_factor * v;
        const float dL_dz_factor = dL_dfactor_normal * ray_len_inv;

        dL_du = dL_du_nJ + dL_du_plane + dL_du_factor + dL_du_sigma;
        dL_dv = dL_dv_nJ + dL_dv_plane + dL_dv_factor + dL_dv_sigma;
        dL_dz = dL_dz_nJ + dL_dz_factor;

        const float dL_dvb_xvb = -aux + dL_drsigma * 0.5f * rsigmat;
        // scale/rotation path is always well conditioned (render_backward.cu:362)
        dL_dVrk = neg(outer(Vrk_inv * W_uvh,
                            Vrk_inv * ((Wm * transpose(nJ_inv)) * dL_dplane_append + W_uvh * dL_dvb_xvb)));
        dL_dVrk = dL_dVrk / vb;
    }

    const float opacity = __ldg(in.opacities + idx);
    const float dL_dcoef = dL_dconic.w * opacity;
    const float dL_dsqrtcoef = dL_dcoef * 0.5f / (coef + 1e-6f);
    const float dL_ddet0 = dL_dsqrtcoef / det_1;
    const float dL_ddet1 = -dL_ddet0 * coef;
    const float dcoef_da = dL_ddet0 * cov2D.c[1][1] + dL_ddet1 * (cov2D.c[1][1] + kernel_size);
    const float dcoef_db = (-2.f * cov2D.c[0][1]) * (dL_ddet0 + dL_ddet1);
    const float dcoef_dc = dL_ddet0 * cov2D.c[0][0] + dL_ddet1 * (cov2D.c[0][0] + kernel_size);
    const float a = cov2D.c[0][0] + kernel_size;
    const float b = cov2D.c[0][1];
    const float c = cov2D.c[1][1] + kernel_size;

    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 1e-7f);

    float dL_dcov_local[6];
    float dL_dopacity = 0.f;
    if (denom2inv != 0) {
        dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
        dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
        dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
        dL_da += dcoef_da;
        dL_dc += dcoef_dc;
        dL_db += dcoef_db;
        dL_dopacity = dL_dconic.w * coef;

        dL_dcov_local[0] =
            (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
        dL_dcov_local[3] =
            (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
        dL_dcov_local[5] =
            (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
        dL_dcov_local[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db +
                           2 * T.c[1][0] * T.c[1][1] * dL_dc;
        dL_dcov_local[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db +
                           2 * T.c[1][0] * T.c[1][2] * dL_dc;
        dL_dcov_local[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db +
                           2 * T.c[1][1] * T.c[1][2] * dL_dc;
    } else {
#pragma unroll
        for (int i = 0; i < 6; i++) dL_dcov_local[i] = 0;
    }
    dL_dcov_local[0] += dL_dVrk.c[0][0];
    dL_dcov_local[3] += dL_dVrk.c[1][1];
    dL_dcov_local[5] += dL_dVrk.c[2][2];
    dL_dcov_local[1] += dL_dVrk.c[0][1] + dL_dVrk.c[1][0];
    dL_dcov_local[2] += dL_dVrk.c[0][2] + dL_dVrk.c[2][0];
    dL_dcov_local[4] += dL_dVrk.c[1][2] + dL_dVrk.c[2][1];

    // ---- scale / rotation chain (render_backward.cu:193-244); scale_local already holds mod*scale
    float3 dL_dscale;
    float4 dL_dq;
    {
        const M3 Ms = S * R;
        const M3 dL_dSigma = m3_cols(dL_dcov_local[0], 0.5f * dL_dcov_local[1], 0.5f * dL_dcov_local[2],
                                     0.5f * dL_dcov_local[1], dL_dcov_local[3], 0.5f * dL_dcov_local[4],
                                     0.5f * dL_dcov_local[2], 0.5f * dL_dcov_local[4], dL_dcov_local[5]);
        const M3 dL_dM = (Ms * 2.0f) * dL_dSigma;
        const M3 Rt = transpose(R);
        M3 dL_dMt = transpose(dL_dM);
        dL_dscale.x = Rt.c[0][0] * dL_dMt.c[0][0] + Rt.c[0][1] * dL_dMt.c[0][1] + Rt.c[0][2] * dL_dMt.c[0][2];
        dL_dscale.y = Rt.c[1][0] * dL_dMt.c[1][0] + Rt.c[1][1] * dL_dMt.c[1][1] + Rt.c[1][2] * dL_dMt.c[1][2];
        dL_dscale.z = Rt.c[2][0] * dL_dMt.c[2][0] + Rt.c[2][1] * dL_dMt.c[2][1] + Rt.c[2][2] * dL_dMt.c[2][2];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            dL_dMt.c[0][k] *= scale_local.x;
            dL_dMt.c[1][k] *= scale_local.y;
            dL_dMt.c[2][k] *= scale_local.z;
        }
        const float r = qr, x = qx, y = qy, z = qz;
        dL_dq.x = 2 * z * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * y * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) +
                  2 * x * (dL_dMt.c[1][2] - dL_dMt.c[2][1]);
        dL_dq.y = 2 * y * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * z * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) +
                  2 * r * (dL_dMt.c[1][2] - dL_dMt.c[2][1]) - 4 * x * (dL_dMt.c[2][2] + dL_dMt.c[1][1]);
        dL_dq.z = 2 * x * (dL_dMt.c[1][0] + dL_dMt.c[0][1]) + 2 * r * (dL_dMt.c[2][0] - dL_dMt.c[0][2]) +
                  2 * z * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * y * (dL_dMt.c[2][2] + dL_dMt.c[0][0]);
        dL_dq.w = 2 * r * (dL_dMt.c[0][1] - dL_dMt.c[1][0]) + 2 * x * (dL_dMt.c[2][0] + dL_dMt.c[0][2]) +
                  2 * y * (dL_dMt.c[1][2] + dL_dMt.c[2][1]) - 4 * z * (dL_dMt.c[1][1] + dL_dMt.c[0][0]);
    }

    // ---- gradient w.r.t. T = W*J, then J, then the view-space point -----------------------------
    const float dL_dT00 = 2 * (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT01 = 2 * (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT02 = 2 * (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_da +
                          (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_db;
    const float dL_dT10 = 2 * (T.c[1][0] * Vrk.c[0][0] + T.c[1][1] * Vrk.c[0][1] + T.c[1][2] * Vrk.c[0][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[0][0] + T.c[0][1] * Vrk.c[0][1] + T.c[0][2] * Vrk.c[0][2]) * dL_db;
    const float dL_dT11 = 2 * (T.c[1][0] * Vrk.c[1][0] + T.c[1][1] * Vrk.c[1][1] + T.c[1][2] * Vrk.c[1][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[1][0] + T.c[0][1] * Vrk.c[1][1] + T.c[0][2] * Vrk.c[1][2]) * dL_db;
    const float dL_dT12 = 2 * (T.c[1][0] * Vrk.c[2][0] + T.c[1][1] * Vrk.c[2][1] + T.c[1][2] * Vrk.c[2][2]) * dL_dc +
                          (T.c[0][0] * Vrk.c[2][0] + T.c[0][1] * Vrk.c[2][1] + T.c[0][2] * Vrk.c[2][2]) * dL_db;

    const float dL_dJ00 = Wm.c[0][0] * dL_dT00 + Wm.c[0][1] * dL_dT01 + Wm.c[0][2] * dL_dT02;
    const float dL_dJ02 = Wm.c[2][0] * dL_dT00 + Wm.c[2][1] * dL_dT01 + Wm.c[2][2] * dL_dT02;
    const float dL_dJ11 = Wm.c[1][0] * dL_dT10 + Wm.c[1][1] * dL_dT11 + Wm.c[1][2] * dL_dT12;
    const float dL_dJ12 = Wm.c[2][0] * dL_dT10 + Wm.c[2][1] * dL_dT11 + Wm.c[2][2] * dL_dT12;

    const float tz = 1.f / t.z;
    const float tz2 = tz * tz;
    const float tz3 = tz2 * tz;
    const float dL_dtx = x_grad_mul * (-h_x * tz2 * dL_dJ02 + dL_du * tz);
    const float dL_dty = y_grad_mul * (-h_y * tz2 * dL_dJ12 + dL_dv * tz);
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + ((1 + x_grad_mul) * h_x * t.x) * tz3 * dL_dJ02 +
                         ((1 + y_grad_mul) * h_y * t.y) * tz3 * dL_dJ12 -
                         (x_grad_mul * dL_du * t.x + y_grad_mul * dL_dv * t.y) * tz2 + dL_dz;
    float3 dL_dmean =
        xform_vec_4x3_transpose({dL_dtx + dL_dt.x, dL_dty + dL_dt.y, dL_dtz + dL_dt.z}, view_matrix);

    // ================= projection part (render_backward.cu:689-705) =============================
    {
        const float* proj = fp.view.proj;
        const float3 m = mean;
        const float4 m_hom = xform_point_4x4(m, proj);
        const float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
        dL_dmean.x += (proj[0] * m_w - proj[3] * mul1) * dL_dmean2D.x + (proj[1] * m_w - proj[3] * mul2) * dL_dmean2D.y;
        dL_dmean.y += (proj[4] * m_w - proj[7] * mul1) * dL_dmean2D.x + (proj[5] * m_w - proj[7] * mul2) * dL_dmean2D.y;
        dL_dmean.z +=
            (proj[8] * m_w - proj[11] * mul1) * dL_dmean2D.x + (proj[9] * m_w - proj[11] * mul2) * dL_dmean2D.y;
    }

    // ================= colour backward (render_backward.cu:56-191) ==============================
    if (has_sh) {
        mbar_wait(&s_rows_landed[warp], 0);  // every lane's share of the warp's SH rows is in shared memory
        const float3 dir_orig = {mean.x - fp.view.campos[0], mean.y - fp.view.campos[1], mean.z - fp.view.campos[2]};
        const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
        const float3 dir = {dir_orig.x / len, dir_orig.y / len, dir_orig.z / len};
        // the thread's SH row lives in shared memory: every coefficient is read before the gradients overwrite it
        auto Csh = [&](int k) -> float3 { return float3{srow[3 * k], srow[3 * k + 1], srow[3 * k + 2]}; };
        const uint8_t mask = clamped[idx];
        float3 dL_dRGB = dL_dcolor;
        dL_dRGB.x *= (mask & 1) ? 0 : 1;
        dL_dRGB.y *= (mask & 2) ? 0 : 1;
        dL_dRGB.z *= (mask & 4) ? 0 : 1;

        float3 dRGBdx = {0, 0, 0}, dRGBdy = {0, 0, 0}, dRGBdz = {0, 0, 0};
        const float x = dir.x, y = dir.y, z = dir.z;
        const float xx = x * x, yy = y * y, zz = z * z;
        const float xy = x * y, yz = y * z, xz = x * z;
        auto axpy = [](float3& acc, float w, const float3& c) {
            acc.x += w * c.x;
            acc.y += w * c.y;
            acc.z += w * c.z;
        };
        // ---- d colour / d direction: same term order as render_backward.cu:83-166 ----
        if (fp.SHD > 0) {
            const float3 s1 = Csh(1), s2 = Csh(2), s3 = Csh(3);
            axpy(dRGBdx, -kSH_C1, s3);
            axpy(dRGBdy, -kSH_C1, s1);
            axpy(dRGBdz, kSH_C1, s2);
            if (fp.SHD > 1) {
                const float3 s4 = Csh(4), s5 = Csh(5), s6 = Csh(6), s7 = Csh(7), s8 = Csh(8);
                axpy(dRGBdx, kSH_C2[0] * y, s4);
                axpy(dRGBdx, kSH_C2[2] * 2.f * -x, s6);
                axpy(dRGBdx, kSH_C2[3] * z, s7);
                axpy(dRGBdx, kSH_C2[4] * 2.f * x, s8);
                axpy(dRGBdy, kSH_C2[0] * x, s4);
                axpy(dRGBdy, kSH_C2[1] * z, s5);
                axpy(dRGBdy, kSH_C2[2] * 2.f * -y, s6);
                axpy(dRGBdy, kSH_C2[4] * 2.f * -y, s8);
                axpy(dRGBdz, kSH_C2[1] * y, s5);
                axpy(dRGBdz, kSH_C2[2] * 2.f * 2.f * z, s6);
                axpy(dRGBdz, kSH_C2[3] * x, s7);
                if (fp.SHD > 2) {
                    const float3 s9 = Csh(9), s10 = Csh(10), s11 = Csh(11), s12 = Csh(12), s13 = Csh(13),
                                 s14 = Csh(14), s15 = Csh(15);
                    axpy(dRGBdx, kSH_C3[0] * 3.f * 2.f * xy, s9);
                    axpy(dRGBdx, kSH_C3[1] * yz, s10);
                    axpy(dRGBdx, kSH_C3[2] * -2.f * xy, s11);
                    axpy(dRGBdx, kSH_C3[3] * -3.f * 2.f * xz, s12);
                    axpy(dRGBdx, kSH_C3[4] * (-3.f * xx + 4.f * zz - yy), s13);
                    axpy(dRGBdx, kSH_C3[5] * 2.f * xz, s14);
                    axpy(dRGBdx, kSH_C3[6] * 3.f * (xx - yy), s15);
                    axpy(dRGBdy, kSH_C3[0] * 3.f * (xx - yy), s9);
                    axpy(dRGBdy, kSH_C3[1] * xz, s10);
                    axpy(dRGBdy, kSH_C3[2] * (-3.f * yy + 4.f * zz - xx), s11);
                    axpy(dRGBdy, kSH_C3[3] * -3.f * 2.f * yz, s12);
                    axpy(dRGBdy, kSH_C3[4] * -2.f * xy, s13);
                    axpy(dRGBdy, kSH_C3[5] * -2.f * yz, s14);
                    axpy(dRGBdy, kSH_C3[6] * -3.f * 2.f * xy, s15);
                    axpy(dRGBdz, kSH_C3[1] * xy, s10);
                    axpy(dRGBdz, kSH_C3[2] * 4.f * 2.f * yz, s11);
                    axpy(dRGBdz, kSH_C3[3] * 3.f * (2.f * zz - xx - yy), s12);
                    axpy(dRGBdz, kSH_C3[4] * 4.f * 2.f * xz, s13);
                    axpy(dRGBdz, kSH_C3[5] * (xx - yy), s14);
                }
            }
        }
        if (g.dL_drgb) st3(g.dL_drgb, idx, dL_dRGB.x, dL_dRGB.y, dL_dRGB.z);
        rgb_push = dL_dRGB;
        // ---- d colour / d SH coefficient: overwrite the row with the gradient (render_backward.cu:76-160) ----
        // (skipped when the caller forms the SH gradient from dL_drgb itself: view-parallel training)
        if (g.dL_dsh) {
        auto put = [&](int k, float w) {
            srow[3 * k] = w * dL_dRGB.x;
            srow[3 * k + 1] = w * dL_dRGB.y;
            srow[3 * k + 2] = w * dL_dRGB.z;
        };
        put(0, kSH_C0);
        int written = 1;
        if (fp.SHD > 0) {
            put(1, -kSH_C1 * y);
            put(2, kSH_C1 * z);
            put(3, -kSH_C1 * x);
            written = 4;
            if (fp.SHD > 1) {
                put(4, kSH_C2[0] * xy);
                put(5, kSH_C2[1] * yz);
                put(6, kSH_C2[2] * (2.f * zz - xx - yy));
                put(7, kSH_C2[3] * xz);
                put(8, kSH_C2[4] * (xx - yy));
                written = 9;
                if (fp.SHD > 2) {
                    put(9, kSH_C3[0] * y * (3.f * xx - yy));
                    put(10, kSH_C3[1] * xy * z);
                    put(11, kSH_C3[2] * y * (4.f * zz - xx - yy));
                    put(12, kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                    put(13, kSH_C3[4] * x * (4.f * zz - xx - yy));
                    put(14, kSH_C3[5] * z * (xx - yy));
                    put(15, kSH_C3[6] * x * (xx - 3.f * yy));
                    written = 16;
                }
            }
        }
        for (int k = written * 3; k < F; k++) srow[k] = 0.f;
        }

        float3 dL_ddir = {dRGBdx.x * dL_dRGB.x + dRGBdx.y * dL_dRGB.y + dRGBdx.z * dL_dRGB.z,
                          dRGBdy.x * dL_dRGB.x + dRGBdy.y * dL_dRGB.y + dRGBdy.z * dL_dRGB.z,
                          dRGBdz.x * dL_dRGB.x + dRGBdz.y * dL_dRGB.y + dRGBdz.z * dL_dRGB.z};
        for (int k = 0; k < SGM; k++) {
            const size_t o = idx * (size_t)SGM + k;
            if (k < fp.SGD) {
                const float3 axis = ld3g(in.sg_axis, o);
                const float sharpness = __ldg(in.sg_sharpness + o);
                const float3 gcolor = ld3g(in.sg_color, o);
                const float aux = (axis.x * dir.x + axis.y * dir.y + axis.z * dir.z) - 1.0f;
                const float gaussian = expf(sharpness * aux);
                st3(g.dL_dsg_color, o, dL_dRGB.x * gaussian, dL_dRGB.y * gaussian, dL_dRGB.z * gaussian);
                const float dL_dgaussian = gcolor.x * dL_dRGB.x + gcolor.y * dL_dRGB.y + gcolor.z * dL_dRGB.z;
                const float dL_dexponent = dL_dgaussian * gaussian;
                g.dL_dsg_sharpness[o] = dL_dexponent * aux;
                const float dL_daux = dL_dexponent * sharpness;
                st3(g.dL_dsg_axis, o, dL_daux * dir.x, dL_daux * dir.y, dL_daux * dir.z);
                dL_ddir.x += dL_daux * axis.x;
                dL_ddir.y += dL_daux * axis.y;
                dL_ddir.z += dL_daux * axis.z;
            } else {
                st3(g.dL_dsg_color, o, 0.f, 0.f, 0.f);
                st3(g.dL_dsg_axis, o, 0.f, 0.f, 0.f);
                g.dL_dsg_sharpness[o] = 0.f;
            }
        }
        const float3 dm = dnormvdv(dir_orig, dL_ddir);
        dL_dmean.x += dm.x;
        dL_dmean.y += dm.y;
        dL_dmean.z += dm.z;
    }

    st3(g.dL_dmeans3D, idx, dL_dmean.x, dL_dmean.y, dL_dmean.z);
    st3(g.dL_dmeans2D, idx, dL_dmean2D.x, dL_dmean2D.y, dL_dmean2D.z);
    st3(g.dL_dcolors, idx, dL_dcolor.x, dL_dcolor.y, dL_dcolor.z);
    g.dL_dopacity[idx] = dL_dopacity;
    st3(g.dL_dscales, idx, dL_dscale.x, dL_dscale.y, dL_dscale.z);
    reinterpret_cast<float4*>(g.dL_drotations)[idx] = dL_dq;
    }  // vis

    // ---- view-parallel exchange: this view's dL/dRGB straight into slot `rank` of EVERY rank's colour stage ------
    // (ggs_rasterize_backward_peer with GgsPeerExchange.fuse_rgb_push: the warp's 32 rows are one 384-byte segment,
    // staged in the slab -- free here, the SH gradient is formed elsewhere in this mode -- and stored as 24 x 16 bytes
    // through the NVSwitch multicast mapping; the camera position goes with block 0)
    if (g.mc_rgb) {
        __syncwarp();  // every lane is done reading its SH row
        swarp[3 * lane] = rgb_push.x;
        swarp[3 * lane + 1] = rgb_push.y;
        swarp[3 * lane + 2] = rgb_push.z;
        __syncwarp();
        if (ng == 32) {
            if (lane < 24) {
                const float4 v = {swarp[4 * lane], swarp[4 * lane + 1], swarp[4 * lane + 2], swarp[4 * lane + 3]};
                float4* dst = reinterpret_cast<float4*>(g.mc_rgb) + g0 * 3 / 4 + lane;  // g0 % 32 == 0: aligned
                asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(v.x),
                             "f"(v.y), "f"(v.z), "f"(v.w)
                             : "memory");
            }
        } else {
            for (int e = lane; e < ng * 3; e += 32)
                asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(g.mc_rgb + g0 * 3 + e), "f"(swarp[e])
                             : "memory");
        }
        if (block == 0 && threadIdx.x < 3)
            asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(g.mc_campos + threadIdx.x),
                         "f"(fp.view.campos[threadIdx.x])
                         : "memory");
    }

    // ---- gradient slab of the warp's 32 Gaussians: shared memory -> global, coalesced ---------------
    if (g.dL_dsh) {
        __syncwarp();
        if (vec4) {
            store_rows_from_shared(reinterpret_cast<float4*>(g.dL_dsh + g0 * F), F >> 2, ng, swarp, RS, lane);
        } else {
            float* dst = g.dL_dsh + g0 * F;
            int gi = 0, rem = lane;
            for (int e = lane; e < ng * F; e += 32, rem += 32) {
                while (rem >= F) {
                    rem -= F;
                    gi++;
                }
                dst[e] = swarp[gi * RS + rem];
            }
        }
    }
}

void launch_gaussian_backward(const FrameParams& fp, const GaussianInputs& in, const int* radii,
                              const uint8_t* clamped, const double* grad_acc, const GaussianGrads& g,
                              cudaStream_t stream, int64_t first, int64_t last) {
    if (last < 0 || last > fp.P) last = fp.P;
    if (fp.P == 0 || first >= last) return;
    const size_t smem = (size_t)(GB_THREADS / 32) * 32 * ((fp.SHM * 3) | 1) * sizeof(float);
    static SmemOptIn opted;
    opt_in_shared_memory(gaussian_backward_kernel, smem, opted);
    // `first` is a multiple of GB_THREADS; a piece that ends inside a block ends the model (the kernel checks fp.P)
    const uint32_t first_block = (uint32_t)(first / GB_THREADS);
    const uint32_t blocks = (uint32_t)((last - first + GB_THREADS - 1) / GB_THREADS);
    gaussian_backward_kernel<<<blocks, GB_THREADS, smem, stream>>>(fp, in, radii, clamped, grad_acc, g, first_block);
}

}  // namespace ggs
