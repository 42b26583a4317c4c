// Patch NCC between a reference view and a neighbour view under the plane-induced homography of one pixel
// (SURVEY.md 8(f) N4).  Follows the arithmetic of forward_mode_differentiation_CUDA<3, true>
// (submodules/warp-patch-ncc/cuda_warp_patch_ncc/warp_patch_ncc_impl.cu:18-262): 7x7 samples at half-pixel steps,
// forward-mode derivative of the NCC with respect to the pixel's depth and normal accumulated next to the value.
//
// What is ours: the reference-view pass comes first (25 pixels loaded once into registers, its 49 samples parked in a
// per-thread scratch column, its two sums), so that a patch whose reference variance is too small stops before any
// neighbour pixel is touched; the neighbour pass is a compact rolled loop over rows that stops at the first sample
// warped outside the image instead of sampling all 49 positions and discarding them.  The accumulation order of
// every sum is the reference's (dv outer, du inner, one accumulator each), so results round alike.
//
// The routine is host/device so that the very code the kernel runs can be checked on a CPU by the test suite
// (tests/test_patch_ncc_cpu.py builds it with g++); the product only ever calls it from the kernel.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define GGS_NCC_HD __host__ __device__ __forceinline__
#else
#define GGS_NCC_HD inline
#endif

namespace ggs {
namespace ncc {

constexpr int RADIUS = 3;                         // samples per side of the centre
constexpr int LINE = 2 * RADIUS + 1;              // 7
constexpr int SAMPLES = LINE * LINE;              // 49
constexpr int REACH = 2;                          // ceil(RADIUS / 2) pixels of the reference image per side
constexpr int PATCH = 2 * REACH + 1;              // 5
constexpr float HALF_EXTENT = 0.5f * RADIUS;      // 1.5 px

struct Pair {  // both views: rotation (column-major) and translation reference -> neighbour, intrinsics, sizes
    float R[9];
    float T[3];
    float fx_r, fy_r, cx_r, cy_r;
    float fx_n, fy_n, cx_n, cy_n;
    int h_r, w_r, h_n, w_n;
};

struct PointResult {
    float ncc, d_depth, d_nx, d_ny, d_nz;
    bool valid;
};

GGS_NCC_HD float load_px(const float* p) {
#if defined(__CUDA_ARCH__)
    return __ldg(p);
#else
    return *p;
#endif
}

GGS_NCC_HD int floor_to_int(float x) {
#if defined(__CUDA_ARCH__)
    return __float2int_rd(x);
#else
    return (int)floorf(x);
#endif
}

GGS_NCC_HD int ceil_to_int(float x) {
#if defined(__CUDA_ARCH__)
    return __float2int_ru(x);
#else
    return (int)ceilf(x);
#endif
}

GGS_NCC_HD int clampi(int x, int lo, int hi) { return x < lo ? lo : (x > hi ? hi : x); }

struct V3 {
    float x, y, z;
};
GGS_NCC_HD V3 v3(float x, float y, float z) { return V3{x, y, z}; }
GGS_NCC_HD float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GGS_NCC_HD V3 add3(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
GGS_NCC_HD V3 sub3(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
GGS_NCC_HD V3 mul3(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
GGS_NCC_HD V3 smul3(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
GGS_NCC_HD V3 div3(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }


// Reference-image sample (du, dv) in half-pixel steps out of the 5x5 register patch: a pixel, the mean of two, or
// the mean of four, rounded as warp_patch_ncc_impl.cu:139-157 rounds it (products by 0.5 are exact, one rounding per
// sum of two).
GGS_NCC_HD float reference_sample(const float (&px)[PATCH][PATCH], int du, int dv) {
    const int u0 = (du >> 1) + REACH, u1 = ((du + 1) >> 1) + REACH;  // floor / ceil of du / 2
    const int v0 = (dv >> 1) + REACH, v1 = ((dv + 1) >> 1) + REACH;
    const bool odd_u = du & 1, odd_v = dv & 1;
    const float top = odd_u ? px[v0][u0] * 0.5f + px[v0][u1] * 0.5f : px[v0][u0];
    if (!odd_v) return top;
    const float bottom = odd_u ? px[v1][u0] * 0.5f + px[v1][u1] * 0.5f : px[v1][u0];
    return top * 0.5f + bottom * 0.5f;
}

GGS_NCC_HD bool window_inside(float x, float y, int w, int h) {
    return x - HALF_EXTENT > 0 && x + HALF_EXTENT < w - 1 && y - HALF_EXTENT > 0 && y + HALF_EXTENT < h - 1;
}

// One pixel (u, v) of the reference view with its depth and unit normal.  Invalid patches return all zeros, which is
// what the reference writes for them (warp_patch_ncc_impl.cu:257-261).
// `scratch`: SAMPLES floats for this thread, STRIDE apart (a shared-memory column on the GPU: [sample][thread]).
template <int STRIDE>
GGS_NCC_HD PointResult evaluate_point(const Pair& k, int u, int v, float depth, V3 n, const float* image_r,
                                      const float* image_n, float* scratch) {
    PointResult out = {0.f, 0.f, 0.f, 0.f, 0.f, false};
    constexpr float INV_SAMPLES = 1.f / static_cast<float>(SAMPLES);
    const float uf = static_cast<float>(u), vf = static_cast<float>(v);
    if (!window_inside(uf, vf, k.w_r, k.h_r)) return out;
    const V3 ray = v3((uf - k.cx_r) / k.fx_r, (vf - k.cy_r) / k.fy_r, 1.f);
    const float distance = -dot3(ray, n) * depth;

    // ---- reference view: 25 pixels, 49 samples, its two sums ----
    float px[PATCH][PATCH];
#pragma unroll
    for (int j = 0; j < PATCH; j++)
#pragma unroll
        for (int i = 0; i < PATCH; i++) px[j][i] = load_px(image_r + (v - REACH + j) * k.w_r + u - REACH + i);
    float sum_r = 0.f, sum_r2 = 0.f;
#pragma unroll
    for (int dv = -RADIUS; dv <= RADIUS; dv++)
#pragma unroll
        for (int du = -RADIUS; du <= RADIUS; du++) {
            const float c_r = reference_sample(px, du, dv);
            scratch[((dv + RADIUS) * LINE + du + RADIUS) * STRIDE] = c_r;
            sum_r += c_r;
            sum_r2 += c_r * c_r;
        }
    const float variance_r = sum_r2 - sum_r * sum_r * INV_SAMPLES;
    if (!(variance_r > 5e-6f)) return out;

    // ---- homography K_n (R - T n^T / distance) K_r^-1, columns H0 H1 H2 (warp_patch_ncc_impl.cu:63-84) ----
    const V3 T = v3(k.T[0], k.T[1], k.T[2]);
    const V3 hn0 = sub3(v3(k.R[0], k.R[1], k.R[2]), smul3(n.x / distance, T));
    const V3 hn1 = sub3(v3(k.R[3], k.R[4], k.R[5]), smul3(n.y / distance, T));
    const V3 hn2 = sub3(v3(k.R[6], k.R[7], k.R[8]), smul3(n.z / distance, T));
    V3 H0 = v3(k.fx_n * hn0.x + k.cx_n * hn0.z, k.fy_n * hn0.y + k.cy_n * hn0.z, hn0.z);
    V3 H1 = v3(k.fx_n * hn1.x + k.cx_n * hn1.z, k.fy_n * hn1.y + k.cy_n * hn1.z, hn1.z);
    V3 H2 = v3(k.fx_n * hn2.x + k.cx_n * hn2.z, k.fy_n * hn2.y + k.cy_n * hn2.z, hn2.z);
    H2 = add3(add3(mul3(H0, -k.cx_r / k.fx_r), mul3(H1, -k.cy_r / k.fy_r)), H2);
    H0 = div3(H0, k.fx_r);
    H1 = div3(H1, k.fy_r);
    const V3 centre = v3(uf, vf, 1.f);
    const V3 H_c = v3(dot3(v3(H0.x, H1.x, H2.x), centre), dot3(v3(H0.y, H1.y, H2.y), centre),
                      dot3(v3(H0.z, H1.z, H2.z), centre));
    const V3 aux = div3(T, distance);

    // ---- neighbour view: bilinear samples along the warped grid + their derivative along aux ----
    float sum_n = 0.f, sum_n2 = 0.f, sum_rn = 0.f;
    V3 g_n = v3(0.f, 0.f, 0.f), g_n2 = g_n, g_rn = g_n;
#pragma unroll 1
    for (int dv = -RADIUS; dv <= RADIUS; dv++) {
        const float dv_f = 0.5f * dv;
        const V3 H_row = add3(H_c, smul3(dv_f, H1));
#pragma unroll
        for (int du = -RADIUS; du <= RADIUS; du++) {
            const float du_f = 0.5f * du;
            const float c_r = scratch[((dv + RADIUS) * LINE + du + RADIUS) * STRIDE];
            const V3 Hp = add3(H_row, smul3(du_f, H0));
            const float xn = Hp.x / Hp.z, yn = Hp.y / Hp.z;
            if (!window_inside(xn, yn, k.w_n, k.h_n)) return out;  // the whole patch is invalid from here on
            const int u0 = clampi(floor_to_int(xn), 0, k.w_n - 1);
            const int v0 = clampi(floor_to_int(yn), 0, k.h_n - 1);
            const int u1 = clampi(ceil_to_int(nextafterf(xn, INFINITY)), 0, k.w_n - 1);
            const int v1 = clampi(ceil_to_int(nextafterf(yn, INFINITY)), 0, k.h_n - 1);
            const float c00 = load_px(image_n + v0 * k.w_n + u0), c01 = load_px(image_n + v0 * k.w_n + u1);
            const float c10 = load_px(image_n + v1 * k.w_n + u0), c11 = load_px(image_n + v1 * k.w_n + u1);
            const float wv0 = v1 - yn, wv1 = yn - v0, wu0 = u1 - xn, wu1 = xn - u0;
            const float c_n = wv0 * (wu0 * c00 + wu1 * c01) + wv1 * (wu0 * c10 + wu1 * c11);
            sum_n += c_n;
            sum_n2 += c_n * c_n;
            sum_rn += c_r * c_n;

            const float dc_dx = -c00 * wv0 + c01 * wv0 - c10 * wv1 + c11 * wv1;
            const float dc_dy = -c00 * wu0 - c01 * wu1 + c10 * wu0 + c11 * wu1;
            const V3 dH = v3(dc_dx / Hp.z, dc_dy / Hp.z, (-dc_dx * xn - dc_dy * yn) / Hp.z);
            const V3 left = v3(dH.x * k.fx_n, dH.y * k.fy_n, dH.x * k.cx_n + dH.y * k.cy_n + dH.z);
            const V3 right = v3(((uf + du_f) - k.cx_r) / k.fx_r, ((vf + dv_f) - k.cy_r) / k.fy_r, 1.f);
            const V3 g = mul3(right, dot3(left, aux));
            g_n = add3(g_n, g);
            g_n2 = add3(g_n2, smul3(2.f * c_n, g));
            g_rn = add3(g_rn, smul3(c_r, g));
        }
    }

    // ---- NCC^2 and its derivative (warp_patch_ncc_impl.cu:236-256) ----
    const float cross = sum_rn - sum_r * sum_n * INV_SAMPLES;
    const float variance_n = sum_n2 - sum_n * sum_n * INV_SAMPLES;
    const float value = cross * cross / (variance_r * variance_n + 1e-8f);
    if (!(variance_n > 5e-6f)) return out;
    const float d_cross = 2.f * cross / (variance_r * variance_n + 1e-8f);
    const float d_variance_n = -value / (variance_n + 1e-8f);
    const float d_sum_n = (-d_cross * sum_r - d_variance_n * 2.f * sum_n) * INV_SAMPLES;
    const V3 ga = add3(add3(smul3(d_sum_n, g_n), smul3(d_variance_n, g_n2)), smul3(d_cross, g_rn));
    const float d_distance = dot3(ga, n) / distance;
    const V3 d_normal = add3(v3(-ga.x, -ga.y, -ga.z), smul3(-depth * d_distance, ray));
    out.ncc = value;
    out.d_depth = -dot3(ray, n) * d_distance;
    out.d_nx = d_normal.x;
    out.d_ny = d_normal.y;
    out.d_nz = d_normal.z;
    out.valid = true;
    return out;
}
}  // namespace ncc
}  // namespace ggs
