// Small device-side math used by the per-Gaussian kernels.
//
// M3 is a column-major 3x3 (c[col][row]) whose product formulas are written in the same term
// order as the reference's matrix library so that, built with the reference's nvcc flags
// (-O3 --use_fast_math, DGR/setup.py:22-33), the float pipelines that feed the bit-exact
// integer outputs (radii, tile rects, depth keys) round identically.
#pragma once
#include "ggs_common.cuh"

namespace ggs {

struct V3 {
    float x, y, z;
};

struct M3 {
    float c[3][3];  // c[col][row]
};

__device__ __forceinline__ M3 m3_cols(float a0, float a1, float a2, float b0, float b1, float b2, float c0, float c1,
                                      float c2) {
    M3 m;
    m.c[0][0] = a0; m.c[0][1] = a1; m.c[0][2] = a2;
    m.c[1][0] = b0; m.c[1][1] = b1; m.c[1][2] = b2;
    m.c[2][0] = c0; m.c[2][1] = c1; m.c[2][2] = c2;
    return m;
}

__device__ __forceinline__ M3 m3_diag(float a, float b, float c) {
    return m3_cols(a, 0.f, 0.f, 0.f, b, 0.f, 0.f, 0.f, c);
}

// r = a * b  (column-major): r[j][i] = a[0][i]*b[j][0] + a[1][i]*b[j][1] + a[2][i]*b[j][2]
__device__ __forceinline__ M3 operator*(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++) {
#pragma unroll
        for (int i = 0; i < 3; i++) {
            r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
        }
    }
    return r;
}

__device__ __forceinline__ V3 operator*(const M3& m, const V3& v) {
    V3 r;
    r.x = m.c[0][0] * v.x + m.c[1][0] * v.y + m.c[2][0] * v.z;
    r.y = m.c[0][1] * v.x + m.c[1][1] * v.y + m.c[2][1] * v.z;
    r.z = m.c[0][2] * v.x + m.c[1][2] * v.y + m.c[2][2] * v.z;
    return r;
}

// row-vector * matrix
__device__ __forceinline__ V3 vec_mul_mat(const V3& v, const M3& m) {
    V3 r;
    r.x = m.c[0][0] * v.x + m.c[0][1] * v.y + m.c[0][2] * v.z;
    r.y = m.c[1][0] * v.x + m.c[1][1] * v.y + m.c[1][2] * v.z;
    r.z = m.c[2][0] * v.x + m.c[2][1] * v.y + m.c[2][2] * v.z;
    return r;
}

__device__ __forceinline__ M3 transpose(const M3& a) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) r.c[j][i] = a.c[i][j];
    return r;
}

// outer(c, r): column i = c * r[i]
__device__ __forceinline__ M3 outer(const V3& c, const V3& r) {
    M3 m;
    m.c[0][0] = c.x * r.x; m.c[0][1] = c.y * r.x; m.c[0][2] = c.z * r.x;
    m.c[1][0] = c.x * r.y; m.c[1][1] = c.y * r.y; m.c[1][2] = c.z * r.y;
    m.c[2][0] = c.x * r.z; m.c[2][1] = c.y * r.z; m.c[2][2] = c.z * r.z;
    return m;
}

__device__ __forceinline__ M3 operator+(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) r.c[j][i] = a.c[j][i] + b.c[j][i];
    return r;
}

__device__ __forceinline__ M3 operator*(const M3& a, float s) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) r.c[j][i] = a.c[j][i] * s;
    return r;
}

__device__ __forceinline__ M3 operator/(const M3& a, float s) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) r.c[j][i] = a.c[j][i] / s;
    return r;
}

__device__ __forceinline__ M3 neg(const M3& a) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) r.c[j][i] = -a.c[j][i];
    return r;
}

__device__ __forceinline__ V3 operator+(const V3& a, const V3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(const V3& a, const V3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(const V3& a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 operator*(float s, const V3& a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ V3 operator/(const V3& a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ float dot(const V3& a, const V3& b) {
    float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
    return tx + ty + tz;
}
__device__ __forceinline__ V3 normalize(const V3& v) { return v * rsqrtf(dot(v, v)); }

// ---- explicitly rounded float steps --------------------------------------------------------------
// The intrinsics below are never contracted, reassociated or re-fused by nvcc / ptxas, so an expression
// written with them evaluates in exactly one way.  The per-Gaussian preprocess uses them to follow the
// evaluation order of the reference binary (learned with tools/sassflow.py) and so reproduce its geometry
// records bit for bit; plain `a * b + c` would leave the FMA contraction pattern to the compiler.
__device__ __forceinline__ float MUL(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float ADD(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float FMA(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float sqrt_approx(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rsqrt_approx(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// a0*b0 + a1*b1 + a2*b2 as the compiler contracts a three-term dot product:
// the middle product is rounded first, then two fused steps.
__device__ __forceinline__ float sop3(float a0, float b0, float a1, float b1, float a2, float b2) {
    return FMA(a2, b2, FMA(a0, b0, MUL(a1, b1)));
}
__device__ __forceinline__ float twice(float w) { return ADD(w, w); }
__device__ __forceinline__ float one_minus_twice(float w) { return ADD(-ADD(w, w), 1.f); }

// world/view transforms with the reference's term order (auxiliary.h:51-85)
__device__ __forceinline__ float3 xform_point_4x3(const float3& p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
__device__ __forceinline__ float4 xform_point_4x4(const float3& p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
}
__device__ __forceinline__ float3 xform_vec_4x3_transpose(const float3& p, const float* m) {
    return {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
            m[8] * p.x + m[9] * p.y + m[10] * p.z};
}

// pixel coordinate from NDC; evaluated in double like the reference (auxiliary.h:38-40)
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// tile rectangle of a splat (auxiliary.h:42-49)
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, uint2& rect_min, uint2& rect_max,
                                          unsigned gx, unsigned gy) {
    rect_min = {min(gx, max((int)0, (int)((p.x - max_radius) / TILE_X))),
                min(gy, max((int)0, (int)((p.y - max_radius) / TILE_Y)))};
    rect_max = {min(gx, max((int)0, (int)((p.x + max_radius + TILE_X - 1) / TILE_X))),
                min(gy, max((int)0, (int)((p.y + max_radius + TILE_Y - 1) / TILE_Y)))};
}

__constant__ const float kSH_C0 = 0.28209479177387814f;
__constant__ const float kSH_C1 = 0.4886025119029199f;
__constant__ const float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                      -1.0925484305920792f, 0.5462742152960396f};
__constant__ const float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f,  -0.4570457994644658f,
                                      0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

}  // namespace ggs
