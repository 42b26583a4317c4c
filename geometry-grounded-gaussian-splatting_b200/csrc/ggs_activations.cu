// Caller-side activations of the Gaussian parameters, fused: what the reference computes with
// ~10 elementwise torch kernels per iteration before it calls the rasterizer, and ~20 more in autograd on the way back,
//     scene/gaussian_model.py:203-212  get_scaling_n_opacity_with_3D_filter   (exp, 3D filter, sigmoid, coefficient)
//     scene/gaussian_model.py:158-160  get_rotation = torch.nn.functional.normalize(_rotation)
// as ONE forward and ONE backward kernel, each a single pass over 36 + 32 bytes per Gaussian:
//     s = exp(raw_scaling);  a = s^2 + f^2;  scales = sqrt(a)
//     coef = sqrt(prod s^2) * rsqrt(prod a);  opacity = sigmoid(raw_opacity) * coef
//     rotation = q / max(|q|, 1e-12)
// filter_3D (f) is a buffer without gradient in the reference (gaussian_model.py:223,261-262).
// This file is compiled WITHOUT --use_fast_math (csrc/Makefile): the outputs feed the rasterizer's integer decisions
// (radii, tile rectangles), so they should sit within an ulp or two of torch's IEEE results.
#include "ggs_kernels.cuh"

namespace ggs {

constexpr int ACT_THREADS = 256;

__global__ void __launch_bounds__(ACT_THREADS)
    activate_forward_kernel(int P, const float* __restrict__ raw_scaling, const float* __restrict__ raw_opacity,
                            const float* __restrict__ raw_rotation, const float* __restrict__ filter_3D,
                            float* __restrict__ scales, float* __restrict__ opacities, float* __restrict__ rotations) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float f = __ldg(filter_3D + idx);
    const float f2 = f * f;
    float det1 = 1.f, det2 = 1.f;
    float out[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float s = expf(__ldg(raw_scaling + 3 * (size_t)idx + k));
        const float u = s * s;
        const float a = u + f2;
        det1 *= u;  // prod(dim=1) multiplies left to right
        det2 *= a;
        out[k] = sqrtf(a);
    }
#pragma unroll
    for (int k = 0; k < 3; k++) scales[3 * (size_t)idx + k] = out[k];
    const float coef = sqrtf(det1) * rsqrtf(det2);  // det1.sqrt() * det2.rsqrt()
    const float x = __ldg(raw_opacity + idx);
    const float o = 1.0f / (1.0f + expf(-x));
    opacities[idx] = o * coef;
    const float4 q = __ldg(reinterpret_cast<const float4*>(raw_rotation) + idx);
    const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);  // F.normalize eps
    reinterpret_cast<float4*>(rotations)[idx] = {q.x / n, q.y / n, q.z / n, q.w / n};
}

// With S_k = scales_k, u_k = s_k^2, a_k = u_k + f^2, o = sigmoid(x), O = o * coef:
//   dO/dx = coef * o (1 - o)
//   dS_k/dr_k = u_k / S_k                     (r = raw_scaling)
//   d ln(coef)/dr_k = 1 - u_k / a_k = f^2 / a_k
//   d(q/n)/dq . g = (g - R (R . g)) / n       (R = q / n; the clamp is inactive for any real quaternion)
__global__ void __launch_bounds__(ACT_THREADS)
    activate_backward_kernel(int P, const float* __restrict__ raw_scaling, const float* __restrict__ raw_opacity,
                             const float* __restrict__ raw_rotation, const float* __restrict__ filter_3D,
                             const float* __restrict__ dL_dscales, const float* __restrict__ dL_dopacities,
                             const float* __restrict__ dL_drotations, float* __restrict__ dL_draw_scaling,
                             float* __restrict__ dL_draw_opacity, float* __restrict__ dL_draw_rotation) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float f = __ldg(filter_3D + idx);
    const float f2 = f * f;
    float u[3], a[3];
    float det1 = 1.f, det2 = 1.f;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float s = expf(__ldg(raw_scaling + 3 * (size_t)idx + k));
        u[k] = s * s;
        a[k] = u[k] + f2;
        det1 *= u[k];
        det2 *= a[k];
    }
    const float coef = sqrtf(det1) * rsqrtf(det2);
    const float x = __ldg(raw_opacity + idx);
    const float o = 1.0f / (1.0f + expf(-x));
    const float gO = __ldg(dL_dopacities + idx);
    dL_draw_opacity[idx] = gO * coef * o * (1.0f - o);
    const float gcoef = gO * o * coef;  // dL/d ln(coef)
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float gS = __ldg(dL_dscales + 3 * (size_t)idx + k);
        dL_draw_scaling[3 * (size_t)idx + k] = gS * u[k] / sqrtf(a[k]) + gcoef * f2 / a[k];
    }
    const float4 q = __ldg(reinterpret_cast<const float4*>(raw_rotation) + idx);
    const float4 g = __ldg(reinterpret_cast<const float4*>(dL_drotations) + idx);
    const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    const float n = fmaxf(norm, 1e-12f);
    const float4 R = {q.x / n, q.y / n, q.z / n, q.w / n};
    // below the clamp the output is q / eps: plain scaling (what autograd gives for clamp_min)
    const float dot = norm > 1e-12f ? (R.x * g.x + R.y * g.y + R.z * g.z + R.w * g.w) : 0.f;
    reinterpret_cast<float4*>(dL_draw_rotation)[idx] = {(g.x - R.x * dot) / n, (g.y - R.y * dot) / n,
                                                        (g.z - R.z * dot) / n, (g.w - R.w * dot) / n};
}

void launch_activate_forward(int P, const float* raw_scaling, const float* raw_opacity, const float* raw_rotation,
                             const float* filter_3D, float* scales, float* opacities, float* rotations,
                             cudaStream_t stream) {
    if (P == 0) return;
    activate_forward_kernel<<<(P + ACT_THREADS - 1) / ACT_THREADS, ACT_THREADS, 0, stream>>>(
        P, raw_scaling, raw_opacity, raw_rotation, filter_3D, scales, opacities, rotations);
}

void launch_activate_backward(int P, const float* raw_scaling, const float* raw_opacity, const float* raw_rotation,
                              const float* filter_3D, const float* dL_dscales, const float* dL_dopacities,
                              const float* dL_drotations, float* dL_draw_scaling, float* dL_draw_opacity,
                              float* dL_draw_rotation, cudaStream_t stream) {
    if (P == 0) return;
    activate_backward_kernel<<<(P + ACT_THREADS - 1) / ACT_THREADS, ACT_THREADS, 0, stream>>>(
        P, raw_scaling, raw_opacity, raw_rotation, filter_3D, dL_dscales, dL_dopacities, dL_drotations,
        dL_draw_scaling, dL_draw_opacity, dL_draw_rotation);
}

}  // namespace ggs
