// libggs_ssim.so: fused SSIM loss (include/ggs_ssim.h).  Kernels in ggs_ssim_kernel.cuh.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ggs_ssim.h"
#include "ggs_ssim_kernel.cuh"

#define GGS_SSIM_EXPORT extern "C" __attribute__((visibility("default")))

static thread_local char g_error[256] = "";

static int fail(int code, const char* what) {
    snprintf(g_error, sizeof(g_error), "%s", what);
    return code;
}

static int launched(const char* what) {
    const cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return 0;
    snprintf(g_error, sizeof(g_error), "%s: launch failed: %s", what, cudaGetErrorString(e));
    return GGS_SSIM_ECUDA;
}

static bool to_plan(const GgsSsimPlan* in, ggs::ssim::Plan* out) {
    if (!in || in->planes <= 0 || in->height <= 0 || in->width <= 0 || in->planes > 65535) return false;
    if (in->valid && (in->height <= 2 * ggs::ssim::HALO || in->width <= 2 * ggs::ssim::HALO)) return false;
    out->planes = in->planes; out->height = in->height; out->width = in->width; out->valid = in->valid ? 1 : 0;
    out->c1 = in->c1; out->c2 = in->c2;
    for (int k = 0; k < ggs::ssim::TAPS; k++) out->window[k] = in->window[k];
    return true;
}

static dim3 map_grid(const ggs::ssim::Plan& p) {
    const ggs::ssim::Region r = ggs::ssim::map_region(p);
    return dim3(ggs::ssim::tiles_along(r.x1 - r.x0), ggs::ssim::tiles_along(r.y1 - r.y0), p.planes);
}

GGS_SSIM_EXPORT int ggs_ssim_abi_version(void) { return GGS_SSIM_ABI_VERSION; }
GGS_SSIM_EXPORT const char* ggs_ssim_last_error(void) { return g_error; }

GGS_SSIM_EXPORT int64_t ggs_ssim_partials(const GgsSsimPlan* plan) {
    ggs::ssim::Plan p;
    if (!to_plan(plan, &p)) return fail(GGS_SSIM_EINVAL, "ggs_ssim_partials: bad plan");
    const dim3 g = map_grid(p);
    return static_cast<int64_t>(g.x) * g.y * g.z;
}

GGS_SSIM_EXPORT int ggs_ssim_forward(const GgsSsimPlan* plan, const float* img1, const float* img2, float* dm_dmu1,
                                     float* dm_dsigma1_sq, float* dm_dsigma12, float* partials, float* mean,
                                     void* stream) {
    ggs::ssim::Plan p;
    if (!to_plan(plan, &p)) return fail(GGS_SSIM_EINVAL, "ggs_ssim_forward: bad plan");
    if (!img1 || !img2 || !partials || !mean) return fail(GGS_SSIM_EINVAL, "ggs_ssim_forward: null buffer");
    const int maps = (dm_dmu1 != nullptr) + (dm_dsigma1_sq != nullptr) + (dm_dsigma12 != nullptr);
    if (maps != 0 && maps != 3) return fail(GGS_SSIM_EINVAL, "ggs_ssim_forward: pass all three derivative maps or none");
    const dim3 g = map_grid(p);
    const ggs::ssim::Region r = ggs::ssim::map_region(p);
    const long long n = static_cast<long long>(g.x) * g.y * g.z;
    const double count = static_cast<double>(r.y1 - r.y0) * (r.x1 - r.x0) * p.planes;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    ggs::ssim::forward_kernel<<<g, ggs::ssim::THREADS, 0, s>>>(p, img1, img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12,
                                                               partials);
    if (int rc = launched("ggs_ssim_forward")) return rc;
    ggs::ssim::mean_kernel<<<1, ggs::ssim::THREADS, 0, s>>>(partials, n, 1.0 / count, mean);
    return launched("ggs_ssim_forward (mean)");
}

GGS_SSIM_EXPORT int ggs_ssim_backward(const GgsSsimPlan* plan, const float* img1, const float* img2,
                                      const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                                      const float* grad_mean, float* dL_dimg1, void* stream) {
    ggs::ssim::Plan p;
    if (!to_plan(plan, &p)) return fail(GGS_SSIM_EINVAL, "ggs_ssim_backward: bad plan");
    if (!img1 || !img2 || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !grad_mean || !dL_dimg1)
        return fail(GGS_SSIM_EINVAL, "ggs_ssim_backward: null buffer");
    const ggs::ssim::Region r = ggs::ssim::map_region(p);
    const double count = static_cast<double>(r.y1 - r.y0) * (r.x1 - r.x0) * p.planes;
    const dim3 g(ggs::ssim::tiles_along(p.width), ggs::ssim::tiles_along(p.height), p.planes);
    ggs::ssim::backward_kernel<<<g, ggs::ssim::THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
        p, img1, img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, grad_mean, static_cast<float>(1.0 / count), dL_dimg1);
    return launched("ggs_ssim_backward");
}
