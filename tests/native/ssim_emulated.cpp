// Test harness (not product code): runs the SOURCE of the SSIM kernels (ggs_ssim_kernel.cuh) on the host under
// cuda_emu.h with the launch geometry of ggs_ssim.cu.
// Build: g++ -O1 -ffp-contract=off -pthread -shared -fPIC -I<csrc> ssim_emulated.cpp
#include "cuda_emu.h"
// after the shim
#include "ggs_ssim_kernel.cuh"

using namespace ggs::ssim;

static Plan make_plan(int planes, int h, int w, int valid, const float* window) {
    Plan p;
    p.planes = planes; p.height = h; p.width = w; p.valid = valid;
    p.c1 = 0.01f * 0.01f; p.c2 = 0.03f * 0.03f;
    for (int k = 0; k < TAPS; k++) p.window[k] = window[k];
    return p;
}

extern "C" void ggs_test_ssim_forward(int planes, int h, int w, int valid, const float* window, const float* img1,
                                      const float* img2, float* dm0, float* dm1, float* dm2, float* partials,
                                      float* mean) {
    const Plan p = make_plan(planes, h, w, valid, window);
    const Region r = map_region(p);
    const dim3 g(tiles_along(r.x1 - r.x0), tiles_along(r.y1 - r.y0), planes);
    emu_launch(g, dim3(THREADS), [=] { forward_kernel(p, img1, img2, dm0, dm1, dm2, partials); });
    const long long n = (long long)g.x * g.y * g.z;
    const double count = (double)(r.y1 - r.y0) * (r.x1 - r.x0) * planes;
    emu_launch(dim3(1), dim3(THREADS), [=] { mean_kernel(partials, n, 1.0 / count, mean); });
}

extern "C" void ggs_test_ssim_backward(int planes, int h, int w, int valid, const float* window, const float* img1,
                                       const float* img2, const float* dm0, const float* dm1, const float* dm2,
                                       const float* grad_mean, float* dL_dimg1) {
    const Plan p = make_plan(planes, h, w, valid, window);
    const Region r = map_region(p);
    const double count = (double)(r.y1 - r.y0) * (r.x1 - r.x0) * planes;
    const dim3 g(tiles_along(w), tiles_along(h), planes);
    emu_launch(g, dim3(THREADS), [=] {
        backward_kernel(p, img1, img2, dm0, dm1, dm2, grad_mean, (float)(1.0 / count), dL_dimg1);
    });
}
