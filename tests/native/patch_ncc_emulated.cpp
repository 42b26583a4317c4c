// Test harness (not product code): runs the SOURCE of warp_patch_kernel (ggs_patch_ncc_kernel.cuh) on the host under
// cuda_emu.h, block by block with real barriers, the way ggs_ncc_warp_patch launches it.
// Build: g++ -O1 -ffp-contract=off -pthread -shared -fPIC -I<csrc> patch_ncc_emulated.cpp
#include "cuda_emu.h"
// after the shim
#include "ggs_patch_ncc_kernel.cuh"

extern "C" void ggs_test_ncc_kernel(long long P, const float* depths, const float* normals, const int32_t* uvs,
                                    const float* R, const float* T, const float* image_r, const float* image_n,
                                    const float* intr, int h_r, int w_r, int h_n, int w_n, float* ncc,
                                    float* grad_depths, float* grad_normals, uint8_t* valid) {
    ggs::ncc::Pair k;
    for (int i = 0; i < 9; i++) k.R[i] = 0.f;
    for (int i = 0; i < 3; i++) k.T[i] = 0.f;
    k.fx_r = intr[0]; k.fy_r = intr[1]; k.cx_r = intr[2]; k.cy_r = intr[3];
    k.fx_n = intr[4]; k.fy_n = intr[5]; k.cx_n = intr[6]; k.cy_n = intr[7];
    k.h_r = h_r; k.w_r = w_r; k.h_n = h_n; k.w_n = w_n;
    const unsigned blocks = (unsigned)((P + ggs::ncc::THREADS - 1) / ggs::ncc::THREADS);
    emu_launch(dim3(blocks), dim3(ggs::ncc::THREADS), [=] {
        ggs::ncc::warp_patch_kernel(k, R, T, P, depths, normals, reinterpret_cast<const int2*>(uvs), image_r, image_n,
                                    ncc, grad_depths, grad_normals, valid);
    });
}
