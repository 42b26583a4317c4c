// Test harness (not product code): compiles the kernel's per-point routine, ggs_patch_ncc_point.cuh, for the host so
// that tests/test_patch_ncc_cpu.py can run the very code the GPU runs against the C oracle and a float64 model.
// Build: g++ -O2 -ffp-contract=off -shared -fPIC -I<csrc> patch_ncc_host.cpp
#include <stdint.h>

#include "ggs_patch_ncc_point.cuh"

extern "C" void ggs_test_ncc_points(int P, const float* depths, const float* normals, const int32_t* uvs,
                                    const float* R, const float* T, const float* image_r, const float* image_n,
                                    const float* intr /* fx_r fy_r cx_r cy_r fx_n fy_n cx_n cy_n */, int h_r, int w_r,
                                    int h_n, int w_n, float* ncc, float* grad_depths, float* grad_normals,
                                    uint8_t* valid) {
    ggs::ncc::Pair k;
    for (int i = 0; i < 9; i++) k.R[i] = R[i];
    for (int i = 0; i < 3; i++) k.T[i] = T[i];
    k.fx_r = intr[0]; k.fy_r = intr[1]; k.cx_r = intr[2]; k.cy_r = intr[3];
    k.fx_n = intr[4]; k.fy_n = intr[5]; k.cx_n = intr[6]; k.cy_n = intr[7];
    k.h_r = h_r; k.w_r = w_r; k.h_n = h_n; k.w_n = w_n;
    for (int i = 0; i < P; i++) {
        const ggs::ncc::V3 n = ggs::ncc::v3(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]);
        float scratch[ggs::ncc::SAMPLES];
        const ggs::ncc::PointResult r =
            ggs::ncc::evaluate_point<1>(k, uvs[2 * i], uvs[2 * i + 1], depths[i], n, image_r, image_n, scratch);
        ncc[i] = r.ncc;
        grad_depths[i] = r.d_depth;
        grad_normals[3 * i] = r.d_nx;
        grad_normals[3 * i + 1] = r.d_ny;
        grad_normals[3 * i + 2] = r.d_nz;
        valid[i] = r.valid ? 1 : 0;
    }
}
