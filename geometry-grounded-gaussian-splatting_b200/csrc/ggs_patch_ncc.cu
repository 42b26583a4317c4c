// libggs_patch_ncc.so: patch NCC under the plane-induced homography (include/ggs_patch_ncc.h).
// One thread per pixel, sums in the reference's order (see ggs_patch_ncc_point.cuh for what differs from
// warp_patch_ncc_impl.cu).  The intrinsics travel as a kernel parameter (constant bank); the pose stays a device
// pointer, as in the reference, so that the caller needs no host synchronisation; [P,3] rows go through shared memory
// so that global traffic is coalesced; invalid patches stop early.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ggs_patch_ncc.h"
#include "ggs_patch_ncc_kernel.cuh"

#define GGS_NCC_EXPORT extern "C" __attribute__((visibility("default")))

static thread_local char g_error[256] = "";

static int fail(int code, const char* what) {
    snprintf(g_error, sizeof(g_error), "%s", what);
    return code;
}

GGS_NCC_EXPORT int ggs_ncc_abi_version(void) { return GGS_NCC_ABI_VERSION; }
GGS_NCC_EXPORT const char* ggs_ncc_last_error(void) { return g_error; }

GGS_NCC_EXPORT int ggs_ncc_warp_patch(const GgsNccPair* pair, int64_t P, const float* depths, const float* normals,
                                      const int32_t* uvs, const float* image_r, const float* image_n, float* ncc,
                                      float* grad_depths, float* grad_normals, uint8_t* valid, void* stream) {
    if (!pair) return fail(GGS_NCC_EINVAL, "ggs_ncc_warp_patch: pair is null");
    if (!pair->R || !pair->T) return fail(GGS_NCC_EINVAL, "ggs_ncc_warp_patch: null pose");
    if (P < 0) return fail(GGS_NCC_EINVAL, "ggs_ncc_warp_patch: negative point count");
    if (P == 0) return 0;
    if (!depths || !normals || !uvs || !image_r || !image_n || !ncc || !grad_depths || !grad_normals || !valid)
        return fail(GGS_NCC_EINVAL, "ggs_ncc_warp_patch: null buffer");
    if (pair->height_r <= 0 || pair->width_r <= 0 || pair->height_n <= 0 || pair->width_n <= 0)
        return fail(GGS_NCC_EINVAL, "ggs_ncc_warp_patch: empty image");
    ggs::ncc::Pair k;
    for (int i = 0; i < 9; i++) k.R[i] = 0.f;  // read on the device
    for (int i = 0; i < 3; i++) k.T[i] = 0.f;
    k.fx_r = pair->fx_r; k.fy_r = pair->fy_r; k.cx_r = pair->cx_r; k.cy_r = pair->cy_r;
    k.fx_n = pair->fx_n; k.fy_n = pair->fy_n; k.cx_n = pair->cx_n; k.cy_n = pair->cy_n;
    k.h_r = pair->height_r; k.w_r = pair->width_r; k.h_n = pair->height_n; k.w_n = pair->width_n;
    const long long blocks = (P + ggs::ncc::THREADS - 1) / ggs::ncc::THREADS;
    if (blocks > 0x7fffffffLL) return fail(GGS_NCC_EINVAL, "ggs_ncc_warp_patch: too many points for one launch");
    ggs::ncc::warp_patch_kernel<<<static_cast<unsigned>(blocks), ggs::ncc::THREADS, 0,
                                  static_cast<cudaStream_t>(stream)>>>(
        k, pair->R, pair->T, P, depths, normals, reinterpret_cast<const int2*>(uvs), image_r, image_n, ncc, grad_depths,
        grad_normals, valid);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(g_error, sizeof(g_error), "ggs_ncc_warp_patch: launch failed: %s", cudaGetErrorString(e));
        return GGS_NCC_ECUDA;
    }
    return 0;
}
