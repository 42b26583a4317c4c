/* C ABI of the patch-NCC kernel: libggs_patch_ncc.so.
 *
 * Replaces the reference's pybind entry `warp_patch_ncc._C.warp_patch_ncc`
 * (submodules/warp-patch-ncc/ext.cpp:4-6 -> WarpPatchNCC, warp_patch_ncc.cu:5-57 -> forward_mode_differentiation,
 * cuda_warp_patch_ncc/warp_patch_ncc_impl.cu:264-302).  Plain pointers and sizes only; every buffer is device
 * memory owned by the caller; all work is enqueued on `stream`.  Functions return 0 on success and a negative
 * GGS_NCC_E* code otherwise; ggs_ncc_last_error() describes the last failure on the calling thread.
 */
#ifndef GGS_PATCH_NCC_H
#define GGS_PATCH_NCC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGS_NCC_ABI_VERSION 1
#define GGS_NCC_EINVAL (-1) /* bad argument (null pointer, negative size, empty image) */
#define GGS_NCC_ECUDA (-2)  /* a CUDA call failed */

/* Relative pose reference -> neighbour and both pinhole intrinsics.
 * R: DEVICE pointer to 9 floats, COLUMN-major (what the reference reads from the row-major tensor
 * `ref_to_nearest_r.T`, utils/loss_utils.py:243); T: DEVICE pointer to 3 floats.  They stay on the device, as in the
 * reference, so that a training step needs no host synchronisation to call this.  Sizes in pixels. */
typedef struct GgsNccPair {
    const float* R;
    const float* T;
    float fx_r, fy_r, cx_r, cy_r;
    float fx_n, fy_n, cx_n, cy_n;
    int32_t height_r, width_r, height_n, width_n;
} GgsNccPair;

int ggs_ncc_abi_version(void);
const char* ggs_ncc_last_error(void);

/* Squared NCC of the 7x7 half-pixel patch around each pixel and its derivative with respect to the pixel's depth and
 * normal (forward-mode, computed together with the value as the reference does).
 *   depths [P] f32, normals [P,3] f32 (unit), uvs [P,2] i32 (x, y) pixel coordinates in the reference image,
 *   image_r [height_r, width_r] f32, image_n [height_n, width_n] f32 (grey),
 *   ncc [P] f32, grad_depths [P] f32, grad_normals [P,3] f32, valid [P] u8 (0/1).
 * Every output element is written (zeros for invalid patches), so the buffers need no clearing.  P == 0 is a no-op. */
int ggs_ncc_warp_patch(const GgsNccPair* pair, int64_t P, const float* depths, const float* normals,
                       const int32_t* uvs, const float* image_r, const float* image_n, float* ncc,
                       float* grad_depths, float* grad_normals, uint8_t* valid, void* stream);

#ifdef __cplusplus
}
#endif
#endif
