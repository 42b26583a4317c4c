/* C ABI of the fused SSIM loss kernels: libggs_ssim.so.
 *
 * Replaces the third-party `fused_ssim` extension the reference imports (utils/loss_utils.py:18,48-49:
 * `fused_ssim(img1, img2, padding="valid")`; rahul-goel/fused-ssim, installed from git HEAD by README.md:37 and not
 * under /root/reference).  The algorithm is the reference's own statement of it, utils/loss_utils.py:34-73.
 * Plain pointers and sizes; every buffer is device memory owned by the caller; all work is enqueued on `stream`.
 * Functions return 0 on success and a negative GGS_SSIM_E* code otherwise.
 */
#ifndef GGS_SSIM_H
#define GGS_SSIM_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGS_SSIM_ABI_VERSION 1
#define GGS_SSIM_EINVAL (-1)
#define GGS_SSIM_ECUDA (-2)
#define GGS_SSIM_TAPS 11

typedef struct GgsSsimPlan {
    int32_t planes, height, width; /* planes = batch * channels of the [B,C,H,W] images */
    int32_t valid;                 /* 1: padding="valid" (map = (H-10) x (W-10)); 0: padding="same" (zero padding) */
    float c1, c2;                  /* 0.01^2, 0.03^2 */
    float window[GGS_SSIM_TAPS];   /* normalised 1-D Gaussian, sigma 1.5 (loss_utils.py:34-37) */
} GgsSsimPlan;

int ggs_ssim_abi_version(void);
const char* ggs_ssim_last_error(void);

/* Number of float partial sums ggs_ssim_forward needs for this plan (host query, no GPU work); < 0 on a bad plan
 * (valid padding needs height and width > 10). */
int64_t ggs_ssim_partials(const GgsSsimPlan* plan);

/* mean[0] = mean of the SSIM map of img1 vs img2 ([planes, H, W] f32 each).  When the three derivative maps are
 * non-null ([planes, H, W] f32 each, all three or none) they receive d map / d mu1, d map / d sigma1^2 and
 * d map / d sigma12 on the map region for ggs_ssim_backward; elements outside the region are not written and not
 * read.  `partials`: scratch of ggs_ssim_partials(plan) floats. */
int ggs_ssim_forward(const GgsSsimPlan* plan, const float* img1, const float* img2, float* dm_dmu1,
                     float* dm_dsigma1_sq, float* dm_dsigma12, float* partials, float* mean, void* stream);

/* dL_dimg1 [planes, H, W] (every element written) from the derivative maps and the scalar gradient of the mean,
 * `grad_mean` (one float in DEVICE memory: no host synchronisation). */
int ggs_ssim_backward(const GgsSsimPlan* plan, const float* img1, const float* img2, const float* dm_dmu1,
                      const float* dm_dsigma1_sq, const float* dm_dsigma12, const float* grad_mean, float* dL_dimg1,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif
