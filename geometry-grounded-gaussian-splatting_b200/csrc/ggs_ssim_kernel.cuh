// Fused SSIM loss kernels (the `fused_ssim` package the reference imports at
// utils/loss_utils.py:18,48-49 is third-party and not under /root/reference: rahul-goel/fused-ssim, unpinned in
// README.md:37).  The algorithm is the reference's own torch statement of it, utils/loss_utils.py:34-73 (11-tap
// Gaussian window, sigma 1.5, C1 = 0.01^2, C2 = 0.03^2), with zero padding ("same") or none ("valid"), the mean of the
// map as the result, and the gradient with respect to the first image.
//
// Forward, one CTA per 32x32 tile of one image plane: both images' 42x42 halo tiles -> shared memory; separable
// window: a horizontal pass for the five moments (x, y, xx, yy, xy), a vertical pass; SSIM and the three partial
// derivatives of the map (d/dmu1, d/dsigma1^2, d/dsigma12) per pixel; the tile's sum of the map reduced in shared
// memory to one partial per CTA (no float atomics: the mean is bit-reproducible).  The map itself is never written.
// Backward: dL/dimg1[p] = sum_q w(q-p) g(q) (dm_dmu1(q) + 2 img1(p) dm_dsigma1_sq(q) + img2(p) dm_dsigma12(q)),
// three separable convolutions of the stored derivative maps in the same tile structure; the incoming scalar gradient
// is read from device memory, so nothing synchronises with the host.
//
// No CUDA runtime includes and no warp intrinsics: the test suite also runs this source on the host under
// tests/native/cuda_emu.h.
#pragma once
#include <stdint.h>

namespace ggs {
namespace ssim {

constexpr int TAPS = 11;
constexpr int HALO = TAPS / 2;           // 5
constexpr int TILE = 32;                 // output tile edge
constexpr int SPAN = TILE + 2 * HALO;    // 42: input tile edge
constexpr int THREADS = 256;
constexpr int ROWS = TILE * TILE / THREADS;  // 4: rows of one column a thread owns in the vertical pass
static_assert(ROWS * THREADS == TILE * TILE && THREADS % TILE == 0, "one column strip per thread");
static_assert(sizeof(float) * (2 * SPAN * SPAN + 5 * SPAN * TILE + THREADS) <= 48 * 1024, "static shared memory");

struct Plan {
    int planes, height, width;  // planes = batch * channels
    int valid;                  // 1: no padding, map = centres [HALO, size - HALO); 0: zero padding, map = whole image
    float c1, c2;
    float window[TAPS];
};

struct Region {
    int y0, y1, x0, x1;  // centres of the map
};

__host__ __device__ inline Region map_region(const Plan& p) {
    Region r;
    r.y0 = p.valid ? HALO : 0;
    r.x0 = p.valid ? HALO : 0;
    r.y1 = p.valid ? p.height - HALO : p.height;
    r.x1 = p.valid ? p.width - HALO : p.width;
    return r;
}

// tiles cover the map region (forward) or the whole image (backward)
__host__ __device__ inline int tiles_along(int extent) { return extent > 0 ? (extent + TILE - 1) / TILE : 0; }


// Sum of `v` over the CTA, valid in thread 0 (shared-memory tree: fixed order, bit-reproducible).
template <class T>
__device__ inline T block_sum(T v, T* red) {
    const int tid = threadIdx.x;
    red[tid] = v;
    __syncthreads();
    for (int s = THREADS / 2; s > 0; s >>= 1) {
        if (tid < s) red[tid] += red[tid + s];
        __syncthreads();
    }
    return red[0];
}

__global__ void __launch_bounds__(THREADS)
forward_kernel(const Plan plan, const float* __restrict__ img1, const float* __restrict__ img2,
               float* __restrict__ dm_dmu1, float* __restrict__ dm_dsigma1_sq, float* __restrict__ dm_dsigma12,
               float* __restrict__ partials) {
    __shared__ float a[SPAN][SPAN];
    __shared__ float b[SPAN][SPAN];
    __shared__ float h[5][SPAN][TILE];
    __shared__ float red[THREADS];
    const Region r = map_region(plan);
    const int tid = threadIdx.x, H = plan.height, W = plan.width;
    const int oy0 = r.y0 + static_cast<int>(blockIdx.y) * TILE, ox0 = r.x0 + static_cast<int>(blockIdx.x) * TILE;
    const size_t base = static_cast<size_t>(blockIdx.z) * H * W;

    for (int i = tid; i < SPAN * SPAN; i += THREADS) {
        const int y = i / SPAN, x = i - y * SPAN;
        const int gy = oy0 - HALO + y, gx = ox0 - HALO + x;
        const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;
        const size_t at = base + static_cast<size_t>(in ? gy : 0) * W + (in ? gx : 0);
        a[y][x] = in ? __ldg(img1 + at) : 0.f;
        b[y][x] = in ? __ldg(img2 + at) : 0.f;
    }
    __syncthreads();

    for (int i = tid; i < SPAN * TILE; i += THREADS) {
        const int y = i / TILE, x = i - y * TILE;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f;
#pragma unroll
        for (int k = 0; k < TAPS; k++) {
            const float w = plan.window[k], p = a[y][x + k], q = b[y][x + k];
            s0 += w * p;
            s1 += w * q;
            s2 += w * p * p;
            s3 += w * q * q;
            s4 += w * p * q;
        }
        h[0][y][x] = s0; h[1][y][x] = s1; h[2][y][x] = s2; h[3][y][x] = s3; h[4][y][x] = s4;
    }
    __syncthreads();

    // vertical pass: each thread owns ROWS consecutive rows of one column, so the ROWS + 10 inputs of a moment are
    // read from shared memory once and reused by the ROWS windows (sums still run k = 0..10, so nothing rounds anew)
    float local = 0.f;
    {
        const int x = tid % TILE, y0 = (tid / TILE) * ROWS;
        float m[5][ROWS];
#pragma unroll
        for (int q = 0; q < 5; q++) {
            float v[ROWS + TAPS - 1];
#pragma unroll
            for (int j = 0; j < ROWS + TAPS - 1; j++) v[j] = h[q][y0 + j][x];
#pragma unroll
            for (int r2 = 0; r2 < ROWS; r2++) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < TAPS; k++) acc += plan.window[k] * v[r2 + k];
                m[q][r2] = acc;
            }
        }
#pragma unroll
        for (int r2 = 0; r2 < ROWS; r2++) {
            const int cy = oy0 + y0 + r2, cx = ox0 + x;
            if (cy >= r.y1 || cx >= r.x1) continue;
            const float mu1 = m[0][r2], mu2 = m[1][r2], e11 = m[2][r2], e22 = m[3][r2], e12 = m[4][r2];
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2;
            const float sigma1_sq = e11 - mu1_sq, sigma2_sq = e22 - mu2_sq, sigma12 = e12 - mu1 * mu2;
            const float A = mu1_sq + mu2_sq + plan.c1, B = sigma1_sq + sigma2_sq + plan.c2;
            const float C = 2.f * mu1 * mu2 + plan.c1, D = 2.f * sigma12 + plan.c2;
            // two reciprocals per pixel instead of seven divisions:
            //   m = C D / (A B);  dm/dmu1 = 2 [mu2 (D - C) + mu1 C D (1/B - 1/A)] / (A B);
            //   dm/dsigma1^2 = -C D / (A B^2);  dm/dsigma12 = 2 C / (A B)
            const float inv_a = 1.f / A, inv_b = 1.f / B, inv_ab = inv_a * inv_b, cd = C * D;
            local += cd * inv_ab;
            if (dm_dmu1) {
                const size_t at = base + static_cast<size_t>(cy) * W + cx;
                dm_dmu1[at] = 2.f * inv_ab * (mu2 * (D - C) + mu1 * cd * (inv_b - inv_a));
                dm_dsigma1_sq[at] = -cd * inv_ab * inv_b;
                dm_dsigma12[at] = 2.f * C * inv_ab;
            }
        }
    }
    const float total = block_sum(local, red);
    if (tid == 0)
        partials[(static_cast<size_t>(blockIdx.z) * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = total;
}

// mean = sum(partials) / count, accumulated in double by one CTA
__global__ void __launch_bounds__(THREADS)
mean_kernel(const float* __restrict__ partials, const long long n, const double inv_count, float* __restrict__ mean) {
    __shared__ double red[THREADS];
    double v = 0.0;
    for (long long i = threadIdx.x; i < n; i += THREADS) v += static_cast<double>(partials[i]);
    const double total = block_sum(v, red);
    if (threadIdx.x == 0) mean[0] = static_cast<float>(total * inv_count);
}

// dL/dimg1 over the whole image; `grad_mean` is the scalar gradient of the mean (device memory), `inv_count` = 1 / size
// of the map.  Derivative maps are only defined on the map region; everything outside it reads as zero.
__global__ void __launch_bounds__(THREADS)
backward_kernel(const Plan plan, const float* __restrict__ img1, const float* __restrict__ img2,
                const float* __restrict__ dm_dmu1, const float* __restrict__ dm_dsigma1_sq,
                const float* __restrict__ dm_dsigma12, const float* __restrict__ grad_mean, const float inv_count,
                float* __restrict__ dL_dimg1) {
    __shared__ float d[3][SPAN][SPAN];
    __shared__ float h[3][SPAN][TILE];
    const Region r = map_region(plan);
    const int tid = threadIdx.x, H = plan.height, W = plan.width;
    const int oy0 = static_cast<int>(blockIdx.y) * TILE, ox0 = static_cast<int>(blockIdx.x) * TILE;
    const size_t base = static_cast<size_t>(blockIdx.z) * H * W;
    const float scale = __ldg(grad_mean) * inv_count;

    for (int i = tid; i < SPAN * SPAN; i += THREADS) {
        const int y = i / SPAN, x = i - y * SPAN;
        const int gy = oy0 - HALO + y, gx = ox0 - HALO + x;
        const bool in = gy >= r.y0 && gy < r.y1 && gx >= r.x0 && gx < r.x1;
        const size_t at = base + static_cast<size_t>(in ? gy : 0) * W + (in ? gx : 0);
        d[0][y][x] = in ? scale * __ldg(dm_dmu1 + at) : 0.f;
        d[1][y][x] = in ? scale * __ldg(dm_dsigma1_sq + at) : 0.f;
        d[2][y][x] = in ? scale * __ldg(dm_dsigma12 + at) : 0.f;
    }
    __syncthreads();

    for (int i = tid; i < SPAN * TILE; i += THREADS) {
        const int y = i / TILE, x = i - y * TILE;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < TAPS; k++) {
            const float w = plan.window[k];
            s0 += w * d[0][y][x + k];
            s1 += w * d[1][y][x + k];
            s2 += w * d[2][y][x + k];
        }
        h[0][y][x] = s0; h[1][y][x] = s1; h[2][y][x] = s2;
    }
    __syncthreads();

    {
        const int x = tid % TILE, y0 = (tid / TILE) * ROWS;
        float m[3][ROWS];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            float v[ROWS + TAPS - 1];
#pragma unroll
            for (int j = 0; j < ROWS + TAPS - 1; j++) v[j] = h[q][y0 + j][x];
#pragma unroll
            for (int r2 = 0; r2 < ROWS; r2++) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < TAPS; k++) acc += plan.window[k] * v[r2 + k];
                m[q][r2] = acc;
            }
        }
#pragma unroll
        for (int r2 = 0; r2 < ROWS; r2++) {
            const int py = oy0 + y0 + r2, px = ox0 + x;
            if (py >= H || px >= W) continue;
            const float v0 = m[0][r2], v1 = m[1][r2], v2 = m[2][r2];
            const size_t at = base + static_cast<size_t>(py) * W + px;
            dL_dimg1[at] = v0 + 2.f * __ldg(img1 + at) * v1 + __ldg(img2 + at) * v2;
        }
    }
}
}  // namespace ssim
}  // namespace ggs
