// View-parallel training: the SH-coefficient gradient without an all-reduce of [P, SHM, 3], and the sum of the
// other parameter gradients over views through peer memory (NVLink / NVSwitch) in the same launch.
//
// For one view the SH gradient of Gaussian p is rank one (render_backward.cu:76-160):
//     dL/dsh[p][k][c] = basis_k(dir(p, campos)) * dL/dRGB[p][c]          (after the colour clamp mask)
// so a rank only needs the 3 floats dL/dRGB[p] of every other view (and that view's camera position) to form the
// SUM over views locally.  At SH degree 3 that replaces an all-reduce of 192 bytes per Gaussian by 12 bytes per
// Gaussian and view -- 81 % of the parameter-gradient bytes leave the wire.
// One thread per Gaussian; the [P][SHM*3] output rows are written as warp slabs through shared memory
// (coalesced 16-byte stores), as in ggs_gaussian_backward.cu.
//
// view_sum_kernel is the fused form: every rank's per-view gradients live in a symmetric exchange buffer that all
// ranks of the node map (torch symmetric memory: cuMem handles exchanged once, at set-up).  The leading blocks of
// the launch own 1/world of the flat region of the remaining 11 floats per Gaussian: they sum it over the views
// (peer loads, or ONE multimem.ld_reduce per 16 bytes when the buffer has an NVSwitch multicast mapping: the switch
// adds) and store the sum into every rank's result region (peer stores, or one multimem.st).  The other blocks pull
// dL/dRGB of every view straight out of the peers' buffers and write the summed SH gradient locally.  No NCCL
// call, no staging copy: the per-Gaussian backward writes its gradients directly into the exchange buffer.
#include "ggs_kernels.cuh"
#include "ggs_math.cuh"

namespace ggs {

#ifndef GGS_MV_THREADS
#define GGS_MV_THREADS 128
#endif
#ifndef GGS_MV_MIN_BLOCKS
#define GGS_MV_MIN_BLOCKS 6  // 80 registers: 102 -> 88 us for 8 views of 1 M Gaussians (tools/sh_sum_bench.py); 8 blocks (64 registers) spill
#endif
constexpr int MV_THREADS = GGS_MV_THREADS;

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float4* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* mc, const float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}

// Leading blocks of view_sum_kernel: out[i] = sum over ranks of send_r[i] for this rank's slice of the flat region,
// stored into every rank's recv region.  16 bytes per thread and step, grid-stride.
__device__ __forceinline__ float4 sum_over_ranks(const PeerJob& job, size_t i) {
    if (job.mc_send) return multimem_ld_reduce_add(job.mc_send + i);
    // four loads in flight per step (the slots are local memory in the push design; this kernel's register budget
    // is set by its SH half)
    float4 acc = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
    for (int r0 = 0; r0 < job.world; r0 += 4) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            v[k] = (r0 + k < job.world) ? __ldcg(job.send[r0 + k] + i) : float4{0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            acc.x += v[k].x;
            acc.y += v[k].y;
            acc.z += v[k].z;
            acc.w += v[k].w;
        }
    }
    return acc;
}
__device__ __forceinline__ void store_to_ranks(const PeerJob& job, size_t i, const float4 acc) {
    if (job.mc_recv) {
        multimem_st(job.mc_recv + i, acc);
        return;
    }
#pragma unroll
    for (int r = 0; r < MAX_VIEWS; r++)
        if (r < job.world) job.recv[r][i] = acc;
}
__device__ __forceinline__ void reduce_slice(const PeerJob& job, int block, int blocks) {
    const size_t stride = (size_t)blocks * blockDim.x;
    size_t i = job.lo4 + (size_t)block * blockDim.x + threadIdx.x;
    // two independent 16-byte columns per step: the link needs ~2 MB in flight (900 GB/s x ~2 us)
    for (; i + stride < job.hi4; i += 2 * stride) {
        const float4 a = sum_over_ranks(job, i);
        const float4 b = sum_over_ranks(job, i + stride);
        store_to_ranks(job, i, a);
        store_to_ranks(job, i + stride, b);
    }
    if (i < job.hi4) store_to_ranks(job, i, sum_over_ranks(job, i));
}

__global__ void __launch_bounds__(MV_THREADS, GGS_MV_MIN_BLOCKS)
    view_sum_kernel(int P, int SHD, int SHM, int V, const float* __restrict__ means3D, const ViewSources src,
                    float* __restrict__ dL_dsh, const PeerJob job) {
    extern __shared__ float s_slab[];  // [warps][32][RS]
    if ((int)blockIdx.x < job.blocks) {
        reduce_slice(job, blockIdx.x, job.blocks);
        return;
    }
    const uint32_t block = blockIdx.x - job.blocks;
    const size_t idx = (size_t)block * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int F = SHM * 3;
    const int RS = F | 1;
    float* const swarp = s_slab + (size_t)warp * 32 * RS;
    float* const srow = swarp + lane * RS;
    const size_t g0 = (size_t)block * blockDim.x + warp * 32;
    const int ng = g0 < (size_t)P ? (int)min((size_t)32, (size_t)P - g0) : 0;

    // The warp's 32 colour-gradient rows of a view are one contiguous 384-byte segment: 24 lanes copy 16 bytes each
    // into the warp's stage, for ALL views before any is consumed (independent loads, issued back to back), and rows
    // are then handed out from shared memory.  Misaligned sources (the all-gathered [V,P,3] form with P % 4 != 0)
    // and the last, partial warp take the per-lane path.
    float* const s_stage = s_slab + (size_t)(MV_THREADS / 32) * 32 * RS + (size_t)warp * V * 96;  // [V][96]
    // camera positions once per block (every warp of the grid asking the L2 for the same sector serialises there)
    __shared__ float s_cam[MAX_VIEWS * 3];
    if (threadIdx.x < 3 * V) s_cam[threadIdx.x] = __ldcg(src.campos[threadIdx.x / 3] + threadIdx.x % 3);
    __syncthreads();
    uint32_t vec_ok = (ng == 32);
#pragma unroll
    for (int v = 0; v < MAX_VIEWS; v++)
        if (v < V) vec_ok &= ((reinterpret_cast<uintptr_t>(src.rgb[v] + g0 * 3) & 15) == 0);
    if (vec_ok && lane < 24) {
#pragma unroll 4
        for (int v = 0; v < V; v++)
            reinterpret_cast<float4*>(s_stage + v * 96)[lane] =
                __ldcg(reinterpret_cast<const float4*>(src.rgb[v] + g0 * 3) + lane);
    }
    __syncwarp();
    float3 mean = {0.f, 0.f, 0.f};
    if (idx < (size_t)P) mean = {__ldg(means3D + 3 * idx), __ldg(means3D + 3 * idx + 1), __ldg(means3D + 3 * idx + 2)};
    float acc[48];  // the row's sums over views stay in registers (every index below is a compile-time constant)
#pragma unroll
    for (int k = 0; k < 48; k++) acc[k] = 0.f;
#pragma unroll 1
    for (int v = 0; v < V; v++) {
        float3 d;
        if (vec_ok) {
            const float* sv = s_stage + v * 96 + 3 * lane;
            d = {sv[0], sv[1], sv[2]};
        } else if (idx < (size_t)P) {
            const float* gr = src.rgb[v] + idx * 3;  // possibly a peer's memory: L2-coherent loads
            d = {__ldcg(gr), __ldcg(gr + 1), __ldcg(gr + 2)};
        } else {
            d = {0.f, 0.f, 0.f};
        }
        if (idx < (size_t)P) {
            if (d.x == 0.f && d.y == 0.f && d.z == 0.f) continue;  // culled (or untouched) in this view
            const float3 dir_orig = {mean.x - s_cam[3 * v], mean.y - s_cam[3 * v + 1], mean.z - s_cam[3 * v + 2]};
            // same normalisation as the per-view kernels (render_forward.cu:31-32)
            const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
            const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
            auto add = [&](int k, float w) {
                acc[3 * k] += w * d.x;
                acc[3 * k + 1] += w * d.y;
                acc[3 * k + 2] += w * d.z;
            };
            add(0, kSH_C0);
            if (SHD > 0) {
                add(1, -kSH_C1 * y);
                add(2, kSH_C1 * z);
                add(3, -kSH_C1 * x);
                if (SHD > 1) {
                    const float xx = x * x, yy = y * y, zz = z * z;
                    const float xy = x * y, yz = y * z, xz = x * z;
                    add(4, kSH_C2[0] * xy);
                    add(5, kSH_C2[1] * yz);
                    add(6, kSH_C2[2] * (2.f * zz - xx - yy));
                    add(7, kSH_C2[3] * xz);
                    add(8, kSH_C2[4] * (xx - yy));
                    if (SHD > 2) {
                        add(9, kSH_C3[0] * y * (3.f * xx - yy));
                        add(10, kSH_C3[1] * xy * z);
                        add(11, kSH_C3[2] * y * (4.f * zz - xx - yy));
                        add(12, kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
                        add(13, kSH_C3[4] * x * (4.f * zz - xx - yy));
                        add(14, kSH_C3[5] * z * (xx - yy));
                        add(15, kSH_C3[6] * x * (xx - 3.f * yy));
                    }
                }
            }
        }
    }
    if (idx < (size_t)P) {
#pragma unroll
        for (int k = 0; k < 48; k++)
            if (k < F) srow[k] = acc[k];
    }
    __syncwarp();
    if ((F & 3) == 0) {
        store_rows_from_shared(reinterpret_cast<float4*>(dL_dsh + g0 * F), F >> 2, ng, swarp, RS, lane);
    } else {
        float* dst = dL_dsh + g0 * F;
        int gi = 0, rem = lane;
        for (int e = lane; e < ng * F; e += 32, rem += 32) {
            while (rem >= F) {
                rem -= F;
                gi++;
            }
            dst[e] = swarp[gi * RS + rem];
        }
    }
}

// Sum over ranks of up to eight ranges of the gradient region through the switch: out[i] = multimem.ld_reduce(send[i]),
// multimem.st(recv[i]), 16 bytes per unit.  Used for the pieces of the region a chunk of the per-Gaussian backward has
// just written (ggs_peer_sum_ranges): the links sum chunk k while the SMs compute chunk k + 1.  The ranges are walked
// as one concatenated index space, two columns in flight per thread (more in flight did not help: the in-switch
// reduction sustains ~350 GB/s per GPU whatever the issue pattern -- 5.5 MB per rank in 114 us with this shape, 128 us
// with four columns and twice the threads).
constexpr int RANGE_THREADS = 128;
__global__ void __launch_bounds__(RANGE_THREADS) sum_ranges_multimem_kernel(const float4* __restrict__ mc_send,
                                                                           float4* __restrict__ mc_recv,
                                                                           const SumRanges r, const size_t total) {
    const size_t stride = (size_t)gridDim.x * RANGE_THREADS;
    auto locate = [&](size_t j) -> size_t {  // unit j of the concatenation -> unit of the region
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (k < r.count) {
                const size_t n = r.hi4[k] - r.lo4[k];
                if (j < n) return r.lo4[k] + j;
                j -= n;
            }
        }
        return r.lo4[0];
    };
    size_t j = (size_t)blockIdx.x * RANGE_THREADS + threadIdx.x;
    for (; j + stride < total; j += 2 * stride) {
        const size_t ia = locate(j), ib = locate(j + stride);
        const float4 a = multimem_ld_reduce_add(mc_send + ia);
        const float4 b = multimem_ld_reduce_add(mc_send + ib);
        multimem_st(mc_recv + ia, a);
        multimem_st(mc_recv + ib, b);
    }
    if (j < total) {
        const size_t ia = locate(j);
        multimem_st(mc_recv + ia, multimem_ld_reduce_add(mc_send + ia));
    }
}

void launch_sum_ranges_multimem(const float4* mc_send, float4* mc_recv, const SumRanges& r, cudaStream_t stream) {
    size_t total = 0;
    for (int k = 0; k < r.count; k++) total += r.hi4[k] - r.lo4[k];
    if (total == 0) return;
    const size_t want = (total + 2 * RANGE_THREADS - 1) / (2 * RANGE_THREADS);
    const unsigned blocks = (unsigned)(want < (size_t)(148 * 4) ? want : (size_t)(148 * 4));
    sum_ranges_multimem_kernel<<<blocks, RANGE_THREADS, 0, stream>>>(mc_send, mc_recv, r, total);
}

static void launch_view_sum(int P, int SHD, int SHM, int V, const float* means3D, const ViewSources& src,
                            float* dL_dsh, const PeerJob& job, cudaStream_t stream) {
    // output slab [warps][32][RS] + colour stage [warps][V][96]; RS = 3 SHM | 1 keeps the slab a multiple of 16 bytes
    const size_t smem = (size_t)(MV_THREADS / 32) * (32 * ((SHM * 3) | 1) + (size_t)V * 96) * sizeof(float);
    static SmemOptIn opted;
    opt_in_shared_memory(view_sum_kernel, smem, opted);
    const int sh_blocks = (P + MV_THREADS - 1) / MV_THREADS;
    if (sh_blocks + job.blocks == 0) return;
    view_sum_kernel<<<sh_blocks + job.blocks, MV_THREADS, smem, stream>>>(P, SHD, SHM, V, means3D, src, dL_dsh, job);
}

void launch_sh_gradient_from_views(int P, int SHD, int SHM, int V, const float* means3D, const float* campos,
                                   const float* dL_drgb, float* dL_dsh, cudaStream_t stream) {
    if (P == 0) return;
    ViewSources src = {};
    for (int v = 0; v < V; v++) {
        src.rgb[v] = dL_drgb + (size_t)v * P * 3;
        src.campos[v] = campos + 3 * v;
    }
    PeerJob job = {};
    launch_view_sum(P, SHD, SHM, V, means3D, src, dL_dsh, job, stream);
}

void launch_view_sum_peer(int P, int SHD, int SHM, const float* means3D, const ViewSources& src, float* dL_dsh,
                          const PeerJob& job, cudaStream_t stream) {
    launch_view_sum(P, SHD, SHM, job.world, means3D, src, dL_dsh, job, stream);
}

}  // namespace ggs
