// Staging of per-tile Gaussian batches into shared memory, shared by the tile kernels (forward and backward).
// A list entry is an index into the 64-byte SplatRecord array, so a batch is a gather: every thread copies the
// quads of one record with cp.async (16 bytes each, L1 bypassed: neighbouring tiles reuse records through the
// L2) into an 80-byte slot of a double buffer, so the gather of batch i+1 overlaps the arithmetic on batch i.
#pragma once
#include "ggs_common.cuh"

namespace ggs {

constexpr int TILE_BATCH = TILE_PIXELS;  // list entries per staged batch (one per thread)
constexpr int TILE_WARPS = TILE_PIXELS / 32;
constexpr int REC_QUADS = 5;  // 80-byte shared-memory stride: conflict-free 16-byte stores, broadcast reads
constexpr int STAGE_QUADS = TILE_BATCH * REC_QUADS;  // float4s per staging buffer

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Gathers the first `quads` (3 or 4) 16-byte pieces of the record of Gaussian `id` into slot `slot`.
__device__ __forceinline__ void stage_record(float4* buf, int slot, const SplatRecord* records, uint32_t id,
                                             int quads) {
    const float4* src = reinterpret_cast<const float4*>(records + id);
    float4* dst = buf + slot * REC_QUADS;
    cp_async16(dst, src);
    cp_async16(dst + 1, src + 1);
    cp_async16(dst + 2, src + 2);
    if (quads > 3) cp_async16(dst + 3, src + 3);
}

// Staged records are read through 32-bit shared-memory addresses: one multiply-add per record instead of the 64-bit
// generic pointer arithmetic the compiler emits for `buf + slot * REC_QUADS`.
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}

}  // namespace ggs
