// Test infrastructure: just enough of the CUDA execution model to run a kernel's SOURCE on the host, one block at a
// time with one OS thread per CUDA thread and a real barrier, so that indexing, shared-memory staging, partial blocks
// and barrier placement of a kernel can be checked without a GPU.  Not a simulator of warps: no shuffles, no votes,
// no atomics ordering, no performance meaning.  Include BEFORE the kernel header; build with g++ -pthread.
#pragma once
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <string.h>

#include <thread>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static  // one block runs at a time, all its threads see the same function-local static

struct uint3 {
    unsigned x, y, z;
};
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct int2 {
    int x, y;
};
struct float2 {
    float x, y;
};
struct float4 {
    float x, y, z, w;
};

static thread_local uint3 threadIdx, blockIdx;
static dim3 blockDim, gridDim;
static pthread_barrier_t emu_block_barrier;

inline void __syncthreads() { pthread_barrier_wait(&emu_block_barrier); }
template <class T>
inline T __ldg(const T* p) {
    return *p;
}
inline int __float2int_rd(float x) { return (int)floorf(x); }
inline int __float2int_ru(float x) { return (int)ceilf(x); }
inline long long min(long long a, long long b) { return a < b ? a : b; }
inline long long max(long long a, long long b) { return a > b ? a : b; }
inline int min(int a, int b) { return a < b ? a : b; }
inline int max(int a, int b) { return a > b ? a : b; }

// emu_launch(grid, block, [&] { kernel(args...); });
template <class Body>
void emu_launch(dim3 grid, dim3 block, Body body) {
    gridDim = grid;
    blockDim = block;
    const unsigned per_block = block.x * block.y * block.z;
    for (unsigned bz = 0; bz < grid.z; bz++)
        for (unsigned by = 0; by < grid.y; by++)
            for (unsigned bx = 0; bx < grid.x; bx++) {
                pthread_barrier_init(&emu_block_barrier, nullptr, per_block);
                std::vector<std::thread> threads;
                threads.reserve(per_block);
                for (unsigned t = 0; t < per_block; t++)
                    threads.emplace_back([=] {
                        blockIdx = uint3{bx, by, bz};
                        threadIdx = uint3{t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
                        body();
                    });
                for (auto& th : threads) th.join();
                pthread_barrier_destroy(&emu_block_barrier);
            }
}
