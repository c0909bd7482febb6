/*
 * cuda_host_shim.h — lets g++ compile the reference's .cu files for the HOST so that
 * the reference's own kernels can be run on CPU cores as the parity pin / CPU baseline.
 * TEST INFRASTRUCTURE ONLY (oracle/); force-included by oracle/build.py with -include.
 *
 * What it supplies: empty CUDA qualifiers, thread-local threadIdx/blockIdx/blockDim,
 * a host launcher that walks the grid, CUDA's float min/max (fminf/fmaxf semantics)
 * and the bit-cast intrinsics.  float3/dim3/__half come from the CUDA toolkit's own
 * host-side headers.
 */
#pragma once
#ifndef __CUDACC__

#include <math.h>   /* libstdc++ wrapper: pulls std::abs/exp/sqrt float overloads into :: */
#include <stdlib.h>
#include <stdint.h>
#include <string.h>
#include <sys/types.h> /* uint */

#include <cuda_runtime_api.h>
#include <vector_functions.h>

#define __global__
#define __device__
#define __host__
#define __constant__
#define __forceinline__ inline
#define __restrict__

struct rtx_host_idx {
    unsigned x, y, z;
};
static thread_local rtx_host_idx threadIdx, blockIdx, blockDim, gridDim;

/* CUDA's overloads of min/max for float compile to PTX min.f32/max.f32. */
static inline float min(float a, float b) { return fminf(a, b); }
static inline float max(float a, float b) { return fmaxf(a, b); }

static inline unsigned __float_as_uint(float f)
{
    unsigned u;
    memcpy(&u, &f, 4);
    return u;
}
static inline float __uint_as_float(unsigned u)
{
    float f;
    memcpy(&f, &u, 4);
    return f;
}

/* kernel<<<grid, block, ...>>>(args) is rewritten by oracle/build.py (on the fly, nothing
 * stored) to rtx_host_launch(grid, block, kernel, args): one host call per CUDA thread. */
template <typename K, typename... A>
static void rtx_host_launch(dim3 grid, dim3 block, K kernel, A... args)
{
    const long long nblocks = (long long)grid.x * grid.y * grid.z;
#pragma omp parallel for schedule(dynamic, 1)
    for (long long b = 0; b < nblocks; b++) {
        gridDim = {grid.x, grid.y, grid.z};
        blockDim = {block.x, block.y, block.z};
        blockIdx = {(unsigned)(b % grid.x), (unsigned)((b / grid.x) % grid.y),
                    (unsigned)(b / ((long long)grid.x * grid.y))};
        for (unsigned tz = 0; tz < block.z; tz++)
            for (unsigned ty = 0; ty < block.y; ty++)
                for (unsigned tx = 0; tx < block.x; tx++) {
                    threadIdx = {tx, ty, tz};
                    kernel(args...);
                }
    }
}
template <typename K, typename... A>
static void rtx_host_launch(int grid, int block, K kernel, A... args)
{
    rtx_host_launch(dim3(grid), dim3(block), kernel, args...);
}

#endif /* !__CUDACC__ */
