// common.cuh — shared helpers for the sm_100a kernels of librtxnerf_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdlib.h>

#include <atomic>

#include "../../include/rtxnerf.h"

#define RTX_NUM_SMS 148  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

// the occupancy grid behind the opaque rtxnerf_grid handle (march.cu, grid_update.cu)
struct rtxnerf_grid {
    int G;
    size_t n_words;
    uint32_t* d_bits;   // G^3 bits, bit (p&31) of word (p>>5), p = (x*G+y)*G+z
    unsigned long long* d_bricks;   // the same bits, one 64-bit word per 4x4x4 cells (march.cu: GridBits)
    unsigned gb;                    // bricks per axis = ceil(G / 4)
    float* d_aabb;      // lazily built make_grid list (compat with Params.aabb)
};

// re-derive d_bricks from d_bits on `stream` (march.cu); every writer of d_bits calls it
int rtx_grid_rebuild_bricks(const rtxnerf_grid* g, cudaStream_t stream);

#define RTX_CUDA_TRY(expr)                      \
    do {                                        \
        cudaError_t _e = (expr);                \
        if (_e != cudaSuccess) return (int)_e;  \
    } while (0)

// Kernels launched by this library in this process (rtxnerf_kernel_launches(), general.cu):
// bench.py's gpu_launches claim is read from here, not counted by the caller.
extern unsigned long long g_rtx_kernel_launches;
static inline int rtx_launch_status(int n_launched = 1)
{
    __atomic_fetch_add(&g_rtx_kernel_launches, (unsigned long long)n_launched, __ATOMIC_RELAXED);
    return (int)cudaGetLastError();
}

// Function attributes (dynamic shared memory above 48 KB) are per device and per context: remember
// which devices of this process have them, so that a host driving several GPUs from several
// threads (SURVEY 8b "Threading") gets them on each.  Racing threads may both set the attribute;
// that is idempotent.
struct RtxPerDeviceOnce {
    std::atomic<unsigned long long> done[4];   // 256 device ordinals
};
// 1: the current device is already marked, 0: not yet, < 0 never (errors are returned as positive CUDA codes by
// the callers through RTX_CUDA_TRY on cudaGetDevice)
static inline bool rtx_device_done(RtxPerDeviceOnce& st)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    return (st.done[(dev >> 6) & 3].load(std::memory_order_acquire) >> (dev & 63)) & 1ull;
}
static inline void rtx_device_mark(RtxPerDeviceOnce& st)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return;
    st.done[(dev >> 6) & 3].fetch_or(1ull << (dev & 63), std::memory_order_release);
}
template <class Kernel>
static inline int rtx_set_max_smem(RtxPerDeviceOnce& st, Kernel kernel, int bytes)
{
    if (rtx_device_done(st)) return 0;
    RTX_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    rtx_device_mark(st);
    return 0;
}
// development knobs from the environment: read once per process (C++11 static initialisation is
// thread-safe), never per device
static inline int rtx_env_int(const char* name, int dflt)
{
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

// ---------------------------------------------------------------------------------------
// Bit-defined fp32 arithmetic.  The hit-cell lists must equal the reference's bit for bit,
// so everything that feeds a comparison is spelled with the _rn intrinsics: nvcc never
// contracts or reorders them, and the CPU oracle spells the same sequence with fmaf().
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float rn_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float rn_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float rn_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float rn_div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float rn_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// 256-bit global accesses (sm_100: LDG.E.256 / STG.E.256), 32-byte aligned addresses.  When the lanes of a warp
// read records that lie far apart, what a load costs the L1 is one tag lookup per lane and per instruction,
// whatever its width: a 32-byte record is one instruction instead of two (or of eight scalar ones).
__device__ __forceinline__ void ldg_v8(const void* p, uint32_t (&r)[8])
{
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}
__device__ __forceinline__ void stg_v8(void* p, const uint32_t (&r)[8])
{
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}

// streaming (read-once / write-once) global accesses: keep them out of L1
__device__ __forceinline__ float4 ld_stream_f4(const float4* p)
{
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p)
{
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ float ld_stream_f1(const float* p)
{
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream_f4(float4* p, float4 v)
{
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ void st_stream_f1(float* p, float v)
{
    asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

__device__ __forceinline__ int warp_incl_scan_i32(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}
