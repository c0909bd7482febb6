/*
 * optix.h — stand-in for the OptiX 7.7 SDK header, which is not in this image.
 * TEST INFRASTRUCTURE ONLY (oracle/).
 *
 *  - under nvcc it is EMPTY: the reference's sampler.cu / vol_render.cu include
 *    <optix.h> without using it, so an empty header is all they need to compile
 *    for sm_100a (oracle/_ref/libref_cuda.so).
 *  - under g++ it is a minimal EMULATION of the OptiX device API — just the calls
 *    rtx/src/optixPrograms.cu makes — so that the reference's own raygen /
 *    intersection / closest-hit / miss programs run unmodified on the host.  One
 *    optixTrace = test every primitive with the reference's intersection program,
 *    keep the smallest reported t inside [tmin, tmax] (ties: lowest primitive index,
 *    a declared convention — OptiX leaves it undefined), then run closest-hit or miss.
 *    This is our model of a closed runtime, not OptiX.
 */
#pragma once
#ifndef __CUDACC__

#include "cuda_host_shim.h"

typedef unsigned long long OptixTraversableHandle;
typedef unsigned int OptixVisibilityMask;
struct OptixAabb {
    float minX, minY, minZ, maxX, maxY, maxZ;
};
enum { OPTIX_RAY_FLAG_NONE = 0, OPTIX_RAY_FLAG_DISABLE_ANYHIT = 1 };

struct rtx_optix_ctx {
    uint3 launch_index;
    float3 origin, direction;
    float tmin, tmax;        /* current ray interval; tmax shrinks to the closest hit */
    unsigned prim;           /* primitive under test / closest primitive */
    unsigned payload[4];
    int n_prims;             /* the "GAS": primitives 0..n_prims-1 */
    int have_hit;
    float best_t;
    unsigned best_prim;
    /* guard: the reference has no bound on hits per ray (Q11); the emulator forces a
     * miss once a ray has filled its slots so a host run cannot write out of bounds */
    const int* num_hits;
    unsigned ray_stride_w;
    int arr_size;
    int guard_tripped;
};
static thread_local rtx_optix_ctx rtx_ctx;

extern "C" void __intersection__ray_march();
extern "C" void __closesthit__ray_march();
extern "C" void __miss__ray_march();

static inline uint3 optixGetLaunchIndex() { return rtx_ctx.launch_index; }
static inline float3 optixGetWorldRayOrigin() { return rtx_ctx.origin; }
static inline float3 optixGetWorldRayDirection() { return rtx_ctx.direction; }
static inline unsigned optixGetPrimitiveIndex() { return rtx_ctx.prim; }
static inline float optixGetRayTmax() { return rtx_ctx.tmax; }
static inline float optixGetRayTmin() { return rtx_ctx.tmin; }
static inline void optixSetPayload_0(unsigned v) { rtx_ctx.payload[0] = v; }
static inline void optixSetPayload_1(unsigned v) { rtx_ctx.payload[1] = v; }
static inline void optixSetPayload_2(unsigned v) { rtx_ctx.payload[2] = v; }
static inline void optixSetPayload_3(unsigned v) { rtx_ctx.payload[3] = v; }
static inline unsigned optixGetPayload_0() { return rtx_ctx.payload[0]; }
static inline void optixIgnoreIntersection() {}

static inline bool optixReportIntersection(float t, unsigned /*kind*/)
{
    /* accepted iff rayTmin <= t <= rayTmax; an equal-t later primitive does not replace */
    if (!(t >= rtx_ctx.tmin && t <= rtx_ctx.tmax)) return false;
    if (rtx_ctx.have_hit && !(t < rtx_ctx.best_t)) return false;
    rtx_ctx.have_hit = 1;
    rtx_ctx.best_t = t;
    rtx_ctx.best_prim = rtx_ctx.prim;
    return true;
}

static inline void optixTrace(OptixTraversableHandle, float3 o, float3 d, float tmin, float tmax,
                              float, OptixVisibilityMask, unsigned, unsigned, unsigned, unsigned,
                              unsigned& p0, unsigned& p1, unsigned& p2, unsigned& p3)
{
    rtx_optix_ctx& c = rtx_ctx;
    c.origin = o;
    c.direction = d;
    c.tmin = tmin;
    c.tmax = tmax;
    c.payload[0] = p0;
    c.payload[1] = p1;
    c.payload[2] = p2;
    c.payload[3] = p3;
    c.have_hit = 0;
    unsigned ray = c.launch_index.x + c.launch_index.y * c.ray_stride_w;
    if (c.num_hits && c.num_hits[ray] >= c.arr_size) {
        c.guard_tripped = 1;
    } else {
        for (int p = 0; p < c.n_prims; p++) {
            c.prim = (unsigned)p;
            __intersection__ray_march();
        }
    }
    if (c.have_hit) {
        c.tmax = c.best_t;
        c.prim = c.best_prim;
        __closesthit__ray_march();
    } else {
        __miss__ray_march();
    }
    p0 = c.payload[0];
    p1 = c.payload[1];
    p2 = c.payload[2];
    p3 = c.payload[3];
}

/* payload-free overload used only by the never-launched ray_sample pipeline */
static inline void optixTrace(OptixTraversableHandle, float3, float3, float, float, float,
                              OptixVisibilityMask, unsigned, unsigned, unsigned, unsigned)
{
}

#endif /* !__CUDACC__ */
