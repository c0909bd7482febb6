// compat/optix.h — NOT OptiX.  The handful of OptiX names the reference's driver touches
// (main.cu:13-16 includes, :154-167 OptixAabb, :491 gas_handle, :504-508 optixLaunch), defined
// over librtxnerf_b200.so so that main.cu's ray-march call sequence compiles and links against
// the sm_100a occupancy-grid march (B200 has no RT cores).  Header-only; include/rtxnerf.h is
// the real interface.
#pragma once
#include <cuda_runtime_api.h>
#include <stddef.h>
#include <stdint.h>

#include "../include/rtxnerf.h"

typedef struct OptixAabb { float minX, minY, minZ, maxX, maxY, maxZ; } OptixAabb;   // main.cu:154-167
typedef unsigned long long OptixTraversableHandle;   // carries the rtxnerf_grid* (params.h:16)
typedef unsigned long long CUdeviceptr_compat;
#ifndef __cuda_cuda_h__
typedef unsigned long long CUdeviceptr;
#endif
typedef struct OptixDeviceContext_t* OptixDeviceContext;
typedef struct OptixModule_t* OptixModule;
typedef struct OptixProgramGroup_t* OptixProgramGroup;
typedef struct OptixPipeline_t { int which; }* OptixPipeline;
typedef struct OptixPipelineCompileOptions { int unused; } OptixPipelineCompileOptions;
typedef struct OptixShaderBindingTable { CUdeviceptr raygenRecord, missRecordBase, hitgroupRecordBase; } OptixShaderBindingTable;
typedef int OptixResult;
#define OPTIX_SUCCESS 0
#define OPTIX_SBT_RECORD_ALIGNMENT 16
#define OPTIX_SBT_RECORD_HEADER_SIZE 32

static inline const char* optixGetErrorString(OptixResult r) { return rtxnerf_error_string(r); }

// main.cu:506-508: optixLaunch(pipeline, stream, (CUdeviceptr)d_param, sizeof(Params), &sbt, W, H, 1).
// The reference passes a DEVICE copy of Params (main.cu:503); the march takes its arguments by
// value, so the block is read back here (132 bytes) before the launch.
static inline OptixResult optixLaunch(OptixPipeline, cudaStream_t stream, CUdeviceptr d_params, size_t params_size,
                                      const OptixShaderBindingTable*, unsigned int width, unsigned int height,
                                      unsigned int)
{
    rtxnerf_params p;
    if (params_size != sizeof(p)) return RTXNERF_EINVAL;
    cudaError_t e = cudaMemcpyAsync(&p, (const void*)d_params, sizeof(p), cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) return (int)e;
    e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return (int)e;
    p.width = width;
    p.height = height;
    return rtxnerf_ray_march_dense(&p, (void*)stream);
}
