/*
 * ref_cuda_glue.cu — C entry points over the reference's own CUDA launchers, linked with
 * the reference's sampler.cu / vol_render.cu compiled in place for sm_100a
 * (oracle/_ref/libref_cuda.so).  TEST INFRASTRUCTURE ONLY: second, independent GPU
 * oracle and "the kernels to beat" timing.  All pointers are device pointers.
 */
#include <cuda_runtime_api.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "sampler.h"
#include "vol_render.h"

extern "C" int ref_cuda_launchSampler(float* start_points, float* end_points, float* view_dirs,
                                      float* t_vals, float* samples, int batch_size, int grid_res,
                                      int* num_hits, int* indices, int sample_type, void* stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    launchSampler((float3*)start_points, (float3*)end_points, (float2*)view_dirs, t_vals, samples,
                  batch_size, grid_res, num_hits, indices, (SAMPLING_TYPE)sample_type, s);
    return (int)cudaGetLastError();
}
extern "C" int ref_cuda_volrender(float* network_outputs, int* num_hits, int* indices,
                                  float* ray_hit, int batch_size, int S, float* pixels)
{
    launch_volrender_cuda(nullptr, network_outputs, num_hits, indices, ray_hit, batch_size, S, pixels);
    return (int)cudaGetLastError();
}
extern "C" int ref_cuda_volrender_backward(uint16_t* loss_gradients, float* radiance, float* t_hit,
                                           int* num_hits, int* indices, int batch_size, int S,
                                           uint16_t* radiance_gradients)
{
    launch_volrender_backward_cuda(nullptr, (__half*)loss_gradients, radiance, t_hit, num_hits,
                                   indices, batch_size, S, (__half*)radiance_gradients);
    return (int)cudaGetLastError();
}
