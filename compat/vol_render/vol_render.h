// compat/vol_render/vol_render.h — the reference's compositing interface
// (vol_render/vol_render.h:5-25) over librtxnerf_b200.so.  The reference launches on the legacy
// default stream (vol_render.cu:155,179): so do these (stream 0).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime_api.h>

#include "../../include/rtxnerf.h"

inline int& rtxnerf_volrender_status() { static thread_local int s = 0; return s; }

inline void launch_volrender_cuda(float* network_inputs, float* network_outputs, int* num_hits, int* indices,
                                  float* ray_hit, int batch_size, int num_samples_per_hit, float* pixels)
{
    rtxnerf_volrender_status() = rtxnerf_launch_volrender(network_inputs, network_outputs, num_hits, indices, ray_hit,
                                                          batch_size, num_samples_per_hit, pixels, nullptr);
}

inline void launch_volrender_backward_cuda(float* loss_values, __half* loss_gradients, float* sampled_points_radiance,
                                           float* t_hit, int* num_hits, int* indices, int batch_size,
                                           int num_samples_per_hit, __half* radiance_gradients)
{
    rtxnerf_volrender_status() = rtxnerf_launch_volrender_backward(
        loss_values, reinterpret_cast<const uint16_t*>(loss_gradients), sampled_points_radiance, t_hit, num_hits,
        indices, batch_size, num_samples_per_hit, reinterpret_cast<uint16_t*>(radiance_gradients), nullptr);
}
