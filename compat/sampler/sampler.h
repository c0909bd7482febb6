// compat/sampler/sampler.h — the reference's sampler interface (sampler/sampler.h:4-30) over
// librtxnerf_b200.so.  Same names, argument order and meaning; forwards to
// rtxnerf_launch_sampler with the reference's compile-time 32 samples per segment.
#pragma once
#include <cuda_runtime_api.h>

#include "../../include/rtxnerf.h"

#define NUM_SAMPLES_PER_SEGMENT 32
enum SAMPLING_TYPE { SAMPLING_REGULAR, SAMPLING_STRATIFIED_JITTERING, SAMPLING_UNIFORM };

struct float5 { float x, y, z, theta, phi; };   // declared by the reference, unused there too

// Asynchronous on `stream` (passed by reference, as in the reference).  Unlike the reference's
// void launcher the status is available: rtxnerf_last_status() after the call.
inline int& rtxnerf_last_status() { static thread_local int s = 0; return s; }

inline void launchSampler(float3* d_start_points, float3* d_end_points, float2* d_view_dirs, float* d_t_vals,
                          float* d_sampled_points, int batch_size, int grid_res, int* d_num_hits, int* d_indices,
                          SAMPLING_TYPE sample_type, cudaStream_t& stream)
{
    rtxnerf_last_status() = rtxnerf_launch_sampler(
        reinterpret_cast<const float*>(d_start_points), reinterpret_cast<const float*>(d_end_points),
        reinterpret_cast<const float*>(d_view_dirs), d_t_vals, d_sampled_points, batch_size, grid_res, d_num_hits,
        d_indices, (int)sample_type, NUM_SAMPLES_PER_SEGMENT, (void*)stream);
}
