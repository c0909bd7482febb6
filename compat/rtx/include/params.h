// compat/rtx/include/params.h — the reference's launch-parameter block
// (rtx/include/params.h:14-42), field for field; layout-identical to rtxnerf_params
// (static_asserts below), which is what rtxnerf_ray_march_dense reads.
#pragma once
#include <cuda_runtime_api.h>
#include <stddef.h>

#include "../../optix.h"

struct RayGenData {};
struct HitGroupData {};
struct MissData {};

struct Params {
    OptixTraversableHandle handle;
    float* look_at;
    int num_primitives;
    int intersection_arr_size;
    OptixAabb* aabb;
    float3* start_points;
    float3* end_points;
    float3* ray_origins;
    float* t_start;
    float* t_end;
    int* num_hits;
    float2* viewing_direction;
    float focal_length;
    float aspect_ratio;
    float3 delta;
    float3 min_point;
    float3 max_point;
    unsigned int width, height;
};

static_assert(sizeof(Params) == sizeof(rtxnerf_params), "Params must mirror rtxnerf_params");
static_assert(offsetof(Params, num_hits) == offsetof(rtxnerf_params, num_hits), "Params layout");
static_assert(offsetof(Params, focal_length) == offsetof(rtxnerf_params, focal_length), "Params layout");
static_assert(offsetof(Params, width) == offsetof(rtxnerf_params, width), "Params layout");
