// compat/rtx/include/rtxFunctions.h — the reference's OptiX host wrapper
// (rtx/include/rtxFunctions.h:56-100) as a thin holder of an rtxnerf_grid.  Context, module,
// program groups, pipeline and SBT have nothing to build without OptiX and are no-ops, so the
// reference's call sequence (main.cu:381-399) compiles and runs unchanged;
// initAccelerationStructure builds the occupancy bitfield from the AABB list and returns the
// DEVICE AABB array the caller stores in Params.aabb (main.cu:399,492).
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../../common/common.h"
#include "params.h"

struct RTXDataHolder {
    Params params = {};
    OptixTraversableHandle gas_handle = 0;               // read by the caller (main.cu:491)
    OptixShaderBindingTable sbt_ray_march = {};          // passed to optixLaunch (main.cu:504)
    OptixShaderBindingTable sbt_ray_sample = {};
    OptixPipeline pipeline_ray_march = nullptr;          // passed to optixLaunch (main.cu:506)
    OptixPipeline pipeline_ray_sample = nullptr;
    cudaStream_t stream = nullptr;
    rtxnerf_grid* grid = nullptr;

    void initContext() {}
    void createModule(const std::string) {}
    void createProgramGroups() {}
    void linkPipeline(bool) {}
    void buildSBT() {}
    void setStream(const cudaStream_t& stream_in) { stream = stream_in; }

    // The grid resolution is recovered from the cell size of the first AABB (make_grid cells,
    // main.cu:154-174).  Throws std::runtime_error like OPTIX_ERROR_CHECK (rtxFunctions.cpp:41-49).
    OptixAabb* initAccelerationStructure(const std::vector<OptixAabb>& cells)
    {
        if (cells.empty()) throw std::runtime_error("initAccelerationStructure: empty grid");
        const float L = cells[0].maxX - cells[0].minX;
        const int G = (int)(2.0f / L + 0.5f);
        if (grid) rtxnerf_grid_destroy(grid);
        int rc = rtxnerf_grid_create_from_aabbs(&cells[0].minX, (int)cells.size(), G, &grid);
        if (rc) throw std::runtime_error(std::string("initAccelerationStructure: ") + rtxnerf_error_string(rc));
        gas_handle = (OptixTraversableHandle)(uintptr_t)grid;
        float* d_aabb = nullptr;
        rc = rtxnerf_grid_device_aabbs(grid, &d_aabb);
        if (rc) throw std::runtime_error(std::string("initAccelerationStructure: ") + rtxnerf_error_string(rc));
        return reinterpret_cast<OptixAabb*>(d_aabb);
    }

    ~RTXDataHolder()
    {
        if (grid) rtxnerf_grid_destroy(grid);
    }
};
