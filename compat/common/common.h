// compat/common/common.h — the error macros and SBT record template the reference's driver uses
// (common/common.h:38-82), print-and-continue like the reference's.
#pragma once
#include <iostream>
#include <string>

#include "../optix.h"

#define OPTIX_CHECK(error)                                                                                         \
    {                                                                                                              \
        int _rtx_e = (error);                                                                                      \
        if (_rtx_e != OPTIX_SUCCESS)                                                                               \
            std::cerr << __FILE__ << ":" << __LINE__ << " march error: '" << optixGetErrorString(_rtx_e) << "'\n"; \
    }
#define CUDA_CHECK(error)                                                                                         \
    {                                                                                                             \
        cudaError_t _rtx_c = (error);                                                                             \
        if (_rtx_c != cudaSuccess)                                                                                \
            std::cerr << __FILE__ << ":" << __LINE__ << " CUDA Error: '" << cudaGetErrorString(_rtx_c) << "'\n";  \
    }
#define BUILD_DIR "./build/"

template <typename T> struct SbtRecord {
    alignas(OPTIX_SBT_RECORD_ALIGNMENT) char header[OPTIX_SBT_RECORD_HEADER_SIZE];
    T data;
};
