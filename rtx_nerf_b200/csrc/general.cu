// general.cu — version and error reporting of librtxnerf_b200.so.
#include "common.cuh"

unsigned long long g_rtx_kernel_launches = 0;

extern "C" unsigned long long rtxnerf_kernel_launches(void)
{
    return __atomic_load_n(&g_rtx_kernel_launches, __ATOMIC_RELAXED);
}

extern "C" int rtxnerf_abi_version(void) { return RTXNERF_ABI_VERSION; }

extern "C" const char* rtxnerf_error_string(int code)
{
    switch (code) {
        case RTXNERF_OK: return "ok";
        case RTXNERF_EINVAL: return "invalid argument";
        case RTXNERF_EGRID: return "grid resolution must be a power of two in [2, 1024]";
        case RTXNERF_ECAPACITY: return "caller-provided capacity too small";
        case RTXNERF_ENOMEM: return "out of host memory";
        case RTXNERF_ECONFIG: return "unsupported network / encoding configuration";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown rtxnerf error";
}
