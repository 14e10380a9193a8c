// capi.cu -- ABI identification and status strings of libgs_b200.so (see include/gs_b200.h).
#include "gs_common.cuh"

extern "C" int gs_abi_version(void) { return 113; }

extern "C" const char *gs_status_string(int status)
{
    if (status == GS_OK) return "ok";
    if (status == GS_E_INVALID) return "invalid argument";
    if (status == GS_E_CAPACITY) return "count exceeds the int32 / 30-bit index contract";
    if (status == GS_E_STATIC_LIMIT)
        return "size exceeds the fixed scratch of a *_static entry point (see GS_PROJECT_STATIC_MAX_POINTS)";
    if (status > 0) return cudaGetErrorString((cudaError_t)status);
    return "unknown status";
}
