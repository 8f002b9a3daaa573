// lp_api.cu -- error state, device probes and version of liblp_b200.so.
#include <cuda_runtime.h>

#include <string>

#include "lp_host.h"

namespace lpb {

static thread_local std::string g_last_error;

void set_error(const std::string &msg) { g_last_error = msg; }

int cuda_fail(cudaError_t e, const char *what)
{
    g_last_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) return LPB200_E_NO_DEVICE;
    return LPB200_E_CUDA;
}

int get_device_info(DeviceInfo *info)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) {
        (void)cudaGetLastError();
        set_error("no CUDA device is visible: liblp_b200 has no CPU fallback");
        return LPB200_E_NO_DEVICE;
    }
    LPB_CUDA(cudaGetDevice(&info->device));
    LPB_CUDA(cudaDeviceGetAttribute(&info->sm_count, cudaDevAttrMultiProcessorCount, info->device));
    LPB_CUDA(cudaDeviceGetAttribute(&info->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, info->device));
    return 0;
}

}  // namespace lpb

extern "C" const char *lpb200_version(void) { return "lp_b200 0.1.0 (sm_100a)"; }

extern "C" const char *lpb200_last_error(void) { return lpb::g_last_error.c_str(); }

extern "C" int lpb200_device_count(void)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    return count;
}
