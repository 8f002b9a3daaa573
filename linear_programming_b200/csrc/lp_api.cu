// lp_api.cu -- error state, device probes and version of liblp_b200.so.
#include <cuda_runtime.h>

#include <mutex>
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
    LPB_CUDA(cudaDeviceGetAttribute(&info->max_smem_per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, info->device));
    return 0;
}

// Stream-ordered scratch (the batched kernel's 4-byte work counter) comes from a pool of our own
// that keeps its memory across synchronisation points: the device's default pool hands unused
// memory back at every sync (release threshold 0), which made each call issued on an idle GPU pay
// ~240 us for a fresh physical allocation.
int scratch_alloc(void **ptr, size_t bytes, int device, cudaStream_t stream)
{
    static std::mutex mu;
    static cudaMemPool_t pools[64] = {};
    if (device < 0 || device >= 64) {
        set_error("scratch_alloc: unsupported device ordinal");
        return LPB200_E_INVALID;
    }
    cudaMemPool_t pool;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!pools[device]) {
            cudaMemPoolProps props = {};
            props.allocType = cudaMemAllocationTypePinned;
            props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice;
            props.location.id = device;
            LPB_CUDA(cudaMemPoolCreate(&pools[device], &props));
            unsigned long long keep = ~0ull;
            LPB_CUDA(cudaMemPoolSetAttribute(pools[device], cudaMemPoolAttrReleaseThreshold, &keep));
        }
        pool = pools[device];
    }
    LPB_CUDA(cudaMallocFromPoolAsync(ptr, bytes, pool, stream));
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
