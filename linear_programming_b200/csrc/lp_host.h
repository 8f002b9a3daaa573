// lp_host.h -- host-side helpers shared by the translation units of liblp_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/lp_b200.h"

namespace lpb {

void set_error(const std::string &msg);
int cuda_fail(cudaError_t e, const char *what);

#define LPB_CUDA(call)                                        \
    do {                                                      \
        cudaError_t _e = (call);                              \
        if (_e != cudaSuccess) return lpb::cuda_fail(_e, #call); \
    } while (0)

// Size of the table CPython ends with after set(range(lo, lo+cnt)) (grown 8 -> ... one add at a
// time, resize when fill*5 >= mask*3 to the first power of two > 4*used).
inline int pyset_initial_size(int cnt)
{
    int size = 8;
    for (int fill = 1; fill <= cnt; ++fill) {
        if ((long long)fill * 5 >= (long long)(size - 1) * 3) {
            int ns = 8;
            while (ns <= fill * 4) ns <<= 1;
            size = ns;
        }
    }
    return size;
}

// Largest table a set that never holds more than cnt keys can be rebuilt into.
inline int pyset_capacity(int cnt)
{
    int ns = 8;
    while (ns <= cnt * 4) ns <<= 1;
    const int init = pyset_initial_size(cnt);
    return ns > init ? ns : init;
}

// Bytes per element of an LPB200_DTYPE_* code; 0 = not a dtype.
inline int dtype_bytes(int dtype)
{
    switch (dtype) {
    case LPB200_DTYPE_F64: return 8;
    case LPB200_DTYPE_F32: return 4;
    case LPB200_DTYPE_I32: return 4;
    case LPB200_DTYPE_I16: return 2;
    case LPB200_DTYPE_I8: return 1;
    default: return 0;
    }
}

struct DeviceInfo {
    int device;
    int sm_count;
    int max_smem_optin;   // per CTA (opt-in)
    int max_smem_per_sm;  // per SM
};
int get_device_info(DeviceInfo *info);
// stream-ordered scratch from the library's own retaining pool; release with cudaFreeAsync
int scratch_alloc(void **ptr, size_t bytes, int device, cudaStream_t stream);

}  // namespace lpb
