// simplex_batched_variants.h -- the kernel variants compiled in simplex_batched_variants.cu, as seen from the
// host code in simplex_batched.cu.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "lp_host.h"

enum {
    LPB_V_SMEM8 = 0,  // the product kernel (simplex_batched.cu)
    LPB_V_REG15,
    LPB_V_TINY1,
    LPB_V_SMALL1,
    LPB_V_MID4,
    LPB_V_BLAND8,
    LPB_V_DANTZIG8,
    LPB_V_DEVEX8,
    LPB_V_DEVEX_TINY1,
    LPB_V_BLAND_TINY1,
    LPB_V_DANTZIG_TINY1,
};

int lpb_batch_variant_launch(int variant, const lpb::DeviceInfo &di, int B, int n, int m, const void *A, int a_dtype,
                             const double *b, const double *c, double tol, int max_pivots, int32_t *status,
                             double *x, double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                             cudaStream_t stream);
long long lpb_batch_variant_smem_bytes(int variant, int n, int m);  // dynamic shared memory the variant needs
bool lpb_batch_variant_eligible(int variant, int n, int m);
