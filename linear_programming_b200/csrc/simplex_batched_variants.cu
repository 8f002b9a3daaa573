// simplex_batched_variants.cu -- the non-headline instantiations of the one-CTA-per-LP simplex kernel
// (simplex_batched_impl.cuh): smaller CTAs for smaller problems, the register-resident experiment and the
// corrected pivoting rules.  Kept in their own translation unit so that the product kernel (smem8, in
// simplex_batched.cu) builds in parallel with them.  simplex_batched.cu reaches them through the three
// dispatchers at the bottom (declared in simplex_batched_variants.h).
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>

#include <string>

#include "lp_common.cuh"
#include "lp_host.h"
#include "simplex_batched_variants.h"

#define LPB_NS reg15
#define LPB_SWEEP_WARPS 15
#define LPB_REG_ROWS 12
#define LPB_MIN_BLOCKS 1
#define LPB_RULE 0
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

// Small problems: one sweep warp + the bookkeeping warp per LP.  A 64-thread CTA needs a fraction of
// the registers of the 288-thread one, so 8-16 LPs are resident per SM instead of 2 (measured on
// B200, tools/gpu/variant_grid.py: 3x the pivots/s at 30x30, 1.8x at 50x50).  Two register budgets:
// tiny1 (64 registers, 16 CTAs/SM) for diet-size problems, small1 (120 registers, 8 CTAs/SM) where
// shared memory limits the residency anyway.
#define LPB_NS tiny1
#define LPB_SWEEP_WARPS 1
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 16
#define LPB_RULE 0
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

#define LPB_NS small1
#define LPB_SWEEP_WARPS 1
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 8
#define LPB_RULE 0
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

// course-size problems (ad_allocation: up to 50 x 100): four sweep warps + the bookkeeping warp
#define LPB_NS mid4
#define LPB_SWEEP_WARPS 4
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 4
#define LPB_RULE 0
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

// the corrected (non-reference) mode, LPB200_RULE_BLAND: same kernel, other pivoting decisions
#define LPB_NS bland8
#define LPB_SWEEP_WARPS 8
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 2
#define LPB_RULE 1
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE


// LPB200_RULE_DANTZIG: largest-coefficient pricing with a Bland fallback on degenerate streaks
#define LPB_NS dantzig8
#define LPB_SWEEP_WARPS 8
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 2
#define LPB_RULE 2
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

// LPB200_RULE_DEVEX: Devex reference-framework weights (approximate steepest edge), same Bland fallback
#define LPB_NS devex8
#define LPB_SWEEP_WARPS 8
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 2
#define LPB_RULE 3
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

#define LPB_NS devex_tiny1
#define LPB_SWEEP_WARPS 1
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 16
#define LPB_RULE 3
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

// the corrected rules on diet-size problems (n, m <= 30): the two-warp CTA
#define LPB_NS bland_tiny1
#define LPB_SWEEP_WARPS 1
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 16
#define LPB_RULE 1
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

#define LPB_NS dantzig_tiny1
#define LPB_SWEEP_WARPS 1
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 16
#define LPB_RULE 2
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE


using namespace lpb;

#define LPB_FOR_EACH_VARIANT(X)                                                                                  \
    X(LPB_V_REG15, reg15) X(LPB_V_TINY1, tiny1) X(LPB_V_SMALL1, small1) X(LPB_V_MID4, mid4) X(LPB_V_BLAND8, bland8) \
    X(LPB_V_DANTZIG8, dantzig8) X(LPB_V_DEVEX8, devex8) X(LPB_V_DEVEX_TINY1, devex_tiny1)                        \
    X(LPB_V_BLAND_TINY1, bland_tiny1) X(LPB_V_DANTZIG_TINY1, dantzig_tiny1)

int lpb_batch_variant_launch(int variant, const DeviceInfo &di, int B, int n, int m, const void *A, int a_dtype,
                             const double *b, const double *c, double tol, int max_pivots, int32_t *status,
                             double *x, double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                             cudaStream_t stream)
{
    switch (variant) {
#define X(id, ns)                                                                                                 \
    case id:                                                                                                      \
        return ns::launch(di, B, n, m, A, a_dtype, b, c, tol, max_pivots, status, x, objective, npivots, pivot_log, \
                          log_cap, stream);
        LPB_FOR_EACH_VARIANT(X)
#undef X
    default: break;
    }
    set_error("lpb200_simplex_batch: unknown kernel variant");
    return LPB200_E_INVALID;
}

long long lpb_batch_variant_smem_bytes(int variant, int n, int m)
{
    switch (variant) {
#define X(id, ns) \
    case id: return (long long)ns::make_batch_shape(n, m).bytes;
        LPB_FOR_EACH_VARIANT(X)
#undef X
    default: break;
    }
    return -1;
}

bool lpb_batch_variant_eligible(int variant, int n, int m)
{
    switch (variant) {
#define X(id, ns) \
    case id: return ns::eligible(n, m);
        LPB_FOR_EACH_VARIANT(X)
#undef X
    default: break;
    }
    return false;
}
