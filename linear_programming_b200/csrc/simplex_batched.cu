// simplex_batched.cu -- one-CTA-per-LP two-phase simplex for many small dense LPs (sm_100a).
//
// Replaces, per problem, the reference's Simplex.solve() (week-2/diet/simplex.py:586-630)
// with its SlackForm pivot loop (:442-491).  B200 design:
//   * the condensed tableau ((n+2) x (1+m+1) fp64; 83 KB for 100x100) lives in SHARED MEMORY
//     for the whole solve; HBM is touched once to read (A,b,c) and once to write (status,x);
//   * persistent CTAs (2 per SM at 100x100) pull problems off a device-side work queue, so
//     LPs with 10 pivots and LPs with 1600 pivots balance across the 148 SMs;
//   * 8 sweep warps + 1 bookkeeping warp, TWO block barriers per pivot:
//       step 1  one thread per row reads its pivot-column entry, stages it in fc[], zeroes it
//               in place (the column now belongs to the leaving variable), and runs the
//               min-ratio test: warp-shuffle lexicographic argmin over (ratio, variable id);
//       step 2  the rank-1 sweep: every lane derives its double2 of the scaled pivot row from
//               the (read-only) old pivot row, keeps it in registers and walks down a block of
//               rows, 4 at a time, with 128-bit shared-memory loads/stores (conflict-free).
//               The new pivot row is written into a SPARE physical row and the old one becomes
//               the next spare (rows are slots: rowvar[] maps slot -> variable), so no thread
//               ever reads a location another thread writes within a step.  The lanes that
//               update the objective row also pick the NEXT entering column (redux.sync min
//               over set-slot priority), and the 9th warp runs the CPython-set emulation that
//               defines that priority, both overlapped with the sweep;
//   * no FMA anywhere on the data path (__dmul_rn/__dadd_rn), IEEE division.
//
// Kernel <-> reference map (file:line in /root/reference/week-2/diet/simplex.py)
//   select_entering()   objective_has_positive_coefficients :387-388, pick_entering_idx :390-392
//   stage_and_ratio()   pick_leaving_idx :394-436, list_wrapper.min_index :157-161
//   sweep()             make_pivot_constraint :325-345,
//                       SlackForm.pivot :348-385 -> Constraint.substitute/sexp :185-214
//   bookkeeping()       set_replace :369-381 (CPython set order)
//   solve_one()         Simplex.solve :586-630, find_basic_feasible :514-584,
//                       make_auxiliary_form :632-660, SlackForm.solve epilogue :473-491
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "lp_common.cuh"
#include "lp_host.h"


// ---- kernel variants (same source, different parameters) ----
#define LPB_NS smem8
#define LPB_SWEEP_WARPS 8
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 2
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS

#define LPB_NS reg15
#define LPB_SWEEP_WARPS 15
#define LPB_REG_ROWS 12
#define LPB_MIN_BLOCKS 1
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS


// ======================================================================================
// Host side: C ABI
// ======================================================================================
using namespace lpb;

static int check_batch_args(int B, int n, int m, double tol, int max_pivots, int log_cap)
{
    if (B < 0 || n < 1 || m < 1 || !(tol >= 0.0) || max_pivots < 0 || log_cap < 0) {
        set_error("lpb200_simplex_batch: need B >= 0, n >= 1, m >= 1, tol >= 0, max_pivots >= 0");
        return LPB200_E_INVALID;
    }
    if (n + m + 1 >= 32000) {
        set_error("lpb200_simplex_batch: n + m too large for the one-CTA-per-LP kernel; use lpb200_simplex_single");
        return LPB200_E_TOO_LARGE;
    }
    return 0;
}

extern "C" int64_t lpb200_simplex_batch_smem_bytes(int n, int m)
{
    if (n < 1 || m < 1 || n + m + 1 >= 32000) return INT64_MAX;
    return (int64_t)smem8::make_batch_shape(n, m).bytes;
}

// Which kernel variant serves a shape: the register-resident one when it is eligible (one
// column-pair pass, at most LPB_REG_ROWS rows per thread) and fits, else the shared-memory one.
// LPB_BATCH_KERNEL=smem|reg overrides the choice (A/B measurements, tests).
// The shared-memory kernel is the product path.  The register-resident variant is bit-exact but
// slower on B200 (32 M vs 81 M pivots/s on Config 4 D2: with one CTA per SM nothing overlaps the
// serial stage/ratio chain); LPB_BATCH_KERNEL=reg selects it for experiments.
static int pick_variant(int n, int m, const DeviceInfo &di)
{
    const char *force = getenv("LPB_BATCH_KERNEL");
    if (force && force[0] == 'r')
        return reg15::eligible(n, m) && reg15::make_batch_shape(n, m).bytes <= di.max_smem_optin ? 1 : 0;
    return 0;
}

extern "C" int lpb200_simplex_batch_device(int B, int n, int m, const double *A, const double *b,
                                           const double *c, double tol, int max_pivots, int32_t *status,
                                           double *x, double *objective, int32_t *npivots,
                                           int32_t *pivot_log, int log_cap, void *stream_v)
{
    int rc = check_batch_args(B, n, m, tol, max_pivots, log_cap);
    if (rc) return rc;
    if (B == 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_v;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    if (pick_variant(n, m, di) == 1)
        return reg15::launch(di, B, n, m, A, b, c, tol, max_pivots, status, x, objective, npivots, pivot_log,
                             log_cap, stream);
    return smem8::launch(di, B, n, m, A, b, c, tol, max_pivots, status, x, objective, npivots, pivot_log, log_cap,
                         stream);
}

// Host-pointer entry point: the batch is cut into chunks that are copied in, solved and
// copied out on two alternating streams, so the PCIe copy of chunk k+1 overlaps the solve of
// chunk k and device memory stays bounded (two chunk-sized buffer sets).
extern "C" int lpb200_simplex_batch(int B, int n, int m, const double *A, const double *b, const double *c,
                                    double tol, int max_pivots, int32_t *status, double *x,
                                    double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap)
{
    int rc = check_batch_args(B, n, m, tol, max_pivots, log_cap);
    if (rc) return rc;
    if (B == 0) return 0;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    const size_t lpA = (size_t)n * m, lpb = (size_t)n, lpc = (size_t)m;
    const size_t lplog = pivot_log ? (size_t)log_cap * 2 : 0;
    const size_t lp_bytes = (lpA + lpb + 2 * lpc + 1) * 8 + 8 + lplog * 4;
    long long chunk = (long long)((size_t)(96u << 20) / lp_bytes);  // ~96 MB per chunk
    if (chunk < 1) chunk = 1;
    if (chunk > B) chunk = B;
    const int nset = (chunk < B) ? 2 : 1;
    struct Set {
        double *A, *b, *c, *x, *obj;
        int32_t *st, *np, *log;
        cudaStream_t stream;
    } set[2] = {};
    rc = 0;
#define LPB_TRY(call)                                                 \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess && rc == 0) rc = cuda_fail(_e, #call);  \
    } while (0)
    for (int s = 0; s < nset; ++s) {
        LPB_TRY(cudaStreamCreateWithFlags(&set[s].stream, cudaStreamNonBlocking));
        LPB_TRY(cudaMalloc((void **)&set[s].A, chunk * lpA * 8));
        LPB_TRY(cudaMalloc((void **)&set[s].b, chunk * lpb * 8));
        LPB_TRY(cudaMalloc((void **)&set[s].c, chunk * lpc * 8));
        LPB_TRY(cudaMalloc((void **)&set[s].x, chunk * lpc * 8));
        LPB_TRY(cudaMalloc((void **)&set[s].obj, chunk * 8));
        LPB_TRY(cudaMalloc((void **)&set[s].st, chunk * 4));
        LPB_TRY(cudaMalloc((void **)&set[s].np, chunk * 4));
        if (lplog) LPB_TRY(cudaMalloc((void **)&set[s].log, chunk * lplog * 4));
    }
    int k = 0;
    for (long long lo = 0; lo < B && rc == 0; lo += chunk, ++k) {
        const long long cnt = (B - lo < chunk) ? (B - lo) : chunk;
        Set &q = set[k % nset];
        LPB_TRY(cudaMemcpyAsync(q.A, A + lo * lpA, cnt * lpA * 8, cudaMemcpyHostToDevice, q.stream));
        LPB_TRY(cudaMemcpyAsync(q.b, b + lo * lpb, cnt * lpb * 8, cudaMemcpyHostToDevice, q.stream));
        LPB_TRY(cudaMemcpyAsync(q.c, c + lo * lpc, cnt * lpc * 8, cudaMemcpyHostToDevice, q.stream));
        if (lplog) LPB_TRY(cudaMemsetAsync(q.log, 0xff, cnt * lplog * 4, q.stream));
        if (rc == 0)
            rc = lpb200_simplex_batch_device((int)cnt, n, m, q.A, q.b, q.c, tol, max_pivots, q.st, q.x, q.obj,
                                             q.np, q.log, log_cap, q.stream);
        LPB_TRY(cudaMemcpyAsync(status + lo, q.st, cnt * 4, cudaMemcpyDeviceToHost, q.stream));
        LPB_TRY(cudaMemcpyAsync(x + lo * lpc, q.x, cnt * lpc * 8, cudaMemcpyDeviceToHost, q.stream));
        LPB_TRY(cudaMemcpyAsync(objective + lo, q.obj, cnt * 8, cudaMemcpyDeviceToHost, q.stream));
        LPB_TRY(cudaMemcpyAsync(npivots + lo, q.np, cnt * 4, cudaMemcpyDeviceToHost, q.stream));
        if (lplog)
            LPB_TRY(cudaMemcpyAsync(pivot_log + lo * lplog, q.log, cnt * lplog * 4, cudaMemcpyDeviceToHost, q.stream));
    }
    for (int s = 0; s < nset; ++s) {
        if (set[s].stream) {
            cudaError_t e = cudaStreamSynchronize(set[s].stream);
            if (e != cudaSuccess && rc == 0) rc = cuda_fail(e, "cudaStreamSynchronize");
        }
    }
#undef LPB_TRY
    for (int s = 0; s < nset; ++s) {
        cudaFree(set[s].A); cudaFree(set[s].b); cudaFree(set[s].c); cudaFree(set[s].x); cudaFree(set[s].obj);
        cudaFree(set[s].st); cudaFree(set[s].np); cudaFree(set[s].log);
        if (set[s].stream) cudaStreamDestroy(set[s].stream);
    }
    return rc;
}
