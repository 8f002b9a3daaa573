// verify_bounds.cu -- batched Simplex.verify_bounds (week-2/diet/simplex.py:663-671) for sm_100a.
//
// The reference checks an answer row by row: sum = 0; for x in [row[i] * ansx[i] ...]: sum += x;
// assert(sum <= b[r] + 1e-3) -- products and the running sum rounded separately, left to right, the
// literal 1e-3 whatever `tolerance` was passed.  stress.py:21-87 runs it after every "Bounded solution";
// here it is the acceptance test of the GPU-batched fuzzer: one CTA per LP, A read
// from HBM exactly once (the kernel is HBM-bound: 8*n*m bytes per LP), staged in shared memory with
// coalesced 16-byte loads, then one thread per constraint walks its row in the reference's order against
// x (broadcast reads).  Output per LP: the first row whose assertion would fire, or -1.
// LPs whose status is not LPB200_STATUS_BOUNDED have no x and are skipped (-1).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "lp_common.cuh"
#include "lp_host.h"

namespace lpb {

constexpr int kVbThreads = 256;
constexpr int kVbRows = 8;  // rows per warp and tile of the staging loop: 8 x 4 loads in flight per lane

// kStaged: the LP's A fits the CTA's shared memory (row stride padded to an odd number of doubles, so
// the threads of a warp -- one row each -- hit distinct banks); otherwise rows are read straight from
// global memory (each thread streams its own row; L1 turns the 8-byte steps into line reuse).
template <bool kStaged>
__global__ void __launch_bounds__(kVbThreads)
verify_bounds_kernel(int B, int n, int m, const double *__restrict__ A, const double *__restrict__ b,
                     const double *__restrict__ x, const int32_t *__restrict__ status, double tolerance,
                     int32_t *__restrict__ first_bad)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x;
    const int sm_stride = m | 1;
    double *xs = reinterpret_cast<double *>(smem);                     // [m]
    double *As = xs + ((m + 1) & ~1);                                   // [n][sm_stride]   (staged only)
    __shared__ int bad;
    for (int lp = blockIdx.x; lp < B; lp += gridDim.x) {
        if (status != nullptr && status[lp] != LPB200_STATUS_BOUNDED) {  // uniform across the CTA
            if (tid == 0) first_bad[lp] = -1;
            continue;
        }
        const double *Alp = A + (size_t)lp * n * m;
        if (tid == 0) bad = 0x7fffffff;
        for (int j = tid; j < m; j += kVbThreads) xs[j] = x[(size_t)lp * m + j];
        if (kStaged) {
            // one warp per row, lanes along the row: coalesced 8-byte loads, no index division.  All 32 loads of
            // an 8-row x 128-column tile are issued before the first shared-memory store (a store right behind
            // its load would leave each warp with ONE load in flight: 0.8 TB/s instead of the HBM rate)
            const int warp = tid >> 5, lane = tid & 31;
            constexpr int kWarps = kVbThreads / 32;
            for (int r0 = warp * kVbRows; r0 < n; r0 += kWarps * kVbRows) {
                for (int j0 = 0; j0 < m; j0 += 128) {
                    double v[kVbRows][4];
#pragma unroll
                    for (int u = 0; u < kVbRows; ++u)
#pragma unroll
                        for (int w = 0; w < 4; ++w) {
                            const int r = r0 + u, j = j0 + 32 * w + lane;
                            if (r < n && j < m) v[u][w] = __ldcs(Alp + (size_t)r * m + j);
                        }
#pragma unroll
                    for (int u = 0; u < kVbRows; ++u)
#pragma unroll
                        for (int w = 0; w < 4; ++w) {
                            const int r = r0 + u, j = j0 + 32 * w + lane;
                            if (r < n && j < m) As[(size_t)r * sm_stride + j] = v[u][w];
                        }
                }
            }
        }
        __syncthreads();
        {
            // pull the next LP of this CTA towards L2 while the rows are summed (a latency chain of m dependent
            // additions on n of the CTA's threads): its staging loop then starts from L2 instead of DRAM
            const int nxt = lp + (int)gridDim.x;
            if (nxt < B && (status == nullptr || status[nxt] == LPB200_STATUS_BOUNDED)) {
                const char *pa = reinterpret_cast<const char *>(A + (size_t)nxt * n * m);
                const size_t bytes = (size_t)n * m * 8;
                for (size_t o = (size_t)(kVbThreads - 1 - tid) * 128; o < bytes; o += (size_t)kVbThreads * 128)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pa + o));
            }
        }
        int mine = 0x7fffffff;
        for (int r = tid; r < n; r += kVbThreads) {
            const double *row = kStaged ? (As + (size_t)r * sm_stride) : (Alp + (size_t)r * m);
            double sum = 0.0;  // int 0 + float in the reference: the same value
            int i = 0;
            for (; i + 8 <= m; i += 8) {  // eight products ahead of the (inherently serial) chain of additions
                double pr[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) pr[k] = __dmul_rn(row[i + k], xs[i + k]);
#pragma unroll
                for (int k = 0; k < 8; ++k) sum = __dadd_rn(sum, pr[k]);
            }
            for (; i < m; ++i) sum = __dadd_rn(sum, __dmul_rn(row[i], xs[i]));
            if (!(sum <= __dadd_rn(b[(size_t)lp * n + r], tolerance))) mine = min(mine, r);
        }
        mine = __reduce_min_sync(0xffffffffu, mine);
        if ((tid & 31) == 0 && mine != 0x7fffffff) atomicMin(&bad, mine);
        __syncthreads();
        if (tid == 0) first_bad[lp] = (bad == 0x7fffffff) ? -1 : bad;
        __syncthreads();  // xs / As / bad are rewritten by the next LP
    }
}


// ---- pipelined variant (n <= 256): the default for course / Config-4 shapes ----
// One thread per constraint keeps its running sum in a register while A streams through shared memory in
// BLOCKS OF 32 COLUMNS (n x 32 doubles, double-buffered, filled with cp.async straight from global memory): the
// next block -- of this LP or of the CTA's next one -- is in flight while the current one is summed, so a CTA never
// alternates between a load phase and a latency-bound summing phase, and at 2 x 26 KB per CTA four CTAs share an SM
// (the staged kernel above holds a whole LP, 81 KB, two CTAs per SM: 0.48 of the HBM peak).  The chain of additions
// of a row is walked strictly left to right, product and sum rounded separately: the reference's bits.
// kVec16 (m even, 16-byte aligned inputs): 16-byte cp.async.cg (L2 only: with 218 KB of the SM's 256 KB carved out as
// shared memory the L1 left over is too small to stream through), rows 32 doubles apart with the 16-byte granules of
// row r XOR-swizzled by r & 15, so that both the copies and the one-row-per-thread 128-bit reads are conflict-free.
// Otherwise 8-byte cp.async.ca into rows 33 doubles apart.
constexpr int kVpThreads = 256;
constexpr int kVpCols = 32;  // columns per block

__device__ __forceinline__ void cp_async8(void *dst_smem, const void *src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory"); }

template <bool kVec16>
__global__ void __launch_bounds__(kVpThreads)
verify_bounds_pipelined_kernel(int B, int n, int m, const double *__restrict__ A, const double *__restrict__ b,
                               const double *__restrict__ x, const int32_t *__restrict__ status, double tolerance,
                               int32_t *__restrict__ first_bad)
{
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int kStride = kVec16 ? kVpCols : kVpCols + 1;  // doubles per buffered row
    const int tid = threadIdx.x;
    const int mpad = (m + 1) & ~1;
    double *xs = reinterpret_cast<double *>(smem);  // [2][mpad]: x of the LP being summed / of the next one
    double *buf = xs + 2 * mpad;                     // [2][n][kStride]
    __shared__ int bad;
    const int nblk = (m + kVpCols - 1) / kVpCols;
    const size_t bufsz = (size_t)n * kStride;

    // the CTA's stream of work units: (LP, column block), LPs blockIdx.x, blockIdx.x + gridDim.x, ...; an LP without
    // an answer (status != bounded) reports -1 and is skipped when the stream is built
    auto next_lp = [&](int lp) {
        lp += (int)gridDim.x;
        while (lp < B && status != nullptr && status[lp] != LPB200_STATUS_BOUNDED) {
            if (tid == 0) first_bad[lp] = -1;
            lp += (int)gridDim.x;
        }
        return lp;
    };
    // block k of LP lp -> buf[slot]; with block 0 also the LP's x -> xs[xslot]; one cp.async group
    auto issue = [&](int lp, int k, int slot, int xslot) {
        const int j0 = k * kVpCols;
        const int cb = min(kVpCols, m - j0);
        const double *src = A + (size_t)lp * n * m + j0;
        double *dst = buf + (size_t)slot * bufsz;
        if (kVec16) {
            const int ng = (cb + 1) >> 1;  // 16-byte granules per row (cb is even: m is)
            if (ng == 16) {
                for (int e = tid; e < n * 16; e += kVpThreads) {
                    const int r = e >> 4, g = e & 15;
                    cp_async16(dst + (size_t)r * kStride + ((g ^ (r & 15)) << 1), src + (size_t)r * m + 2 * g);
                }
            } else {
                for (int e = tid; e < n * ng; e += kVpThreads) {
                    const int r = e / ng, g = e - r * ng;
                    cp_async16(dst + (size_t)r * kStride + ((g ^ (r & 15)) << 1), src + (size_t)r * m + 2 * g);
                }
            }
            if (k == 0)
                for (int g = tid; g < (m >> 1); g += kVpThreads)
                    cp_async16(xs + (size_t)xslot * mpad + 2 * g, x + (size_t)lp * m + 2 * g);
        } else {
            for (int e = tid; e < n * cb; e += kVpThreads) {
                const int r = e / cb, j = e - r * cb;
                cp_async8(dst + (size_t)r * kStride + j, src + (size_t)r * m + j);
            }
            if (k == 0)
                for (int j = tid; j < m; j += kVpThreads) cp_async8(xs + (size_t)xslot * mpad + j, x + (size_t)lp * m + j);
        }
        cp_async_commit();
    };

    int lp = next_lp((int)blockIdx.x - (int)gridDim.x);
    if (lp >= B) return;
    int slot = 0, xpar = 0;
    issue(lp, 0, 0, 0);
    while (lp < B) {
        const int lp_next = next_lp(lp);
        double sum = 0.0;  // int 0 + float in the reference: the same value
        if (tid == 0) bad = 0x7fffffff;
        for (int k = 0; k < nblk; ++k) {
            // put the following unit in flight, then wait for this one
            bool more = true;
            if (k + 1 < nblk) issue(lp, k + 1, slot ^ 1, xpar);
            else if (lp_next < B) issue(lp_next, 0, slot ^ 1, xpar ^ 1);
            else more = false;
            if (more) cp_async_wait<1>(); else cp_async_wait<0>();
            __syncthreads();
            if (tid < n) {
                const int j0 = k * kVpCols;
                const int cb = min(kVpCols, m - j0);
                const double *row = buf + (size_t)slot * bufsz + (size_t)tid * kStride;
                const double *xv = xs + (size_t)xpar * mpad + j0;
                if (kVec16) {
                    const double2 *row2 = reinterpret_cast<const double2 *>(row);
                    const double2 *xv2 = reinterpret_cast<const double2 *>(xv);
                    const int sw = tid & 15;
                    int g = 0;
                    for (; g + 4 <= (cb >> 1); g += 4) {  // eight products ahead of the (inherently serial) additions
                        double pr[8];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const double2 a2 = row2[(g + q) ^ sw];
                            const double2 x2 = xv2[g + q];
                            pr[2 * q] = __dmul_rn(a2.x, x2.x);
                            pr[2 * q + 1] = __dmul_rn(a2.y, x2.y);
                        }
#pragma unroll
                        for (int q = 0; q < 8; ++q) sum = __dadd_rn(sum, pr[q]);
                    }
                    for (; g < (cb >> 1); ++g) {
                        const double2 a2 = row2[g ^ sw];
                        const double2 x2 = xv2[g];
                        sum = __dadd_rn(sum, __dmul_rn(a2.x, x2.x));
                        sum = __dadd_rn(sum, __dmul_rn(a2.y, x2.y));
                    }
                } else {
                    int i = 0;
                    for (; i + 8 <= cb; i += 8) {
                        double pr[8];
#pragma unroll
                        for (int q = 0; q < 8; ++q) pr[q] = __dmul_rn(row[i + q], xv[i + q]);
#pragma unroll
                        for (int q = 0; q < 8; ++q) sum = __dadd_rn(sum, pr[q]);
                    }
                    for (; i < cb; ++i) sum = __dadd_rn(sum, __dmul_rn(row[i], xv[i]));
                }
            }
            __syncthreads();  // buf[slot] is refilled two units from now
            slot ^= 1;
        }
        int mine = 0x7fffffff;
        if (tid < n && !(sum <= __dadd_rn(b[(size_t)lp * n + tid], tolerance))) mine = tid;
        mine = __reduce_min_sync(0xffffffffu, mine);
        if ((tid & 31) == 0 && mine != 0x7fffffff) atomicMin(&bad, mine);
        __syncthreads();
        if (tid == 0) first_bad[lp] = (bad == 0x7fffffff) ? -1 : bad;
        __syncthreads();
        lp = lp_next;
        xpar ^= 1;
    }
}

}  // namespace lpb

using namespace lpb;

extern "C" int lpb200_verify_bounds_batch_device(int B, int n, int m, const double *A, const double *b,
                                                 const double *x, const int32_t *status, double tolerance,
                                                 int32_t *first_bad_row, void *stream_v)
{
    if (B < 0 || n < 1 || m < 1 || !(tolerance >= 0.0)) {
        set_error("lpb200_verify_bounds_batch: need B >= 0, n >= 1, m >= 1, tolerance >= 0");
        return LPB200_E_INVALID;
    }
    if (B == 0) return 0;
    if (A == nullptr || b == nullptr || x == nullptr || first_bad_row == nullptr) {
        set_error("lpb200_verify_bounds_batch: null argument");
        return LPB200_E_INVALID;
    }
    DeviceInfo di;
    int rc = get_device_info(&di);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_v;
    const size_t xs_bytes = (size_t)((m + 1) & ~1) * 8;
    // n <= 256: the pipelined kernel (column blocks through a double buffer, several CTAs per SM)
    const bool vec16 = (m & 1) == 0 && ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(x)) & 15) == 0;
    const size_t piped_bytes = 2 * xs_bytes + 2 * (size_t)n * (vec16 ? kVpCols : kVpCols + 1) * 8;
    const char *force = getenv("LPB_DEBUG_VERIFY");  // "staged" / "direct": the round-1 kernels (A/B, tests)
    if (n <= kVpThreads && piped_bytes <= (size_t)di.max_smem_optin && force == nullptr) {
        int per_sm = (int)((size_t)di.max_smem_per_sm / (piped_bytes + 1024));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 8) per_sm = 8;
        long long grid = (long long)di.sm_count * per_sm;
        if (grid > B) grid = B;
        if (vec16) {
            LPB_CUDA(cudaFuncSetAttribute(verify_bounds_pipelined_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)piped_bytes));
            verify_bounds_pipelined_kernel<true><<<(int)grid, kVpThreads, piped_bytes, stream>>>(
                B, n, m, A, b, x, status, tolerance, first_bad_row);
        } else {
            LPB_CUDA(cudaFuncSetAttribute(verify_bounds_pipelined_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)piped_bytes));
            verify_bounds_pipelined_kernel<false><<<(int)grid, kVpThreads, piped_bytes, stream>>>(
                B, n, m, A, b, x, status, tolerance, first_bad_row);
        }
        LPB_CUDA(cudaGetLastError());
        return 0;
    }
    const size_t staged_bytes = xs_bytes + (size_t)n * (m | 1) * 8;
    const bool staged = staged_bytes <= (size_t)di.max_smem_optin && !(force && force[0] == 'd');
    if (!staged && xs_bytes > (size_t)di.max_smem_optin) {
        set_error("lpb200_verify_bounds_batch: m too large for the shared-memory copy of x");
        return LPB200_E_TOO_LARGE;
    }
    const size_t smem = staged ? staged_bytes : xs_bytes;
    int per_sm = (int)((size_t)di.max_smem_per_sm / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 16) per_sm = 16;
    long long grid = (long long)di.sm_count * per_sm;
    if (grid > B) grid = B;
    if (staged) {
        LPB_CUDA(cudaFuncSetAttribute(verify_bounds_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        verify_bounds_kernel<true><<<(int)grid, kVbThreads, smem, stream>>>(B, n, m, A, b, x, status, tolerance,
                                                                            first_bad_row);
    } else {
        LPB_CUDA(cudaFuncSetAttribute(verify_bounds_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        verify_bounds_kernel<false><<<(int)grid, kVbThreads, smem, stream>>>(B, n, m, A, b, x, status, tolerance,
                                                                             first_bad_row);
    }
    LPB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int lpb200_verify_bounds_batch(int B, int n, int m, const double *A, const double *b, const double *x,
                                          const int32_t *status, double tolerance, int32_t *first_bad_row)
{
    if (B < 0 || n < 1 || m < 1 || !(tolerance >= 0.0)) {
        set_error("lpb200_verify_bounds_batch: need B >= 0, n >= 1, m >= 1, tolerance >= 0");
        return LPB200_E_INVALID;
    }
    if (B == 0) return 0;
    DeviceInfo di;
    int rc = get_device_info(&di);
    if (rc) return rc;
    const size_t nA = (size_t)B * n * m * 8, nb = (size_t)B * n * 8, nx = (size_t)B * m * 8, ns = (size_t)B * 4;
    char *d = nullptr;
    const size_t oA = 0, ob = oA + (nA + 255) / 256 * 256, ox = ob + (nb + 255) / 256 * 256,
                 os = ox + (nx + 255) / 256 * 256, of = os + (ns + 255) / 256 * 256, total = of + ns;
    cudaError_t e = cudaMalloc((void **)&d, total);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    rc = 0;
#define LPB_TRY(call)                                                 \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess && rc == 0) rc = cuda_fail(_e, #call);  \
    } while (0)
    LPB_TRY(cudaMemcpy(d + oA, A, nA, cudaMemcpyHostToDevice));
    LPB_TRY(cudaMemcpy(d + ob, b, nb, cudaMemcpyHostToDevice));
    LPB_TRY(cudaMemcpy(d + ox, x, nx, cudaMemcpyHostToDevice));
    if (status) LPB_TRY(cudaMemcpy(d + os, status, ns, cudaMemcpyHostToDevice));
    if (rc == 0)
        rc = lpb200_verify_bounds_batch_device(B, n, m, reinterpret_cast<const double *>(d + oA),
                                               reinterpret_cast<const double *>(d + ob),
                                               reinterpret_cast<const double *>(d + ox),
                                               status ? reinterpret_cast<const int32_t *>(d + os) : nullptr, tolerance,
                                               reinterpret_cast<int32_t *>(d + of), nullptr);
    LPB_TRY(cudaMemcpy(first_bad_row, d + of, ns, cudaMemcpyDeviceToHost));
#undef LPB_TRY
    cudaFree(d);
    return rc;
}
