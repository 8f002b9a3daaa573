// gauss_jordan_batched.cu -- batched Gauss-Jordan elimination, one CTA per system (sm_100a).
//
// Replaces SolveEquation(Equation{a,b}) of week-2/energy_values/energy_values.cc:166-197:
//   SelectPivotElement :85-114  step k pivots on (k,k); on an EXACT zero swap with the first
//                               lower row holding a non-zero in column k; none -> carry on
//   ProcessPivotElement :132-159  row k /= a[k][k] (division per element); every other row i:
//                               a[i][j] += (-a[i][k]) * a[k][j], b likewise (mul and add
//                               rounded separately -- the reference is built without FMA)
//
// B200 design.  A step-by-step sweep would stream the whole 2 MB system through HBM 512
// times (1.08 GB/system).  Instead the elimination is PANEL-BLOCKED without
// changing a single rounding: for a panel of K = 16 consecutive pivots
//   1. the n x K panel is factorised in shared memory (K tiny steps), leaving in place the
//      scale s[i][k] = -a[i][k] every row used at step k and the pivot value of row k;
//   2. every trailing column j (and b) is then read ONCE, taken through the K updates in the
//      reference's order in registers -- a = a + s[i][k]*r[k][j], k ascending, product and sum
//      rounded separately, the pivot rows resolved first by a K x K triangular pass -- and
//      written ONCE.
// Each element sees exactly the operations, operands and order of the reference, so x is
// bit-identical, while HBM traffic drops K-fold (67 MB/system at n = 512) and the kernel is
// bound by the fp64 pipe instead of HBM.  Columns left of the current pivot are never
// touched (they are exact zeros the reference only rewrites with zeros).
// A zero pivot met inside a panel truncates the panel there; the swap happens in global
// memory at the start of the next panel.
//
// kPartial = true is the PARTIAL-PIVOTING mode (not reference behaviour:
// energy_values.cc:100-110 only ever swaps on an exact zero).  Step k pivots on the row i >= k
// with the largest |a[i][k]| (first maximum).  The search and the row exchange happen inside the
// shared-memory panel -- the whole panel row moves, scales of the earlier steps included, so the
// trailing update still sees, per row, the operations of the step-by-step algorithm in order --
// and the K exchanges are replayed on each trailing column by the thread that owns it at the
// start of pass 1.  No panel is ever truncated; the arithmetic per element is unchanged.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <string>

#include "lp_common.cuh"
#include "lp_host.h"

namespace lpb {

#ifndef LPB_GJ_K
#define LPB_GJ_K 16
#endif
constexpr int kGjK = LPB_GJ_K;      // pivots per panel
constexpr int kGjPS = kGjK + 2;     // panel row stride in shared memory (16-byte aligned rows)
constexpr int kGjMaxThreads = 512;  // CTA size for large systems; small ones get 64 or 128 threads (more systems per SM)
#ifndef LPB_GJ_ROWS
#define LPB_GJ_ROWS 2
#endif
#ifndef LPB_GJ_AHEAD
#define LPB_GJ_AHEAD 6
#endif
constexpr int kGjAhead = LPB_GJ_AHEAD;  // prefetch distance in chunks of kGjRows rows
constexpr int kGjRows = LPB_GJ_ROWS;
#ifndef LPB_GJ_BLOCK
#define LPB_GJ_BLOCK 32
#endif
#ifndef LPB_GJ_PAIR_BLOCK
#define LPB_GJ_PAIR_BLOCK 32
#endif
constexpr int kGjPairBlock = LPB_GJ_PAIR_BLOCK;  // rows per work item of the column-pair trailing update
constexpr int kGjBlockRows = LPB_GJ_BLOCK;  // rows per work item of the trailing update   // rows in flight per thread in the trailing update

// Step Q of the panel factorisation for the rows of one thread: whole panel rows go through registers
// with 16-byte shared-memory accesses (conflict-free at a row stride of 9 x 16 bytes).  Element by
// element, every load would wait for the store of the previous element (the compiler cannot tell
// P[i][t] from P[k][t+1]) and 8-byte accesses at this stride are 4-way bank conflicts.  Columns >= K of
// a short last panel are padding: updated, never used.
template <int Q, int kThreads>
__device__ __forceinline__ void gj_factor_rows(double *P, int n, int k, int tid)
{
    constexpr int T0 = Q & ~1;  // first 16-byte piece that holds a column >= Q
    double pk[kGjK];
    {
        const double2 *Pk2 = reinterpret_cast<const double2 *>(P + (size_t)k * kGjPS);
#pragma unroll
        for (int t = T0; t < kGjK; t += 2) {
            const double2 v = Pk2[t >> 1];
            pk[t] = v.x;
            pk[t + 1] = v.y;
        }
    }
    for (int i = tid; i < n; i += kThreads) {
        if (i == k) continue;
        double2 *Pi2 = reinterpret_cast<double2 *>(P + (size_t)i * kGjPS);
        double r[kGjK];
#pragma unroll
        for (int t = T0; t < kGjK; t += 2) {
            const double2 v = Pi2[t >> 1];
            r[t] = v.x;
            r[t + 1] = v.y;
        }
        const double s = -r[Q];  // scale = -1 * a[i][k]  (:151)
#pragma unroll
        for (int t = Q + 1; t < kGjK; ++t) r[t] = __dadd_rn(r[t], __dmul_rn(s, pk[t]));
        r[Q] = s;  // keep the scale for the trailing update
#pragma unroll
        for (int t = T0; t < kGjK; t += 2) Pi2[t >> 1] = make_double2(r[t], r[t + 1]);
    }
}

// a: [B][n][n] working copy (destroyed), x: [B][n] holds b on entry and the solution on exit.
#ifndef LPB_GJ_MINB
#define LPB_GJ_MINB 2
#endif
// kPairs: the trailing update takes TWO adjacent columns per thread (one 16-byte load/store per row, the
// row's K scales read from shared memory once for both columns): half the shared-memory and global
// instructions per fp64 operation.  Needs 2 x K pivot-row values in registers, hence 256-thread CTAs at
// 128 registers, two per SM.  Same operations per element in the same order: bit-identical.
template <bool kPartial, int kGjThreads, bool kPairs = false>
#ifndef LPB_GJ_MINB256
#define LPB_GJ_MINB256 4
#endif
__global__ void __launch_bounds__(kGjThreads, (kPairs ? 2 : (kGjThreads == kGjMaxThreads ? LPB_GJ_MINB : (kGjThreads == 256 ? LPB_GJ_MINB256 : 1024 / kGjThreads))))
gauss_jordan_batched_kernel(int n, double *__restrict__ a_all, double *__restrict__ x_all,
                            int32_t *__restrict__ status_all)
{
    extern __shared__ __align__(16) unsigned char smem[];
    double *P = reinterpret_cast<double *>(smem);                 // [n][kGjPS]
    int *flag = reinterpret_cast<int *>(P + (size_t)n * kGjPS);   // [2]
    // partial pivoting only: perm[q] = row exchanged with row k0+q at step q; per-warp argmax slots
    int *perm = flag + 4;                                                       // [kGjK]
    unsigned long long *wkey = reinterpret_cast<unsigned long long *>(perm + kGjK);  // [kGjThreads / 32]
    int *widx = reinterpret_cast<int *>(wkey + kGjThreads / 32);                // [kGjThreads / 32]
    const int tid = threadIdx.x;
    double *a = a_all + (size_t)blockIdx.x * n * n;
    double *bx = x_all + (size_t)blockIdx.x * n;
    // first step whose pivot column holds no non-zero at or below the diagonal (the reference then divides by
    // zero and returns inf/nan, energy_values.cc:100-110 finds no row to swap with); -1: every pivot was non-zero
    int first_singular = -1;

#ifdef LPB_GJ_TIMING
    long long tacc[5] = {0, 0, 0, 0, 0}, tlast = clock64();
#define LPB_GJ_T(ph)                       \
    do {                                   \
        const long long _t = clock64();    \
        tacc[ph] += _t - tlast;            \
        tlast = _t;                        \
    } while (0)
#else
#define LPB_GJ_T(ph) do { } while (0)
#endif
    for (int k0 = 0; k0 < n;) {
        const int K = min(kGjK, n - k0);
        // ---- SelectPivotElement for step k0, on the up-to-date matrix in global memory ----
        if (!kPartial && tid == 0) {
            int sw = -1;
            if (a[(size_t)k0 * n + k0] == 0.0) {
                sw = k0;
                while (sw < n && a[(size_t)sw * n + k0] == 0.0) ++sw;
                if (sw >= n) {  // the reference carries on and divides by zero
                    sw = -1;
                    if (first_singular < 0) first_singular = k0;
                }
            }
            flag[0] = sw;
        }
        __syncthreads();
        const int sw = kPartial ? -1 : flag[0];
        if (sw > k0) {  // std::swap(a[k0], a[sw]); std::swap(b[k0], b[sw])  (:106-109)
            for (int j = k0 + tid; j <= n; j += kGjThreads) {
                double *p0 = (j < n) ? &a[(size_t)k0 * n + j] : &bx[k0];
                double *p1 = (j < n) ? &a[(size_t)sw * n + j] : &bx[sw];
                const double t = *p0;
                *p0 = *p1;
                *p1 = t;
            }
            __syncthreads();
        }
        // ---- load the n x K panel ----
        // (batching these loads through registers, 8 x 16 bytes per thread, was measured: 73.3 ms against 70.2 ms)
        for (int idx = tid; idx < n * K; idx += kGjThreads) {
            const int i = idx / K, q = idx - i * K;
            P[(size_t)i * kGjPS + q] = a[(size_t)i * n + k0 + q];
        }
        __syncthreads();
        LPB_GJ_T(0);
        // ---- factorise the panel in shared memory ----
        int Kp = K;
        for (int q = 0; q < K; ++q) {
            const int k = k0 + q;
            if (kPartial) {
                // r = first row >= k with the largest |a[i][k]|; key = (|v| bits, ~i): max wins
                unsigned long long key = 0ull;
                int best = k;
                for (int i = k + tid; i < n; i += kGjThreads) {
                    const unsigned long long m =
                        (unsigned long long)__double_as_longlong(fabs(P[(size_t)i * kGjPS + q]));
                    if (m > key || (m == key && i < best)) {
                        key = m;
                        best = i;
                    }
                }
                if (k + tid >= n) best = 0x7fffffff;
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
                    const unsigned long long ok = __shfl_down_sync(0xffffffffu, key, off);
                    const int ob = __shfl_down_sync(0xffffffffu, best, off);
                    if (ok > key || (ok == key && ob < best)) {
                        key = ok;
                        best = ob;
                    }
                }
                if ((tid & 31) == 0) {
                    wkey[tid >> 5] = key;
                    widx[tid >> 5] = best;
                }
                __syncthreads();
                int r = widx[0];
                {
                    unsigned long long bk = wkey[0];
                    for (int w = 1; w < kGjThreads / 32; ++w) {
                        const unsigned long long ok = wkey[w];
                        const int ob = widx[w];
                        if (ok > bk || (ok == bk && ob < r)) {
                            bk = ok;
                            r = ob;
                        }
                    }
                }
                if (tid == 0) {
                    perm[q] = r;
                    if (first_singular < 0) {
                        unsigned long long bk = wkey[0];
                        for (int w = 1; w < kGjThreads / 32; ++w) bk = wkey[w] > bk ? wkey[w] : bk;
                        if (bk == 0ull) first_singular = k;  // the whole column is zero: no pivot exists
                    }
                }
                if (r != k && tid < K) {  // the whole panel row moves, earlier scales included
                    const double t = P[(size_t)k * kGjPS + tid];
                    P[(size_t)k * kGjPS + tid] = P[(size_t)r * kGjPS + tid];
                    P[(size_t)r * kGjPS + tid] = t;
                }
                __syncthreads();
            }
            const double value = P[(size_t)k * kGjPS + q];
            if (!kPartial && q > 0 && value == 0.0) {  // needs a row swap: finish this panel at q pivots
                Kp = q;
                break;
            }
            // scale the pivot row (division per element, :143-145)
            if (tid > q && tid < K) P[(size_t)k * kGjPS + tid] = __ddiv_rn(P[(size_t)k * kGjPS + tid], value);
            __syncthreads();
            if (kPairs) {
                switch (q) {  // compile-time step index: no predicated-off fp64 work, no selects
#define LPB_GJ_CASE(Q) case Q: gj_factor_rows<Q, kGjThreads>(P, n, k, tid); break;
                    LPB_GJ_CASE(0) LPB_GJ_CASE(1) LPB_GJ_CASE(2) LPB_GJ_CASE(3) LPB_GJ_CASE(4) LPB_GJ_CASE(5)
                    LPB_GJ_CASE(6) LPB_GJ_CASE(7) LPB_GJ_CASE(8) LPB_GJ_CASE(9) LPB_GJ_CASE(10) LPB_GJ_CASE(11)
                    LPB_GJ_CASE(12) LPB_GJ_CASE(13) LPB_GJ_CASE(14) LPB_GJ_CASE(15)
#undef LPB_GJ_CASE
                }
            } else {
                for (int i = tid; i < n; i += kGjThreads) {
                    if (i == k) continue;
                    double *Pi = P + (size_t)i * kGjPS;
                    const double s = -Pi[q];  // scale = -1 * a[i][k]  (:151)
                    for (int t = q + 1; t < K; ++t)
                        Pi[t] = __dadd_rn(Pi[t], __dmul_rn(s, P[(size_t)k * kGjPS + t]));
                    Pi[q] = s;  // keep the scale for the trailing update
                }
            }
            __syncthreads();
        }
        // ---- trailing update: columns k0+Kp .. n-1 and b (column index n) ----
        // Three passes with a block barrier between them, so that the non-pivot rows can be
        // split over SEVERAL threads per column once the trailing matrix is narrower than the
        // CTA (otherwise most threads idle through the second half of the elimination):
        //   1. one thread per column: the Kp pivot rows are taken through the updates of the
        //      earlier pivots of the panel and the division, R[q]; R is parked in the pivot rows;
        //   2. (column, row block) threads: every other row, Kp updates in the reference's
        //      order, one read and one write per element, four rows in flight;
        //   3. one thread per column: the pivot rows are eliminated by the later pivots of the
        //      panel (Gauss-JORDAN) and take their final values.
        LPB_GJ_T(1);
        const int jfirst = k0 + Kp;
        const int C = n + 1 - jfirst;
        for (int j = jfirst + tid; j <= n; j += kGjThreads) {
            double *col = (j < n) ? (a + j) : bx;
            const size_t cs = (j < n) ? (size_t)n : 1;  // element stride down the column
            if (kPartial) {  // replay the panel's row exchanges on this column, in order
                for (int q = 0; q < Kp; ++q) {
                    const int r = perm[q];
                    if (r != k0 + q) {
                        const double t = col[(size_t)(k0 + q) * cs];
                        col[(size_t)(k0 + q) * cs] = col[(size_t)r * cs];
                        col[(size_t)r * cs] = t;
                    }
                }
            }
            // all Kp loads first, all stores last: interleaved, each load would wait for the store before it
            // (the compiler cannot tell the rows apart) and the chain would pay 16 memory round trips
            double R[kGjK];
#pragma unroll
            for (int q = 0; q < kGjK; ++q) R[q] = (q < Kp) ? col[(size_t)(k0 + q) * cs] : 0.0;
#pragma unroll
            for (int q = 0; q < kGjK; ++q) {
                if (q < Kp) {
                    const double *Pk = P + (size_t)(k0 + q) * kGjPS;
                    double v = R[q];
#pragma unroll
                    for (int t = 0; t < q; ++t) v = __dadd_rn(v, __dmul_rn(Pk[t], R[t]));
                    R[q] = __ddiv_rn(v, Pk[q]);
                }
            }
#pragma unroll
            for (int q = 0; q < kGjK; ++q)
                if (q < Kp) col[(size_t)(k0 + q) * cs] = R[q];
        }
        __syncthreads();
        LPB_GJ_T(2);
        // column pairs: every row of `a` starts 16-byte aligned (n even) and the trailing block starts on an
        // even column; the b column (and everything, when the conditions fail) goes through the one-column loop
        const bool pairs = kPairs && Kp == kGjK && (n & 1) == 0 && (jfirst & 1) == 0;
        if (kPairs && pairs) {
            const int nblk = (n + kGjPairBlock - 1) / kGjPairBlock;
            const int CP = (n - jfirst) >> 1;
            const int items = CP * nblk;
            const size_t cs2 = (size_t)(n >> 1);  // row stride in double2
            for (int item = tid; item < items; item += kGjThreads) {
                const int blk = item / CP, jp = item - blk * CP;
                const int i0 = blk * kGjPairBlock, i1 = min(n, i0 + kGjPairBlock);
                double2 *col = reinterpret_cast<double2 *>(a + jfirst + 2 * jp);
                double2 R[kGjK];
#pragma unroll
                for (int q = 0; q < kGjK; ++q) R[q] = __ldcg(col + (size_t)(k0 + q) * cs2);
                // the block's rows minus the panel's pivot rows [k0, k0 + K): at most two contiguous runs
#pragma unroll 1
                for (int run = 0; run < 2; ++run) {
                    const int ra = run == 0 ? i0 : max(i0, k0 + kGjK);
                    const int rb = run == 0 ? min(i1, k0) : i1;
                    if (ra >= rb) continue;
                    double2 *g = col + (size_t)ra * cs2;
                    const double *Pr = P + (size_t)ra * kGjPS;
                    const int npair = (rb - ra) >> 1;
                    // software pipeline: the loads of the next two rows are in flight while these two are worked on
                    double2 cur0, cur1, nxt0, nxt1;
                    if (npair > 0) {
                        cur0 = g[0];
                        cur1 = g[cs2];
                    }
#pragma unroll 1
                    for (int c = 0; c < npair; ++c) {
                        double2 *gn = g + 2 * cs2;
                        if (c + 1 < npair) {
                            nxt0 = gn[0];
                            nxt1 = gn[cs2];
                        }
                        if (c + kGjAhead < npair) {
                            asm volatile("prefetch.global.L1 [%0];" ::"l"(g + 2 * kGjAhead * cs2));
                            asm volatile("prefetch.global.L1 [%0];" ::"l"(g + (2 * kGjAhead + 1) * cs2));
                        }
#pragma unroll
                        for (int t = 0; t < kGjK; t += 2) {
                            const double2 s0 = *reinterpret_cast<const double2 *>(Pr + t);
                            const double2 s1 = *reinterpret_cast<const double2 *>(Pr + kGjPS + t);
                            cur0.x = __dadd_rn(cur0.x, __dmul_rn(s0.x, R[t].x));
                            cur0.y = __dadd_rn(cur0.y, __dmul_rn(s0.x, R[t].y));
                            cur1.x = __dadd_rn(cur1.x, __dmul_rn(s1.x, R[t].x));
                            cur1.y = __dadd_rn(cur1.y, __dmul_rn(s1.x, R[t].y));
                            cur0.x = __dadd_rn(cur0.x, __dmul_rn(s0.y, R[t + 1].x));
                            cur0.y = __dadd_rn(cur0.y, __dmul_rn(s0.y, R[t + 1].y));
                            cur1.x = __dadd_rn(cur1.x, __dmul_rn(s1.y, R[t + 1].x));
                            cur1.y = __dadd_rn(cur1.y, __dmul_rn(s1.y, R[t + 1].y));
                        }
                        g[0] = cur0;
                        g[cs2] = cur1;
                        cur0 = nxt0;
                        cur1 = nxt1;
                        g = gn;
                        Pr += 2 * kGjPS;
                    }
                    if ((rb - ra) & 1) {  // odd run: the last row on its own
                        double2 v = g[0];
#pragma unroll
                        for (int t = 0; t < kGjK; ++t) {
                            v.x = __dadd_rn(v.x, __dmul_rn(Pr[t], R[t].x));
                            v.y = __dadd_rn(v.y, __dmul_rn(Pr[t], R[t].y));
                        }
                        g[0] = v;
                    }
                }
            }
        }
        if (kPairs && pairs) {
            // the b column, one row per thread and round (as one 32-row item per thread it would keep 16
            // threads busy and the rest of the CTA waiting at the barrier)
            double Rb[kGjK];
#pragma unroll
            for (int q = 0; q < kGjK; ++q) Rb[q] = __ldcg(bx + k0 + q);
            for (int i = tid; i < n; i += kGjThreads) {
                if (i >= k0 && i < k0 + kGjK) continue;  // pivot rows are handled by passes 1 and 3
                const double *Pi = P + (size_t)i * kGjPS;
                double v = bx[i];
#pragma unroll
                for (int t = 0; t < kGjK; ++t) v = __dadd_rn(v, __dmul_rn(Pi[t], Rb[t]));
                bx[i] = v;
            }
        } else {
            // work items = (column, block of kGjBlockRows rows), dealt round-robin: consecutive
            // threads take consecutive columns (coalesced) and every thread gets the same number
            // of items to within one, however narrow the trailing matrix has become
            const int nblk = (n + kGjBlockRows - 1) / kGjBlockRows;
            const int jskip = 0;
            const int C1 = C;
            const int items = C1 * nblk;
            for (int item = tid; item < items; item += kGjThreads) {
                const int blk = item / C1, jj = item - blk * C1;
                const int i0 = blk * kGjBlockRows, i1 = min(n, i0 + kGjBlockRows);
                const int j = jfirst + jskip + jj;
                double *col = (j < n) ? (a + j) : bx;
                const size_t cs = (j < n) ? (size_t)n : 1;
                double R[kGjK];
#pragma unroll
                for (int q = 0; q < kGjK; ++q) R[q] = (q < Kp) ? __ldcg(col + (size_t)(k0 + q) * cs) : 0.0;
                int i = i0;
                while (i < i1) {
                    if (Kp == kGjK && i + kGjRows <= i1 && !(i + kGjRows > k0 && i < k0 + Kp)) {
                        // kGjRows rows in flight, none of them a pivot row of this panel
                        double v[kGjRows];
#pragma unroll
                        for (int r = 0; r < kGjRows; ++r) v[r] = col[(size_t)(i + r) * cs];
                        // pull the rows kGjAhead chunks further down towards L1 while these are worked on
#pragma unroll
                        for (int r = 0; r < kGjRows; ++r) {
                            const int ip = i + kGjAhead * kGjRows + r;
                            if (ip < i1) asm volatile("prefetch.global.L1 [%0];" ::"l"(col + (size_t)ip * cs));
                        }
                        const double *Pi = P + (size_t)i * kGjPS;
#pragma unroll
                        for (int t = 0; t < kGjK; t += 2) {
#pragma unroll
                            for (int r = 0; r < kGjRows; ++r) {
                                const double2 s2 = *reinterpret_cast<const double2 *>(Pi + r * kGjPS + t);
                                v[r] = __dadd_rn(v[r], __dmul_rn(s2.x, R[t]));
                                v[r] = __dadd_rn(v[r], __dmul_rn(s2.y, R[t + 1]));
                            }
                        }
#pragma unroll
                        for (int r = 0; r < kGjRows; ++r) col[(size_t)(i + r) * cs] = v[r];
                        i += kGjRows;
                    } else {
                        if (!(i >= k0 && i < k0 + Kp)) {  // pivot rows are handled by passes 1 and 3
                            const double *Pi = P + (size_t)i * kGjPS;
                            double v = col[(size_t)i * cs];
#pragma unroll
                            for (int t = 0; t < kGjK; ++t)
                                if (t < Kp) v = __dadd_rn(v, __dmul_rn(Pi[t], R[t]));
                            col[(size_t)i * cs] = v;
                        }
                        ++i;
                    }
                }
            }
        }
        __syncthreads();
        LPB_GJ_T(3);
        for (int j = jfirst + tid; j <= n; j += kGjThreads) {
            double *col = (j < n) ? (a + j) : bx;
            const size_t cs = (j < n) ? (size_t)n : 1;
            double R[kGjK];
#pragma unroll
            for (int q = 0; q < kGjK; ++q) R[q] = (q < Kp) ? __ldcg(col + (size_t)(k0 + q) * cs) : 0.0;
#pragma unroll
            for (int q = 0; q < kGjK; ++q) {
                if (q < Kp) {
                    const double *Pk = P + (size_t)(k0 + q) * kGjPS;
                    double v = R[q];
#pragma unroll
                    for (int t = q + 1; t < kGjK; ++t)
                        if (t < Kp) v = __dadd_rn(v, __dmul_rn(Pk[t], R[t]));
                    col[(size_t)(k0 + q) * cs] = v;
                }
            }
        }
        __syncthreads();
        LPB_GJ_T(4);
        k0 += Kp;
    }
    if (tid == 0 && status_all != nullptr) status_all[blockIdx.x] = first_singular;
#ifdef LPB_GJ_TIMING
    if (tid == 0 && (blockIdx.x == 0 || blockIdx.x == 300))
        printf("gj timing cta %d: select+load %lld  factor %lld  pass1 %lld  pass2 %lld  pass3 %lld cycles\n", blockIdx.x,
               tacc[0], tacc[1], tacc[2], tacc[3], tacc[4]);
#endif
#undef LPB_GJ_T
}

}  // namespace lpb

using namespace lpb;

static inline int gj_smem_bytes(int n) { return n * kGjPS * 8 + 224 + 4 * kGjK; }  // panel + flag, perm, argmax slots

static int gj_launch(int B, int n, double *wa, double *x, int pivoting, int32_t *status, const DeviceInfo &di,
                     cudaStream_t stream)
{
    const int smem = gj_smem_bytes(n);
    if (smem > di.max_smem_optin) {
        set_error("lpb200_gauss_jordan_batch: n too large for the shared-memory panel (n <= " +
                  std::to_string((di.max_smem_optin - 224 - 4 * kGjK) / (kGjPS * 8)) + ")");
        return LPB200_E_TOO_LARGE;
    }
    // CTA size by system size: a 512-thread CTA idles on the small systems the course feeds this solver
    // (energy_values: n <= 20), and two of them fill an SM; 64- and 128-thread CTAs put 16 / 8 there.
#define LPB_GJ_LAUNCH(PP, T, ...)                                                                                    \
    do {                                                                                                             \
        LPB_CUDA(cudaFuncSetAttribute(gauss_jordan_batched_kernel<PP, T, ##__VA_ARGS__>,                             \
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, smem));                           \
        gauss_jordan_batched_kernel<PP, T, ##__VA_ARGS__><<<B, T, smem, stream>>>(n, wa, x, status);                \
    } while (0)
    const bool pp = (pivoting == LPB200_PIVOT_PARTIAL);
    int threads = n <= 64 ? 64 : (n <= 128 ? 128 : (n <= 256 ? 256 : kGjMaxThreads));  // measured: tools/gpu/gj_small_rates.py
    if (const char *g = getenv("LPB_DEBUG_GJ_THREADS")) threads = atoi(g);
    // n > 192, even: the column-pair trailing update on 256-thread CTAs at 128 registers (two per SM) --
    // 70.2 ms against 88.8 ms of the one-column 512-thread kernel on 4096 x 512 x 512 (tools/bench_gj.py),
    // 0.381 M against 0.299 M systems/s at n = 256, but 1.60 M against 1.93 M at n = 128 (tools/gpu/gj_small_rates.py)
    bool use_pairs = (n > 192) && (n & 1) == 0;
    if (const char *g = getenv("LPB_DEBUG_GJ_PAIRS")) use_pairs = atoi(g) != 0;
    if (use_pairs) {
        if (pp) LPB_GJ_LAUNCH(true, 256, true); else LPB_GJ_LAUNCH(false, 256, true);
    } else if (threads == 64) {
        if (pp) LPB_GJ_LAUNCH(true, 64); else LPB_GJ_LAUNCH(false, 64);
    } else if (threads == 128) {
        if (pp) LPB_GJ_LAUNCH(true, 128); else LPB_GJ_LAUNCH(false, 128);
    } else if (threads == 256) {
        if (pp) LPB_GJ_LAUNCH(true, 256); else LPB_GJ_LAUNCH(false, 256);
    } else {
        if (pp) LPB_GJ_LAUNCH(true, kGjMaxThreads); else LPB_GJ_LAUNCH(false, kGjMaxThreads);
    }
#undef LPB_GJ_LAUNCH
    LPB_CUDA(cudaGetLastError());
    return 0;
}

static int gj_check(int B, int n, int pivoting)
{
    if (B < 0 || n < 0) {
        set_error("lpb200_gauss_jordan_batch: need B >= 0 and n >= 0");
        return LPB200_E_INVALID;
    }
    if (pivoting != LPB200_PIVOT_REFERENCE && pivoting != LPB200_PIVOT_PARTIAL) {
        set_error("lpb200_gauss_jordan_batch: pivoting must be LPB200_PIVOT_REFERENCE or LPB200_PIVOT_PARTIAL");
        return LPB200_E_INVALID;
    }
    return 0;
}

extern "C" int64_t lpb200_gauss_jordan_workspace_bytes(int B, int n)
{
    if (B < 0 || n < 0) return 0;
    return (int64_t)B * n * n * 8;
}

extern "C" int lpb200_gauss_jordan_batch_ex_device(int B, int n, const double *a, const double *b, double *x,
                                                   int pivoting, void *workspace, int64_t workspace_bytes,
                                                   void *stream_v)
{
    return lpb200_gauss_jordan_batch_status_device(B, n, a, b, x, pivoting, nullptr, workspace, workspace_bytes,
                                                   stream_v);
}

extern "C" int lpb200_gauss_jordan_batch_status_device(int B, int n, const double *a, const double *b, double *x,
                                                       int pivoting, int32_t *status, void *workspace,
                                                       int64_t workspace_bytes, void *stream_v)
{
    int rc = gj_check(B, n, pivoting);
    if (rc) return rc;
    if (B == 0 || n == 0) return 0;  // size 0 prints nothing (week-2/energy_values/tests/01)
    cudaStream_t stream = (cudaStream_t)stream_v;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    if (workspace == nullptr || workspace_bytes < lpb200_gauss_jordan_workspace_bytes(B, n)) {
        set_error("lpb200_gauss_jordan_batch: workspace too small");
        return LPB200_E_INVALID;
    }
    double *wa = static_cast<double *>(workspace);
    LPB_CUDA(cudaMemcpyAsync(wa, a, (size_t)B * n * n * 8, cudaMemcpyDeviceToDevice, stream));
    if (x != b) LPB_CUDA(cudaMemcpyAsync(x, b, (size_t)B * n * 8, cudaMemcpyDeviceToDevice, stream));
    return gj_launch(B, n, wa, x, pivoting, status, di, stream);
}

extern "C" int lpb200_gauss_jordan_batch_device(int B, int n, const double *a, const double *b, double *x,
                                                void *workspace, int64_t workspace_bytes, void *stream_v)
{
    return lpb200_gauss_jordan_batch_ex_device(B, n, a, b, x, LPB200_PIVOT_REFERENCE, workspace, workspace_bytes,
                                               stream_v);
}

extern "C" int lpb200_gauss_jordan_batch_ex(int B, int n, const double *a, const double *b, double *x,
                                            int pivoting)
{
    return lpb200_gauss_jordan_batch_status(B, n, a, b, x, pivoting, nullptr);
}

extern "C" int lpb200_gauss_jordan_batch_status(int B, int n, const double *a, const double *b, double *x,
                                                int pivoting, int32_t *status)
{
    int rc = gj_check(B, n, pivoting);
    if (rc) return rc;
    if (B == 0 || n == 0) return 0;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    double *da = nullptr, *dx = nullptr;
    int32_t *dst = nullptr;
    const size_t na = (size_t)B * n * n * 8, nx = (size_t)B * n * 8;
    cudaError_t e;
    if (status != nullptr && (e = cudaMalloc((void **)&dst, (size_t)B * 4)) != cudaSuccess)
        return cuda_fail(e, "cudaMalloc(status)");
    // the host copy of `a` lands directly in the working buffer: no device-to-device copy
    if ((e = cudaMalloc((void **)&da, na)) != cudaSuccess) { cudaFree(dst); return cuda_fail(e, "cudaMalloc(a)"); }
    if ((e = cudaMalloc((void **)&dx, nx)) != cudaSuccess) { cudaFree(da); cudaFree(dst); return cuda_fail(e, "cudaMalloc(x)"); }
    rc = 0;
    if ((e = cudaMemcpy(da, a, na, cudaMemcpyHostToDevice)) != cudaSuccess) rc = cuda_fail(e, "cudaMemcpy(a)");
    if (!rc && (e = cudaMemcpy(dx, b, nx, cudaMemcpyHostToDevice)) != cudaSuccess) rc = cuda_fail(e, "cudaMemcpy(b)");
    if (!rc) rc = gj_launch(B, n, da, dx, pivoting, dst, di, 0);
    if (!rc && (e = cudaMemcpy(x, dx, nx, cudaMemcpyDeviceToHost)) != cudaSuccess) rc = cuda_fail(e, "cudaMemcpy(x)");
    if (!rc && dst && (e = cudaMemcpy(status, dst, (size_t)B * 4, cudaMemcpyDeviceToHost)) != cudaSuccess)
        rc = cuda_fail(e, "cudaMemcpy(status)");
    cudaFree(da);
    cudaFree(dx);
    cudaFree(dst);
    return rc;
}

extern "C" int lpb200_gauss_jordan_batch(int B, int n, const double *a, const double *b, double *x)
{
    return lpb200_gauss_jordan_batch_ex(B, n, a, b, x, LPB200_PIVOT_REFERENCE);
}

// ======================================================================================
// lpb200_gauss_jordan_batch_static_device: the same kernels with NO allocation anywhere and no caller workspace.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The elimination destroys its working copy of `a`; the other device entry points take that copy as a caller-owned
// `workspace`.  Here it lives in a file-scope __device__ array of LPB200_GJ_STATIC_SCRATCH_BYTES (32 MiB), which every
// GPU that loads the library holds until the process exits.  In HOST code the array's name is only a host-side
// stand-in for the symbol, not a device address (a kernel handed it would fault, or silently use the wrong memory where
// the GPU can reach host memory): the device address is looked up with cudaGetSymbolAddress once per device, with that
// device current, and cached.  Every call shares the array: it keeps the last call's contents (nothing here has to
// start at zero -- the whole region a call uses is overwritten by its copy of `a` before the kernel reads it), and two
// calls that overlap on the GPU (different streams, or different host threads on one device) RACE on it; calls on one
// stream are ordered and safe.
// ======================================================================================
constexpr size_t kGjStaticBytes = (size_t)LPB200_GJ_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) double g_gj_static_scratch[kGjStaticBytes / sizeof(double)];

// Device address of g_gj_static_scratch on the CURRENT device (must be `device`).  Once per device.
static int gj_static_scratch_address(int device, double **out)
{
    constexpr int kMaxDevices = 64;
    static std::atomic<double *> cache[kMaxDevices];
    if (device < 0 || device >= kMaxDevices) {
        set_error("lpb200_gauss_jordan_batch_static_device: unsupported device ordinal");
        return LPB200_E_INVALID;
    }
    double *p = cache[device].load(std::memory_order_acquire);
    if (p == nullptr) {
        void *sym = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&sym, g_gj_static_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_gj_static_scratch)");
        if (sym == nullptr || (reinterpret_cast<uintptr_t>(sym) & 255) != 0) {
            set_error("lpb200_gauss_jordan_batch_static_device: the static scratch has no usable device address");
            return LPB200_E_CUDA;
        }
        p = static_cast<double *>(sym);
        cache[device].store(p, std::memory_order_release);  // racing first calls look up the same address
    }
    *out = p;
    return 0;
}

extern "C" int lpb200_gauss_jordan_batch_static_device(int B, int n, const double *a, const double *b, double *x,
                                                       int pivoting, int32_t *status, void *stream_v)
{
    int rc = gj_check(B, n, pivoting);
    if (rc) return rc;
    if (B == 0 || n == 0) return 0;  // size 0 prints nothing (week-2/energy_values/tests/01)
    if (a == nullptr || b == nullptr || x == nullptr) {
        set_error("lpb200_gauss_jordan_batch_static_device: null argument");
        return LPB200_E_INVALID;
    }
    const unsigned long long need = (unsigned long long)B * (unsigned long long)n * (unsigned long long)n * 8ull;
    if (need > (unsigned long long)kGjStaticBytes) {
        set_error("lpb200_gauss_jordan_batch_static_device: B*n*n*8 = " + std::to_string(need) +
                  " bytes of working copy exceed the static scratch of " + std::to_string(kGjStaticBytes) +
                  " bytes (LPB200_GJ_STATIC_SCRATCH_BYTES); split the batch, or use lpb200_gauss_jordan_batch_status_device "
                  "with a workspace of your own");
        return LPB200_E_TOO_LARGE;
    }
    cudaStream_t stream = (cudaStream_t)stream_v;
    DeviceInfo di;
    rc = get_device_info(&di);  // the current device
    if (rc) return rc;
    double *wa = nullptr;
    rc = gj_static_scratch_address(di.device, &wa);
    if (rc) return rc;
    LPB_CUDA(cudaMemcpyAsync(wa, a, (size_t)need, cudaMemcpyDeviceToDevice, stream));
    if (x != b) LPB_CUDA(cudaMemcpyAsync(x, b, (size_t)B * n * 8, cudaMemcpyDeviceToDevice, stream));
    return gj_launch(B, n, wa, x, pivoting, status, di, stream);
}
