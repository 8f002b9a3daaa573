// simplex_single.cu -- grid-wide two-phase simplex for ONE large dense LP (sm_100a).
//
// Same semantics as simplex_batched.cu (the reference's Simplex.solve(),
// week-2/diet/simplex.py:586-630) but the condensed tableau ((n+1) x (m+1) fp64, 4.3 GB at
// 16384 x 32768) lives in HBM and every pivot is ONE streaming read-modify-write sweep over
// it by all 148 SMs.  One cooperative persistent kernel runs the whole solve; per pivot:
//
//   P1  entering column   every CTA scans a slice of the objective row, block-reduces the
//                         packed key (set slot << 32 | position), writes one partial     (:390-392)
//   P2  leaving row       every CTA reduces the partials (-> p), gathers a slice of pivot
//                         column p into the contiguous buffer fc[], does the min-ratio
//                         test on it (ratio, variable id) and writes one partial          (:394-436)
//   P3  pivot row         partials -> (l, r); nr[] = snap(pinv * row r), column p zeroed  (:325-345)
//   P4  rank-1 sweep      tiles of 64 rows x 512 columns: the tile's slice of the pivot
//                         row nr[] and of the pivot column fc[] are staged in shared memory
//                         by TMA bulk copies (cp.async.bulk + mbarrier, double-buffered
//                         across tiles); each thread keeps its double2 of nr in registers
//                         and streams the rows with 128-bit loads/stores, 8 rows in flight:
//                         T[i][k] = snap(T[i][k] + fc[i]*nr[k]), no FMA                  (:348-362, 185-214)
//   grid.sync() between phases (4 per pivot, ~1 % of a 1.3 ms sweep).
// The constant column b is kept contiguous (bcol[]) so the ratio test reads it coalesced.
// Set bookkeeping (CPython set order, :369-381) is done by one thread at the start of P4,
// overlapped with the sweep.  Tensor cores are not used: the update is a rank-1 AXPY at
// 0.125 flop/byte, HBM-bound by construction.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "lp_common.cuh"
#include "lp_host.h"

namespace cg = cooperative_groups;

namespace lpb {

constexpr int kSgThreads = 256;
constexpr int kTileCols = 2 * kSgThreads;  // 512 doubles: one double2 per thread
constexpr int kTileRows = 64;
constexpr int kRowUnroll = 8;

enum {
    ST_STATUS = 0, ST_NPIV, ST_ISIZE, ST_IFILL, ST_IUSED, ST_DSIZE, ST_DFILL, ST_DUSED, ST_NLIST,
    ST_COUNT = 16
};

// Device-side view of the workspace (all pointers into one caller-owned buffer).
struct SingleWS {
    int n, m, S;      // S: tableau row stride in doubles (multiple of 16)
    int size_i0, size_i1, size_d0, cap_i, cap_d, cap_s;
    double *T;        // [(n+1)][S]  non-basic columns only; row n = objective
    double *bcol;     // [n+1]       constant column (b, then v at index n)
    double *nr;       // [S]         scaled pivot row
    double *fc;       // [n+1]       pivot column factors
    double *cstage;   // [m]         scratch: original c / rebuilt objective
    double *scal;     // [4]         nr constant, misc scalars
    int *rowvar, *nonecol, *untouched, *rowslot;  // [n]
    int *colvar, *colslot;                        // [m+1]
    int *varpos;                                  // [n+m+1]
    unsigned *basis0;                             // bitmask over variables
    int *tab_i, *tab_d, *tab_s;
    int *list;                                    // [n]
    unsigned long long *part_key;                 // [2][grid]  set 0: entering, set 1: leaving
    int *part_var, *part_crash;                   // [2][grid]
    int *state;                                   // [ST_COUNT]
};

struct SingleArgs {
    const double *A, *b, *c;
    double tol;
    int max_pivots;
    int32_t *status;
    double *x, *objective;
    int32_t *npivots;
    int32_t *pivot_log;
    int log_cap;
};

// ---- small PTX wrappers: mbarrier + TMA bulk copy (global -> shared) ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// ---- block-level lexicographic argmin; result broadcast to every thread ----
struct BlockRed {
    unsigned long long key[kSgThreads / 32];
    int var[kSgThreads / 32];
    int flag[kSgThreads / 32];
};

__device__ __forceinline__ void block_argmin(BlockRed &br, unsigned long long &key, int &var, int &flag)
{
    const int tid = threadIdx.x;
    warp_argmin(key, var);
    flag = __any_sync(0xffffffffu, flag);
    __syncthreads();  // previous readers of br are done
    if ((tid & 31) == 0) {
        br.key[tid >> 5] = key;
        br.var[tid >> 5] = var;
        br.flag[tid >> 5] = flag;
    }
    __syncthreads();
    key = br.key[0];
    var = br.var[0];
    flag = br.flag[0];
#pragma unroll
    for (int w = 1; w < kSgThreads / 32; ++w) {
        const unsigned long long k = br.key[w];
        const int v = br.var[w];
        if (k < key || (k == key && v < var)) {
            key = k;
            var = v;
        }
        flag |= br.flag[w];
    }
}

// Reduce the per-CTA partials written before the last grid.sync().
__device__ __forceinline__ void reduce_partials(BlockRed &br, const SingleWS &w, int set,
                                                unsigned long long &key, int &var, int &flag)
{
    key = kKeyNone;
    var = 0x7fffffff;
    flag = 0;
    const int lo = set * (int)gridDim.x;
    for (int q = lo + threadIdx.x; q < lo + (int)gridDim.x; q += kSgThreads) {
        const unsigned long long k = w.part_key[q];
        const int v = w.part_var[q];
        if (k < key || (k == key && v < var)) {
            key = k;
            var = v;
        }
        flag |= w.part_crash[q];
    }
    block_argmin(br, key, var, flag);
}

__device__ __forceinline__ void write_partial(const SingleWS &w, int set, unsigned long long key, int var,
                                              int flag)
{
    if (threadIdx.x == 0) {
        const int q = set * (int)gridDim.x + (int)blockIdx.x;
        w.part_key[q] = key;
        w.part_var[q] = var;
        w.part_crash[q] = flag;
    }
}

// ---- P1: entering column (pick_entering_idx :390-392) ----
__device__ __forceinline__ void phase_entering(BlockRed &br, const SingleWS &w, int ncols, double tol)
{
    unsigned long long key = kKeyNone;
    int var = 0, flag = 0;
    const double *obj = w.T + (size_t)w.n * w.S;
    for (int k = blockIdx.x * kSgThreads + threadIdx.x; k < ncols; k += gridDim.x * kSgThreads) {
        if (obj[k] > tol) {
            const unsigned long long kk = ((unsigned long long)(unsigned)w.colslot[k] << 32) | (unsigned)k;
            key = kk < key ? kk : key;
        }
    }
    block_argmin(br, key, var, flag);
    write_partial(w, 0, key, var, flag);
}

// ---- P2: gather pivot column p into fc[], min-ratio test on it (pick_leaving_idx :394-436) ----
__device__ __forceinline__ void phase_ratio(BlockRed &br, const SingleWS &w, int p, int e, double tol)
{
    unsigned long long key = kKeyNone;
    int var = 0x7fffffff, crash = 0;
    const int n = w.n;
    const bool e_in_basis0 = (w.basis0[e >> 5] >> (e & 31)) & 1u;
    for (int i = blockIdx.x * kSgThreads + threadIdx.x; i <= n; i += gridDim.x * kSgThreads) {
        const double f = w.T[(size_t)i * w.S + p];
        w.fc[i] = (fabs(f) <= tol) ? 0.0 : f;  // substitute() skips a zero coefficient (:203-205)
        if (i < n) {
            const int nc = w.nonecol[i];
            crash |= (nc == kNoneInitial && e < n) || (nc == e) || (w.untouched[i] && e_in_basis0);
            const double coeff = -f;
            if (coeff > tol) {
                const double ratio = snap(__ddiv_rn(w.bcol[i], coeff), tol);
                if (ratio != 0.0) {
                    const unsigned long long k = order_key(ratio);
                    const int v = w.rowvar[i];
                    if (k < key || (k == key && v < var)) {
                        key = k;
                        var = v;
                    }
                }
            }
        }
    }
    block_argmin(br, key, var, crash);
    write_partial(w, 1, key, var, crash);
}

// ---- P3: scaled pivot row (make_pivot_constraint :325-345), zero column p, None bookkeeping ----
__device__ __forceinline__ void phase_pivot_row(const SingleWS &w, int ncols, int p, int r, int e, double tol)
{
    const int gtid = blockIdx.x * kSgThreads + threadIdx.x, gsize = gridDim.x * kSgThreads;
    const double pinv = pivot_inverse(w.fc[r], tol);
    const double *Tr = w.T + (size_t)r * w.S;
    for (int k = gtid; k < w.S; k += gsize) {
        double v = 0.0;
        if (k < ncols) v = (k == p) ? snap(-pinv, tol) : snap(__dmul_rn(pinv, Tr[k]), tol);
        w.nr[k] = v;
    }
    if (gtid == 0) w.scal[0] = snap(__dmul_rn(pinv, w.bcol[r]), tol);
    for (int i = gtid; i <= w.n; i += gsize) {
        if (i != r) w.T[(size_t)i * w.S + p] = 0.0;
        if (i < w.n && (w.fc[i] != 0.0 || i == r)) {
            w.nonecol[i] = e;
            w.untouched[i] = 0;
        }
    }
    fence_proxy_async();  // nr[] / fc[] are read back through the async proxy (TMA) in P4
}

// ---- set bookkeeping (set_replace :369-381), one thread, tables in global memory ----
__device__ void single_bookkeeping(const SingleWS &w, bool track_dep, int p, int r, int e, int l)
{
    const int eslot = w.colslot[p];
    w.rowvar[r] = e;
    w.colvar[p] = l;
    w.varpos[e] = ~r;
    w.varpos[l] = p;
    bool rebuild;
    if (track_dep) {
        PySet d = {w.state[ST_DSIZE], w.state[ST_DFILL], w.state[ST_DUSED]};
        pyset_remove_slot(w.tab_d, d, w.rowslot[r]);
        w.rowslot[r] = pyset_add(w.tab_d, d, e, &rebuild);
        if (rebuild) {
            pyset_rebuild(w.tab_d, w.tab_s, d);
            for (int s = 0; s < d.size; ++s) {
                const int v = w.tab_d[s];
                if (v >= 0) w.rowslot[~w.varpos[v]] = s;
            }
        }
        w.state[ST_DSIZE] = d.size;
        w.state[ST_DFILL] = d.fill;
        w.state[ST_DUSED] = d.used;
    }
    PySet s = {w.state[ST_ISIZE], w.state[ST_IFILL], w.state[ST_IUSED]};
    pyset_remove_slot(w.tab_i, s, eslot);
    w.colslot[p] = pyset_add(w.tab_i, s, l, &rebuild);
    if (rebuild) {
        pyset_rebuild(w.tab_i, w.tab_s, s);
        for (int t = 0; t < s.size; ++t) {
            const int v = w.tab_i[t];
            if (v >= 0) w.colslot[w.varpos[v]] = t;
        }
    }
    w.state[ST_ISIZE] = s.size;
    w.state[ST_IFILL] = s.fill;
    w.state[ST_IUSED] = s.used;
}

// ---- P4: the rank-1 sweep (SlackForm.pivot :348-362 -> sexp :185-194) ----
struct SweepSmem {
    double nr[2][kTileCols];
    double fc[2][kTileRows];
    uint64_t bar[2];
};

__device__ __forceinline__ void sweep_issue(SweepSmem &sm, const SingleWS &w, int stage, int tile,
                                            int nstrips)
{
    const int strip = tile % nstrips, rb = tile / nstrips;
    const int c0 = strip * kTileCols, r0 = rb * kTileRows;
    const int cols = min(kTileCols, w.S - c0);                  // even (S is a multiple of 16)
    const int rows = (min(kTileRows, w.n + 1 - r0) + 1) & ~1;   // fc[] is padded to an even length
    const uint32_t bnr = (uint32_t)cols * 8u, bfc = (uint32_t)rows * 8u;
    mbar_expect_tx(&sm.bar[stage], bnr + bfc);
    tma_load_1d(sm.nr[stage], w.nr + c0, bnr, &sm.bar[stage]);
    tma_load_1d(sm.fc[stage], w.fc + r0, bfc, &sm.bar[stage]);
}

__device__ __forceinline__ void phase_sweep(SweepSmem &sm, const SingleWS &w, int ncols, int r, double tol,
                                            uint32_t &parity)
{
    const int tid = threadIdx.x;
    const int n = w.n, S = w.S;
    const int nstrips = (ncols + kTileCols - 1) / kTileCols;
    const int nrb = (n + 1 + kTileRows - 1) / kTileRows;
    const int ntiles = nstrips * nrb;
    // constant column: b_i = snap(b_i + f_i * nr_const); the pivot row's constant is nr_const
    {
        const double nrc = w.scal[0];
        for (int i = blockIdx.x * kSgThreads + tid; i <= n; i += gridDim.x * kSgThreads) {
            if (i == r) {
                w.bcol[i] = nrc;
            } else {
                const double f = w.fc[i];
                if (f != 0.0) w.bcol[i] = axpy_snap(w.bcol[i], f, nrc, tol);
            }
        }
    }
    int tile = blockIdx.x;
    if (tile < ntiles && tid == 0) {
        fence_proxy_async();
        sweep_issue(sm, w, 0, tile, nstrips);
    }
    for (int it = 0; tile < ntiles; tile += gridDim.x, ++it) {
        const int stage = it & 1;
        if (tid == 0 && tile + (int)gridDim.x < ntiles) sweep_issue(sm, w, stage ^ 1, tile + gridDim.x, nstrips);
        mbar_wait(&sm.bar[stage], (parity >> stage) & 1u);
        parity ^= 1u << stage;
        const int strip = tile % nstrips, rb = tile / nstrips;
        const int c0 = strip * kTileCols, r0 = rb * kTileRows;
        const int rows = min(kTileRows, n + 1 - r0);
        const int col = c0 + 2 * tid;
        if (col < S) {
            const double2 r2 = reinterpret_cast<const double2 *>(sm.nr[stage])[tid];
            double2 *base = reinterpret_cast<double2 *>(w.T + (size_t)r0 * S + col);
            const size_t rs = (size_t)S / 2;  // row stride in double2
            for (int ii = 0; ii < rows; ii += kRowUnroll) {
                double2 v[kRowUnroll];
                double f[kRowUnroll];
#pragma unroll
                for (int j = 0; j < kRowUnroll; ++j) {
                    const int q = ii + j;
                    f[j] = (q < rows && r0 + q != r) ? sm.fc[stage][q] : 0.0;
                    if (f[j] != 0.0) v[j] = __ldcs(base + (size_t)q * rs);
                }
#pragma unroll
                for (int j = 0; j < kRowUnroll; ++j) {
                    const int q = ii + j;
                    if (f[j] != 0.0) {
                        v[j].x = axpy_snap(v[j].x, f[j], r2.x, tol);
                        v[j].y = axpy_snap(v[j].y, f[j], r2.y, tol);
                        __stcs(base + (size_t)q * rs, v[j]);
                    } else if (q < rows && r0 + q == r) {
                        __stcs(base + (size_t)q * rs, r2);  // store(entering_idx, row) (:352)
                    }
                }
            }
        }
        __syncthreads();  // everyone is done with sm.*[stage] before it is refilled
    }
}

struct LoopCtl {
    int ncols;
    int npiv;
    bool track_dep;
    uint32_t parity;
};

// P3 + P4 of one pivot: e sits at column position p, l owns row r.
__device__ __forceinline__ void single_pivot(cg::grid_group &grid, SweepSmem &sm, const SingleWS &w,
                                             const SingleArgs &a, LoopCtl &ctl, int p, int r, int e, int l)
{
    phase_pivot_row(w, ctl.ncols, p, r, e, a.tol);
    grid.sync();
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kSgThreads - 32) {
        if (a.pivot_log != nullptr && ctl.npiv < a.log_cap) {
            a.pivot_log[2 * ctl.npiv] = e;
            a.pivot_log[2 * ctl.npiv + 1] = l;
        }
        single_bookkeeping(w, ctl.track_dep, p, r, e, l);
    }
    phase_sweep(sm, w, ctl.ncols, r, a.tol, ctl.parity);
    grid.sync();
    ++ctl.npiv;
}

// SlackForm.solve loop (:442-466); every CTA takes the same branch (all reduce the same partials).
__device__ __forceinline__ int single_run_loop(cg::grid_group &grid, BlockRed &br, SweepSmem &sm,
                                               const SingleWS &w, const SingleArgs &a, LoopCtl &ctl)
{
    for (;;) {
        unsigned long long key;
        int var, flag;
        phase_entering(br, w, ctl.ncols, a.tol);
        grid.sync();
        reduce_partials(br, w, 0, key, var, flag);
        if (key == kKeyNone) return 0;
        const int p = (int)(key & 0xffffffffull);
        const int e = w.colvar[p];
        phase_ratio(br, w, p, e, a.tol);
        grid.sync();
        reduce_partials(br, w, 1, key, var, flag);
        if (flag) return LPB200_STATUS_REF_TYPE_ERROR;
        if (key == kKeyNone) return LPB200_STATUS_INFINITY;
        if (ctl.npiv >= a.max_pivots) return LPB200_STATUS_ITERATION_LIMIT;
        const int l = var;
        const int r = ~w.varpos[l];
        single_pivot(grid, sm, w, a, ctl, p, r, e, l);
    }
}

__global__ void __launch_bounds__(kSgThreads)
simplex_single_kernel(const SingleWS w, const SingleArgs a)
{
    cg::grid_group grid = cg::this_grid();
    __shared__ BlockRed br;
    __shared__ __align__(128) SweepSmem sm;
    const int tid = threadIdx.x;
    const int gtid = blockIdx.x * kSgThreads + tid, gsize = gridDim.x * kSgThreads;
    const int n = w.n, m = w.m, S = w.S, nv = n + m + 1;
    const double tol = a.tol;
    if (tid == 0) {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ---- min_constant, idx = min_index(b) (:591): zeros count as +inf, first minimum wins ----
    int idx0;
    {
        unsigned long long key = kKeyNone;
        int var = 0x7fffffff, flag = 0;
        for (int i = gtid; i < n; i += gsize) {
            const double bi = a.b[i];
            const unsigned long long k = order_key(bi != 0.0 ? bi : CUDART_INF);
            if (k < key || (k == key && i < var)) {
                key = k;
                var = i;
            }
        }
        block_argmin(br, key, var, flag);
        write_partial(w, 1, key, var, flag);
        grid.sync();
        reduce_partials(br, w, 1, key, var, flag);
        idx0 = var;
    }
    const double min_c = a.b[idx0];
    const bool phase1 = (min_c < 0.0);
    int status = 0;
    if (!phase1 && !(min_c > tol)) status = LPB200_STATUS_REF_ATTRIBUTE_ERROR;  // :626
    LoopCtl ctl;
    ctl.ncols = phase1 ? m + 1 : m;
    ctl.npiv = 0;
    ctl.track_dep = phase1;
    ctl.parity = 0u;

    if (status == 0) {
        // ---- SlackForm.__init__ (:230-302) (+ the x_aux column of :632-660 in phase 1) ----
        for (int i = blockIdx.x; i <= n; i += gridDim.x) {
            double *Ti = w.T + (size_t)i * S;
            if (i < n) {
                const double *Ai = a.A + (size_t)i * m;
                for (int k = tid; k < S; k += kSgThreads) {
                    double v = 0.0;
                    if (k < m) v = -Ai[k];
                    else if (k == m && phase1) v = 1.0;  // -(-1), :659 then :262
                    Ti[k] = v;
                }
            } else {
                for (int k = tid; k < S; k += kSgThreads) {
                    double v = 0.0;
                    if (k < m) v = phase1 ? 0.0 : a.c[k];
                    else if (k == m && phase1) v = -1.0;  // objective -x_aux (:644-645)
                    Ti[k] = v;
                }
            }
        }
        const int isize = phase1 ? w.size_i1 : w.size_i0;
        for (int i = gtid; i < n + 2; i += gsize) {
            if (i <= n) w.bcol[i] = (i < n) ? a.b[i] : 0.0;
            w.fc[i] = 0.0;
            if (i < n) {
                w.rowvar[i] = i;
                w.varpos[i] = ~i;
                w.nonecol[i] = kNoneInitial;
                w.untouched[i] = 0;
                w.rowslot[i] = i & (w.size_d0 - 1);
            }
        }
        for (int k = gtid; k <= m; k += gsize) {
            w.colvar[k] = n + k;
            w.varpos[n + k] = k;
            w.colslot[k] = (n + k) & (isize - 1);
        }
        for (int q = gtid; q < (nv + 31) / 32; q += gsize) w.basis0[q] = 0u;
        // set(range(n, n+ncols)) / set(range(0, n)) (:289-292): every key in its home slot
        for (int s = gtid; s < isize; s += gsize) {
            const int k = n + ((s - n) & (isize - 1));
            w.tab_i[s] = (k < n + ctl.ncols) ? k : kSetUnused;
        }
        if (phase1)
            for (int s = gtid; s < w.size_d0; s += gsize) w.tab_d[s] = (s < n) ? s : kSetUnused;
        if (gtid == 0) {
            w.state[ST_ISIZE] = isize;
            w.state[ST_IFILL] = ctl.ncols;
            w.state[ST_IUSED] = ctl.ncols;
            w.state[ST_DSIZE] = w.size_d0;
            w.state[ST_DFILL] = n;
            w.state[ST_DUSED] = n;
        }
        grid.sync();

        if (phase1) {
            const int aux = n + m;
            if (ctl.npiv >= a.max_pivots) {
                status = LPB200_STATUS_ITERATION_LIMIT;
            } else {
                // forced pivot: x_aux enters, the most negative row leaves (:524)
                phase_ratio(br, w, m, aux, tol);  // gathers column m into fc[]; its partial is unused
                grid.sync();
                single_pivot(grid, sm, w, a, ctl, m, idx0, aux, idx0);
                status = single_run_loop(grid, br, sm, w, a, ctl);  // :526
            }
            if (status == 0) {
                if (!(fabs(w.bcol[n]) <= tol)) status = LPB200_STATUS_NO_SOLUTION;       // :528-530
                else if (w.varpos[aux] < 0) status = LPB200_STATUS_REF_INDEX_ERROR;      // :541 -> :183
            }
            if (status == 0) {
                // ---- rebuild the original objective (:537-541) ----
                // (a) basic ORIGINAL variables with a non-zero objective coefficient, in
                //     iteration order of the `dependent` set: rows -> list[], factors -> cstage[]
                if (blockIdx.x == 0 && tid < 32) {
                    const int dsize = w.state[ST_DSIZE];
                    int cnt = 0;
                    for (int s0 = 0; s0 < dsize; s0 += 32) {
                        const int s = s0 + tid;
                        const int v = (s < dsize) ? w.tab_d[s] : -1;
                        double f = 0.0;
                        if (v >= n) f = a.c[v - n];
                        const bool keep = (v >= n) && !(fabs(f) <= tol);
                        const unsigned msk = __ballot_sync(0xffffffffu, keep);
                        if (keep) {
                            const int o = cnt + __popc(msk & ((1u << tid) - 1u));
                            w.list[o] = ~w.varpos[v];
                            w.cstage[o] = f;
                        }
                        cnt += __popc(msk);
                    }
                    if (tid == 0) w.state[ST_NLIST] = cnt;
                }
                grid.sync();
                // (b) (v, c) = (0, [None]*n + float(c)); substitute each listed row in order
                {
                    const int nlist = w.state[ST_NLIST];
                    const int kaux = w.varpos[aux];
                    double *obj = w.T + (size_t)n * S;
                    for (int k = gtid; k <= ctl.ncols; k += gsize) {
                        if (k == kaux) continue;  // zip() truncates the x_aux column (:190)
                        double o = 0.0;
                        if (k < ctl.ncols) {
                            const int v = w.colvar[k];
                            o = (v >= n) ? a.c[v - n] : 0.0;
                            for (int t = 0; t < nlist; ++t)
                                o = axpy_snap(o, w.cstage[t], w.T[(size_t)w.list[t] * S + k], tol);
                            obj[k] = o;
                        } else {  // k == ncols: the constant term
                            for (int t = 0; t < nlist; ++t) o = axpy_snap(o, w.cstage[t], w.bcol[w.list[t]], tol);
                            w.bcol[n] = o;
                        }
                    }
                    // clean-up (:547-551): every then-basic column becomes None in every row
                    for (int q = gtid; q < (nv + 31) / 32; q += gsize) {
                        unsigned bits = 0u;
                        for (int bb = 0; bb < 32; ++bb) {
                            const int v = q * 32 + bb;
                            if (v < nv && w.varpos[v] < 0) bits |= 1u << bb;
                        }
                        w.basis0[q] = bits;
                    }
                    for (int i = gtid; i < n; i += gsize) w.untouched[i] = 1;
                }
                grid.sync();
                // (c) drop x_aux (:563-575); the same two set objects live on into phase 2
                {
                    const int kaux = w.varpos[aux], last = ctl.ncols - 1;
                    for (int i = gtid; i <= n; i += gsize) {
                        double *Ti = w.T + (size_t)i * S;
                        if (kaux != last) Ti[kaux] = Ti[last];
                        Ti[last] = 0.0;
                    }
                    grid.sync();  // everyone has read varpos[aux] / colvar before they change
                    if (gtid == 0) {
                        PySet s = {w.state[ST_ISIZE], w.state[ST_IFILL], w.state[ST_IUSED]};
                        pyset_remove_slot(w.tab_i, s, w.colslot[kaux]);  // independent.remove(nvars) (:574)
                        w.state[ST_IUSED] = s.used;
                        if (kaux != last) {
                            const int v = w.colvar[last];
                            w.colvar[kaux] = v;
                            w.colslot[kaux] = w.colslot[last];
                            w.varpos[v] = kaux;
                        }
                    }
                    ctl.ncols = m;
                    grid.sync();
                }
                status = single_run_loop(grid, br, sm, w, a, ctl);  // :599
            }
        } else {
            status = single_run_loop(grid, br, sm, w, a, ctl);  // :619
        }
    }

    // ---- results (:473-491, 603, 622) ----
    for (int j = gtid; j < m; j += gsize) {
        double v = 0.0;
        if (status == 0) {
            const int pos = w.varpos[n + j];
            if (pos < 0) v = w.bcol[~pos];
        }
        a.x[j] = v;
    }
    if (gtid == 0) {
        a.status[0] = status;
        a.npivots[0] = ctl.npiv;
        a.objective[0] = (status == 0) ? w.bcol[n] : 0.0;
    }
}

}  // namespace lpb

// ======================================================================================
// Host side: C ABI
// ======================================================================================
using namespace lpb;

static inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }

// Carves the workspace; returns the bytes needed.  base == nullptr just measures.
static size_t carve_single(SingleWS *w, int n, int m, unsigned char *base, int grid)
{
    const int nv = n + m + 1;
    w->n = n;
    w->m = m;
    w->S = (m + 1 + 15) / 16 * 16;
    w->size_i0 = pyset_initial_size(m);
    w->size_i1 = pyset_initial_size(m + 1);
    w->size_d0 = pyset_initial_size(n);
    w->cap_i = pyset_capacity(m + 1);
    w->cap_d = pyset_capacity(n);
    w->cap_s = w->cap_i > w->cap_d ? w->cap_i : w->cap_d;
    size_t o = 0;
#define CARVE(field, type, count)                         \
    w->field = reinterpret_cast<type *>(base + o);        \
    o += up256(sizeof(type) * (size_t)(count));
    CARVE(T, double, (size_t)(n + 1) * w->S)
    CARVE(bcol, double, n + 2)
    CARVE(nr, double, w->S)
    CARVE(fc, double, n + 2 + kTileRows)
    CARVE(cstage, double, (n > m ? n : m) + 1)
    CARVE(scal, double, 4)
    CARVE(rowvar, int, n)
    CARVE(nonecol, int, n)
    CARVE(untouched, int, n)
    CARVE(rowslot, int, n)
    CARVE(colvar, int, m + 1)
    CARVE(colslot, int, m + 1)
    CARVE(varpos, int, nv)
    CARVE(basis0, unsigned, (nv + 31) / 32)
    CARVE(tab_i, int, w->cap_i)
    CARVE(tab_d, int, w->cap_d)
    CARVE(tab_s, int, w->cap_s)
    CARVE(list, int, n + 1)
    CARVE(part_key, unsigned long long, 2 * grid)
    CARVE(part_var, int, 2 * grid)
    CARVE(part_crash, int, 2 * grid)
    CARVE(state, int, ST_COUNT)
#undef CARVE
    return o;
}

static const int kMaxSingleGrid = 148 * 8;  // workspace is sized for any grid up to this

static int check_single_args(int n, int m, double tol, int max_pivots, int log_cap)
{
    if (n < 1 || m < 1 || !(tol >= 0.0) || max_pivots < 0 || log_cap < 0) {
        set_error("lpb200_simplex_single: need n >= 1, m >= 1, tol >= 0, max_pivots >= 0");
        return LPB200_E_INVALID;
    }
    if ((long long)n + m + 1 >= (1ll << 30)) {
        set_error("lpb200_simplex_single: n + m too large");
        return LPB200_E_TOO_LARGE;
    }
    return 0;
}

extern "C" int64_t lpb200_simplex_single_workspace_bytes(int n, int m)
{
    if (n < 1 || m < 1) return 0;
    SingleWS w;
    return (int64_t)carve_single(&w, n, m, nullptr, kMaxSingleGrid);
}

extern "C" int lpb200_simplex_single_device(int n, int m, const double *A, const double *b, const double *c,
                                            double tol, int max_pivots, int32_t *status, double *x,
                                            double *objective, int32_t *npivots, int32_t *pivot_log,
                                            int log_cap, void *workspace, int64_t workspace_bytes,
                                            void *stream_v)
{
    int rc = check_single_args(n, m, tol, max_pivots, log_cap);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_v;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    int coop = 0;
    LPB_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, di.device));
    if (!coop) {
        set_error("lpb200_simplex_single: device does not support cooperative launch");
        return LPB200_E_CUDA;
    }
    int per_sm = 0;
    LPB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, simplex_single_kernel, kSgThreads, 0));
    if (per_sm < 1) {
        set_error("lpb200_simplex_single: kernel does not fit on an SM");
        return LPB200_E_CUDA;
    }
    if (per_sm > 4) per_sm = 4;
    int grid = di.sm_count * per_sm;
    if (grid > kMaxSingleGrid) grid = kMaxSingleGrid;
    // small problems: no point in more CTAs than there is work
    const long long work = ((long long)(n + 1) * (m + 1) + kSgThreads * 8 - 1) / (kSgThreads * 8);
    if (grid > work) grid = (int)(work < 1 ? 1 : work);
    SingleWS w;
    const size_t need = carve_single(&w, n, m, static_cast<unsigned char *>(workspace), kMaxSingleGrid);
    if (workspace == nullptr || (size_t)workspace_bytes < need) {
        set_error("lpb200_simplex_single: workspace too small (need " + std::to_string(need) + " bytes)");
        return LPB200_E_INVALID;
    }
    if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0) {
        set_error("lpb200_simplex_single: workspace must be 256-byte aligned");
        return LPB200_E_INVALID;
    }
    SingleArgs a;
    a.A = A;
    a.b = b;
    a.c = c;
    a.tol = tol;
    a.max_pivots = max_pivots;
    a.status = status;
    a.x = x;
    a.objective = objective;
    a.npivots = npivots;
    a.pivot_log = pivot_log;
    a.log_cap = log_cap;
    void *params[] = {(void *)&w, (void *)&a};
    LPB_CUDA(cudaLaunchCooperativeKernel((const void *)simplex_single_kernel, dim3(grid), dim3(kSgThreads),
                                         params, 0, stream));
    return 0;
}

extern "C" int lpb200_simplex_single(int n, int m, const double *A, const double *b, const double *c,
                                     double tol, int max_pivots, int32_t *status, double *x,
                                     double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                                     double *stats)
{
    int rc = check_single_args(n, m, tol, max_pivots, log_cap);
    if (rc) return rc;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    const size_t nA = (size_t)n * m * 8, nb = (size_t)n * 8, nc = (size_t)m * 8;
    const size_t nlog = pivot_log ? (size_t)log_cap * 2 * 4 : 0;
    const int64_t wsb = lpb200_simplex_single_workspace_bytes(n, m);
    double *dA = nullptr, *db = nullptr, *dc = nullptr, *dx = nullptr, *dobj = nullptr;
    int32_t *dst = nullptr, *dnp = nullptr, *dlog = nullptr;
    void *ws = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    rc = 0;
#define LPB_TRY(call)                                                 \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess && rc == 0) rc = cuda_fail(_e, #call);  \
    } while (0)
    LPB_TRY(cudaMalloc((void **)&dA, nA));
    LPB_TRY(cudaMalloc((void **)&db, nb));
    LPB_TRY(cudaMalloc((void **)&dc, nc));
    LPB_TRY(cudaMalloc((void **)&dx, nc));
    LPB_TRY(cudaMalloc((void **)&dobj, 8));
    LPB_TRY(cudaMalloc((void **)&dst, 4));
    LPB_TRY(cudaMalloc((void **)&dnp, 4));
    LPB_TRY(cudaMalloc(&ws, (size_t)wsb));
    if (nlog) LPB_TRY(cudaMalloc((void **)&dlog, nlog));
    LPB_TRY(cudaEventCreate(&ev0));
    LPB_TRY(cudaEventCreate(&ev1));
    if (rc == 0) {
        LPB_TRY(cudaMemcpy(dA, A, nA, cudaMemcpyHostToDevice));
        LPB_TRY(cudaMemcpy(db, b, nb, cudaMemcpyHostToDevice));
        LPB_TRY(cudaMemcpy(dc, c, nc, cudaMemcpyHostToDevice));
        if (nlog) LPB_TRY(cudaMemset(dlog, 0xff, nlog));
    }
    if (rc == 0) {
        LPB_TRY(cudaEventRecord(ev0, 0));
        if (rc == 0)
            rc = lpb200_simplex_single_device(n, m, dA, db, dc, tol, max_pivots, dst, dx, dobj, dnp, dlog,
                                              log_cap, ws, wsb, nullptr);
        LPB_TRY(cudaEventRecord(ev1, 0));
    }
    if (rc == 0) {
        LPB_TRY(cudaMemcpy(status, dst, 4, cudaMemcpyDeviceToHost));
        LPB_TRY(cudaMemcpy(x, dx, nc, cudaMemcpyDeviceToHost));
        LPB_TRY(cudaMemcpy(objective, dobj, 8, cudaMemcpyDeviceToHost));
        LPB_TRY(cudaMemcpy(npivots, dnp, 4, cudaMemcpyDeviceToHost));
        if (nlog) LPB_TRY(cudaMemcpy(pivot_log, dlog, nlog, cudaMemcpyDeviceToHost));
        if (stats && rc == 0) {
            float ms = 0.f;
            LPB_TRY(cudaEventElapsedTime(&ms, ev0, ev1));
            stats[0] = ms;
            stats[1] = (double)*npivots;
        }
    }
#undef LPB_TRY
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    cudaFree(dA); cudaFree(db); cudaFree(dc); cudaFree(dx); cudaFree(dobj);
    cudaFree(dst); cudaFree(dnp); cudaFree(dlog); cudaFree(ws);
    return rc;
}
