// simplex_batched_impl.cuh -- device code of the one-CTA-per-LP simplex kernel.
//
// Included once per kernel variant by simplex_batched.cu with
//   LPB_NS            namespace of the variant
//   LPB_SWEEP_WARPS   sweep warps per CTA (a bookkeeping warp is added)
//   LPB_REG_ROWS      0: tableau swept in shared memory (2 CTAs per SM);
//                     R > 0: tableau held in REGISTERS during the pivot loop, R rows x one
//                     column pair per thread (1 CTA per SM), shared memory keeps the vectors
//   LPB_MIN_BLOCKS    __launch_bounds__ minimum resident CTAs per SM
//   LPB_WARP_ROWS     1: every sweep warp OWNS a block of whole tableau rows and the next pivot's ratio test is
//                     fused into the sweep -- ONE block barrier per pivot (see "warp-owns-rows" below);
//                     0 (default): column-pair lanes x row groups, two block barriers per pivot
// See simplex_batched.cu for the design notes and the reference map.
#ifndef LPB_WARP_ROWS
#define LPB_WARP_ROWS 0
#endif
#ifndef LPB_LANE_REDUCE
#define LPB_LANE_REDUCE 1  // per-warp partials / entering keys reduced lane-parallel with redux.sync (0: every thread reads all)
#endif

namespace lpb {
namespace LPB_NS {

// LPB_RULE: 0 = the reference's decisions (parity), 1 = Bland, 2 = Dantzig with Bland fallback,
// 3 = Devex reference-framework weights with the same fallback.
// Everything that differs from the reference is compiled in only when LPB_CORRECTED.
#undef LPB_CORRECTED
#define LPB_CORRECTED (LPB_RULE != 0)
constexpr unsigned long long kKeyZeroRatio = 0x8000000000000000ull;  // order_key(+0.0)
constexpr int kDegenerateStreak = 8;  // Dantzig: degenerate pivots in a row before Bland takes over


constexpr int kSweepWarps = LPB_SWEEP_WARPS;
constexpr int kSweepThreads = kSweepWarps * 32;  // 256
constexpr int kThreads = kSweepThreads + 32;     // + bookkeeping warp
constexpr int kRowChunk = 4;                     // rows per unrolled sweep step

typedef int16_t tab_t;

// Shared-memory carve-up, computed on the host and passed by value.
struct BatchShape {
    int n, m;
    int S;        // tableau row stride in doubles (even): 1 constant + m + 1 auxiliary, rounded up, + padding
    int npairs;   // column pairs the sweep covers = ceil((m + 2) / 2); the padding is never touched
    int cx;       // lanes across column pairs in the sweep (<= 64)
    int groups;   // row groups = 256 / cx
    int rpg;      // physical rows per row group (multiple of kRowChunk)
    int nfc;      // doubles in fc[] (groups * rpg + one chunk of zero padding)
    int rpt;      // rows per thread = ceil((n+2) / groups) (register-resident variants)
    int size_i0;  // initial table size of `independent` (phase 2: m keys / phase 1: m+1 keys)
    int size_i1;
    int size_d0;  // initial table size of `dependent` (n keys)
    int cap_i, cap_d, cap_s;
    // byte offsets from the (16-byte aligned) dynamic smem base
    int off_T, off_nr, off_fc, off_pinvs, off_bcol, off_redkey, off_rowvar, off_nonecol, off_untouched, off_rowslot,
        off_colvar, off_colslot, off_varpos, off_basis0, off_redvar, off_redu32, off_misc,
        off_tab_i, off_tab_d, off_tab_s;
    int bytes;
};

static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

static BatchShape make_batch_shape(int n, int m)
{
    BatchShape s;
    s.n = n;
    s.m = m;
    const int W = align_up(m + 2, 2);
    const int npairs = W / 2;
    s.npairs = npairs;
    s.S = W;
    const int passes = (npairs + 63) / 64;          // column pairs per lane (wide tableaus loop)
    s.cx = (npairs + passes - 1) / passes;           // lanes across column pairs: 51 at m = 100
    const int nphys = n + 2;  // n basic rows + objective row + one spare slot
    s.groups = kSweepThreads / s.cx;                 // row groups: 5 at m = 100 (255 of 256 threads busy)
    s.rpg = align_up((nphys + s.groups - 1) / s.groups, kRowChunk);
    s.rpt = (nphys + s.groups - 1) / s.groups;
    s.nfc = s.groups * s.rpg + kRowChunk;
    if (s.nfc < align_up(nphys, kRowChunk) + kRowChunk) s.nfc = align_up(nphys, kRowChunk) + kRowChunk;
    if (const char *g = getenv("LPB_DEBUG_SPAD")) s.S = W + 2 * (atoi(g) / 2);
    s.size_i0 = pyset_initial_size(m);
    s.size_i1 = pyset_initial_size(m + 1);
    s.size_d0 = pyset_initial_size(n);
    s.cap_i = pyset_capacity(m + 1);
    s.cap_d = pyset_capacity(n);
    s.cap_s = s.cap_i > s.cap_d ? s.cap_i : s.cap_d;
    const int nv = n + m + 1;
    int o = 0;
    s.off_T = o;          o += nphys * s.S * 8;
    s.off_nr = o;         o += s.S * 8;
    s.off_fc = o;         o += s.nfc * 8;
    s.off_pinvs = o;      o += s.nfc * 8;
    s.off_bcol = o;       o += s.nfc * 8;
    s.off_redkey = o;     o += 3 * kSweepWarps * 16;  // sets of per-warp ratio partials: [0],[1] fused (by pivot parity), [2] stand-alone
    s.off_rowvar = o;     o += align_up(nphys, 4) * 4;
    s.off_nonecol = o;    o += align_up(nphys, 4) * 4;
    s.off_untouched = o;  o += align_up(nphys, 4) * 4;
    s.off_rowslot = o;    o += align_up(nphys, 4) * 4;
    s.off_colvar = o;     o += align_up(m + 1, 4) * 4;
    s.off_colslot = o;    o += align_up(m + 1, 4) * 4;
    s.off_varpos = o;     o += align_up(nv, 4) * 4;
    s.off_basis0 = o;     o += align_up((nv + 31) / 32, 4) * 4;
    s.off_redvar = o;     o += align_up(kSweepWarps, 4) * 4;
    s.off_redu32 = o;     o += ((kSweepWarps + 2 + 3) / 4) * 16;
    s.off_misc = o;       o += 16 * 4 + 16;  // 16 ints + two mbarriers (LPB_WARP_ROWS)
    s.off_tab_i = o;      o += align_up(s.cap_i * (int)sizeof(tab_t), 16);
    s.off_tab_d = o;      o += align_up(s.cap_d * (int)sizeof(tab_t), 16);
    s.off_tab_s = o;      o += align_up(s.cap_s * (int)sizeof(tab_t), 16);
    s.bytes = align_up(o, 16);
    return s;
}

struct BatchArgs {
    int B;
    const void *A;  // [B][n][m] elements of a_dtype (LPB200_DTYPE_*), a_esz bytes each
    int a_dtype, a_esz;
    const double *b, *c;
    double tol;
    int max_pivots;
    int32_t *status;
    double *x, *objective;
    int32_t *npivots;
    int32_t *pivot_log;
    int log_cap;
    unsigned int *work_counter;
    unsigned long long *timing;  // debug (LPB_DEBUG_TIMING): per-phase cycle counters, else null
};

// misc[] slots
enum { MI_LP = 0, MI_ISIZE, MI_IFILL, MI_IUSED, MI_DSIZE, MI_DFILL, MI_DUSED, MI_NLIST, MI_RESELECT };

// Per-CTA view of the carved shared memory (pointers live in registers).
struct Ctx {
    double *T, *nr, *fc, *pinvs, *bcol;
    unsigned long long *redkey;
    int *rowvar, *nonecol, *untouched, *rowslot, *colvar, *colslot, *varpos;
    unsigned *basis0;
    int *redvar;
    unsigned *redu32;
    int *misc;
    tab_t *tab_i, *tab_d, *tab_s;
    int n, m, S, npairs, cx, rpg, rpt;
    int cxi, row0, row1;  // this thread's column-pair lane and physical row block [row0, row1)
    int spare;  // physical row slot currently unused (uniform across the CTA)
    // LPB_WARP_ROWS: rows [r_lo, r_hi) belong to this thread's warp (rpw per warp, nw warps own rows), warp `ow`
    // owns the objective row; nfused counts the fused pivots of this CTA (parity of the two mbarriers below)
    int rpw, r_lo, r_hi, ow, nw;
    unsigned nfused;
    unsigned long long *mbar;  // [0] "objective key published" (warp ow -> 9th warp), [1] "next entering column published"
    double tol;
    bool track_dep;  // the `dependent` set order only matters when phase 1 runs (:538)
    unsigned long long *timing;
};

#if LPB_CORRECTED
// Column to pivot a still-basic x_aux out on after phase 1: the non-basic variable with the lowest
// id whose entry in physical row r is non-zero; -1 when the row is all zero.
__device__ __forceinline__ int driveout_column(const Ctx &c, int ncols, int r)
{
    const int tid = threadIdx.x;
    if (tid < kSweepThreads) {
        unsigned key = 0xffffffffu;
        const double *row = c.T + (size_t)r * c.S + 1;
        for (int k = tid; k < ncols; k += kSweepThreads)
            if (fabs(row[k]) > c.tol) key = min(key, ((unsigned)c.colvar[k] << 16) | (unsigned)k);
        key = __reduce_min_sync(0xffffffffu, key);
        if ((tid & 31) == 0) c.redu32[tid >> 5] = key;
    }
    __syncthreads();
    unsigned best = c.redu32[0];
#pragma unroll
    for (int w = 1; w < kSweepWarps; ++w) best = min(best, c.redu32[w]);
    __syncthreads();
    return best == 0xffffffffu ? -1 : (int)(best & 0xffffu);
}
#endif

// ---- entering column: first positive objective coefficient in set-slot order (:390-392) ----
// Stand-alone pass (loop entry, and after a set-table rebuild).  Returns the column position,
// or -1 when no coefficient is > tol (loop ends, :447).
__device__ __forceinline__ int select_entering(const Ctx &c, int ncols)
{
    const int tid = threadIdx.x;
    if (tid < kSweepThreads) {
        unsigned key = 0xffffffffu;
        const double *obj = c.T + (size_t)c.n * c.S + 1;
        for (int k = tid; k < ncols; k += kSweepThreads)
            if (obj[k] > c.tol) key = min(key, ((unsigned)c.colslot[k] << 16) | (unsigned)k);
        key = __reduce_min_sync(0xffffffffu, key);
        if ((tid & 31) == 0) c.redu32[tid >> 5] = key;
    }
    __syncthreads();
    unsigned best = c.redu32[0];
#pragma unroll
    for (int w = 1; w < kSweepWarps; ++w) best = min(best, c.redu32[w]);
    __syncthreads();  // redu32 is rewritten by the next sweep
    return best == 0xffffffffu ? -1 : (int)(best & 0xffffu);
}

#if LPB_RULE == 2
// Dantzig pricing: the column with the LARGEST objective coefficient > tol enters, ties -> lowest
// variable id; with bland_now (a run of degenerate pivots) the lowest id with a positive
// coefficient, which cannot cycle.  Stand-alone pass over the objective row after every sweep.
__device__ __forceinline__ int select_entering_dantzig(const Ctx &c, int ncols, bool bland_now)
{
    const int tid = threadIdx.x;
    if (tid < kSweepThreads) {
        unsigned long long key = kKeyNone;
        int id = 0x7fffffff;
        const double *obj = c.T + (size_t)c.n * c.S + 1;
        for (int k = tid; k < ncols; k += kSweepThreads) {
            const double v = obj[k];
            if (v > c.tol) {
                // positive doubles order like their bit patterns: ~bits makes the largest the minimum
                const unsigned long long kk = bland_now ? 0ull : ~(unsigned long long)__double_as_longlong(v);
                const int ii = (c.colvar[k] << 16) | k;
                if (kk < key || (kk == key && ii < id)) {
                    key = kk;
                    id = ii;
                }
            }
        }
        warp_argmin(key, id);
        if ((tid & 31) == 0) {
            c.redkey[2 * (tid >> 5)] = key;  // redkey[] is 16 bytes per warp (RatioPartial)
            c.redvar[tid >> 5] = id;
        }
    }
    __syncthreads();
    unsigned long long best = c.redkey[0];
    int bid = c.redvar[0];
#pragma unroll
    for (int w = 1; w < kSweepWarps; ++w) {
        const unsigned long long k = c.redkey[2 * w];
        const int v = c.redvar[w];
        if (k < best || (k == best && v < bid)) {
            best = k;
            bid = v;
        }
    }
    __syncthreads();  // redkey/redvar are rewritten by the next ratio test
    return bid == 0x7fffffff ? -1 : (bid & 0xffff);
}
#endif

#if LPB_RULE == 3
// Devex pricing (Forrest & Goldfarb's approximation of steepest edge): every column position k carries a
// reference-framework weight w[k] (kept in nr[], which the shared-memory variants do not use inside the pivot
// loop); the column with the largest (c*c)/w among c > tol enters, ties -> lowest variable id, Bland's choice on
// a degenerate streak.  upd_row >= 0: the pivot just performed (entering position p_prev, weight wq read before
// the sweep, new pivot row in physical row upd_row) first updates the weights, element-wise:
//   w[k] = max(w[k], (nr[k]*nr[k])*wq) for k != p_prev,   w[p_prev] = max((nr[p_prev]*nr[p_prev])*wq, 1).
// One thread owns a column in this pass, so the update needs no reduction and gives the same bits whatever
// the thread mapping -- the grid-wide kernel and the CPU restatement compute exactly these expressions.
__device__ __forceinline__ int select_entering_devex(const Ctx &c, int ncols, bool bland_now, int upd_row, int p_prev,
                                                     double wq)
{
    const int tid = threadIdx.x;
    if (tid < kSweepThreads) {
        unsigned long long key = kKeyNone;
        int id = 0x7fffffff;
        const double *obj = c.T + (size_t)c.n * c.S + 1;
        const double *nrow = c.T + (size_t)(upd_row >= 0 ? upd_row : 0) * c.S + 1;
        for (int k = tid; k < ncols; k += kSweepThreads) {
            double w = c.nr[k];
            if (upd_row >= 0) {
                const double a = nrow[k];
                const double cand = __dmul_rn(__dmul_rn(a, a), wq);
                if (k == p_prev) w = cand > 1.0 ? cand : 1.0;
                else if (cand > w) w = cand;
                c.nr[k] = w;
            }
            const double v = obj[k];
            if (v > c.tol) {
                const double d = __ddiv_rn(__dmul_rn(v, v), w);
                const unsigned long long kk = bland_now ? 0ull : ~(unsigned long long)__double_as_longlong(d);
                const int ii = (c.colvar[k] << 16) | k;
                if (kk < key || (kk == key && ii < id)) {
                    key = kk;
                    id = ii;
                }
            }
        }
        warp_argmin(key, id);
        if ((tid & 31) == 0) {
            c.redkey[2 * (tid >> 5)] = key;
            c.redvar[tid >> 5] = id;
        }
    }
    __syncthreads();
    unsigned long long best = c.redkey[0];
    int bid = c.redvar[0];
#pragma unroll
    for (int w = 1; w < kSweepWarps; ++w) {
        const unsigned long long k = c.redkey[2 * w];
        const int v = c.redvar[w];
        if (k < best || (k == best && v < bid)) {
            best = k;
            bid = v;
        }
    }
    __syncthreads();  // redkey/redvar are rewritten by the next ratio test
    return bid == 0x7fffffff ? -1 : (bid & 0xffff);
}
#endif

// ---- step 1: stage the pivot column, min-ratio test with the reference's degenerate skip (:394-436) ----
// One thread per physical row: fc[i] = snapped T[i][1+p]; the entry is zeroed in place (the
// column now belongs to the leaving variable, whose coefficient starts from None, :187 opt).
// Every candidate row also leaves its scaled reciprocal pivot in pinvs[i] (:328-330), so the
// sweep starts without a division on its critical path.
// Returns (leaving VARIABLE id << 16 | its physical row), -1 if unbounded, -2 if the reference
// would raise TypeError.  The row travels with the id through the reduction because varpos[]
// is rewritten by the bookkeeping lane as soon as the barrier below is passed.
struct __align__(16) RatioPartial {
    unsigned long long key;
    int id;
    int pad;
};

__device__ __forceinline__ int stage_and_ratio(const Ctx &c, int p, int e, bool forced,
                                               unsigned long long *best_key = nullptr)
{
    const int tid = threadIdx.x;
    const int nphys = c.n + 2;
    int crash = 0;
    // only the warps that own rows take part (4 of 8 at n = 100)
    const int nsw = min(kSweepWarps, (nphys + 31) >> 5);
    RatioPartial *part = reinterpret_cast<RatioPartial *>(c.redkey) + 2 * kSweepWarps;  // the stand-alone set
    if (tid < nsw * 32) {
        unsigned long long key = kKeyNone;
        int var = 0x7fffffff;
#if !LPB_CORRECTED
        const bool e_in_basis0 = (c.basis0[e >> 5] >> (e & 31)) & 1u;
        const bool e_is_slack = e < c.n;
#endif
        for (int i = tid; i < nphys; i += nsw * 32) {
            const int v = c.rowvar[i];
            double f = 0.0;
            if (v >= 0 || i == c.n) {
                double *q = c.T + (size_t)i * c.S + 1 + p;
                f = *q;
                *q = 0.0;
                if (fabs(f) <= c.tol) f = 0.0;  // substitute() returns the row unchanged (:203-205)
            }
            c.fc[i] = f;
            if (v >= 0) {
#if !LPB_CORRECTED
                const int nc = c.nonecol[i];
                crash |= (nc == kNoneInitial && e_is_slack) || (nc == e) || (c.untouched[i] && e_in_basis0);
#endif
                if (f != 0.0) {  // row is rewritten by this pivot: its only None is column e
                    c.nonecol[i] = e;
                    c.untouched[i] = 0;
                }
                const double coeff = -f;
                const bool cand = coeff > c.tol;
                if (cand || (forced && f != 0.0)) c.pinvs[i] = pivot_inverse(f, c.tol);
                if (cand) {
                    const double ratio = snap(__ddiv_rn(c.T[(size_t)i * c.S], coeff), c.tol);
                    if (ratio != 0.0 || LPB_CORRECTED) {  // zero ratios count as +inf in min_index (:161)
                        const unsigned long long k = order_key(ratio);
                        const int vi = (v << 16) | i;  // ties -> lowest variable id (rows are unique per id)
                        if (k < key || (k == key && vi < var)) {
                            key = k;
                            var = vi;
                        }
                    }
                }
            }
        }
        // lexicographic (key, id) minimum across the warp with three redux.sync steps
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
        const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
        const unsigned mi = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? (unsigned)var : 0x7fffffffu);
        if ((tid & 31) == 0) {
            RatioPartial r;
            r.key = ((unsigned long long)mh << 32) | ml;
            r.id = (int)mi;
            r.pad = 0;
            part[tid >> 5] = r;
        }
    }
    if (forced) crash = 0;
    if (__syncthreads_or(crash)) return -2;
#if LPB_LANE_REDUCE
    // lane w of every warp takes partial w (ONE shared-memory wavefront per warp instead of nsw broadcast loads per
    // thread: 126 of the 2 480 wavefronts of a pivot went here), three redux.sync steps give the lexicographic minimum
    {
        const int lane = tid & 31;
        unsigned long long key = kKeyNone;
        unsigned id = 0x7fffffffu;
        if (lane < nsw) {
            const RatioPartial r = part[lane];
            key = r.key;
            id = (unsigned)r.id;
        }
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
        const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
        const unsigned mi = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? id : 0x7fffffffu);
        if (best_key) *best_key = ((unsigned long long)mh << 32) | ml;
        return (mh == 0xffffffffu && ml == 0xffffffffu) ? -1 : (int)mi;
    }
#else
    RatioPartial best = part[0];
    for (int w = 1; w < nsw; ++w) {
        const RatioPartial r = part[w];
        if (r.key < best.key || (r.key == best.key && r.id < best.id)) best = r;
    }
    if (best_key) *best_key = best.key;
    return best.key == kKeyNone ? -1 : best.id;
#endif
}

// ---- set bookkeeping by one lane of the 9th warp, overlapped with the sweep (:369-381) ----
// e sat at column position p; l owned physical row `src`; the new row of e is written to `dest`.
// Also publishes the entering-key candidate of column p (now variable l) for the next pivot.
__device__ __forceinline__ void bookkeeping(const Ctx &c, int p, int src, int dest, int e, int l, double pinv)
{
    const int eslot = c.colslot[p];
    const int lrowslot = c.rowslot[src];
    c.rowvar[src] = -1;  // becomes the spare slot
    c.rowvar[dest] = e;
    c.nonecol[dest] = e;
    c.untouched[dest] = 0;
    c.colvar[p] = l;
    c.varpos[e] = ~dest;
    c.varpos[l] = p;
    bool rebuild;
    int reselect = 0;
    if (c.track_dep) {  // dependent.remove(l); dependent.add(e)
        PySet d = {c.misc[MI_DSIZE], c.misc[MI_DFILL], c.misc[MI_DUSED]};
        pyset_remove_slot(c.tab_d, d, lrowslot);
        c.rowslot[dest] = pyset_add(c.tab_d, d, e, &rebuild);
        if (rebuild) {
            pyset_rebuild(c.tab_d, c.tab_s, d);
            for (int s = 0; s < d.size; ++s) {
                const int v = (int)c.tab_d[s];
                if (v >= 0) c.rowslot[~c.varpos[v]] = s;
            }
        }
        c.misc[MI_DSIZE] = d.size;
        c.misc[MI_DFILL] = d.fill;
        c.misc[MI_DUSED] = d.used;
    }
#if LPB_CORRECTED
    (void)eslot;
    c.colslot[p] = l;  // Bland: the entering priority of a column is the id of its variable
#else
    {  // independent.remove(e); independent.add(l)
        PySet s = {c.misc[MI_ISIZE], c.misc[MI_IFILL], c.misc[MI_IUSED]};
        pyset_remove_slot(c.tab_i, s, eslot);
        c.colslot[p] = pyset_add(c.tab_i, s, l, &rebuild);
        if (rebuild) {  // every slot moved: the keys computed by the sweep lanes are stale
            pyset_rebuild(c.tab_i, c.tab_s, s);
            for (int t = 0; t < s.size; ++t) {
                const int v = (int)c.tab_i[t];
                if (v >= 0) c.colslot[c.varpos[v]] = t;
            }
            reselect = 1;
        }
        c.misc[MI_ISIZE] = s.size;
        c.misc[MI_IFILL] = s.fill;
        c.misc[MI_IUSED] = s.used;
    }
#endif
    // new objective coefficient of column p: snap(None + c_e * (1/pivot)) (:314-318, :343-344)
    unsigned key = 0xffffffffu;
    const double fobj = c.fc[c.n];
    if (fobj != 0.0) {
        const double v = axpy_snap(0.0, fobj, snap(-pinv, c.tol), c.tol);
        if (v > c.tol) key = ((unsigned)c.colslot[p] << 16) | (unsigned)p;
    }
    c.redu32[kSweepWarps] = key;
    c.redu32[kSweepWarps + 1] = (unsigned)reselect;
}

// ---- shared-memory helpers for the sweep (32-bit shared addresses, explicit instruction forms) ----
__device__ __forceinline__ void lds_f64x2(uint32_t addr, double &x, double &y)
{
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(addr) : "memory");
}
__device__ __forceinline__ void sts_f64x2(uint32_t addr, double x, double y)
{
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(x), "d"(y) : "memory");
}
// snap-to-zero as a predicated 8-byte store of +0.0 over a value already stored (|v| <= tol is
// rare; this replaces a compare + two selects + register shuffling per element by compare + store)
__device__ __forceinline__ void sts_zero_if_snapped(uint32_t addr, double v, double tol)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .f64 t;\n"
        "abs.f64 t, %1;\n"
        "setp.le.f64 p, t, %2;\n"
        "@p st.shared.u64 [%0], 0;\n"
        "}\n" ::"r"(addr), "d"(v), "d"(tol) : "memory");
}

// ---- step 2: SlackForm.pivot(e, l) (:348-385) as one sweep ----
// e sits at column position p, l owns physical row src, c.spare receives the new row of e.
__device__ __forceinline__ void sweep(Ctx &c, int ncols, int p, int src, int e, int l, int32_t *log,
                                      int log_cap, int npiv, long long *t_own = nullptr)
{
    const int tid = threadIdx.x;
    const int n = c.n, S = c.S;
    const int nphys = n + 2;
    const int dest = c.spare;
    const double tol = c.tol;
    const double pinv = c.pinvs[src];  // snap(-1 * (1.0 / pivot)), left by the row's own thread in step 1
#ifdef LPB_TIMING
    const long long ts_in = clock64();
    long long ts_pro = ts_in, ts_slow = 0;
    int n_slow = 0;
#endif
    if (tid < kSweepThreads) {
        unsigned mykey = 0xffffffffu;
        if (c.row0 < c.row1) {
            const int npairs = c.npairs;
            const double2 *Tsrc = reinterpret_cast<const double2 *>(c.T + (size_t)src * S);
            const uint32_t tbase = (uint32_t)__cvta_generic_to_shared(c.T);
            const uint32_t rsb = (uint32_t)S * 8u;  // row stride in bytes
            const int src4 = src & ~(kRowChunk - 1), dest4 = dest & ~(kRowChunk - 1),
                      n4 = n & ~(kRowChunk - 1);
            for (int cp = c.cxi; cp < npairs; cp += c.cx) {
                // this lane's two entries of the scaled pivot row (make_pivot_constraint :325-345)
                double2 r2 = Tsrc[cp];
                r2.x = (2 * cp == 1 + p) ? snap(-pinv, tol) : snap(__dmul_rn(pinv, r2.x), tol);
                r2.y = (2 * cp + 1 == 1 + p) ? snap(-pinv, tol) : snap(__dmul_rn(pinv, r2.y), tol);
                uint32_t qa = tbase + (uint32_t)c.row0 * rsb + (uint32_t)cp * 16u;
                const double *fcp = c.fc + c.row0;
#ifdef LPB_TIMING
                ts_pro = clock64();
#endif
                for (int i = c.row0; i < c.row1; i += kRowChunk, qa += kRowChunk * rsb, fcp += kRowChunk) {
                    const double2 fa = *reinterpret_cast<const double2 *>(fcp);
                    const double2 fb = *reinterpret_cast<const double2 *>(fcp + 2);
                    const bool special = (i == src4) | (i == dest4) | (i == n4);
                    if (!special && fa.x != 0.0 && fa.y != 0.0 && fb.x != 0.0 && fb.y != 0.0) {
                        double a0x, a0y, a1x, a1y, a2x, a2y, a3x, a3y;
                        lds_f64x2(qa, a0x, a0y);
                        lds_f64x2(qa + rsb, a1x, a1y);
                        lds_f64x2(qa + 2 * rsb, a2x, a2y);
                        lds_f64x2(qa + 3 * rsb, a3x, a3y);
                        a0x = __dadd_rn(a0x, __dmul_rn(fa.x, r2.x));
                        a0y = __dadd_rn(a0y, __dmul_rn(fa.x, r2.y));
                        a1x = __dadd_rn(a1x, __dmul_rn(fa.y, r2.x));
                        a1y = __dadd_rn(a1y, __dmul_rn(fa.y, r2.y));
                        a2x = __dadd_rn(a2x, __dmul_rn(fb.x, r2.x));
                        a2y = __dadd_rn(a2y, __dmul_rn(fb.x, r2.y));
                        a3x = __dadd_rn(a3x, __dmul_rn(fb.y, r2.x));
                        a3y = __dadd_rn(a3y, __dmul_rn(fb.y, r2.y));
#ifndef LPB_EXP_NOSTORE
                        sts_f64x2(qa, a0x, a0y);
                        sts_f64x2(qa + rsb, a1x, a1y);
                        sts_f64x2(qa + 2 * rsb, a2x, a2y);
                        sts_f64x2(qa + 3 * rsb, a3x, a3y);
#endif
#ifndef LPB_EXP_NOSNAP
                        sts_zero_if_snapped(qa, a0x, tol);
                        sts_zero_if_snapped(qa + 8, a0y, tol);
                        sts_zero_if_snapped(qa + rsb, a1x, tol);
                        sts_zero_if_snapped(qa + rsb + 8, a1y, tol);
                        sts_zero_if_snapped(qa + 2 * rsb, a2x, tol);
                        sts_zero_if_snapped(qa + 2 * rsb + 8, a2y, tol);
                        sts_zero_if_snapped(qa + 3 * rsb, a3x, tol);
                        sts_zero_if_snapped(qa + 3 * rsb + 8, a3y, tol);
#endif
                    } else {
#ifdef LPB_TIMING
                        const long long ts0 = clock64();
                        ++n_slow;
#endif
                        const double f4[kRowChunk] = {fa.x, fa.y, fb.x, fb.y};
                        double2 *col = reinterpret_cast<double2 *>(c.T) + cp;
                        const int rs = S >> 1;  // row stride in double2
#pragma unroll
                        for (int j = 0; j < kRowChunk; ++j) {
                            const int ii = i + j;
                            if (ii >= nphys || ii == src) continue;
                            double2 *q = col + (size_t)ii * rs;
                            if (ii == dest) {  // store(entering_idx, entering_constraint) (:352)
                                *q = r2;
                                continue;
                            }
                            double2 a = *q;
                            const double f = f4[j];
                            if (f != 0.0) {
                                a.x = axpy_snap(a.x, f, r2.x, tol);
                                a.y = axpy_snap(a.y, f, r2.y, tol);
                                *q = a;
                            }
                            if (ii == n) {  // objective row: candidates for the NEXT entering column (:390-392)
                                const int k0 = 2 * cp - 1;  // column position of a.x (-1 is the constant)
                                if (k0 >= 0 && k0 < ncols && k0 != p && a.x > tol)
                                    mykey = min(mykey, ((unsigned)c.colslot[k0] << 16) | (unsigned)k0);
                                if (k0 + 1 < ncols && k0 + 1 != p && a.y > tol)
                                    mykey = min(mykey, ((unsigned)c.colslot[k0 + 1] << 16) | (unsigned)(k0 + 1));
                            }
                        }
#ifdef LPB_TIMING
                        ts_slow += clock64() - ts0;
#endif
                    }
                }
            }
        }
        mykey = __reduce_min_sync(0xffffffffu, mykey);
        if ((tid & 31) == 0) c.redu32[tid >> 5] = mykey;
    } else if (tid == kSweepThreads) {
        if (log != nullptr && npiv < log_cap) {
            log[2 * npiv] = e;
            log[2 * npiv + 1] = l;
        }
        bookkeeping(c, p, src, dest, e, l, pinv);
    }
    c.spare = src;
    if (t_own) *t_own = clock64();
#ifdef LPB_TIMING
    if (c.timing != nullptr && (tid & 31) == 0 && tid < kSweepThreads) {
        unsigned long long *t = c.timing + 32 * 4 + (tid >> 5) * 4;  // [prologue, general-path cycles, general-path chunks, -]
        atomicAdd(t + 0, (unsigned long long)(ts_pro - ts_in));
        atomicAdd(t + 1, (unsigned long long)ts_slow);
        atomicAdd(t + 2, (unsigned long long)n_slow);
    }
#endif
    __syncthreads();
}

// Next entering position after a sweep: the partial keys it left in redu32[0..8] (three
// 128-bit loads), or a fresh pass when the set table was rebuilt and every slot moved
// (redu32[9] != 0; redu32[10..11] stay 0xffffffff).
constexpr int kKeySlots = kSweepWarps + 1;               // sweep-warp partial keys + the bookkeeping lane's key
constexpr int kKeyVecs = (kKeySlots + 1 + 3) / 4;        // + the reselect flag, read as uint4
__device__ __forceinline__ int next_entering(const Ctx &c, int ncols)
{
#if LPB_LANE_REDUCE
    static_assert(kKeySlots + 1 <= 32, "one lane per key slot");
    const int lane = threadIdx.x & 31;
    const unsigned v = (lane <= kKeySlots) ? c.redu32[lane] : 0xffffffffu;  // one wavefront per warp
    if (__shfl_sync(0xffffffffu, v, kKeySlots) != 0u) {
        __syncthreads();  // everyone has read the flag before the bookkeeping lane can rewrite it
        return select_entering(c, ncols);
    }
    const unsigned best = __reduce_min_sync(0xffffffffu, lane < kKeySlots ? v : 0xffffffffu);
    return best == 0xffffffffu ? -1 : (int)(best & 0xffffu);
#else
    unsigned u[kKeyVecs * 4];
#pragma unroll
    for (int q = 0; q < kKeyVecs; ++q) {
        const uint4 v = reinterpret_cast<const uint4 *>(c.redu32)[q];
        u[4 * q] = v.x;
        u[4 * q + 1] = v.y;
        u[4 * q + 2] = v.z;
        u[4 * q + 3] = v.w;
    }
    if (u[kKeySlots] != 0u) {
        __syncthreads();  // everyone has read the flag before the bookkeeping lane can rewrite it
        return select_entering(c, ncols);
    }
    unsigned best = u[0];
#pragma unroll
    for (int q = 1; q < kKeySlots; ++q) best = min(best, u[q]);
    return best == 0xffffffffu ? -1 : (int)(best & 0xffffu);
#endif
}

#if LPB_WARP_ROWS
// =========================================================================================
// Warp-owns-rows pivot loop: ONE block barrier per pivot.
//
// Every sweep warp owns a contiguous block of whole physical rows (13 of the 102 at 100x100); its 32 lanes
// cover the column pairs in passes of 32 (two passes at m = 100: 32 + 19 lanes).  Consequences:
//   * every per-row decision (pivot row / spare slot / zero factor) is warp-uniform: no divergence;
//   * a row is only ever written by its own warp, so the NEXT pivot's ratio test of a row -- read its entry
//     in the next entering column, zero it in place, divide the row's constant by it (pick_leaving_idx,
//     simplex.py:394-436) -- can be done by that warp right behind its own part of the sweep, ordered by a
//     __syncwarp() alone.  The stage_and_ratio() pass between two block barriers disappears;
//   * the warp that owns the objective row updates it FIRST and publishes its entering key; the 9th warp
//     (set bookkeeping, simplex.py:369-381) merges in the key of the column that changed hands and publishes
//     the next entering column.  Both hand-overs are mbarriers in shared memory: the producer arrives once,
//     the consumers only wait -- they never wait for each other.
// Per pivot and warp:  [block barrier] reduce 8 ratio partials -> sweep own rows -> wait(next column) ->
// ratio test of own rows -> partial -> [block barrier].
// Same operations on the same operands as the two-barrier path: bit-identical results (the parity tests run
// both).
// =========================================================================================
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}

// Ratio test of the rows this warp owns for the NEXT pivot (entering column position p2), fused behind the
// warp's part of the sweep.  Same per-row work as stage_and_ratio(); leaves the warp's partial in part[w].
__device__ __forceinline__ void stage_rows(const Ctx &c, int p2, RatioPartial *part)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int e2 = c.colvar[p2];
    unsigned long long key = kKeyNone;
    int var = 0x7fffffff;
    int crash = 0;
#if !LPB_CORRECTED
    const bool e_in_basis0 = (c.basis0[e2 >> 5] >> (e2 & 31)) & 1u;
    const bool e_is_slack = e2 < c.n;
#endif
    for (int i = c.r_lo + lane; i < c.r_hi; i += 32) {
        const int v = c.rowvar[i];
        double f = 0.0;
        if (v >= 0 || i == c.n) {
            double *q = c.T + (size_t)i * c.S + 1 + p2;
            f = *q;
            *q = 0.0;
            if (fabs(f) <= c.tol) f = 0.0;  // substitute() returns the row unchanged (:203-205)
        }
        c.fc[i] = f;
        if (v >= 0) {
#if !LPB_CORRECTED
            const int nc = c.nonecol[i];
            crash |= (nc == kNoneInitial && e_is_slack) || (nc == e2) || (c.untouched[i] && e_in_basis0);
#endif
            if (f != 0.0) {  // row is rewritten by this pivot: its only None is column e
                c.nonecol[i] = e2;
                c.untouched[i] = 0;
            }
            const double coeff = -f;
            if (coeff > c.tol) {
                c.pinvs[i] = pivot_inverse(f, c.tol);
                const double ratio = snap(__ddiv_rn(c.T[(size_t)i * c.S], coeff), c.tol);
                if (ratio != 0.0 || LPB_CORRECTED) {  // zero ratios count as +inf in min_index (:161)
                    const unsigned long long k = order_key(ratio);
                    const int vi = (v << 16) | i;
                    if (k < key || (k == key && vi < var)) {
                        key = k;
                        var = vi;
                    }
                }
            }
        }
    }
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
    const unsigned mi = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? (unsigned)var : 0x7fffffffu);
    crash = __any_sync(0xffffffffu, crash);
    if (lane == 0) {
        RatioPartial r;
        r.key = ((unsigned long long)mh << 32) | ml;
        r.id = (int)mi;
        r.pad = crash;
        part[w] = r;
    }
}

// After the block barrier: the leaving row from the warps' partials -- lane w of every warp takes part[w], three
// redux.sync steps give the lexicographic (ratio key, id) minimum.  Same return value as stage_and_ratio().
__device__ __forceinline__ int reduce_rows(const Ctx &c, const RatioPartial *part, unsigned long long *best_key = nullptr)
{
    const int lane = threadIdx.x & 31;
    unsigned long long key = kKeyNone;
    unsigned var = 0x7fffffffu;
    int crash = 0;
    if (lane < c.nw) {
        const RatioPartial r = part[lane];
        key = r.key;
        var = (unsigned)r.id;
        crash = r.pad;
    }
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
    const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
    const unsigned mi = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? var : 0x7fffffffu);
    crash = __any_sync(0xffffffffu, crash);
    if (best_key) *best_key = ((unsigned long long)mh << 32) | ml;
    if (crash) return -2;
    return (mh == 0xffffffffu && ml == 0xffffffffu) ? -1 : (int)mi;
}

// R rows (warp-uniform indices ri[] relative to the block base q0) x the two column-pair passes of this lane:
// all loads first, then T[i][k] = snap(T[i][k] + f_i * r_k) with two roundings (sexp, simplex.py:185-194).
// The factor of row ri[j] sits in lane ri[j]'s register.
template <int R>
__device__ __forceinline__ void sweep_chunk(uint32_t q0, uint32_t rsb, const int (&ri)[kRowChunk], double fmine,
                                            const double2 rA, const double2 rB, bool hasA, bool hasB, double tol)
{
    double f[R];
#pragma unroll
    for (int j = 0; j < R; ++j) f[j] = __shfl_sync(0xffffffffu, fmine, ri[j]);
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
        const bool has = pass ? hasB : hasA;
        const double2 r2 = pass ? rB : rA;
        const uint32_t qb = q0 + (pass ? 512u : 0u);
        if (has) {
            double vx[R], vy[R];
#pragma unroll
            for (int j = 0; j < R; ++j) lds_f64x2(qb + (uint32_t)ri[j] * rsb, vx[j], vy[j]);
            double px[R], py[R];
#pragma unroll
            for (int j = 0; j < R; ++j) {  // the products do not depend on the loads: they run while those are in flight
                px[j] = __dmul_rn(f[j], r2.x);
                py[j] = __dmul_rn(f[j], r2.y);
            }
#ifndef LPB_EXP_NOMATH  // (LPB_EXP_*: timing experiments only, tools/gpu/pivot_rate.py -- the answers are wrong)
#pragma unroll
            for (int j = 0; j < R; ++j) {
                vx[j] = __dadd_rn(vx[j], px[j]);
                vy[j] = __dadd_rn(vy[j], py[j]);
            }
#endif
#ifndef LPB_EXP_NOSTORE
#pragma unroll
            for (int j = 0; j < R; ++j) sts_f64x2(qb + (uint32_t)ri[j] * rsb, vx[j], vy[j]);
#endif
#ifndef LPB_EXP_NOSNAP
#pragma unroll
            for (int j = 0; j < R; ++j) {
                sts_zero_if_snapped(qb + (uint32_t)ri[j] * rsb, vx[j], tol);
                sts_zero_if_snapped(qb + (uint32_t)ri[j] * rsb + 8u, vy[j], tol);
            }
#endif
        }
    }
}

// SlackForm.pivot(e, l) (:348-385) with every warp sweeping the rows it owns; with do_next also the selection of
// the next entering column and the ratio test on it.  Returns (uniform across the CTA) the published next key:
// 0xffffffff = no candidate, bit 31 of *resel set = the set table was rebuilt (every slot moved): reselect.
__device__ __forceinline__ unsigned sweep_rows(Ctx &c, int ncols, int p, int src, int e, int l, int32_t *log,
                                               int log_cap, int npiv, bool do_next, unsigned *resel)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int n = c.n, S = c.S;
    const int dest = c.spare;
    const double tol = c.tol;
    const double pinv = c.pinvs[src];  // snap(-1 * (1.0 / pivot)), left by the row's own ratio test
    const uint32_t par = c.nfused & 1u;
    RatioPartial *part = reinterpret_cast<RatioPartial *>(c.redkey) + (do_next ? par * kSweepWarps : 0u);
    unsigned nk = 0xffffffffu, rs = 0u;
#ifdef LPB_TIMING
    const long long tt0 = clock64();
    long long tt1 = tt0, tt2 = tt0, tt3 = tt0, tt4 = tt0;
#endif
    if (tid < kSweepThreads) {
        const int npairs = c.npairs;
        const double2 *Tsrc = reinterpret_cast<const double2 *>(c.T + (size_t)src * S);
        const double dpiv = snap(-pinv, tol);  // coefficient of l in the new row of e (:343-344)
        const int pp = 1 + p;
        // ---- the objective row first: its owner publishes the entering key of every column but p ----
        if (w == c.ow) {
            const double fobj = c.fc[n];
            unsigned mykey = 0xffffffffu;
            double2 *Tobj = reinterpret_cast<double2 *>(c.T + (size_t)n * S);
            for (int cp = lane; cp < npairs; cp += 32) {
                double2 r2 = Tsrc[cp];
                r2.x = (2 * cp == pp) ? dpiv : snap(__dmul_rn(pinv, r2.x), tol);
                r2.y = (2 * cp + 1 == pp) ? dpiv : snap(__dmul_rn(pinv, r2.y), tol);
                double2 a = Tobj[cp];
                if (fobj != 0.0) {
                    a.x = axpy_snap(a.x, fobj, r2.x, tol);
                    a.y = axpy_snap(a.y, fobj, r2.y, tol);
                    Tobj[cp] = a;
                }
                const int k0 = 2 * cp - 1;  // column position of a.x (-1 is the constant)
                if (k0 >= 0 && k0 < ncols && k0 != p && a.x > tol)
                    mykey = min(mykey, ((unsigned)c.colslot[k0] << 16) | (unsigned)k0);
                if (k0 + 1 < ncols && k0 + 1 != p && a.y > tol)
                    mykey = min(mykey, ((unsigned)c.colslot[k0 + 1] << 16) | (unsigned)(k0 + 1));
            }
            if (do_next) {
                mykey = __reduce_min_sync(0xffffffffu, mykey);
                if (lane == 0) {
                    c.redu32[0] = mykey;
                    mbar_arrive(&c.mbar[0]);  // release: the key is visible to whoever sees the phase flip
                }
            }
        }
#ifdef LPB_TIMING
        tt1 = clock64();
#endif
        // ---- the rows of this warp, 64 column pairs (two passes of the 32 lanes) at a time ----
        // Blocks of up to 32 rows: lane j keeps the factor of row rb + j in a register; the rows that take the
        // plain update (non-zero factor, not the old pivot row, not the slot of the new one, not the objective
        // row) are a bit mask, walked four at a time -- every decision is warp-uniform, there is no slow path.
        const uint32_t tbase = smem_addr(c.T);
        const uint32_t rsb = (uint32_t)S * 8u;  // row stride in bytes
        for (int rb = c.r_lo; rb < c.r_hi; rb += 32) {
            const int nrows = min(32, c.r_hi - rb);
            const double fmine = (lane < nrows) ? c.fc[rb + lane] : 0.0;
            unsigned live = __ballot_sync(0xffffffffu, fmine != 0.0);
            if ((unsigned)(src - rb) < 32u) live &= ~(1u << (src - rb));
            if ((unsigned)(dest - rb) < 32u) live &= ~(1u << (dest - rb));
            if ((unsigned)(n - rb) < 32u) live &= ~(1u << (n - rb));
            const bool dest_here = (unsigned)(dest - rb) < (unsigned)nrows;
            for (int g0 = 0; g0 < npairs; g0 += 64) {
                const int cpA = g0 + lane, cpB = cpA + 32;
                const bool hasA = cpA < npairs, hasB = cpB < npairs;
                double2 rA = make_double2(0.0, 0.0), rB = make_double2(0.0, 0.0);
                if (hasA) {  // this lane's entries of the scaled pivot row (make_pivot_constraint :325-345)
                    rA = Tsrc[cpA];
                    rA.x = (2 * cpA == pp) ? dpiv : snap(__dmul_rn(pinv, rA.x), tol);
                    rA.y = (2 * cpA + 1 == pp) ? dpiv : snap(__dmul_rn(pinv, rA.y), tol);
                }
                if (hasB) {
                    rB = Tsrc[cpB];
                    rB.x = (2 * cpB == pp) ? dpiv : snap(__dmul_rn(pinv, rB.x), tol);
                    rB.y = (2 * cpB + 1 == pp) ? dpiv : snap(__dmul_rn(pinv, rB.y), tol);
                }
                const uint32_t q0 = tbase + (uint32_t)rb * rsb + (uint32_t)cpA * 16u;  // pass B sits 512 bytes further
                if (dest_here) {  // store(entering_idx, entering_constraint) (:352)
                    const uint32_t qd = q0 + (uint32_t)(dest - rb) * rsb;
                    if (hasA) sts_f64x2(qd, rA.x, rA.y);
                    if (hasB) sts_f64x2(qd + 512u, rB.x, rB.y);
                }
                unsigned m = live;
                while (m) {
                    const int cnt = min(__popc(m), kRowChunk);
                    int ri[kRowChunk];
#pragma unroll
                    for (int j = 0; j < kRowChunk; ++j) {
                        ri[j] = m ? (__ffs(m) - 1) : 0;
                        m &= m - 1u;  // (0 stays 0)
                    }
                    if (cnt == 4) {
                        sweep_chunk<4>(q0, rsb, ri, fmine, rA, rB, hasA, hasB, tol);
                    } else {  // the tail of the list: 1 .. 3 rows, one at a time
#pragma unroll
                        for (int j = 0; j < kRowChunk - 1; ++j) {
                            if (j < cnt) {
                                const int one[kRowChunk] = {ri[j], 0, 0, 0};
                                sweep_chunk<1>(q0, rsb, one, fmine, rA, rB, hasA, hasB, tol);
                            }
                        }
                    }
                }
            }
        }
#ifdef LPB_TIMING
        tt2 = tt3 = tt4 = clock64();
#endif
        if (do_next) {
            __syncwarp();                // the lanes' stores to this warp's rows are ordered before the reads below
            mbar_wait(&c.mbar[1], par);  // acquire: the 9th warp's bookkeeping and the key it published
#ifdef LPB_TIMING
            tt3 = tt4 = clock64();
#endif
            nk = *reinterpret_cast<const volatile unsigned *>(c.redu32 + kSweepWarps);
            rs = *reinterpret_cast<const volatile unsigned *>(c.redu32 + kSweepWarps + 1);
#ifdef LPB_EXP_NOSTAGE
            if (rs == 0u && nk != 0xffffffffu && lane == 0) {  // keep the loop alive: any candidate row of this warp
                RatioPartial r;
                r.key = (c.rowvar[c.r_lo] >= 0 && c.r_lo != c.spare) ? 0x8000000000000001ull : kKeyNone;
                r.id = (c.rowvar[c.r_lo] << 16) | c.r_lo;
                r.pad = 0;
                part[w] = r;
                c.pinvs[c.r_lo] = -0.5;
            }
#else
            if (rs == 0u && nk != 0xffffffffu) stage_rows(c, (int)(nk & 0xffffu), part);
#endif
#ifdef LPB_TIMING
            tt4 = clock64();
#endif
        }
    } else {
        if (lane == 0) {
            if (log != nullptr && npiv < log_cap) {
                log[2 * npiv] = e;
                log[2 * npiv + 1] = l;
            }
            bookkeeping(c, p, src, dest, e, l, pinv);  // leaves (key of column p, reselect) in redu32[kSweepWarps..+1]
        }
#ifdef LPB_TIMING
        tt1 = tt2 = tt3 = tt4 = clock64();
#endif
        if (do_next) {
            __syncwarp();
            mbar_wait(&c.mbar[0], par);  // the objective row's key
#ifdef LPB_TIMING
            tt2 = tt3 = tt4 = clock64();
#endif
            if (lane == 0) {
                const unsigned kp = c.redu32[kSweepWarps];
                nk = min(kp, *reinterpret_cast<const volatile unsigned *>(c.redu32));
                rs = c.redu32[kSweepWarps + 1];
                c.redu32[kSweepWarps] = nk;
                mbar_arrive(&c.mbar[1]);  // release: row/column maps, set tables and the key
            }
            nk = __shfl_sync(0xffffffffu, nk, 0);
            rs = __shfl_sync(0xffffffffu, rs, 0);
        }
    }
    if (do_next) ++c.nfused;
    c.spare = src;
    *resel = rs;
    __syncthreads();
#ifdef LPB_TIMING
    if (c.timing != nullptr && lane == 0 && do_next) {
        // [warp][0] objective pre-pass / bookkeeping, [1] own rows / wait for the objective key, [2] wait for the
        // next column, [3] ratio test of own rows, [4] block barrier, [5] pivots
        unsigned long long *t = c.timing + 192 + w * 6;
        const long long tt5 = clock64();
        atomicAdd(t + 0, (unsigned long long)(tt1 - tt0));
        atomicAdd(t + 1, (unsigned long long)(tt2 - tt1));
        atomicAdd(t + 2, (unsigned long long)(tt3 - tt2));
        atomicAdd(t + 3, (unsigned long long)(tt4 - tt3));
        atomicAdd(t + 4, (unsigned long long)(tt5 - tt4));
        atomicAdd(t + 5, 1ull);
    }
#endif
    return nk;
}

__device__ __forceinline__ int run_loop(Ctx &c, int ncols, int &npiv, int max_pivots, int32_t *log,
                                        int log_cap, int forced_p, int forced_lr)
{
    bool forced = forced_lr >= 0;
    int p = forced ? forced_p : select_entering(c, ncols);
    int status = 0;
    int lr = 0;
    bool staged = false;  // the ratio test on column p was done inside the previous sweep
    for (;;) {
        if (p < 0) break;
        const int e = c.colvar[p];
        if (!staged) lr = stage_and_ratio(c, p, e, forced);
        if (forced) lr = forced_lr;
        if (lr == -2) { status = LPB200_STATUS_REF_TYPE_ERROR; break; }
        if (lr < 0) { status = LPB200_STATUS_INFINITY; break; }
        if (npiv >= max_pivots) { status = LPB200_STATUS_ITERATION_LIMIT; break; }
        const int l = lr >> 16, src = lr & 0xffff;
        const RatioPartial *part = reinterpret_cast<const RatioPartial *>(c.redkey) + (c.nfused & 1u) * kSweepWarps;
        unsigned resel = 0u;
        const unsigned nk = sweep_rows(c, ncols, p, src, e, l, log, log_cap, npiv, !forced, &resel);
        ++npiv;
        if (forced || resel != 0u) {
            // after a forced pivot, or when the set table was rebuilt (every slot moved, the keys are stale):
            // a stand-alone pass over the objective row, then the stand-alone ratio test
            forced = false;
            staged = false;
            p = select_entering(c, ncols);
        } else {
            p = (nk == 0xffffffffu) ? -1 : (int)(nk & 0xffffu);
            staged = true;
            if (p >= 0) lr = reduce_rows(c, part);
        }
    }
    return status;
}
#endif  // LPB_WARP_ROWS

#if LPB_REG_ROWS > 0
// =========================================================================================
// Register-resident pivot loop.  During the loop every sweep thread holds its block of the
// tableau -- LPB_REG_ROWS row slots x one column pair -- in REGISTERS; shared memory only
// carries the vectors a pivot exchanges between threads (the staged pivot column fc[], the
// constant column bcol[], the pivot row nr[]) and the objective row (kept fresh for the
// stand-alone entering pass).  A sweep then costs no shared-memory traffic for the tableau,
// which is what bounds the shared-memory variant (128 B/clk/SM).  Four block barriers per
// pivot: stage | ratio | pivot-row publish | sweep.  The tableau is read from shared memory
// when the loop starts and written back when it ends, so set-up, the phase-1 epilogue and
// the results code are shared with the other variant.
// =========================================================================================

// step 2 of a register-resident pivot: min-ratio test on the staged column (fc[]) and the staged
// constant column (bcol[]); same semantics and return value as stage_and_ratio().
__device__ __forceinline__ int ratio_staged(const Ctx &c, int e, bool forced)
{
    const int tid = threadIdx.x;
    const int nphys = c.n + 2;
    int crash = 0;
    const int nsw = min(kSweepWarps, (nphys + 31) >> 5);
    RatioPartial *part = reinterpret_cast<RatioPartial *>(c.redkey);
    if (tid < nsw * 32) {
        unsigned long long key = kKeyNone;
        int var = 0x7fffffff;
#if !LPB_CORRECTED
        const bool e_in_basis0 = (c.basis0[e >> 5] >> (e & 31)) & 1u;
        const bool e_is_slack = e < c.n;
#endif
        for (int i = tid; i < nphys; i += nsw * 32) {
            const int v = c.rowvar[i];
            double f = c.fc[i];
            if (v < 0 && i != c.n) {  // the spare slot holds no row
                f = 0.0;
                c.fc[i] = 0.0;
            }
            if (v >= 0) {
#if !LPB_CORRECTED
                const int nc = c.nonecol[i];
                crash |= (nc == kNoneInitial && e_is_slack) || (nc == e) || (c.untouched[i] && e_in_basis0);
#endif
                if (f != 0.0) {
                    c.nonecol[i] = e;
                    c.untouched[i] = 0;
                }
                const double coeff = -f;
                const bool cand = coeff > c.tol;
                if (cand || (forced && f != 0.0)) c.pinvs[i] = pivot_inverse(f, c.tol);
                if (cand) {
                    const double ratio = snap(__ddiv_rn(c.bcol[i], coeff), c.tol);
                    if (ratio != 0.0 || LPB_CORRECTED) {
                        const unsigned long long k = order_key(ratio);
                        const int vi = (v << 16) | i;
                        if (k < key || (k == key && vi < var)) {
                            key = k;
                            var = vi;
                        }
                    }
                }
            }
        }
        const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
        const unsigned mh = __reduce_min_sync(0xffffffffu, hi);
        const unsigned ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
        const unsigned mi = __reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? (unsigned)var : 0x7fffffffu);
        if ((tid & 31) == 0) {
            RatioPartial r;
            r.key = ((unsigned long long)mh << 32) | ml;
            r.id = (int)mi;
            r.pad = 0;
            part[tid >> 5] = r;
        }
    }
    if (forced) crash = 0;
    if (__syncthreads_or(crash)) return -2;
    RatioPartial best = part[0];
    for (int w = 1; w < nsw; ++w) {
        const RatioPartial r = part[w];
        if (r.key < best.key || (r.key == best.key && r.id < best.id)) best = r;
    }
    return best.key == kKeyNone ? -1 : best.id;
}

__device__ __forceinline__ int run_loop(Ctx &c, int ncols, int &npiv, int max_pivots, int32_t *log,
                                        int log_cap, int forced_p, int forced_lr)
{
    constexpr int R = LPB_REG_ROWS;
    const int tid = threadIdx.x;
    const int n = c.n, S = c.S, nphys = n + 2;
    const double tol = c.tol;
    const bool active = tid < kSweepThreads && c.row0 < c.row1;  // owns a block of the tableau
    const int row0 = c.row0, cxi = c.cxi;
    double2 *T2 = reinterpret_cast<double2 *>(c.T);
    const int rs = S >> 1;
    double2 *nrs = reinterpret_cast<double2 *>(c.nr);

    // ---- tableau block: shared memory -> registers ----
    double2 a[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
        const int row = row0 + j;
        a[j] = (active && j < c.rpt && row < nphys) ? T2[(size_t)row * rs + cxi] : make_double2(0.0, 0.0);
    }

    bool forced = forced_lr >= 0;
    int p = forced ? forced_p : select_entering(c, ncols);
    int status = 0;
#ifdef LPB_TIMING
    const bool tm = (c.timing != nullptr) && ((threadIdx.x & 31) == 0);
    unsigned long long acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
#endif
    for (;;) {
        if (p < 0) break;
#ifdef LPB_TIMING
        const long long t0 = clock64();
#endif
        const int e = c.colvar[p];
        // ---- step 1: the owners of column p stage it (and zero it: it now belongs to the leaving
        //      variable); the owners of the constant column publish b ----
        if (active) {
            const int cpp = (1 + p) >> 1;
            if (cxi == cpp) {
                const bool odd = (1 + p) & 1;
#pragma unroll
                for (int j = 0; j < R; ++j) {
                    double f = odd ? a[j].y : a[j].x;
                    if (odd) a[j].y = 0.0; else a[j].x = 0.0;
                    if (fabs(f) <= tol) f = 0.0;  // substitute() returns the row unchanged (:203-205)
                    if (j < c.rpt) c.fc[row0 + j] = f;
                }
            }
            if (cxi == 0) {
#pragma unroll
                for (int j = 0; j < R; ++j)
                    if (j < c.rpt) c.bcol[row0 + j] = a[j].x;
            }
        }
        __syncthreads();
        // ---- step 2: min-ratio test (:394-436) ----
        int lr = ratio_staged(c, e, forced);
        if (forced) lr = forced_lr;
        if (lr == -2) { status = LPB200_STATUS_REF_TYPE_ERROR; break; }
        if (lr < 0) { status = LPB200_STATUS_INFINITY; break; }
        if (npiv >= max_pivots) { status = LPB200_STATUS_ITERATION_LIMIT; break; }
        const int l = lr >> 16, src = lr & 0xffff;
        const int dest = c.spare;
#ifdef LPB_TIMING
        const long long t1 = clock64();
#endif
        // ---- step 3: the owners of the pivot row publish it (column p already zeroed) ----
        if (active) {
            const int js = src - row0;
#pragma unroll
            for (int j = 0; j < R; ++j)
                if (j == js && j < c.rpt) nrs[cxi] = a[j];
        }
        __syncthreads();
        // ---- step 4: the sweep, in registers (:348-362 -> sexp :185-194) ----
        const double pinv = c.pinvs[src];
        unsigned mykey = 0xffffffffu;
        if (active) {
            double2 r2 = nrs[cxi];  // make_pivot_constraint (:325-345)
            r2.x = (2 * cxi == 1 + p) ? snap(-pinv, tol) : snap(__dmul_rn(pinv, r2.x), tol);
            r2.y = (2 * cxi + 1 == 1 + p) ? snap(-pinv, tol) : snap(__dmul_rn(pinv, r2.y), tol);
            const int jd = dest - row0, jsrc = src - row0, jn = n - row0;
#pragma unroll
            for (int j = 0; j < R; ++j) {
                if (j >= c.rpt) break;  // R is the compile-time capacity, rpt the rows this shape uses
                const double f = c.fc[row0 + j];  // zero beyond the last row slot (padding)
                if (j == jd) {
                    a[j] = r2;  // store(entering_idx, entering_constraint) (:352)
                } else if (j != jsrc && f != 0.0) {
                    a[j].x = axpy_snap(a[j].x, f, r2.x, tol);
                    a[j].y = axpy_snap(a[j].y, f, r2.y, tol);
                }
                if (j == jn) {  // objective row: keep the shared copy fresh, pick the next entering column
                    T2[(size_t)n * rs + cxi] = a[j];
                    const int k0 = 2 * cxi - 1;
                    if (k0 >= 0 && k0 < ncols && k0 != p && a[j].x > tol)
                        mykey = min(mykey, ((unsigned)c.colslot[k0] << 16) | (unsigned)k0);
                    if (k0 + 1 < ncols && k0 + 1 != p && a[j].y > tol)
                        mykey = min(mykey, ((unsigned)c.colslot[k0 + 1] << 16) | (unsigned)(k0 + 1));
                }
            }
        }
        if (tid < kSweepThreads) {
            mykey = __reduce_min_sync(0xffffffffu, mykey);
            if ((tid & 31) == 0) c.redu32[tid >> 5] = mykey;
        } else if (tid == kSweepThreads) {
            if (log != nullptr && npiv < log_cap) {
                log[2 * npiv] = e;
                log[2 * npiv + 1] = l;
            }
            bookkeeping(c, p, src, dest, e, l, pinv);
        }
        c.spare = src;
#ifdef LPB_TIMING
        const long long t2 = clock64();
#endif
        __syncthreads();
#ifdef LPB_TIMING
        const long long t3 = clock64();
        if (tm) { acc0 += t1 - t0; acc1 += t2 - t1; acc2 += t3 - t1; acc3 += 1; }
#endif
        ++npiv;
        if (forced) {
            forced = false;
            p = select_entering(c, ncols);
        } else {
            p = next_entering(c, ncols);
        }
    }
#ifdef LPB_TIMING
    if (tm && acc3) {
        unsigned long long *t = c.timing + (threadIdx.x >> 5) * 4;
        atomicAdd(t + 0, acc0); atomicAdd(t + 1, acc1); atomicAdd(t + 2, acc2); atomicAdd(t + 3, acc3);
    }
#endif
    // ---- tableau block: registers -> shared memory (every exit path; the callers read T) ----
#pragma unroll
    for (int j = 0; j < R; ++j) {
        const int row = row0 + j;
        if (active && j < c.rpt && row < nphys) T2[(size_t)row * rs + cxi] = a[j];
    }
    __syncthreads();
    return status;
}
#elif !LPB_WARP_ROWS
// SlackForm.solve loop (:442-466).  Returns 0 when optimal, else a terminal status code.
// forced_lr >= 0: the first pivot is forced -- the variable at column position forced_p enters,
// the basic variable forced_lr >> 16 in physical row forced_lr & 0xffff leaves (phase 1, :524:
// x_aux enters, the row of min b leaves) -- then the loop runs.
__device__ __forceinline__ int run_loop(Ctx &c, int ncols, int &npiv, int max_pivots, int32_t *log,
                                        int log_cap, int forced_p, int forced_lr)
{
    bool forced = forced_lr >= 0;
#if LPB_RULE == 3
    int streak = 0;  // degenerate pivots in a row (uniform across the CTA)
    unsigned long long best_key = 0ull;
    // a new reference framework wherever a pivot loop is entered (also ahead of a forced pivot)
    for (int k = threadIdx.x; k < c.S; k += kThreads) c.nr[k] = 1.0;
    __syncthreads();
    int p = forced ? forced_p : select_entering_devex(c, ncols, false, -1, 0, 1.0);
#elif LPB_RULE == 2
    int streak = 0;  // degenerate pivots in a row (uniform across the CTA)
    unsigned long long best_key = 0ull;
    int p = forced ? forced_p : select_entering_dantzig(c, ncols, false);
#else
    int p = forced ? forced_p : select_entering(c, ncols);
#endif
#ifdef LPB_TIMING
    const bool tm = (c.timing != nullptr) && ((threadIdx.x & 31) == 0);
    unsigned long long acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
#endif
    int status = 0;
    for (;;) {
        if (p < 0) break;
#ifdef LPB_TIMING
        long long t0 = clock64(), t1, t2 = 0, t3;
#endif
        const int e = c.colvar[p];
#if LPB_RULE == 3
        const double wq = c.nr[p];  // every thread reads it here; the barriers of the ratio test order the later write
        const int dest_row = c.spare;  // physical row the sweep writes the new pivot row to
#endif
#if LPB_RULE >= 2
        int lr = stage_and_ratio(c, p, e, forced, &best_key);
        if (!forced) streak = (best_key == kKeyZeroRatio) ? streak + 1 : 0;
#else
        int lr = stage_and_ratio(c, p, e, forced);
#endif
#ifdef LPB_TIMING
        t1 = clock64();
#endif
        if (forced) lr = forced_lr;
        if (lr == -2) { status = LPB200_STATUS_REF_TYPE_ERROR; break; }
        if (lr < 0) { status = LPB200_STATUS_INFINITY; break; }
        if (npiv >= max_pivots) { status = LPB200_STATUS_ITERATION_LIMIT; break; }
        const int l = lr >> 16, src = lr & 0xffff;
#ifdef LPB_TIMING
        sweep(c, ncols, p, src, e, l, log, log_cap, npiv, &t2);
        t3 = clock64();
        if (tm) { acc0 += t1 - t0; acc1 += t2 - t1; acc2 += t3 - t1; acc3 += 1; }
#else
        sweep(c, ncols, p, src, e, l, log, log_cap, npiv);
#endif
        ++npiv;
#if LPB_RULE == 3
        forced = false;
        p = select_entering_devex(c, ncols, streak > kDegenerateStreak, dest_row, p, wq);
#elif LPB_RULE == 2
        forced = false;
        p = select_entering_dantzig(c, ncols, streak > kDegenerateStreak);
#else
        if (forced) {
            forced = false;
            p = select_entering(c, ncols);
        } else {
            p = next_entering(c, ncols);
        }
#endif
    }
#ifdef LPB_TIMING
    // [warp][0]=stage incl. barrier, [1]=own sweep work, [2]=sweep incl. barrier, [3]=pivots
    if (tm && acc3) {
        unsigned long long *t = c.timing + (threadIdx.x >> 5) * 4;
        atomicAdd(t + 0, acc0); atomicAdd(t + 1, acc1); atomicAdd(t + 2, acc2); atomicAdd(t + 3, acc3);
    }
#endif
    return status;
}
#endif  // LPB_REG_ROWS

// -A into the tableau body (SlackForm.__init__, simplex.py:262: A'[i][n+j] = -A[i][j]).  Kept out
// of line: the pivot loop's register allocation should not depend on this once-per-LP code.
#ifndef LPB_SETUP_ATTR
#define LPB_SETUP_ATTR __noinline__
#endif
// 16-byte loads of A in flight per thread while a problem is read (fp64 path).  The inputs were L2-prefetched
// one queue round ahead, and the CTA that reads shares its SM's load/store path with the other CTA's sweep:
// measured on one box (bench.py, D1): depth 1 3.93 M LPs/s, 2 3.80 M, 4 3.81 M, 8 3.73 M, 12 3.61 M -- a burst
// of loads holds up the neighbour's shared-memory traffic for longer than the latency it hides.
#ifndef LPB_LOAD_DEPTH
#define LPB_LOAD_DEPTH 1
#endif
constexpr int kLoadDepth = LPB_LOAD_DEPTH;
#ifndef LPB_LOAD_DEPTH_TYPED
#define LPB_LOAD_DEPTH_TYPED 4  // typed (int8 .. float32) reader: not yet A/B-measured, see DESIGN.md section 10
#endif
constexpr int kLoadDepthTyped = LPB_LOAD_DEPTH_TYPED;
__device__ LPB_SETUP_ATTR void load_constraints(double *T, int S, const char *gAraw, int a_dtype, int a_esz,
                                                int n, int m)
{
    const int tid = threadIdx.x;
    const double *gA = reinterpret_cast<const double *>(gAraw);  // valid when a_dtype is F64
    const int warp = tid >> 5, lane = tid & 31;
    if (a_dtype != LPB200_DTYPE_F64) {
        // typed A: widened to fp64 here, exactly (ints and float32 are subsets of fp64)
        if ((m & 3) == 0 && ((reinterpret_cast<uintptr_t>(gAraw) & (uintptr_t)(4 * a_esz - 1)) == 0)) {
            // four consecutive elements per load (never straddling a row), kLoadDepthTyped loads in flight
            const int total = (n * m) >> 2, qm = m >> 2;
            for (int base = tid; base < total; base += kLoadDepthTyped * kSweepThreads) {
                double v[kLoadDepthTyped][4];
#pragma unroll
                for (int u = 0; u < kLoadDepthTyped; ++u) {
                    const int idx = base + u * kSweepThreads;
                    if (idx < total) load_typed4(gAraw, a_dtype, idx, v[u]);
                }
#pragma unroll
                for (int u = 0; u < kLoadDepthTyped; ++u) {
                    const int idx = base + u * kSweepThreads;
                    if (idx < total) {
                        const int i = idx / qm, j = (idx - i * qm) * 4;
                        double *Ti = T + (size_t)i * S + 1 + j;
#pragma unroll
                        for (int q = 0; q < 4; ++q) Ti[q] = -v[u][q];
                    }
                }
            }
        } else {
            for (int i = warp; i < n; i += kSweepWarps) {
                double *Ti = T + (size_t)i * S;
                for (int j = lane; j < m; j += 32)
                    Ti[1 + j] = -load_typed1(gAraw, a_dtype, (size_t)i * m + j);
            }
        }
    } else if ((m & 1) == 0 && ((reinterpret_cast<uintptr_t>(gA) & 15) == 0)) {
        // A as a flat stream of 128-bit loads, kLoadDepth in flight per thread; element pairs
        // never straddle a row because m is even
        const double2 *g2 = reinterpret_cast<const double2 *>(gA);
        const int total = (n * m) >> 1, hm = m >> 1;
        for (int base = tid; base < total; base += kLoadDepth * kSweepThreads) {
            double2 v[kLoadDepth];
#pragma unroll
            for (int u = 0; u < kLoadDepth; ++u) {
                const int idx = base + u * kSweepThreads;
                if (idx < total) v[u] = __ldcs(g2 + idx);
            }
#pragma unroll
            for (int u = 0; u < kLoadDepth; ++u) {
                const int idx = base + u * kSweepThreads;
                if (idx < total) {
                    const int i = idx / hm, j = (idx - i * hm) * 2;
                    double *Ti = T + (size_t)i * S + 1 + j;
                    Ti[0] = -v[u].x;
                    Ti[1] = -v[u].y;
                }
            }
        }
    } else {
        for (int i = warp; i < n; i += kSweepWarps) {
            double *Ti = T + (size_t)i * S;
            const double *Ai = gA + (size_t)i * m;
            for (int j = lane; j < m; j += 32) Ti[1 + j] = -Ai[j];
        }
    }
}

#if LPB_REG_ROWS > 0
// 16 warps (15 sweep + 1 bookkeeping) x 128 registers = the SM's whole register file: one CTA per
// SM, the tableau block of every thread (2 x LPB_REG_ROWS doubles) stays in registers
__global__ void __maxnreg__(128)
#else
__global__ void __launch_bounds__(kThreads, LPB_MIN_BLOCKS)
#endif
simplex_batched_kernel(const BatchShape sh, const BatchArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x;
    const int n = sh.n, m = sh.m, S = sh.S;
    const int nv = n + m + 1;
    const int nphys = n + 2;
    Ctx c;
    c.T = reinterpret_cast<double *>(smem + sh.off_T);
    c.nr = reinterpret_cast<double *>(smem + sh.off_nr);
    c.fc = reinterpret_cast<double *>(smem + sh.off_fc);
    c.pinvs = reinterpret_cast<double *>(smem + sh.off_pinvs);
    c.bcol = reinterpret_cast<double *>(smem + sh.off_bcol);
    c.redkey = reinterpret_cast<unsigned long long *>(smem + sh.off_redkey);
    c.rowvar = reinterpret_cast<int *>(smem + sh.off_rowvar);
    c.nonecol = reinterpret_cast<int *>(smem + sh.off_nonecol);
    c.untouched = reinterpret_cast<int *>(smem + sh.off_untouched);
    c.rowslot = reinterpret_cast<int *>(smem + sh.off_rowslot);
    c.colvar = reinterpret_cast<int *>(smem + sh.off_colvar);
    c.colslot = reinterpret_cast<int *>(smem + sh.off_colslot);
    c.varpos = reinterpret_cast<int *>(smem + sh.off_varpos);
    c.basis0 = reinterpret_cast<unsigned *>(smem + sh.off_basis0);
    c.redvar = reinterpret_cast<int *>(smem + sh.off_redvar);
    c.redu32 = reinterpret_cast<unsigned *>(smem + sh.off_redu32);
    c.misc = reinterpret_cast<int *>(smem + sh.off_misc);
    c.tab_i = reinterpret_cast<tab_t *>(smem + sh.off_tab_i);
    c.tab_d = reinterpret_cast<tab_t *>(smem + sh.off_tab_d);
    c.tab_s = reinterpret_cast<tab_t *>(smem + sh.off_tab_s);
    c.n = n;
    c.m = m;
    c.S = S;
    c.npairs = sh.npairs;
    c.cx = sh.cx;
    c.rpg = sh.rpg;
    c.rpt = sh.rpt;
    c.tol = a.tol;
    c.timing = a.timing;
    const double tol = a.tol;
    const int nfc = sh.nfc;
#if LPB_WARP_ROWS
    c.mbar = reinterpret_cast<unsigned long long *>(smem + sh.off_misc + 64);
    {
        const int rpw = (nphys + kSweepWarps - 1) / kSweepWarps;
        const int w = tid >> 5;
        c.rpw = rpw;
        c.r_lo = min(w * rpw, nphys);
        c.r_hi = min(c.r_lo + rpw, nphys);
        c.ow = n / rpw;
        c.nw = (nphys + rpw - 1) / rpw;
    }
    c.nfused = 0u;
    if (tid == 0) {
        mbar_init(&c.mbar[0], 1);
        mbar_init(&c.mbar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
#endif
    {
        const int ry = tid / sh.cx;
        c.cxi = tid - ry * sh.cx;
#if LPB_REG_ROWS > 0
        c.row0 = ry * sh.rpt;
        c.row1 = c.row0 + sh.rpt;
#else
        c.row0 = ry * sh.rpg;
        c.row1 = min(c.row0 + sh.rpg, (nphys + kRowChunk - 1) / kRowChunk * kRowChunk);
#endif
        if (ry >= sh.groups) c.row1 = c.row0;  // leftover threads (groups*cx .. kSweepThreads) have no block
    }

    for (;;) {
        // ---- next problem off the device work queue ----
        __syncthreads();
        if (tid == 0) c.misc[MI_LP] = (int)atomicAdd(a.work_counter, 1u);
        __syncthreads();
        const int lp = c.misc[MI_LP];
        if (lp >= a.B) break;
        if (tid >= kSweepThreads) {
            // The 9th warp pulls the inputs of problem lp + gridDim.x into L2: with the queue
            // handing out indices in order, that is the problem some CTA (usually this one) sets
            // up one solve from now, which then reads A at L2 rather than HBM latency.
            const long long nxt = (long long)lp + gridDim.x;
            if (nxt < a.B) {
                const size_t bytes = (size_t)n * m * a.a_esz;
                const char *pa = reinterpret_cast<const char *>(a.A) + (size_t)nxt * bytes;
                for (size_t o = (size_t)(tid - kSweepThreads) * 128; o < bytes; o += 32 * 128)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pa + o));
            }
        }
        const char *gAraw = reinterpret_cast<const char *>(a.A) + (size_t)lp * n * m * a.a_esz;
        const double *gb = a.b + (size_t)lp * n;
        const double *gc = a.c + (size_t)lp * m;
        int32_t *log = a.pivot_log ? a.pivot_log + (size_t)lp * a.log_cap * 2 : nullptr;

        // ---- min_constant, idx = min_index(b): zeros count as +inf, first minimum wins (:591) ----
        if (tid < kSweepThreads) {
            unsigned long long key = kKeyNone;
            int idx = 0x7fffffff;
            for (int i = tid; i < n; i += kSweepThreads) {
                const double bi = gb[i];
                const unsigned long long k = order_key((bi != 0.0 || LPB_CORRECTED) ? bi : CUDART_INF);
                if (k < key || (k == key && i < idx)) {
                    key = k;
                    idx = i;
                }
            }
            warp_argmin(key, idx);
            if ((tid & 31) == 0) {
                c.redkey[tid >> 5] = key;
                c.redvar[tid >> 5] = idx;
            }
        }
        __syncthreads();
        int idx0 = c.redvar[0];
        {
            unsigned long long key = c.redkey[0];
#pragma unroll
            for (int w = 1; w < kSweepWarps; ++w) {
                const unsigned long long k = c.redkey[w];
                const int v = c.redvar[w];
                if (k < key || (k == key && v < idx0)) {
                    key = k;
                    idx0 = v;
                }
            }
        }
        const double min_c = gb[idx0];
        const bool phase1 = (min_c < 0.0);
        int status = 0;
        int npiv = 0;
        int ncols = phase1 ? m + 1 : m;
        if (!LPB_CORRECTED && !phase1 && !(min_c > tol)) status = LPB200_STATUS_REF_ATTRIBUTE_ERROR;  // :626
        c.track_dep = phase1;
        c.spare = n + 1;

        if (status == 0) {
            // ---- SlackForm.__init__ (:230-302) (+ the x_aux column of :632-660 in phase 1) ----
            if (tid < kSweepThreads) {
                load_constraints(c.T, S, gAraw, a.a_dtype, a.a_esz, n, m);
                for (int i = tid; i < n; i += kSweepThreads) {
                    double *Ti = c.T + (size_t)i * S;
                    Ti[0] = gb[i];
                    Ti[1 + m] = phase1 ? 1.0 : 0.0;  // -(-1), :659 then :262
                    for (int k = m + 2; k < S; ++k) Ti[k] = 0.0;
                }
                double *obj = c.T + (size_t)n * S;
                for (int k = tid; k < S; k += kSweepThreads) {
                    double v = 0.0;
                    if (k >= 1 && k <= m) v = phase1 ? 0.0 : gc[k - 1];
                    if (k == 1 + m && phase1) v = -1.0;  // objective -x_aux (:644-645)
                    obj[k] = v;
                    obj[S + k] = 0.0;  // the spare slot
                    c.nr[k] = 0.0;
                }
                for (int i = tid; i < nfc; i += kSweepThreads) c.fc[i] = 0.0;
                for (int i = tid; i < nphys; i += kSweepThreads) {
                    c.rowvar[i] = (i < n) ? i : -1;
                    c.nonecol[i] = kNoneInitial;
                    c.untouched[i] = 0;
                    c.rowslot[i] = i & (sh.size_d0 - 1);
                    if (i < n) c.varpos[i] = ~i;
                }
                const int isize = phase1 ? sh.size_i1 : sh.size_i0;
                for (int k = tid; k <= m; k += kSweepThreads) {
                    c.colvar[k] = n + k;
                    c.varpos[n + k] = k;
                    c.colslot[k] = LPB_CORRECTED ? (n + k) : ((n + k) & (isize - 1));
                }
                for (int w = tid; w < (nv + 31) / 32; w += kSweepThreads) c.basis0[w] = 0u;
            } else {
                // set(range(n, n+ncols)) and set(range(0, n)) (:289-292), built by the 9th warp
                const int lane = tid - kSweepThreads;
                const int isize = phase1 ? sh.size_i1 : sh.size_i0;
                for (int s = lane; s < isize; s += 32) c.tab_i[s] = (tab_t)kSetUnused;
                __syncwarp();
                for (int k = n + lane; k < n + ncols; k += 32) c.tab_i[k & (isize - 1)] = (tab_t)k;
                if (phase1) {
                    for (int s = lane; s < sh.size_d0; s += 32) c.tab_d[s] = (tab_t)kSetUnused;
                    __syncwarp();
                    for (int k = lane; k < n; k += 32) c.tab_d[k & (sh.size_d0 - 1)] = (tab_t)k;
                }
                if (lane == 0) {
                    c.misc[MI_ISIZE] = isize;
                    c.misc[MI_IFILL] = ncols;
                    c.misc[MI_IUSED] = ncols;
                    c.misc[MI_DSIZE] = sh.size_d0;
                    c.misc[MI_DFILL] = n;
                    c.misc[MI_DUSED] = n;
                    c.redu32[kSweepWarps + 1] = 0u;
                }
            }
            __syncthreads();

            // One instance of the pivot loop serves phase 1 (:526, entered through the forced
            // pivot), phase 2 after it (:599) and phase 2 alone (:619).
            // the forced pivot: x_aux (column position m) enters, the row of min b -- which still
            // holds its own slack variable, id == row -- leaves (:524)
            int forced_p = m, forced_lr = phase1 ? ((idx0 << 16) | idx0) : -1;
            bool in_phase1 = phase1;
            for (;;) {
                status = run_loop(c, ncols, npiv, a.max_pivots, log, a.log_cap, forced_p, forced_lr);
                if (!in_phase1) break;  // phase 2 is done
                forced_lr = -1;
                const int aux = n + m;
                if (status == 0) {
                    if (!(fabs(c.T[(size_t)n * S]) <= tol)) status = LPB200_STATUS_NO_SOLUTION;  // :528-530
#if LPB_CORRECTED
                    else if (c.varpos[aux] < 0) {
                        // x_aux is basic at level 0: pivot it out on the lowest-id column with a
                        // non-zero entry in its row; an all-zero row is a redundant constraint, left inert
                        const int r = ~c.varpos[aux];
                        const int k = driveout_column(c, ncols, r);
                        if (k >= 0) {
                            forced_p = k;
                            forced_lr = (aux << 16) | r;
                            continue;
                        }
                    }
#else
                    else if (c.varpos[aux] < 0) status = LPB200_STATUS_REF_INDEX_ERROR;  // :541 -> :183
#endif
                }
                if (status != 0) break;
                in_phase1 = false;
                {
                    // ---- rebuild the original objective (:537-541) ----
                    // basic ORIGINAL variables in iteration order of the `dependent` set
                    // (slack rows have objective coefficient None: skipped, :203)
                    int *list = reinterpret_cast<int *>(c.fc);
                    if (tid < 32) {
                        const int dsize = c.misc[MI_DSIZE];
                        int cnt = 0;
                        for (int s0 = 0; s0 < dsize; s0 += 32) {
                            const int s = s0 + tid;
                            const int v = (s < dsize) ? (int)c.tab_d[s] : -1;
                            const unsigned msk = __ballot_sync(0xffffffffu, v >= n);
                            if (v >= n) list[cnt + __popc(msk & ((1u << tid) - 1u))] = v;
                            cnt += __popc(msk);
                        }
                        if (tid == 0) c.misc[MI_NLIST] = cnt;
                    }
                    // stage c in nr[], start from (v, c) = (0, [None]*n + float(c)) (:516, :537)
                    if (tid < kSweepThreads) {
                        for (int j = tid; j <= m; j += kSweepThreads) c.nr[j] = (j < m) ? gc[j] : 0.0;  // [m]: x_aux
                    }
                    __syncthreads();
                    const int kaux = c.varpos[aux];
                    if (tid < kSweepThreads) {
                        double *obj = c.T + (size_t)n * S;
                        const int nlist = c.misc[MI_NLIST];
                        for (int k = tid; k < 1 + ncols; k += kSweepThreads) {
                            if (k == 1 + kaux) continue;  // zip() truncates the x_aux column (:190)
                            double o = 0.0;
                            if (k >= 1) {
                                const int w = c.colvar[k - 1];
                                o = (w >= n) ? c.nr[w - n] : 0.0;
                            }
                            for (int t = 0; t < nlist; ++t) {
                                const int v = list[t];
                                const double f = c.nr[v - n];
                                if (fabs(f) <= tol) continue;
                                o = axpy_snap(o, f, c.T[(size_t)(~c.varpos[v]) * S + k], tol);
                            }
                            obj[k] = o;
                        }
                    }
                    __syncthreads();
                    // ---- clean-up (:547-575): None out basic columns, drop x_aux, same sets live on ----
                    const int last = ncols - 1;
                    if (tid < kSweepThreads) {
                        for (int w = tid; w < (nv + 31) / 32; w += kSweepThreads) {
                            unsigned bits = 0u;
                            for (int bb = 0; bb < 32; ++bb) {
                                const int v = w * 32 + bb;
                                if (v < nv && c.varpos[v] < 0) bits |= 1u << bb;
                            }
                            c.basis0[w] = bits;
                        }
                        for (int i = tid; i < nphys; i += kSweepThreads) c.untouched[i] = 1;
                        for (int i = tid; i < nphys && kaux >= 0; i += kSweepThreads) {
                            double *Ti = c.T + (size_t)i * S;
                            if (kaux != last) Ti[1 + kaux] = Ti[1 + last];
                            Ti[1 + last] = 0.0;
                        }
                        for (int k = tid; k < S; k += kSweepThreads) c.nr[k] = 0.0;
                        for (int i = tid; i < nfc; i += kSweepThreads) c.fc[i] = 0.0;  // list[] lived here
                    } else if (tid == kSweepThreads && kaux >= 0) {
#if !LPB_CORRECTED
                        PySet s = {c.misc[MI_ISIZE], c.misc[MI_IFILL], c.misc[MI_IUSED]};
                        pyset_remove_slot(c.tab_i, s, c.colslot[kaux]);  // independent.remove(nvars) (:574)
                        c.misc[MI_IUSED] = s.used;
#endif
                        if (kaux != last) {
                            const int w = c.colvar[last];
                            c.colvar[kaux] = w;
                            c.colslot[kaux] = c.colslot[last];
                            c.varpos[w] = kaux;
                        }
                    }
                    if (kaux >= 0) ncols = m;  // (Bland only: x_aux left basic in an all-zero row keeps m+1 columns)
                    __syncthreads();
                }
            }
        }

        // ---- results (:473-491, 603, 622) ----
        if (tid < kSweepThreads) {
            double *gx = a.x + (size_t)lp * m;
            for (int j = tid; j < m; j += kSweepThreads) {
                double v = 0.0;
                if (status == 0) {
                    const int pos = c.varpos[n + j];
                    if (pos < 0) v = c.T[(size_t)(~pos) * S];
                }
                gx[j] = v;
            }
            if (tid == 0) {
                a.status[lp] = status;
                a.npivots[lp] = npiv;
                a.objective[lp] = (status == 0) ? c.T[(size_t)n * S] : 0.0;
            }
        }
    }
}


// ---- host side of this variant ----
static int launch(const DeviceInfo &di, int B, int n, int m, const void *A, int a_dtype, const double *b,
                  const double *c,
                  double tol, int max_pivots, int32_t *status, double *x, double *objective, int32_t *npivots,
                  int32_t *pivot_log, int log_cap, cudaStream_t stream)
{
    const BatchShape sh = make_batch_shape(n, m);
    if (sh.bytes > di.max_smem_optin) {
        set_error("lpb200_simplex_batch: tableau does not fit shared memory (" + std::to_string(sh.bytes) +
                  " > " + std::to_string(di.max_smem_optin) + " bytes); use lpb200_simplex_single");
        return LPB200_E_TOO_LARGE;
    }
    int dyn_smem = sh.bytes;
    if (const char *g = getenv("LPB_DEBUG_PAD")) dyn_smem += atoi(g);
    LPB_CUDA(cudaFuncSetAttribute(simplex_batched_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_smem));
    int per_sm = 0;
    LPB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, simplex_batched_kernel, kThreads, dyn_smem));
    if (per_sm < 1) per_sm = 1;
    long long grid = (long long)di.sm_count * per_sm;
    if (const char *g = getenv("LPB_DEBUG_GRID")) grid = atoi(g);
    if (grid > B) grid = B;
    unsigned int *counter = nullptr;
    {
        const int rc = scratch_alloc((void **)&counter, sizeof(unsigned int), di.device, stream);
        if (rc) return rc;
    }
    LPB_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned int), stream));
    BatchArgs a;
    a.B = B;
    a.A = A;
    a.a_dtype = a_dtype;
    a.a_esz = dtype_bytes(a_dtype);
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
    a.work_counter = counter;
    a.timing = nullptr;
    const bool dbg_timing = getenv("LPB_DEBUG_TIMING") != nullptr;
    if (dbg_timing) {
        LPB_CUDA(cudaMalloc((void **)&a.timing, 64 * 4 * 8));
        LPB_CUDA(cudaMemset(a.timing, 0, 64 * 4 * 8));
    }
    simplex_batched_kernel<<<(unsigned)grid, kThreads, dyn_smem, stream>>>(sh, a);
    LPB_CUDA(cudaGetLastError());
    LPB_CUDA(cudaFreeAsync(counter, stream));
    if (dbg_timing) {  // debug only: synchronises
        unsigned long long h[64 * 4];
        LPB_CUDA(cudaStreamSynchronize(stream));
        LPB_CUDA(cudaMemcpy(h, a.timing, sizeof(h), cudaMemcpyDeviceToHost));
        for (int w = 0; w <= kSweepWarps; ++w)
            if (h[w * 4 + 3])
                fprintf(stderr, "[lpb timing] warp %d: stage+bar %.0f  own sweep %.0f  sweep+bar %.0f cycles/pivot (%llu pivots)\n",
                        w, (double)h[w * 4] / h[w * 4 + 3], (double)h[w * 4 + 1] / h[w * 4 + 3],
                        (double)h[w * 4 + 2] / h[w * 4 + 3], h[w * 4 + 3]);
#if LPB_WARP_ROWS
        for (int w = 0; w <= kSweepWarps; ++w) {
            const unsigned long long *t = h + 192 + w * 6;
            if (t[5])
                fprintf(stderr, "[lpb timing] warp %d: %s %.0f  %s %.0f  wait-next %.0f  ratio-own-rows %.0f  barrier %.0f  "
                                "cycles/pivot (%llu fused pivots)\n", w, w < kSweepWarps ? "objective-row" : "bookkeeping",
                        (double)t[0] / t[5], w < kSweepWarps ? "own-rows" : "wait-objective-key", (double)t[1] / t[5],
                        (double)t[2] / t[5], (double)t[3] / t[5], (double)t[4] / t[5], t[5]);
        }
#endif
        for (int w = 0; w < kSweepWarps; ++w)
            if (h[w * 4 + 3])
                fprintf(stderr, "[lpb timing] warp %d sweep: prologue %.0f  general path %.0f cycles/pivot in %.2f chunks/pivot\n", w,
                        (double)h[128 + w * 4] / h[w * 4 + 3], (double)h[128 + w * 4 + 1] / h[w * 4 + 3],
                        (double)h[128 + w * 4 + 2] / h[w * 4 + 3]);
        cudaFree(a.timing);
    }
    return 0;
}

// Shapes this variant can serve (register-resident variants need one column-pair pass and at
// most LPB_REG_ROWS rows per thread; the shared-memory variant takes anything that fits).
[[maybe_unused]] static bool eligible(int n, int m)
{
#if LPB_REG_ROWS > 0
    const BatchShape s = make_batch_shape(n, m);
    return s.npairs <= 64 && s.rpt <= LPB_REG_ROWS;
#else
    (void)n;
    (void)m;
    return true;
#endif
}

}  // namespace LPB_NS
}  // namespace lpb
