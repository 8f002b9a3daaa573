// simplex_single.cu -- grid-wide two-phase simplex for ONE large dense LP (sm_100a), on one
// GPU or COLUMN-SHARDED over the GPUs of one NVLink/NVSwitch box.
//
// Same semantics as simplex_batched.cu (the reference's Simplex.solve(),
// week-2/diet/simplex.py:586-630) but the condensed tableau ((n+1) x (m+1) fp64, 4.3 GB at
// 16384 x 32768) lives in HBM and every pivot is ONE streaming read-modify-write sweep over
// it by all SMs.  One cooperative persistent kernel per GPU runs the whole solve; per pivot:
//
//   P1  entering column   every CTA scans a slice of the objective row, block-reduces the
//                         packed key (set slot << 32 | position), writes one partial     (:390-392)
//       [ranks > 1]       KEY EXCHANGE: every rank pushes its 8-byte key into every peer's
//                         exchange buffer over NVLink (all-to-all of 16-byte LL packets,
//                         data+epoch flag in each 8-byte word, no fence) and takes the min
//   P2  leaving row       the rank that owns column p gathers it and BROADCASTS it (n+1
//                         doubles as LL packets) into every rank's buffer; every rank decodes
//                         it into fc[] and runs the min-ratio test redundantly on the
//                         replicated constant column -> identical (l, r) everywhere       (:394-436)
//   P3  pivot row         nr[] = snap(pinv * row r) on the rank's own columns             (:325-345)
//   P4  rank-1 sweep      tiles of 64 rows x 512 columns: the tile's slice of the pivot
//                         row nr[] and of the pivot column fc[] are staged in shared memory
//                         by TMA bulk copies (cp.async.bulk + mbarrier, double-buffered
//                         across tiles); each thread keeps its double2 of nr in registers
//                         and streams the rows with 128-bit loads/stores, 8 rows in flight:
//                         T[i][k] = snap(T[i][k] + fc[i]*nr[k]), no FMA                  (:348-362, 185-214)
//   grid.sync() between phases (4 per pivot, ~1 % of a 1.3 ms sweep).
//
// Column sharding: rank g owns column positions [g*Wc, (g+1)*Wc) of ALL rows
// incl. the objective row; the constant column, the basis/slot tables and the CPython-set
// emulation are replicated and updated identically by every rank, so the only traffic per
// pivot is the 8-byte key all-to-all and the pivot-column broadcast: compute and collective
// are ONE kernel working on peer-mapped memory, no NCCL call and no host round trip on the
// pivot path.  Waits are bounded: a rank that sees no packet for `timeout` traps instead of
// hanging the GPU.  For tests (and boxes with fewer GPUs than ranks) the same kernel runs all
// ranks as CTA groups of one cooperative launch ("emulated"), which makes every wait safe.
//
// Pivoting rule (SingleArgs.rule, a run-time argument: the selection phases are a few microseconds
// of a 1.3 ms pivot, the sweep is rule-independent): LPB200_RULE_REFERENCE = the reference's
// decisions (parity); LPB200_RULE_BLAND / LPB200_RULE_DANTZIG / LPB200_RULE_DEVEX = the corrected
// modes of the batch kernel (lowest-id / largest-coefficient / Devex-weighted entering, zero ratios
// kept, true minimum of b, x_aux driven out after phase 1, no crash classes), checked bit-exact
// against oracle/simplex_oracle.c.  Devex keeps one weight per owned column (wts[]); the weight of
// the entering column rides along with the pivot-column broadcast as element n+1.
//
// Tensor cores are not used: the update is a rank-1 AXPY at 0.125 flop/byte, HBM-bound.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include <stdlib.h>

#include <vector>

#include "lp_common.cuh"
#include "lp_host.h"

namespace cg = cooperative_groups;

namespace lpb {

constexpr int kSgThreads = 256;
constexpr int kTileCols = 2 * kSgThreads;  // 512 doubles: one double2 per thread
#ifndef LPB_TILE_ROWS
#define LPB_TILE_ROWS 64
#endif
constexpr int kTileRows = LPB_TILE_ROWS;
constexpr int kRowUnroll = 8;
constexpr int kMaxRanks = 32;
constexpr int kDegenerateStreak = 8;  // Dantzig: degenerate pivots in a row before Bland takes over (as in the batch kernel)
constexpr unsigned long long kKeyZeroRatio = 0x8000000000000000ull;  // order_key(+0.0)

enum { ST_ISIZE = 0, ST_IFILL, ST_IUSED, ST_DSIZE, ST_DFILL, ST_DUSED, ST_NLIST, ST_NCOLS,
       ST_TCUR,                               // 1: the tableau the solve stopped on is T2 (pipelined loop, see below)
       ST_LA_STATUS,                          // look-ahead hand-over, [2][4] by pivot parity: outcome of the NEXT pivot's
                                              // selection (LA_* code, entering position, leaving variable, its row)
       ST_COUNT = 32 };
enum { LA_PIVOT = 0, LA_OPTIMAL, LA_INFINITY, LA_TYPE_ERROR };

// Exchange buffer of one rank, in 8-byte words.  Every payload word is an "LL" word:
// low 32 bits data, high 32 bits the epoch of the exchange that wrote it.
constexpr int kXbEpoch = 0;  // running epoch, persists across solves on the same communicator
constexpr int kXbKx = 1;     // key exchanges so far (its parity selects the key buffer): persists too, so that the
                             // "at most one exchange ahead" invariant of the double buffer holds ACROSS solves
constexpr int kXbKeys = 8;   // [2][nranks] x (key packet, id packet) of 2 words each (double-buffered by exchange parity)
__host__ __device__ inline int xb_fc(int nranks) { return kXbKeys + 8 * nranks; }  // [(n+2)] packets
static inline size_t xbuf_bytes_for(int n, int nranks)
{
    return ((size_t)(xb_fc(nranks) + 2 * (n + 2)) * 8 + 255) / 256 * 256;
}

// Per-rank device view: workspace carve-up, column range, exchange buffers, outputs.
struct SingleWS {
    int n, m, S;                  // S: LOCAL row stride in doubles (multiple of 16) = positions per rank
    int rank, nranks, c_lo, wloc; // owns global column positions [c_lo, c_lo + wloc)
    int size_i0, size_i1, size_d0, cap_i, cap_d, cap_s;
    const double *A;              // this rank's columns of A: A[i*lda + (k - c_lo)]
    long long lda;
    double *T;                    // [(n+1)][S]  owned non-basic columns; row n = objective
    double *T2;                   // [(n+1)][S]  second tableau buffer: the pipelined loop sweeps T -> T2 -> T ...
    double *bcol;                 // [n+1]       constant column (b, then v at index n), replicated
    double *nr;                   // [S]         scaled pivot row, owned columns
    double *fc;                   // [n+1]       pivot column factors, replicated
    double *fc2;                  // [n+1]       the NEXT pivot's column (written by the look-ahead while fc is in use)
    double *cstage;               // scratch: factors of the objective rebuild
    double *scal;                 // [4]         nr constant
    double *wts;                  // [S]         LPB200_RULE_DEVEX: reference-framework weights of the owned columns
    int *rowvar, *nonecol, *untouched, *rowslot;  // [n]      replicated
    int *colvar, *colslot;                        // [m+1]    replicated (global positions)
    int *varpos;                                  // [n+m+1]  replicated
    unsigned *basis0;
    int *tab_i, *tab_d, *tab_s;
    int *list;
    unsigned long long *part_key;                 // [2][ctas]  set 0: entering, set 1: leaving
    int *part_var, *part_crash;
    int *state;
    unsigned *labar;              // counter of the control barrier (pipelined loop)
    unsigned long long *xbuf;                     // this rank's exchange buffer (nranks > 1)
    unsigned long long *const *peers;             // [nranks] every rank's exchange buffer, peer-mapped
    int32_t *status;
    double *x, *objective;
    int32_t *npivots, *pivot_log;
};

struct SingleArgs {
    const double *b, *c;
    double tol;
    int max_pivots;
    int log_cap;
    int cpr;  // CTAs per rank in this launch
    int rule; // LPB200_RULE_*: 0 = the reference's decisions, 1 = Bland, 2 = Dantzig (corrected modes)
    int nctl_max;   // control CTAs per rank of the pipelined loop, at most (LPB_DEBUG_SINGLE_NCTL)
    int pipelined;  // 1: rules 0/1 run the pipelined loop (one barrier per pivot); 0: step by step (LPB_DEBUG_SINGLE_PIPELINE=0)
    unsigned long long timeout_ns;
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

// ---- LL packets over peer-mapped memory: 16 bytes = two (data32 | epoch32) words ----
// Only 8-byte store atomicity is relied on: the reader accepts a packet when BOTH words
// carry the expected epoch, so no fence and no separate flag round trip is needed.
__device__ __forceinline__ void ll_store(unsigned long long *dst, unsigned long long v, unsigned ep)
{
    const unsigned long long w0 = (v & 0xffffffffull) | ((unsigned long long)ep << 32);
    const unsigned long long w1 = (v >> 32) | ((unsigned long long)ep << 32);
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(w0), "l"(w1) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned long long ll_load(const unsigned long long *src, unsigned ep,
                                                      unsigned long long timeout_ns)
{
    unsigned long long w0, w1, t0 = 0;
    unsigned spins = 0;
    for (;;) {
        asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(src) : "memory");
        if ((unsigned)(w0 >> 32) == ep && (unsigned)(w1 >> 32) == ep) break;
        if ((++spins & 0xfffu) == 0) {  // bounded wait: a dead peer must not hang the GPU
            const unsigned long long t = globaltimer_ns();
            if (t0 == 0) t0 = t;
            else if (t - t0 > timeout_ns) __trap();
        }
    }
    return (w0 & 0xffffffffull) | (w1 << 32);
}

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

// Per-thread loop state (uniform across all threads of all ranks).
struct LoopCtl {
    int bid, nblk;    // CTA index within this rank's group, CTAs per rank
    int ncols;        // column positions alive (m, or m+1 once phase 1 has run)
    int npiv;
    bool track_dep;
    uint32_t parity;  // mbarrier phase bits of the two sweep stages
    unsigned epoch;   // exchange epoch (continues across solves)
    unsigned kx;      // key exchanges so far (parity selects the key buffer)
    int streak;       // LPB200_RULE_DANTZIG / DEVEX: degenerate pivots in a row in the current loop
    int tcur;         // 1: the kernel's w.T is the workspace's T2 (the pipelined loop ended on an odd pivot count)
    int nctl;         // pipelined loop: CTAs 0 .. nctl-1 of the rank are its control CTAs, the others sweep
    unsigned labar_gen;  // control barriers passed so far (control CTAs only)
    // LPB200_RULE_DEVEX: the last pivot (entering position, its weight) still has to update the weights
    bool upd;
    int p_prev;
    double wq;
};

// Reduce the per-CTA partials written before the last grid.sync().
__device__ __forceinline__ void reduce_partials(BlockRed &br, const SingleWS &w, const LoopCtl &ctl, int set,
                                                unsigned long long &key, int &var, int &flag)
{
    key = kKeyNone;
    var = 0x7fffffff;
    flag = 0;
    const int lo = set * ctl.nblk;
    for (int q = lo + threadIdx.x; q < lo + ctl.nblk; q += kSgThreads) {
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

__device__ __forceinline__ void write_partial(const SingleWS &w, const LoopCtl &ctl, int set,
                                              unsigned long long key, int var, int flag)
{
    if (threadIdx.x == 0) {
        const int q = set * ctl.nblk + ctl.bid;
        w.part_key[q] = key;
        w.part_var[q] = var;
        w.part_crash[q] = flag;
    }
}

// ---- P1: entering column (pick_entering_idx :390-392), own columns only ----
// pricing: 0 = first positive coefficient in CPython-set slot order (the reference), 1 = lowest
// variable id (Bland), 2 = largest coefficient, ties -> lowest id (Dantzig).  The partial is the
// lexicographic minimum of (key, id): rules 0/1 pack (priority << 32 | position) into the key,
// rule 2 uses key = ~bits(c_j) (positive doubles order like their bit patterns) and id = variable.
// Devex (rule 3): key = ~bits((c_j*c_j)/w_j), id = variable; the weights of the owned columns are first brought
// up to date with the pivot performed last (ctl.upd; one thread of the grid owns a column in this pass):
// w[k] = max(w[k], (nr[k]*nr[k])*wq), w[p_prev] = max((nr[p_prev]*nr[p_prev])*wq, 1) -- the expressions of
// select_entering_devex() in the batch kernel and of the CPU restatement.
__device__ __forceinline__ void phase_entering(BlockRed &br, const SingleWS &w, const LoopCtl &ctl, double tol,
                                               int pricing, int rule)
{
    unsigned long long key = kKeyNone;
    int var = 0x7fffffff, flag = 0;
    const double *obj = w.T + (size_t)w.n * w.S;
    const int lim = min(w.wloc, ctl.ncols - w.c_lo);
    const int *prio = (pricing == 0) ? w.colslot : w.colvar;
    for (int kl = ctl.bid * kSgThreads + threadIdx.x; kl < lim; kl += ctl.nblk * kSgThreads) {
        const double v = obj[kl];
        double wgt = 1.0;
        if (rule == LPB200_RULE_DEVEX) {
            wgt = w.wts[kl];
            if (ctl.upd) {
                const double a = w.nr[kl];
                const double cand = __dmul_rn(__dmul_rn(a, a), ctl.wq);
                if (w.c_lo + kl == ctl.p_prev) wgt = cand > 1.0 ? cand : 1.0;
                else if (cand > wgt) wgt = cand;
                w.wts[kl] = wgt;
            }
        }
        if (v > tol) {
            const int k = w.c_lo + kl;
            if (pricing == LPB200_RULE_DEVEX) {
                const unsigned long long kk = ~(unsigned long long)__double_as_longlong(__ddiv_rn(__dmul_rn(v, v), wgt));
                const int id = w.colvar[k];
                if (kk < key || (kk == key && id < var)) {
                    key = kk;
                    var = id;
                }
            } else if (pricing == 2) {
                const unsigned long long kk = ~(unsigned long long)__double_as_longlong(v);
                const int id = w.colvar[k];
                if (kk < key || (kk == key && id < var)) {
                    key = kk;
                    var = id;
                }
            } else {
                const unsigned long long kk = ((unsigned long long)(unsigned)prio[k] << 32) | (unsigned)k;
                key = kk < key ? kk : key;
            }
        }
    }
    block_argmin(br, key, var, flag);
    write_partial(w, ctl, 0, key, var, flag);
}

// ---- all-to-all of the ranks' entering keys over NVLink, min on arrival ----
// with_id: the (key, id) pair travels as two packets and the minimum is lexicographic (Dantzig).
__device__ __forceinline__ unsigned long long exchange_key(BlockRed &br, const SingleWS &w, LoopCtl &ctl,
                                                           unsigned long long key, int &var, bool with_id,
                                                           unsigned long long timeout_ns)
{
    if (w.nranks == 1) return key;
    const int tid = threadIdx.x, G = w.nranks;
    const unsigned ep = ++ctl.epoch;
    const int buf = (int)(ctl.kx++ & 1u);
    if (ctl.bid == 0 && tid < G) {
        unsigned long long *dst = w.peers[tid] + kXbKeys + 4 * (buf * G + w.rank);
        ll_store(dst, key, ep);
        if (with_id) ll_store(dst + 2, (unsigned long long)(unsigned)var, ep);
    }
    __syncthreads();  // previous readers of br are done
    if (tid < 32) {
        unsigned long long k = kKeyNone;
        int v = 0x7fffffff;
        if (tid < G) {
            const unsigned long long *src = w.xbuf + kXbKeys + 4 * (buf * G + tid);
            k = ll_load(src, ep, timeout_ns);
            if (with_id) v = (int)(unsigned)ll_load(src + 2, ep, timeout_ns);
        }
        warp_argmin(k, v);
        if (tid == 0) {
            br.key[0] = k;
            br.var[0] = v;
        }
    }
    __syncthreads();
    var = br.var[0];
    return br.key[0];
}

// ---- P2: pivot column p -> fc[] on every rank, min-ratio test on it (pick_leaving_idx :394-436) ----
// corrected (rules 1, 2, 3): zero ratios are kept and the None bookkeeping raises nothing.
// devex: the Devex weight of the entering column travels with the pivot column as element n+1 of fc[].
__device__ __forceinline__ void phase_ratio(BlockRed &br, const SingleWS &w, LoopCtl &ctl, int p, int e,
                                            double tol, bool corrected, bool devex, unsigned long long timeout_ns)
{
    unsigned long long key = kKeyNone;
    int var = 0x7fffffff, crash = 0;
    const int n = w.n, G = w.nranks;
    const int owner = p / w.S, pl = p - owner * w.S;
    const bool mine = (owner == w.rank);
    const bool e_in_basis0 = (w.basis0[e >> 5] >> (e & 31)) & 1u;
    unsigned ep = 0;
    unsigned long long *fcll = nullptr;
    if (G > 1) {
        ep = ++ctl.epoch;
        fcll = w.xbuf + xb_fc(G);
    }
    const int ilast = devex ? n + 1 : n;
    for (int i = ctl.bid * kSgThreads + threadIdx.x; i <= ilast; i += ctl.nblk * kSgThreads) {
        double f = 0.0;
        if (mine) {
            if (i <= n) {
                f = w.T[(size_t)i * w.S + pl];
                if (fabs(f) <= tol) f = 0.0;  // substitute() skips a zero coefficient (:203-205)
            } else {
                f = w.wts[pl];
            }
        }
        if (G > 1) {
            if (mine) {  // broadcast: one LL packet per element into every rank's buffer
                const unsigned long long bits = (unsigned long long)__double_as_longlong(f);
                for (int g = 0; g < G; ++g) ll_store(w.peers[g] + xb_fc(G) + 2 * i, bits, ep);
            }
            f = __longlong_as_double((long long)ll_load(fcll + 2 * i, ep, timeout_ns));
        }
        w.fc[i] = f;
        if (i < n) {
            const int nc = w.nonecol[i];
            crash |= (nc == kNoneInitial && e < n) || (nc == e) || (w.untouched[i] && e_in_basis0);
            const double coeff = -f;
            if (coeff > tol) {
                const double ratio = snap(__ddiv_rn(w.bcol[i], coeff), tol);
                if (ratio != 0.0 || corrected) {
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
    if (corrected) crash = 0;
    block_argmin(br, key, var, crash);
    write_partial(w, ctl, 1, key, var, crash);
}

// ---- P3: scaled pivot row (make_pivot_constraint :325-345), zero column p, None bookkeeping ----
__device__ __forceinline__ void phase_pivot_row(const SingleWS &w, const LoopCtl &ctl, int p, int r, int e,
                                                double tol)
{
    const int gtid = ctl.bid * kSgThreads + threadIdx.x, gsize = ctl.nblk * kSgThreads;
    const double pinv = pivot_inverse(w.fc[r], tol);
    const double *Tr = w.T + (size_t)r * w.S;
    const int lim = min(w.wloc, ctl.ncols - w.c_lo);
    for (int kl = gtid; kl < w.S; kl += gsize) {
        double v = 0.0;
        if (kl < lim) v = (w.c_lo + kl == p) ? snap(-pinv, tol) : snap(__dmul_rn(pinv, Tr[kl]), tol);
        w.nr[kl] = v;
    }
    if (gtid == 0) w.scal[0] = snap(__dmul_rn(pinv, w.bcol[r]), tol);
    const int owner = p / w.S, pl = p - owner * w.S;
    const bool mine = (owner == w.rank);
    for (int i = gtid; i <= w.n; i += gsize) {
        if (mine && i != r) w.T[(size_t)i * w.S + pl] = 0.0;
        if (i < w.n && (w.fc[i] != 0.0 || i == r)) {
            w.nonecol[i] = e;
            w.untouched[i] = 0;
        }
    }
    fence_proxy_async();  // nr[] / fc[] are read back through the async proxy (TMA) in P4
}

// ---- set bookkeeping (set_replace :369-381), one thread per rank, tables in global memory ----
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
    int tile[2];  // pipelined loop: the tile staged in each buffer (tiles are pulled off a counter)
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

__device__ __forceinline__ void phase_sweep(SweepSmem &sm, const SingleWS &w, LoopCtl &ctl, int r, double tol)
{
    const int tid = threadIdx.x;
    const int n = w.n, S = w.S;
    const int lcols = max(0, min(w.wloc, ctl.ncols - w.c_lo));
    const int nstrips = (lcols + kTileCols - 1) / kTileCols;
    const int nrb = (n + 1 + kTileRows - 1) / kTileRows;
    const int ntiles = nstrips * nrb;
    // constant column (replicated): b_i = snap(b_i + f_i * nr_const); the pivot row's constant is nr_const
    {
        const double nrc = w.scal[0];
        for (int i = ctl.bid * kSgThreads + tid; i <= n; i += ctl.nblk * kSgThreads) {
            if (i == r) {
                w.bcol[i] = nrc;
            } else {
                const double f = w.fc[i];
                if (f != 0.0) w.bcol[i] = axpy_snap(w.bcol[i], f, nrc, tol);
            }
        }
    }
    int tile = ctl.bid;
    if (tile < ntiles && tid == 0) {
        fence_proxy_async();
        sweep_issue(sm, w, 0, tile, nstrips);
    }
    for (int it = 0; tile < ntiles; tile += ctl.nblk, ++it) {
        const int stage = it & 1;
        if (tid == 0 && tile + ctl.nblk < ntiles) sweep_issue(sm, w, stage ^ 1, tile + ctl.nblk, nstrips);
        mbar_wait(&sm.bar[stage], (ctl.parity >> stage) & 1u);
        ctl.parity ^= 1u << stage;
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

#ifdef LPB_TIMING
// per-phase wall time of the pivot loop, accumulated by thread 0 of CTA 0 of each rank (globaltimer, ns):
// [0] P1 entering scan  [1] grid.sync  [2] reduce + key exchange  [3] P2 column gather / broadcast / ratio
// [4] grid.sync  [5] reduce  [6] P3 pivot row  [7] grid.sync  [8] bookkeeping + sweep  [9] grid.sync  [10] pivots
__device__ unsigned long long g_single_timing[32][12];
#define LPB_ST(slot)                                                             \
    do {                                                                         \
        if (ctl.bid == 0 && threadIdx.x == 0) {                                  \
            const unsigned long long _t = globaltimer_ns();                      \
            g_single_timing[w.rank][slot] += _t - tlast;                         \
            tlast = _t;                                                          \
        }                                                                        \
    } while (0)
#else
#define LPB_ST(slot) do { } while (0)
#endif

// P3 + P4 of one pivot: e sits at column position p, l owns row r.
__device__ __forceinline__ void single_pivot(cg::grid_group &grid, SweepSmem &sm, const SingleWS &w,
                                             const SingleArgs &a, LoopCtl &ctl, int p, int r, int e, int l,
                                             unsigned long long *tl = nullptr)
{
#ifdef LPB_TIMING
    unsigned long long tlast = tl ? *tl : globaltimer_ns();
#else
    (void)tl;
#endif
    if (a.rule == LPB200_RULE_DEVEX) {  // fc[n+1] carries the entering column's weight (phase_ratio)
        ctl.wq = w.fc[w.n + 1];
        ctl.p_prev = p;
        ctl.upd = true;
    }
    phase_pivot_row(w, ctl, p, r, e, a.tol);
    LPB_ST(6);
    grid.sync();
    LPB_ST(7);
    if (ctl.bid == ctl.nblk - 1 && threadIdx.x == kSgThreads - 32) {
        if (w.pivot_log != nullptr && ctl.npiv < a.log_cap) {
            w.pivot_log[2 * ctl.npiv] = e;
            w.pivot_log[2 * ctl.npiv + 1] = l;
        }
        single_bookkeeping(w, ctl.track_dep, p, r, e, l);
    }
    phase_sweep(sm, w, ctl, r, a.tol);
    LPB_ST(8);
    grid.sync();
    LPB_ST(9);
    ++ctl.npiv;
}

// SlackForm.solve loop (:442-466); every CTA of every rank takes the same branch.
__device__ __forceinline__ int single_run_loop(cg::grid_group &grid, BlockRed &br, SweepSmem &sm,
                                               const SingleWS &w, const SingleArgs &a, LoopCtl &ctl)
{
    const bool corrected = (a.rule != LPB200_RULE_REFERENCE);
    ctl.streak = 0;
    for (;;) {
        unsigned long long key;
        int var, flag;
        // Dantzig prices by coefficient until more than kDegenerateStreak pivots in a row made no progress
        const int pricing = (a.rule >= LPB200_RULE_DANTZIG && ctl.streak > kDegenerateStreak) ? LPB200_RULE_BLAND : a.rule;
        const bool with_id = pricing >= LPB200_RULE_DANTZIG;  // (value key, variable id) instead of (priority, position)
#ifdef LPB_TIMING
        unsigned long long tlast = globaltimer_ns();
#endif
        phase_entering(br, w, ctl, a.tol, pricing, a.rule);
        ctl.upd = false;  // the Devex weights are up to date with the last pivot
        LPB_ST(0);
        grid.sync();
        LPB_ST(1);
        reduce_partials(br, w, ctl, 0, key, var, flag);
        key = exchange_key(br, w, ctl, key, var, with_id, a.timeout_ns);
        LPB_ST(2);
        if (key == kKeyNone) return 0;
        const int p = with_id ? w.varpos[var] : (int)(key & 0xffffffffull);
        const int e = w.colvar[p];
        phase_ratio(br, w, ctl, p, e, a.tol, corrected, a.rule == LPB200_RULE_DEVEX, a.timeout_ns);
        LPB_ST(3);
        grid.sync();
        LPB_ST(4);
        reduce_partials(br, w, ctl, 1, key, var, flag);
        if (flag) return LPB200_STATUS_REF_TYPE_ERROR;
        if (key == kKeyNone) return LPB200_STATUS_INFINITY;
        ctl.streak = (key == kKeyZeroRatio) ? ctl.streak + 1 : 0;
        if (ctl.npiv >= a.max_pivots) return LPB200_STATUS_ITERATION_LIMIT;
        const int l = var;
        const int r = ~w.varpos[l];
        LPB_ST(5);
#ifdef LPB_TIMING
        single_pivot(grid, sm, w, a, ctl, p, r, e, l, &tlast);
        if (ctl.bid == 0 && threadIdx.x == 0) g_single_timing[w.rank][10] += 1;
#else
        single_pivot(grid, sm, w, a, ctl, p, r, e, l);
#endif
    }
}

// Corrected modes: x_aux is still basic (at level 0) after phase 1 -> the non-basic variable with the
// lowest id whose entry in x_aux's row r is non-zero; kKeyNone when the row is all zero (a redundant
// constraint, left inert).  Grid-wide argmin + one key exchange, like P1.
__device__ __forceinline__ unsigned long long driveout_key(cg::grid_group &grid, BlockRed &br, const SingleWS &w,
                                                           const SingleArgs &a, LoopCtl &ctl, int r)
{
    unsigned long long key = kKeyNone;
    int var = 0x7fffffff, flag = 0;
    const double *row = w.T + (size_t)r * w.S;
    const int lim = min(w.wloc, ctl.ncols - w.c_lo);
    for (int kl = ctl.bid * kSgThreads + threadIdx.x; kl < lim; kl += ctl.nblk * kSgThreads) {
        if (fabs(row[kl]) > a.tol) {
            const int k = w.c_lo + kl;
            const unsigned long long kk = ((unsigned long long)(unsigned)w.colvar[k] << 32) | (unsigned)k;
            key = kk < key ? kk : key;
        }
    }
    block_argmin(br, key, var, flag);
    write_partial(w, ctl, 0, key, var, flag);
    grid.sync();
    reduce_partials(br, w, ctl, 0, key, var, flag);
    return exchange_key(br, w, ctl, key, var, false, a.timeout_ns);
}

// ======================================================================================
// Pipelined pivot loop (LPB200_RULE_REFERENCE / LPB200_RULE_BLAND, two or more CTAs per rank): ONE grid barrier
// per pivot instead of four.
//
// The selection of the next pivot needs very little of the tableau the current pivot produces: the objective row
// (entering column, simplex.py:390-392), one column of it (the ratio test, :394-436) and the constant column.  All of
// it can be computed from the OLD tableau, the pivot column fc[] and the old pivot row -- provided the old values are
// still there.  So the sweep is not done in place: pivot k reads tableau buffer A and writes buffer B, pivot k+1
// reads B and writes A (same HBM traffic: every element is read once and written once either way; 2 x 4.3 GB at
// 16384 x 32768 is small change next to 180 GB).  While the other CTAs of a rank sweep A -> B, its first few CTAs --
// the control group (8 by default, LPB_DEBUG_SINGLE_NCTL) -- perform pivot k's set bookkeeping, update the constant column and the objective row, pick the
// next entering column (key exchange with the other ranks' control CTAs), compute that column of B from A (the owner)
// and broadcast it, run the ratio test -- and then join the sweep (tiles are pulled off a counter): by the time the
// sweep ends, pivot k+1 is decided.  The fixed cost of a
// pivot (four grid barriers, three selection phases, two NVLink exchanges: 20 us on one GPU, 23-27 us on eight, measured
// with tools/gpu/single_phases.py) overlaps the sweep; what is left on the critical path is one barrier.
// Every value is produced by the same operations on the same operands as in the step-by-step loop: same bits.
// ======================================================================================

// Sweep of pivot (entering position p, leaving row r) from tableau A into tableau B by the sweeper CTAs (bid >= 1).
// Rows 0 .. n-1 only: the objective row belongs to the control CTA.  The tile's slice of the OLD pivot row (A, row r)
// and of the pivot column are staged by TMA; every thread scales its double2 of the row into the new pivot row
// (make_pivot_constraint :325-345) in registers.  Unlike the in-place sweep every row is written (a row with a zero
// factor is copied) and the old pivot row is never overwritten, so no CTA ever reads what another one writes.
__device__ __forceinline__ void sweep_issue_pp(SweepSmem &sm, const SingleWS &w, const double *A, const double *fcA,
                                               int r, int stage, int tile, int nstrips, int trows)
{
    const int strip = tile % nstrips, rb = tile / nstrips;
    const int c0 = strip * kTileCols, r0 = rb * trows;
    const int cols = min(kTileCols, w.S - c0);                 // even (S is a multiple of 16)
    const int rows = (min(trows, w.n - r0) + 1) & ~1;          // fc[] is padded: reading one element past is harmless
    const uint32_t bnr = (uint32_t)cols * 8u, bfc = (uint32_t)rows * 8u;
    mbar_expect_tx(&sm.bar[stage], bnr + bfc);
    tma_load_1d(sm.nr[stage], A + (size_t)r * w.S + c0, bnr, &sm.bar[stage]);
    tma_load_1d(sm.fc[stage], fcA + r0, bfc, &sm.bar[stage]);
}

__device__ __forceinline__ void phase_sweep_pp(SweepSmem &sm, const SingleWS &w, LoopCtl &ctl, const double *A,
                                               double *B, const double *fcA, int p, int r, double tol)
{
    const int tid = threadIdx.x;
    const int n = w.n, S = w.S;
    const int lcols = max(0, min(w.wloc, ctl.ncols - w.c_lo));
    const int nstrips = (lcols + kTileCols - 1) / kTileCols;
    // tile height: 64 rows when every CTA gets many tiles (one GPU: 55 each), 32 when it gets few (eight GPUs: 7
    // each) -- the end of a sweep idles CTAs for up to one tile.  Measured, 16384 x 32768: one GPU 1.315 ms per pivot
    // with 64 rows against 1.323 with 32; eight GPUs 0.1829 ms against 0.1752 (tools/gpu/pp_tiles.sh)
    const int trows = (nstrips * ((n + kTileRows - 1) / kTileRows) < 16 * ctl.nblk) ? kTileRows / 2 : kTileRows;
    const int nrb = (n + trows - 1) / trows;
    const int ntiles = nstrips * nrb;
    const double pinv = pivot_inverse(fcA[r], tol);
    const double dpiv = snap(-pinv, tol);
    const int pl = p - w.c_lo;  // local position of the entering column; outside [0, S) on the ranks that do not own it
    // Tiles are pulled off a counter (one per pivot parity; the control CTAs zero the other one meanwhile): the
    // control CTAs join the sweep when their look-ahead is done, and nobody waits for a slow CTA at the end.  Thread 0
    // claims two tiles ahead, so the latency of the atomic is never waited for.
    unsigned *ctr = w.labar + 1 + (ctl.npiv & 1);
    int claimed = 0;  // (thread 0) the tile claimed for the iteration after next
    if (tid == 0) {
        const int t0 = (int)atomicAdd(ctr, 1u);
        sm.tile[0] = t0;
        if (t0 < ntiles) {
            fence_proxy_async();
            sweep_issue_pp(sm, w, A, fcA, r, 0, t0, nstrips, trows);
        }
        claimed = (int)atomicAdd(ctr, 1u);
    }
    __syncthreads();
    int tile = sm.tile[0];
    for (int it = 0; tile < ntiles; ++it) {
        const int stage = it & 1;
        if (tid == 0) {
            sm.tile[stage ^ 1] = claimed;
            if (claimed < ntiles) sweep_issue_pp(sm, w, A, fcA, r, stage ^ 1, claimed, nstrips, trows);
            claimed = (claimed < ntiles) ? (int)atomicAdd(ctr, 1u) : ntiles;
        }
        mbar_wait(&sm.bar[stage], (ctl.parity >> stage) & 1u);
        ctl.parity ^= 1u << stage;
        const int strip = tile % nstrips, rb = tile / nstrips;
        const int c0 = strip * kTileCols, r0 = rb * trows;
        const int rows = min(trows, n - r0);
        const int col = c0 + 2 * tid;
        if (col < S) {
            const double2 raw = reinterpret_cast<const double2 *>(sm.nr[stage])[tid];
            const bool zx = (col == pl), zy = (col + 1 == pl);  // column p now belongs to the leaving variable (None -> 0)
            double2 r2;
            r2.x = (col < lcols) ? (zx ? dpiv : snap(__dmul_rn(pinv, raw.x), tol)) : 0.0;
            r2.y = (col + 1 < lcols) ? (zy ? dpiv : snap(__dmul_rn(pinv, raw.y), tol)) : 0.0;
            const double2 *abase = reinterpret_cast<const double2 *>(A + (size_t)r0 * S + col);
            double2 *bbase = reinterpret_cast<double2 *>(B + (size_t)r0 * S + col);
            const size_t rs = (size_t)S / 2;  // row stride in double2
            for (int ii = 0; ii < rows; ii += kRowUnroll) {
                double2 v[kRowUnroll];
                double f[kRowUnroll];
#pragma unroll
                for (int j = 0; j < kRowUnroll; ++j) {
                    const int q = ii + j;
                    f[j] = (q < rows) ? sm.fc[stage][q] : 0.0;
                    if (q < rows && r0 + q != r) v[j] = __ldcs(abase + (size_t)q * rs);
                }
#pragma unroll
                for (int j = 0; j < kRowUnroll; ++j) {
                    const int q = ii + j;
                    if (q < rows) {
                        double2 o;
                        if (r0 + q == r) {
                            o = r2;  // store(entering_idx, row) (:352)
                        } else {
                            o = v[j];
                            if (zx) o.x = 0.0;
                            if (zy) o.y = 0.0;
                            if (f[j] != 0.0) {  // substitute() (:185-214); a zero factor returns the row unchanged
                                o.x = axpy_snap(o.x, f[j], r2.x, tol);
                                o.y = axpy_snap(o.y, f[j], r2.y, tol);
                            }
                        }
                        __stcs(bbase + (size_t)q * rs, o);
                    }
                }
            }
        }
        __syncthreads();  // everyone is done with sm.*[stage] before it is refilled; publishes sm.tile[stage ^ 1]
        tile = sm.tile[stage ^ 1];
    }
    fence_proxy_async();  // B's rows are read through the async proxy (TMA) by the next pivot's sweep
}

#ifdef LPB_TIMING
// inside the look-ahead (pipelined mode; slots 0..5 of g_single_timing): [0] bookkeeping + None records + constant column
// [1] objective row + keys  [2] key exchange  [3] column from A (owner)  [4] broadcast (owner) / receive  [5] ratio test
#define LPB_LA_T(slot)                                                \
    do {                                                              \
        __syncthreads();                                              \
        if (threadIdx.x == 0 && ctl.bid == 0) {                       \
            const unsigned long long _t = globaltimer_ns();           \
            g_single_timing[w.rank][slot] += _t - la_t;               \
            la_t = _t;                                                \
        }                                                             \
    } while (0)
#else
#define LPB_LA_T(slot) do { } while (0)
#endif

// Barrier among the control CTAs of one rank (bid < ctl.nctl): a counter in global memory that only ever grows; the
// k-th barrier since it was zeroed (kernel set-up) completes when it reaches k * nctl.  All control CTAs are resident
// (cooperative launch), so the spin cannot deadlock.
__device__ __forceinline__ void control_barrier(const SingleWS &w, LoopCtl &ctl)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned target = (++ctl.labar_gen) * (unsigned)ctl.nctl;
        __threadfence();
        atomicAdd(w.labar, 1u);
        while (*reinterpret_cast<volatile unsigned *>(w.labar) < target) {
        }
        __threadfence();
    } else {
        ++ctl.labar_gen;  // (every thread keeps the count; only thread 0 uses it)
    }
    __syncthreads();
}

// lexicographic (key, var) minimum + OR of the flags over the control CTAs' partials of `set`
__device__ __forceinline__ void reduce_control(BlockRed &br, const SingleWS &w, const LoopCtl &ctl, int set,
                                               unsigned long long &key, int &var, int &flag)
{
    key = kKeyNone;
    var = 0x7fffffff;
    flag = 0;
    const int lo = set * ctl.nblk;
    for (int q = lo + threadIdx.x; q < lo + ctl.nctl; q += kSgThreads) {
        const unsigned long long k = __ldcg(w.part_key + q);
        const int v = __ldcg(w.part_var + q);
        if (k < key || (k == key && v < var)) {
            key = k;
            var = v;
        }
        flag |= __ldcg(w.part_crash + q);
    }
    block_argmin(br, key, var, flag);
}

// Look-ahead of the CONTROL CTAs (bid < ctl.nctl) while the sweepers apply pivot (p, r, e, l), A -> B.  Leaves the outcome
// of the NEXT pivot's selection in desc[] (LA_* code, entering position, leaving variable, its row) and that pivot's
// column in fcB[].  In the order of the step-by-step loop: the None records of this pivot (P3), its set bookkeeping
// (:369-381), the constant column, the objective row (:314-318) with the entering keys (:390-392), the key exchange,
// the entering column of B computed from A, its broadcast, the ratio test (:394-436).
// While the sweepers saturate HBM a dependent memory access costs several microseconds, so the work is spread over
// nctl x 256 threads with kLaBatch independent accesses in flight each: every phase is one or two memory rounds, and
// three control barriers separate them.  (One control CTA, one access per iteration: 317 us on eight GPUs against a
// 166 us sweep; measured with tools/gpu/single_phases.py.)
__device__ __noinline__ void lookahead(BlockRed &br, const SingleWS &w, const SingleArgs &a, LoopCtl &ctl,
                                       const double *A, double *B, const double *fcA, double *fcB, int *desc, int p,
                                       int r, int e, int l)
{
    constexpr int kLaBatch = 4;  // independent memory accesses in flight per thread
    const int tid = threadIdx.x;
    const int n = w.n, S = w.S, G = w.nranks;
    const int gt = ctl.bid * kSgThreads + tid, gs = ctl.nctl * kSgThreads;  // this thread among the control threads
    const double tol = a.tol;
    const bool corrected = (a.rule != LPB200_RULE_REFERENCE);
    const double pinv = pivot_inverse(fcA[r], tol);
    const double dpiv = snap(-pinv, tol);
    const double nrc = snap(__dmul_rn(pinv, w.bcol[r]), tol);  // the pivot row's constant (:331-336); bcol[r] itself is
                                                               // rewritten only after the first control barrier
#ifdef LPB_TIMING
    unsigned long long la_t = globaltimer_ns();
#endif
    if (gt == 0) w.labar[1 + ((ctl.npiv + 1) & 1)] = 0u;  // the NEXT pivot's tile counter (idle since the last barrier)
    // ---- phase 1: set bookkeeping (one thread), None records and constant column (everyone else) ----
    constexpr int kBookThread = kSgThreads - 32;
    if (gt == kBookThread) {
        if (w.pivot_log != nullptr && ctl.npiv < a.log_cap) {
            w.pivot_log[2 * ctl.npiv] = e;
            w.pivot_log[2 * ctl.npiv + 1] = l;
        }
        single_bookkeeping(w, ctl.track_dep, p, r, e, l);
    } else {
        const int wid = gt < kBookThread ? gt : gt - 1, nw = gs - 1;
        for (int i0 = wid; i0 <= n; i0 += nw * kLaBatch) {
            double f[kLaBatch], bv[kLaBatch];
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int i = i0 + j * nw;
                if (i <= n) {
                    f[j] = fcA[i];
                    bv[j] = w.bcol[i];
                }
            }
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int i = i0 + j * nw;
                if (i <= n) {
                    if (i < n && (f[j] != 0.0 || i == r)) {  // the row is rewritten by this pivot: its only None is column e
                        w.nonecol[i] = e;
                        w.untouched[i] = 0;
                    }
                    if (i != r && f[j] != 0.0) w.bcol[i] = axpy_snap(bv[j], f[j], nrc, tol);
                }
            }
        }
    }
    control_barrier(w, ctl);  // bookkeeping done (slots / ids of every column are final); everyone has read bcol[r]
    if (gt == 0) w.bcol[r] = nrc;
    LPB_LA_T(0);
    // ---- phase 2: objective row of B, and the entering key of every owned column ----
    unsigned long long key = kKeyNone;
    int var = 0x7fffffff, flag = 0;
    {
        const double fobj = fcA[n];
        const int lim = min(w.wloc, ctl.ncols - w.c_lo);
        const int pl = p - w.c_lo;
        const int *prio = corrected ? w.colvar : w.colslot;  // Bland: variable id; reference: CPython-set slot
        const double *Ar = A + (size_t)r * S, *An = A + (size_t)n * S;
        double *Bn = B + (size_t)n * S;
        for (int k0 = gt; k0 < S; k0 += gs * kLaBatch) {
            double rr[kLaBatch], oo[kLaBatch];
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int kl = k0 + j * gs;
                if (kl < S) {
                    rr[j] = __ldcg(Ar + kl);
                    oo[j] = __ldcg(An + kl);
                }
            }
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int kl = k0 + j * gs;
                if (kl < S) {
                    double r2 = 0.0;
                    if (kl < lim) r2 = (kl == pl) ? dpiv : snap(__dmul_rn(pinv, rr[j]), tol);
                    double o = (kl == pl) ? 0.0 : oo[j];
                    if (fobj != 0.0) o = axpy_snap(o, fobj, r2, tol);
                    Bn[kl] = o;
                    if (kl < lim && o > tol) {
                        const int k = w.c_lo + kl;
                        const unsigned long long kk = ((unsigned long long)(unsigned)__ldcg(prio + k) << 32) | (unsigned)k;
                        key = kk < key ? kk : key;
                    }
                }
            }
        }
    }
    block_argmin(br, key, var, flag);
    write_partial(w, ctl, 0, key, var, flag);
    control_barrier(w, ctl);
    reduce_control(br, w, ctl, 0, key, var, flag);
    LPB_LA_T(1);
    key = exchange_key(br, w, ctl, key, var, false, a.timeout_ns);
    LPB_LA_T(2);
    if (key == kKeyNone) {
        if (gt == 0) desc[0] = LA_OPTIMAL;
        return;
    }
    const int p2 = (int)(key & 0xffffffffull);
    const int e2 = __ldcg(w.colvar + p2);
    // ---- phase 3: column p2 of B from A (its owner), broadcast, ratio test on the updated constant column ----
    const int owner = p2 / S, pl2 = p2 - owner * S;
    const bool mine = (owner == w.rank);
    double nr_p2 = 0.0;
    if (mine) nr_p2 = (p2 == p) ? dpiv : snap(__dmul_rn(pinv, __ldcg(A + (size_t)r * S + pl2)), tol);
    const bool e_in_basis0 = (w.basis0[e2 >> 5] >> (e2 & 31)) & 1u;
    unsigned ep = 0;
    unsigned long long *fcll = nullptr;
    if (G > 1) {
        ep = ++ctl.epoch;
        fcll = w.xbuf + xb_fc(G);
    }
    key = kKeyNone;
    var = 0x7fffffff;
    int crash = 0;
    const bool same = (p2 == p);  // the column that changed hands: its old entries count as zero (None)
    for (int i0 = gt; i0 <= n; i0 += gs * kLaBatch) {
        double f[kLaBatch];
        if (mine) {  // pass A (owner): the column, computed from A
            double a0[kLaBatch], fi[kLaBatch];
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int i = i0 + j * gs;
                if (i <= n) {
                    a0[j] = (same || i == r) ? 0.0 : __ldcg(A + (size_t)i * S + pl2);
                    fi[j] = fcA[i];
                }
            }
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int i = i0 + j * gs;
                if (i <= n) {
                    double v;
                    if (i == r) v = nr_p2;
                    else v = (fi[j] != 0.0) ? axpy_snap(a0[j], fi[j], nr_p2, tol) : a0[j];
                    f[j] = (fabs(v) <= tol) ? 0.0 : v;  // substitute() skips a zero coefficient (:203-205)
                }
            }
            if (G > 1) {  // pass B (owner): one LL packet per element into every OTHER rank's buffer
#pragma unroll
                for (int j = 0; j < kLaBatch; ++j) {
                    const int i = i0 + j * gs;
                    if (i <= n) {
                        const unsigned long long bits = (unsigned long long)__double_as_longlong(f[j]);
                        for (int g = 0; g < G; ++g)
                            if (g != w.rank) ll_store(w.peers[g] + xb_fc(G) + 2 * i, bits, ep);
                    }
                }
            }
        } else {  // pass C (the others): kLaBatch packets polled at a time
            unsigned long long w0[kLaBatch], w1[kLaBatch];
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int i = i0 + j * gs;
                if (i <= n)
                    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];"
                                 : "=l"(w0[j]), "=l"(w1[j])
                                 : "l"(fcll + 2 * i)
                                 : "memory");
            }
#pragma unroll
            for (int j = 0; j < kLaBatch; ++j) {
                const int i = i0 + j * gs;
                if (i <= n) {
                    unsigned long long bits;
                    if ((unsigned)(w0[j] >> 32) == ep && (unsigned)(w1[j] >> 32) == ep)
                        bits = (w0[j] & 0xffffffffull) | (w1[j] << 32);
                    else
                        bits = ll_load(fcll + 2 * i, ep, a.timeout_ns);  // not there yet: wait for this one
                    f[j] = __longlong_as_double((long long)bits);
                }
            }
        }
        // pass D: the ratio test (every rank, redundantly, on the replicated constant column)
        double bv[kLaBatch];
        int nc[kLaBatch], ut[kLaBatch], rv[kLaBatch];
#pragma unroll
        for (int j = 0; j < kLaBatch; ++j) {
            const int i = i0 + j * gs;
            if (i <= n) fcB[i] = f[j];
            if (i < n) {
                bv[j] = __ldcg(w.bcol + i);
                nc[j] = __ldcg(w.nonecol + i);
                ut[j] = __ldcg(w.untouched + i);
                rv[j] = __ldcg(w.rowvar + i);
            }
        }
#pragma unroll
        for (int j = 0; j < kLaBatch; ++j) {
            const int i = i0 + j * gs;
            if (i < n) {
                crash |= (nc[j] == kNoneInitial && e2 < n) || (nc[j] == e2) || (ut[j] && e_in_basis0);
                const double coeff = -f[j];
                if (coeff > tol) {
                    const double ratio = snap(__ddiv_rn(bv[j], coeff), tol);
                    if (ratio != 0.0 || corrected) {
                        const unsigned long long k = order_key(ratio);
                        if (k < key || (k == key && rv[j] < var)) {
                            key = k;
                            var = rv[j];
                        }
                    }
                }
            }
        }
    }
    LPB_LA_T(3);
    if (corrected) crash = 0;
    block_argmin(br, key, var, crash);
    write_partial(w, ctl, 1, key, var, crash);
    control_barrier(w, ctl);
    reduce_control(br, w, ctl, 1, key, var, crash);
    LPB_LA_T(5);
    if (gt == 0) {
        if (crash) {
            desc[0] = LA_TYPE_ERROR;
        } else if (key == kKeyNone) {
            desc[0] = LA_INFINITY;
        } else {
            desc[0] = LA_PIVOT;
            desc[1] = p2;
            desc[2] = var;
            desc[3] = ~__ldcg(w.varpos + var);
        }
    }
    fence_proxy_async();  // fcB is read through the async proxy (TMA) by the next sweep
}

// SlackForm.solve loop (:442-466), pipelined.  Same decisions, same arithmetic, same exits as single_run_loop.
__device__ __forceinline__ int single_run_loop_pipelined(cg::grid_group &grid, BlockRed &br, SweepSmem &sm,
                                                         SingleWS &w, const SingleArgs &a, LoopCtl &ctl)
{
    const bool corrected = (a.rule != LPB200_RULE_REFERENCE);
    ctl.streak = 0;
    unsigned long long key;
    int var, flag;
    if (ctl.bid == 0 && threadIdx.x == 0) w.labar[1] = w.labar[2] = 0u;  // tile counters (two grid barriers follow)
    // the first pivot of the loop is selected step by step (there is no sweep to hide the selection behind)
    phase_entering(br, w, ctl, a.tol, a.rule, a.rule);
    grid.sync();
    reduce_partials(br, w, ctl, 0, key, var, flag);
    key = exchange_key(br, w, ctl, key, var, false, a.timeout_ns);
    if (key == kKeyNone) return 0;
    int p = (int)(key & 0xffffffffull);
    int e = w.colvar[p];
    phase_ratio(br, w, ctl, p, e, a.tol, corrected, false, a.timeout_ns);
    grid.sync();
    reduce_partials(br, w, ctl, 1, key, var, flag);
    if (flag) return LPB200_STATUS_REF_TYPE_ERROR;
    if (key == kKeyNone) return LPB200_STATUS_INFINITY;
    int l = var;
    int r = ~w.varpos[l];
    double *A = w.T, *B = w.T2, *fcA = w.fc, *fcB = w.fc2;
    int flips = 0, status = 0;
    for (;;) {
        if (ctl.npiv >= a.max_pivots) {
            status = LPB200_STATUS_ITERATION_LIMIT;
            break;
        }
#ifdef LPB_TIMING
        unsigned long long tlast = globaltimer_ns();
#endif
        int *desc = w.state + ST_LA_STATUS + 4 * (ctl.npiv & 1);  // double-buffered by pivot parity
        if (ctl.bid < ctl.nctl) lookahead(br, w, a, ctl, A, B, fcA, fcB, desc, p, r, e, l);
        LPB_ST(8);
        phase_sweep_pp(sm, w, ctl, A, B, fcA, p, r, a.tol);  // the control CTAs join once the next pivot is decided
        LPB_ST(7);
        grid.sync();
        LPB_ST(9);
#ifdef LPB_TIMING
        if (ctl.bid == 0 && threadIdx.x == 0) g_single_timing[w.rank][10] += 1;
#endif
        ++ctl.npiv;
        {
            double *t = A; A = B; B = t;
            t = fcA; fcA = fcB; fcB = t;
            flips ^= 1;
        }
        const int la = desc[0];
        if (ctl.bid >= ctl.nctl && w.nranks > 1) {  // the exchanges the control CTAs performed: keep the counters uniform
            ctl.epoch += (la == LA_OPTIMAL) ? 1u : 2u;
            ctl.kx += 1u;
        }
        if (la == LA_OPTIMAL) break;
        if (la == LA_INFINITY) { status = LPB200_STATUS_INFINITY; break; }
        if (la == LA_TYPE_ERROR) { status = LPB200_STATUS_REF_TYPE_ERROR; break; }
        p = desc[1];
        l = desc[2];
        r = desc[3];
        e = w.colvar[p];  // only the control CTA uses it, and it reads it before its own bookkeeping rewrites it
    }
    if (flips) {  // the tableau the loop stopped on is what was T2 when it started: make it the kernel's T
        __syncthreads();
        if (threadIdx.x == 0) {
            double *t = w.T; w.T = w.T2; w.T2 = t;
            t = w.fc; w.fc = w.fc2; w.fc2 = t;
        }
        ctl.tcur ^= 1;
        __syncthreads();
    }
    return status;
}

__device__ __forceinline__ int run_loop_auto(cg::grid_group &grid, BlockRed &br, SweepSmem &sm, SingleWS &w,
                                             const SingleArgs &a, LoopCtl &ctl)
{
    if (a.pipelined && a.rule < LPB200_RULE_DANTZIG && ctl.nblk >= 2)
        return single_run_loop_pipelined(grid, br, sm, w, a, ctl);
    return single_run_loop(grid, br, sm, w, a, ctl);
}

__global__ void __launch_bounds__(kSgThreads, 2)
simplex_single_kernel(const SingleWS *__restrict__ ranks, const SingleArgs a)
{
    cg::grid_group grid = cg::this_grid();
    __shared__ SingleWS w;
    __shared__ BlockRed br;
    __shared__ __align__(128) SweepSmem sm;
    const int tid = threadIdx.x;
    LoopCtl ctl;
    {
        const int grp = blockIdx.x / a.cpr;
        ctl.bid = blockIdx.x - grp * a.cpr;
        ctl.nblk = a.cpr;
        const unsigned *src = reinterpret_cast<const unsigned *>(ranks + grp);
        unsigned *dst = reinterpret_cast<unsigned *>(&w);
        for (int i = tid; i < (int)(sizeof(SingleWS) / 4); i += kSgThreads) dst[i] = src[i];
    }
    if (tid == 0) {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int gtid = ctl.bid * kSgThreads + tid, gsize = ctl.nblk * kSgThreads;
    const int n = w.n, m = w.m, S = w.S, nv = n + m + 1;
    const double tol = a.tol;
    const bool corrected = (a.rule != LPB200_RULE_REFERENCE);
    const bool devex = (a.rule == LPB200_RULE_DEVEX);
    ctl.upd = false;
    ctl.p_prev = 0;
    ctl.wq = 1.0;
    ctl.epoch = (w.nranks > 1) ? (unsigned)w.xbuf[kXbEpoch] : 0u;
    ctl.kx = (w.nranks > 1) ? (unsigned)w.xbuf[kXbKx] : 0u;
    ctl.streak = 0;
    ctl.tcur = 0;
    ctl.nctl = max(1, min(a.nctl_max, ctl.nblk / 4));
    ctl.labar_gen = 0u;
    if (gtid == 0) *w.labar = 0u;  // (ordered before any control barrier by the grid.sync()s of the set-up)

    // ---- min_constant, idx = min_index(b) (:591): zeros count as +inf, first minimum wins ----
    int idx0;
    {
        unsigned long long key = kKeyNone;
        int var = 0x7fffffff, flag = 0;
        for (int i = gtid; i < n; i += gsize) {
            const double bi = a.b[i];
            double kb = bi;  // the reference counts zeros as +inf (:161); the corrected modes take the true minimum
            if (bi == 0.0) kb = corrected ? 0.0 : CUDART_INF;
            const unsigned long long k = order_key(kb);
            if (k < key || (k == key && i < var)) {
                key = k;
                var = i;
            }
        }
        block_argmin(br, key, var, flag);
        write_partial(w, ctl, 1, key, var, flag);
        grid.sync();
        reduce_partials(br, w, ctl, 1, key, var, flag);
        idx0 = var;
    }
    const double min_c = a.b[idx0];
    const bool phase1 = (min_c < 0.0);
    int status = 0;
    if (!phase1 && !(min_c > tol) && !corrected) status = LPB200_STATUS_REF_ATTRIBUTE_ERROR;  // :626
    ctl.ncols = phase1 ? m + 1 : m;
    ctl.npiv = 0;
    ctl.track_dep = phase1;
    ctl.parity = 0u;

    if (status == 0) {
        // ---- SlackForm.__init__ (:230-302) (+ the x_aux column of :632-660 in phase 1) ----
        for (int i = ctl.bid; i <= n; i += ctl.nblk) {
            double *Ti = w.T + (size_t)i * S;
            if (i < n) {
                const double *Ai = w.A + (size_t)i * w.lda;
                for (int kl = tid; kl < S; kl += kSgThreads) {
                    const int k = w.c_lo + kl;
                    double v = 0.0;
                    if (kl < w.wloc) {
                        if (k < m) v = -Ai[kl];
                        else if (k == m && phase1) v = 1.0;  // -(-1), :659 then :262
                    }
                    Ti[kl] = v;
                }
            } else {
                for (int kl = tid; kl < S; kl += kSgThreads) {
                    const int k = w.c_lo + kl;
                    double v = 0.0;
                    if (kl < w.wloc) {
                        if (k < m) v = phase1 ? 0.0 : a.c[k];
                        else if (k == m && phase1) v = -1.0;  // objective -x_aux (:644-645)
                    }
                    Ti[kl] = v;
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
        if (devex)
            for (int kl = gtid; kl < S; kl += gsize) w.wts[kl] = 1.0;  // reference framework of the first pivot loop
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
                phase_ratio(br, w, ctl, m, aux, tol, corrected, devex, a.timeout_ns);  // column m -> fc[]; its partial is unused
                grid.sync();
                single_pivot(grid, sm, w, a, ctl, m, idx0, aux, idx0);
                status = run_loop_auto(grid, br, sm, w, a, ctl);  // :526
            }
            if (status == 0) {
                if (!(fabs(w.bcol[n]) <= tol)) status = LPB200_STATUS_NO_SOLUTION;       // :528-530
                else if (w.varpos[aux] < 0) {
                    if (!corrected) {
                        status = LPB200_STATUS_REF_INDEX_ERROR;                          // :541 -> :183
                    } else {
                        // x_aux is basic at level 0: pivot it out on the lowest-id column with a non-zero
                        // entry in its row, then let the loop run (the objective is -x_aux again: it ends at once)
                        const int r = ~w.varpos[aux];
                        if (devex) {  // a new reference framework ahead of the forced pivot (grid.sync inside driveout_key)
                            for (int kl = gtid; kl < S; kl += gsize) w.wts[kl] = 1.0;
                            ctl.upd = false;
                        }
                        const unsigned long long dk = driveout_key(grid, br, w, a, ctl, r);
                        if (dk != kKeyNone) {
                            if (ctl.npiv >= a.max_pivots) {
                                status = LPB200_STATUS_ITERATION_LIMIT;
                            } else {
                                const int p = (int)(dk & 0xffffffffull);
                                const int ev = w.colvar[p];
                                phase_ratio(br, w, ctl, p, ev, tol, corrected, devex, a.timeout_ns);  // column p -> fc[]
                                grid.sync();
                                single_pivot(grid, sm, w, a, ctl, p, r, ev, aux);
                                status = run_loop_auto(grid, br, sm, w, a, ctl);
                            }
                        }
                    }
                }
            }
            if (status == 0) {
                // ---- rebuild the original objective (:537-541) ----
                // (a) basic ORIGINAL variables with a non-zero objective coefficient, in
                //     iteration order of the `dependent` set: rows -> list[], factors -> cstage[]
                if (ctl.bid == 0 && tid < 32) {
                    const int dsize = w.state[ST_DSIZE];
                    int cnt = 0;
                    for (int s0 = 0; s0 < dsize; s0 += 32) {
                        const int s = s0 + tid;
                        const int v = (s < dsize) ? w.tab_d[s] : -1;
                        double f = 0.0;
                        if (v >= n && v < n + m) f = a.c[v - n];  // an inert basic x_aux (corrected modes) has none
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
                // (b) (v, c) = (0, [None]*n + float(c)); substitute each listed row in order.
                //     x_aux's column is dropped (:563-575) by zeroing it in place: a zero column
                //     never enters and stays zero under the sweep, so no data moves between ranks.
                {
                    const int nlist = w.state[ST_NLIST];
                    const int kaux = w.varpos[aux];
                    double *obj = w.T + (size_t)n * S;
                    const int lim = min(w.wloc, ctl.ncols - w.c_lo);
                    for (int kl = gtid; kl < lim; kl += gsize) {
                        const int k = w.c_lo + kl;
                        double o = 0.0;
                        if (k != kaux) {  // zip() truncates the x_aux column (:190)
                            const int v = w.colvar[k];
                            o = (v >= n && v < n + m) ? a.c[v - n] : 0.0;
                            for (int t = 0; t < nlist; ++t)
                                o = axpy_snap(o, w.cstage[t], w.T[(size_t)w.list[t] * S + kl], tol);
                        }
                        obj[kl] = o;
                    }
                    if (gtid == gsize - 1) {  // the constant term (replicated)
                        double o = 0.0;
                        for (int t = 0; t < nlist; ++t) o = axpy_snap(o, w.cstage[t], w.bcol[w.list[t]], tol);
                        w.bcol[n] = o;
                    }
                    if (kaux >= w.c_lo && kaux < w.c_lo + w.wloc)
                        for (int i = gtid; i < n; i += gsize) w.T[(size_t)i * S + (kaux - w.c_lo)] = 0.0;
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
                    if (devex) {  // phase 2 starts from a new reference framework
                        for (int kl = gtid; kl < S; kl += gsize) w.wts[kl] = 1.0;
                        ctl.upd = false;
                    }
                    if (gtid == 0 && kaux >= 0) {  // independent.remove(nvars) (:574); the same sets live on
                        PySet s = {w.state[ST_ISIZE], w.state[ST_IFILL], w.state[ST_IUSED]};
                        pyset_remove_slot(w.tab_i, s, w.colslot[kaux]);
                        w.state[ST_IUSED] = s.used;
                    }
                }
                grid.sync();
                status = run_loop_auto(grid, br, sm, w, a, ctl);  // :599
            }
        } else {
            status = run_loop_auto(grid, br, sm, w, a, ctl);  // :619
        }
    }

    // ---- results (:473-491, 603, 622); every rank holds the full answer ----
    for (int j = gtid; j < m; j += gsize) {
        double v = 0.0;
        if (status == 0) {
            const int pos = w.varpos[n + j];
            if (pos < 0) v = w.bcol[~pos];
        }
        w.x[j] = v;
    }
    if (gtid == 0) {
        w.status[0] = status;
        w.npivots[0] = ctl.npiv;
        w.objective[0] = (status == 0) ? w.bcol[n] : 0.0;
        if (w.nranks > 1) {
            w.xbuf[kXbEpoch] = ctl.epoch;
            w.xbuf[kXbKx] = ctl.kx;
        }
        w.state[ST_NCOLS] = (status == LPB200_STATUS_REF_ATTRIBUTE_ERROR) ? 0 : ctl.ncols;  // 0: no tableau was built
        w.state[ST_TCUR] = ctl.tcur;
    }
}

// ---- tableau export by variable id (lpb200_simplex_single_state_device) ----
__global__ void simplex_single_state_kernel(const SingleWS *__restrict__ hdr, double *__restrict__ obj_by_var,
                                            double *__restrict__ const_by_var)
{
    const SingleWS &w = *hdr;
    const int n = w.n, nv = w.n + w.m + 1;
    const bool built = w.state[ST_NCOLS] != 0;
    const double *obj = (w.state[ST_TCUR] ? w.T2 : w.T) + (size_t)n * w.S;
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v <= nv; v += gridDim.x * blockDim.x) {
        double o = 0.0, cst = 0.0;
        if (v == nv) {
            o = built ? w.bcol[n] : 0.0;
        } else if (built) {
            const int pos = w.varpos[v];
            if (pos < 0) cst = w.bcol[~pos];
            else if (pos >= w.c_lo && pos < w.c_lo + w.wloc) o = obj[pos - w.c_lo];
        }
        if (obj_by_var) obj_by_var[v] = o;
        if (const_by_var && v < nv) const_by_var[v] = cst;
    }
}

}  // namespace lpb

// ======================================================================================
// Host side: C ABI
// ======================================================================================
using namespace lpb;

static inline size_t up256(size_t v) { return (v + 255) / 256 * 256; }
static const int kMaxSingleGrid = 148 * 8;  // per-rank scratch is sized for any grid up to this

static int cols_per_rank(int m, int nranks) { return (((m + 1) + nranks - 1) / nranks + 15) / 16 * 16; }

// Carves one rank's part of the workspace; returns the bytes needed.  base == nullptr just measures.
static size_t carve_single(SingleWS *w, int n, int m, int rank, int nranks, unsigned char *base)
{
    const int nv = n + m + 1;
    w->n = n;
    w->m = m;
    w->S = cols_per_rank(m, nranks);
    w->rank = rank;
    w->nranks = nranks;
    w->c_lo = rank * w->S;
    int wl = (m + 1) - w->c_lo;
    w->wloc = wl < 0 ? 0 : (wl > w->S ? w->S : wl);
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
    CARVE(T2, double, (size_t)(n + 1) * w->S)
    CARVE(bcol, double, n + 2)
    CARVE(nr, double, w->S)
    CARVE(fc, double, n + 2 + kTileRows)
    CARVE(fc2, double, n + 2 + kTileRows)
    CARVE(cstage, double, (n > m ? n : m) + 1)
    CARVE(scal, double, 4)
    CARVE(wts, double, w->S)
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
    CARVE(part_key, unsigned long long, 2 * kMaxSingleGrid)
    CARVE(part_var, int, 2 * kMaxSingleGrid)
    CARVE(part_crash, int, 2 * kMaxSingleGrid)
    CARVE(state, int, ST_COUNT)
    CARVE(labar, unsigned, 4)
#undef CARVE
    return o;
}

static const size_t kWsHeader = 4096;  // device copy of the rank's SingleWS lives at the workspace head

static int check_single_args(int n, int m, double tol, int max_pivots, int log_cap, int rule)
{
    if (n < 1 || m < 1 || !(tol >= 0.0) || max_pivots < 0 || log_cap < 0) {
        set_error("lpb200_simplex_single: need n >= 1, m >= 1, tol >= 0, max_pivots >= 0");
        return LPB200_E_INVALID;
    }
    if (rule != LPB200_RULE_REFERENCE && rule != LPB200_RULE_BLAND && rule != LPB200_RULE_DANTZIG &&
        rule != LPB200_RULE_DEVEX) {
        set_error("lpb200_simplex_single: rule must be one of LPB200_RULE_*");
        return LPB200_E_INVALID;
    }
    if ((long long)n + m + 1 >= (1ll << 30)) {
        set_error("lpb200_simplex_single: n + m too large");
        return LPB200_E_TOO_LARGE;
    }
    return 0;
}

// CTAs per rank for a cooperative launch that holds `groups` ranks on the current device.
static int pick_ctas_per_rank(const DeviceInfo &di, int n, int wcols, int groups, int *out)
{
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
    if (per_sm > 2) per_sm = 2;
    int cpr = di.sm_count * per_sm / groups;
    if (cpr > kMaxSingleGrid) cpr = kMaxSingleGrid;
    // small problems: no point in more CTAs than there is work
    const long long work = ((long long)(n + 1) * (wcols + 1) + kSgThreads * 8 - 1) / (kSgThreads * 8);
    if (cpr > work) cpr = (int)(work < 1 ? 1 : work);
    if (const char *g = getenv("LPB_DEBUG_SINGLE_CPR")) {  // measurements: force the CTAs per rank
        const int v = atoi(g);
        if (v >= 1 && v <= di.sm_count * per_sm / groups) cpr = v;
    }
    if (cpr < 1) {
        set_error("lpb200_simplex_single: too many emulated ranks for one cooperative launch");
        return LPB200_E_TOO_LARGE;
    }
    *out = cpr;
    return 0;
}

static int launch_single(const SingleWS *d_ranks, int groups, int cpr, const double *b, const double *c,
                         double tol, int max_pivots, int rule, int log_cap, double timeout_s, cudaStream_t stream)
{
    SingleArgs a;
    a.rule = rule;
    a.b = b;
    a.c = c;
    a.tol = tol;
    a.max_pivots = max_pivots;
    a.log_cap = log_cap;
    a.cpr = cpr;
    {
        const char *g = getenv("LPB_DEBUG_SINGLE_PIPELINE");
        a.pipelined = !(g && g[0] == '0');
        const char *c = getenv("LPB_DEBUG_SINGLE_NCTL");
        a.nctl_max = c ? atoi(c) : 8;
        if (a.nctl_max < 1) a.nctl_max = 1;
    }
    a.timeout_ns = (unsigned long long)(timeout_s * 1e9);
    void *params[] = {(void *)&d_ranks, (void *)&a};
    LPB_CUDA(cudaLaunchCooperativeKernel((const void *)simplex_single_kernel, dim3(groups * cpr),
                                         dim3(kSgThreads), params, 0, stream));
    return 0;
}

#ifdef LPB_TIMING
// debug builds only (not declared in include/lp_b200.h): read and clear the phase timers of the current device
extern "C" int lpb_debug_single_timing(unsigned long long *out /* [32][12] */)
{
    LPB_CUDA(cudaDeviceSynchronize());
    LPB_CUDA(cudaMemcpyFromSymbol(out, g_single_timing, sizeof(unsigned long long) * 32 * 12));
    static unsigned long long zeros[32 * 12];
    LPB_CUDA(cudaMemcpyToSymbol(g_single_timing, zeros, sizeof(zeros)));
    return 0;
}
#endif

extern "C" int64_t lpb200_simplex_single_workspace_bytes(int n, int m)
{
    if (n < 1 || m < 1) return 0;
    SingleWS w;
    return (int64_t)(kWsHeader + carve_single(&w, n, m, 0, 1, nullptr));
}

extern "C" int lpb200_simplex_single_state_device(int n, int m, const void *workspace, double *obj_by_var,
                                                  double *const_by_var, void *stream_v)
{
    if (n < 1 || m < 1 || workspace == nullptr) {
        set_error("lpb200_simplex_single_state_device: need n >= 1, m >= 1 and the workspace of a finished solve");
        return LPB200_E_INVALID;
    }
    DeviceInfo di;
    int rc = get_device_info(&di);
    if (rc) return rc;
    const int nv = n + m + 1;
    const int blocks = (nv + 1 + 255) / 256 < 1024 ? (nv + 1 + 255) / 256 : 1024;
    simplex_single_state_kernel<<<blocks, 256, 0, (cudaStream_t)stream_v>>>(static_cast<const SingleWS *>(workspace),
                                                                             obj_by_var, const_by_var);
    LPB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int lpb200_simplex_single_device(int n, int m, const double *A, const double *b, const double *c,
                                            double tol, int max_pivots, int32_t *status, double *x,
                                            double *objective, int32_t *npivots, int32_t *pivot_log,
                                            int log_cap, void *workspace, int64_t workspace_bytes,
                                            void *stream_v)
{
    return lpb200_simplex_single_ex_device(n, m, A, b, c, tol, max_pivots, LPB200_RULE_REFERENCE, status, x,
                                           objective, npivots, pivot_log, log_cap, workspace, workspace_bytes,
                                           stream_v);
}

extern "C" int lpb200_simplex_single_ex_device(int n, int m, const double *A, const double *b, const double *c,
                                               double tol, int max_pivots, int rule, int32_t *status, double *x,
                                               double *objective, int32_t *npivots, int32_t *pivot_log,
                                               int log_cap, void *workspace, int64_t workspace_bytes,
                                               void *stream_v)
{
    int rc = check_single_args(n, m, tol, max_pivots, log_cap, rule);
    if (rc) return rc;
    cudaStream_t stream = (cudaStream_t)stream_v;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    SingleWS w;
    unsigned char *base = static_cast<unsigned char *>(workspace);
    const size_t need = kWsHeader + carve_single(&w, n, m, 0, 1, base ? base + kWsHeader : nullptr);
    if (workspace == nullptr || (size_t)workspace_bytes < need) {
        set_error("lpb200_simplex_single: workspace too small (need " + std::to_string(need) + " bytes)");
        return LPB200_E_INVALID;
    }
    if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0) {
        set_error("lpb200_simplex_single: workspace must be 256-byte aligned");
        return LPB200_E_INVALID;
    }
    int cpr = 0;
    rc = pick_ctas_per_rank(di, n, m, 1, &cpr);
    if (rc) return rc;
    w.A = A;
    w.lda = m;
    w.xbuf = nullptr;
    w.peers = nullptr;
    w.status = status;
    w.x = x;
    w.objective = objective;
    w.npivots = npivots;
    w.pivot_log = pivot_log;
    // pageable source: the runtime stages it before returning, so the stack copy is safe
    LPB_CUDA(cudaMemcpyAsync(workspace, &w, sizeof(w), cudaMemcpyHostToDevice, stream));
    return launch_single(static_cast<const SingleWS *>(workspace), 1, cpr, b, c, tol, max_pivots, rule, log_cap,
                         0.0, stream);
}

extern "C" int lpb200_simplex_single(int n, int m, const double *A, const double *b, const double *c,
                                     double tol, int max_pivots, int32_t *status, double *x,
                                     double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                                     double *stats)
{
    return lpb200_simplex_single_ex(n, m, A, b, c, tol, max_pivots, LPB200_RULE_REFERENCE, status, x, objective,
                                    npivots, pivot_log, log_cap, stats);
}

extern "C" int lpb200_simplex_single_ex(int n, int m, const double *A, const double *b, const double *c,
                                        double tol, int max_pivots, int rule, int32_t *status, double *x,
                                        double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap,
                                        double *stats)
{
    int rc = check_single_args(n, m, tol, max_pivots, log_cap, rule);
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
            rc = lpb200_simplex_single_ex_device(n, m, dA, db, dc, tol, max_pivots, rule, dst, dx, dobj, dnp, dlog,
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
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    cudaFree(dA); cudaFree(db); cudaFree(dc); cudaFree(dx); cudaFree(dobj);
    cudaFree(dst); cudaFree(dnp); cudaFree(dlog); cudaFree(ws);
    return rc;
}

// ======================================================================================
// Column-sharded mode: communicator (peer-mapped exchange buffers) and entry points
// ======================================================================================
struct lpb200_comm {
    int rank, nranks, device, max_n;
    size_t xbuf_bytes;
    unsigned long long *xbuf;               // cudaMalloc'ed on this rank's device
    void *peer_host[kMaxRanks];             // peer-mapped pointers (own entry == xbuf)
    unsigned long long **peers_dev;         // device table handed to the kernel
    bool connected;
};

extern "C" int64_t lpb200_comm_handle_bytes(void) { return (int64_t)sizeof(cudaIpcMemHandle_t); }

extern "C" int lpb200_comm_create(int rank, int nranks, int max_n, lpb200_comm_t **out, void *handle_out)
{
    if (out == nullptr || handle_out == nullptr || nranks < 1 || nranks > kMaxRanks || rank < 0 ||
        rank >= nranks || max_n < 1) {
        set_error("lpb200_comm_create: need 0 <= rank < nranks <= 32, max_n >= 1, non-null outputs");
        return LPB200_E_INVALID;
    }
    DeviceInfo di;
    int rc = get_device_info(&di);
    if (rc) return rc;
    lpb200_comm *c = new lpb200_comm();
    c->rank = rank;
    c->nranks = nranks;
    c->device = di.device;
    c->max_n = max_n;
    c->xbuf_bytes = xbuf_bytes_for(max_n, nranks);
    c->xbuf = nullptr;
    c->peers_dev = nullptr;
    c->connected = false;
    for (int g = 0; g < kMaxRanks; ++g) c->peer_host[g] = nullptr;
    cudaError_t e;
    if ((e = cudaMalloc((void **)&c->xbuf, c->xbuf_bytes)) != cudaSuccess ||
        (e = cudaMemset(c->xbuf, 0, c->xbuf_bytes)) != cudaSuccess ||
        (e = cudaDeviceSynchronize()) != cudaSuccess ||
        (e = cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t *>(handle_out), c->xbuf)) != cudaSuccess) {
        if (c->xbuf) cudaFree(c->xbuf);
        delete c;
        return cuda_fail(e, "lpb200_comm_create");
    }
    *out = c;
    return 0;
}

extern "C" int lpb200_comm_connect(lpb200_comm_t *c, const void *all_handles)
{
    if (c == nullptr || all_handles == nullptr) {
        set_error("lpb200_comm_connect: null argument");
        return LPB200_E_INVALID;
    }
    if (c->connected) return 0;
    const cudaIpcMemHandle_t *h = static_cast<const cudaIpcMemHandle_t *>(all_handles);
    for (int g = 0; g < c->nranks; ++g) {
        if (g == c->rank) {
            c->peer_host[g] = c->xbuf;
            continue;
        }
        cudaError_t e = cudaIpcOpenMemHandle(&c->peer_host[g], h[g], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return cuda_fail(e, "cudaIpcOpenMemHandle (is peer access available between the GPUs?)");
    }
    LPB_CUDA(cudaMalloc((void **)&c->peers_dev, sizeof(void *) * kMaxRanks));
    LPB_CUDA(cudaMemcpy(c->peers_dev, c->peer_host, sizeof(void *) * kMaxRanks, cudaMemcpyHostToDevice));
    c->connected = true;
    return 0;
}

extern "C" int lpb200_comm_destroy(lpb200_comm_t *c)
{
    if (c == nullptr) return 0;
    for (int g = 0; g < c->nranks; ++g)
        if (g != c->rank && c->peer_host[g]) cudaIpcCloseMemHandle(c->peer_host[g]);
    if (c->peers_dev) cudaFree(c->peers_dev);
    if (c->xbuf) cudaFree(c->xbuf);
    delete c;
    return 0;
}

extern "C" int lpb200_simplex_sharded_columns(int m, int nranks, int rank, int *c_lo, int *c_count)
{
    if (m < 1 || nranks < 1 || nranks > kMaxRanks || rank < 0 || rank >= nranks) {
        set_error("lpb200_simplex_sharded_columns: bad arguments");
        return LPB200_E_INVALID;
    }
    const int S = cols_per_rank(m, nranks);
    int lo = rank * S;
    int cnt = m - lo;  // columns of A (position m is the auxiliary column, not part of A)
    if (cnt < 0) cnt = 0;
    if (cnt > S) cnt = S;
    if (lo > m) lo = m;
    if (c_lo) *c_lo = lo;
    if (c_count) *c_count = cnt;
    return 0;
}

extern "C" int64_t lpb200_simplex_sharded_workspace_bytes(int n, int m, int nranks)
{
    if (n < 1 || m < 1 || nranks < 1 || nranks > kMaxRanks) return 0;
    SingleWS w;
    return (int64_t)(kWsHeader + carve_single(&w, n, m, 0, nranks, nullptr));
}

extern "C" int lpb200_simplex_sharded_device(lpb200_comm_t *comm, int n, int m, const double *A_cols,
                                             int64_t lda, const double *b, const double *c, double tol,
                                             int max_pivots, int32_t *status, double *x, double *objective,
                                             int32_t *npivots, int32_t *pivot_log, int log_cap,
                                             void *workspace, int64_t workspace_bytes, double timeout_s,
                                             void *stream_v)
{
    return lpb200_simplex_sharded_ex_device(comm, n, m, A_cols, lda, b, c, tol, max_pivots, LPB200_RULE_REFERENCE,
                                            status, x, objective, npivots, pivot_log, log_cap, workspace,
                                            workspace_bytes, timeout_s, stream_v);
}

extern "C" int lpb200_simplex_sharded_ex_device(lpb200_comm_t *comm, int n, int m, const double *A_cols,
                                                int64_t lda, const double *b, const double *c, double tol,
                                                int max_pivots, int rule, int32_t *status, double *x,
                                                double *objective, int32_t *npivots, int32_t *pivot_log,
                                                int log_cap, void *workspace, int64_t workspace_bytes,
                                                double timeout_s, void *stream_v)
{
    int rc = check_single_args(n, m, tol, max_pivots, log_cap, rule);
    if (rc) return rc;
    if (comm == nullptr || !comm->connected) {
        set_error("lpb200_simplex_sharded: communicator is not connected");
        return LPB200_E_INVALID;
    }
    if (n > comm->max_n) {
        set_error("lpb200_simplex_sharded: n exceeds the communicator's max_n");
        return LPB200_E_TOO_LARGE;
    }
    cudaStream_t stream = (cudaStream_t)stream_v;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    SingleWS w;
    unsigned char *base = static_cast<unsigned char *>(workspace);
    const size_t need = kWsHeader + carve_single(&w, n, m, comm->rank, comm->nranks, base ? base + kWsHeader : nullptr);
    if (workspace == nullptr || (size_t)workspace_bytes < need ||
        (reinterpret_cast<uintptr_t>(workspace) & 255) != 0) {
        set_error("lpb200_simplex_sharded: workspace null, misaligned or too small (need " +
                  std::to_string(need) + " bytes)");
        return LPB200_E_INVALID;
    }
    // every rank must use the same grid shape? No: ranks only meet through the exchange buffers.
    int cpr = 0;
    rc = pick_ctas_per_rank(di, n, w.S, 1, &cpr);
    if (rc) return rc;
    w.A = A_cols;
    w.lda = lda;
    w.xbuf = comm->xbuf;
    w.peers = comm->peers_dev;
    w.status = status;
    w.x = x;
    w.objective = objective;
    w.npivots = npivots;
    w.pivot_log = pivot_log;
    LPB_CUDA(cudaMemcpyAsync(workspace, &w, sizeof(w), cudaMemcpyHostToDevice, stream));
    return launch_single(static_cast<const SingleWS *>(workspace), 1, cpr, b, c, tol, max_pivots, rule, log_cap,
                         timeout_s > 0.0 ? timeout_s : 20.0, stream);
}

// All `nranks` ranks of the sharded protocol as CTA groups of ONE cooperative launch on the
// current device (host pointers in, per-rank results out).  Exercises the exact kernel,
// packet format and exchange schedule of the multi-GPU path; used by the parity tests.
extern "C" int lpb200_simplex_sharded_emulated(int nranks, int n, int m, const double *A, const double *b,
                                               const double *c, double tol, int max_pivots, int32_t *status,
                                               double *x, double *objective, int32_t *npivots,
                                               int32_t *pivot_log, int log_cap)
{
    return lpb200_simplex_sharded_emulated_ex(nranks, n, m, A, b, c, tol, max_pivots, LPB200_RULE_REFERENCE, status,
                                              x, objective, npivots, pivot_log, log_cap);
}

extern "C" int lpb200_simplex_sharded_emulated_ex(int nranks, int n, int m, const double *A, const double *b,
                                                  const double *c, double tol, int max_pivots, int rule,
                                                  int32_t *status, double *x, double *objective,
                                                  int32_t *npivots, int32_t *pivot_log, int log_cap)
{
    return lpb200_simplex_sharded_emulated_state(nranks, n, m, A, b, c, tol, max_pivots, rule, status, x, objective,
                                                 npivots, pivot_log, log_cap, nullptr, nullptr);
}

// The emulated launch plus the tableau export of lpb200_simplex_single_state_device, per rank (host memory):
// obj_by_var[nranks][n+m+2] (a rank reports the columns it owns), const_by_var[nranks][n+m+1] (replicated).
extern "C" int lpb200_simplex_sharded_emulated_state(int nranks, int n, int m, const double *A, const double *b,
                                                     const double *c, double tol, int max_pivots, int rule,
                                                     int32_t *status, double *x, double *objective,
                                                     int32_t *npivots, int32_t *pivot_log, int log_cap,
                                                     double *obj_by_var, double *const_by_var)
{
    int rc = check_single_args(n, m, tol, max_pivots, log_cap, rule);
    if (rc) return rc;
    if (nranks < 1 || nranks > kMaxRanks) {
        set_error("lpb200_simplex_sharded_emulated: need 1 <= nranks <= 32");
        return LPB200_E_INVALID;
    }
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    const int G = nranks;
    SingleWS probe;
    const size_t per_rank = up256(carve_single(&probe, n, m, 0, G, nullptr));
    const size_t xb = xbuf_bytes_for(n, G);
    const size_t nA = (size_t)n * m * 8, nb = (size_t)n * 8, nc = (size_t)m * 8;
    const size_t nlog = pivot_log ? (size_t)log_cap * 2 * 4 : 0;
    const size_t out_per_rank = up256(nc) + up256(8) + up256(4) + up256(4) + up256(nlog);
    int cpr = 0;
    rc = pick_ctas_per_rank(di, n, probe.S, G, &cpr);
    if (rc) return rc;
    unsigned char *ws = nullptr, *outs = nullptr, *xbufs = nullptr;
    double *dA = nullptr, *db = nullptr, *dc = nullptr;
    SingleWS *d_ranks = nullptr;
    unsigned long long **d_peers = nullptr;
    rc = 0;
    LPB_TRY(cudaMalloc((void **)&ws, per_rank * G));
    LPB_TRY(cudaMalloc((void **)&outs, out_per_rank * G));
    LPB_TRY(cudaMalloc((void **)&xbufs, xb * G));
    LPB_TRY(cudaMalloc((void **)&dA, nA));
    LPB_TRY(cudaMalloc((void **)&db, nb));
    LPB_TRY(cudaMalloc((void **)&dc, nc));
    LPB_TRY(cudaMalloc((void **)&d_ranks, sizeof(SingleWS) * G));
    LPB_TRY(cudaMalloc((void **)&d_peers, sizeof(void *) * kMaxRanks));
    if (rc == 0) {
        LPB_TRY(cudaMemset(xbufs, 0, xb * G));
        LPB_TRY(cudaMemset(outs, 0xff, out_per_rank * G));
        LPB_TRY(cudaMemcpy(dA, A, nA, cudaMemcpyHostToDevice));
        LPB_TRY(cudaMemcpy(db, b, nb, cudaMemcpyHostToDevice));
        LPB_TRY(cudaMemcpy(dc, c, nc, cudaMemcpyHostToDevice));
    }
    std::vector<SingleWS> hw((size_t)G);
    std::vector<unsigned char *> ho((size_t)G);
    if (rc == 0) {
        void *peers[kMaxRanks] = {};
        for (int g = 0; g < G; ++g) peers[g] = xbufs + xb * g;
        LPB_TRY(cudaMemcpy(d_peers, peers, sizeof(peers), cudaMemcpyHostToDevice));
        for (int g = 0; g < G; ++g) {
            SingleWS &w = hw[g];
            carve_single(&w, n, m, g, G, ws + per_rank * g);
            w.A = dA + (w.c_lo < m ? w.c_lo : 0);
            w.lda = m;
            w.xbuf = reinterpret_cast<unsigned long long *>(xbufs + xb * g);
            w.peers = d_peers;
            unsigned char *o = outs + out_per_rank * g;
            ho[g] = o;
            w.x = reinterpret_cast<double *>(o);               o += up256(nc);
            w.objective = reinterpret_cast<double *>(o);       o += up256(8);
            w.status = reinterpret_cast<int32_t *>(o);         o += up256(4);
            w.npivots = reinterpret_cast<int32_t *>(o);        o += up256(4);
            w.pivot_log = nlog ? reinterpret_cast<int32_t *>(o) : nullptr;
        }
        LPB_TRY(cudaMemcpy(d_ranks, hw.data(), sizeof(SingleWS) * G, cudaMemcpyHostToDevice));
    }
    if (rc == 0) rc = launch_single(d_ranks, G, cpr, db, dc, tol, max_pivots, rule, log_cap, 20.0, nullptr);
    if (rc == 0) LPB_TRY(cudaDeviceSynchronize());
    for (int g = 0; g < G && rc == 0; ++g) {
        const SingleWS &w = hw[g];
        LPB_TRY(cudaMemcpy(x + (size_t)g * m, w.x, nc, cudaMemcpyDeviceToHost));
        LPB_TRY(cudaMemcpy(objective + g, w.objective, 8, cudaMemcpyDeviceToHost));
        LPB_TRY(cudaMemcpy(status + g, w.status, 4, cudaMemcpyDeviceToHost));
        LPB_TRY(cudaMemcpy(npivots + g, w.npivots, 4, cudaMemcpyDeviceToHost));
        if (nlog) LPB_TRY(cudaMemcpy(pivot_log + (size_t)g * log_cap * 2, w.pivot_log, nlog, cudaMemcpyDeviceToHost));
    }
    if (rc == 0 && (obj_by_var || const_by_var)) {
        const size_t nv = (size_t)n + m + 1;
        double *dov = nullptr, *dcv = nullptr;
        LPB_TRY(cudaMalloc((void **)&dov, (nv + 1) * 8));
        LPB_TRY(cudaMalloc((void **)&dcv, nv * 8));
        for (int g = 0; g < G && rc == 0; ++g) {
            rc = lpb200_simplex_single_state_device(n, m, d_ranks + g, dov, dcv, nullptr);
            if (rc) break;
            if (obj_by_var) LPB_TRY(cudaMemcpy(obj_by_var + (size_t)g * (nv + 1), dov, (nv + 1) * 8, cudaMemcpyDeviceToHost));
            if (const_by_var) LPB_TRY(cudaMemcpy(const_by_var + (size_t)g * nv, dcv, nv * 8, cudaMemcpyDeviceToHost));
        }
        cudaFree(dov);
        cudaFree(dcv);
    }
#undef LPB_TRY
    cudaFree(ws); cudaFree(outs); cudaFree(xbufs); cudaFree(dA); cudaFree(db); cudaFree(dc);
    cudaFree(d_ranks); cudaFree(d_peers);
    return rc;
}
