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
#include <string.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include "lp_common.cuh"
#include "lp_host.h"
#include "host_narrow.h"
#include "simplex_batched_variants.h"


// ---- kernel variants (same source, different parameters) ----
#ifndef LPB_SMEM_WARPS
#define LPB_SMEM_WARPS 8  // sweep warps of the product kernel (A/B knob: 10 gives six row groups, measured slower)
#endif
#define LPB_NS smem8
#define LPB_SWEEP_WARPS LPB_SMEM_WARPS
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 2
#define LPB_RULE 0
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE

// The same kernel with every sweep warp owning whole rows and the next pivot's ratio test fused into the sweep:
// one block barrier per pivot (simplex_batched_impl.cuh "warp-owns-rows").  Opt-in (LPB_BATCH_KERNEL=w), bit-exact,
// a measured experiment: 74 M pivots/s against the 83 M of smem8 on the D2 batch -- the barrier it removes was not
// what the pivot waits for (stores and the per-warp fixed work are).
#define LPB_NS rows8
#define LPB_SWEEP_WARPS 8
#define LPB_REG_ROWS 0
#define LPB_MIN_BLOCKS 2
#define LPB_RULE 0
#define LPB_WARP_ROWS 1
#include "simplex_batched_impl.cuh"
#undef LPB_NS
#undef LPB_SWEEP_WARPS
#undef LPB_REG_ROWS
#undef LPB_MIN_BLOCKS
#undef LPB_RULE
#undef LPB_WARP_ROWS

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
// Kernel variant for a shape (reference rule).  smem8 (8+1 warps, 2 CTAs/SM) is the kernel the
// headline 100x100 batch runs on; smaller problems get smaller CTAs so that more of them are resident:
//   tiny1   n, m <= 30                                   1+1 warps, 16 CTAs/SM
//   small1  (n+2) * ceil((m+2)/2) <= 1500, m <= 62        1+1 warps,  8 CTAs/SM
//   mid4    three or more tableaus fit one SM's shared memory   4+1 warps
// (small1 and mid4 only when the batch has more than 2 LPs per SM: see below)
// LPB_BATCH_KERNEL=smem|tiny|small|mid|reg forces one (A/B measurements, tests); reg is the
// register-resident experiment: bit-exact but slower on B200 (32 M vs 81 M pivots/s on Config 4 D2:
// with one CTA per SM nothing overlaps the serial stage/ratio chain).
enum { V_ROWS8 = 100, V_SMEM8 = LPB_V_SMEM8, V_REG15 = LPB_V_REG15, V_TINY1 = LPB_V_TINY1, V_SMALL1 = LPB_V_SMALL1, V_MID4 = LPB_V_MID4 };
static int pick_variant(int B, int n, int m, const DeviceInfo &di)
{
    const int npairs = (m + 3) / 2, nphys = n + 2;
    const bool one_warp_ok = npairs <= 32;
    if (const char *force = getenv("LPB_BATCH_KERNEL")) {
        switch (force[0]) {
        case 'r': return lpb_batch_variant_eligible(LPB_V_REG15, n, m) && lpb_batch_variant_smem_bytes(LPB_V_REG15, n, m) <= di.max_smem_optin ? V_REG15 : V_SMEM8;
        case 't': return one_warp_ok ? V_TINY1 : V_SMEM8;
        case 'm': return V_MID4;
        case 'w': return V_ROWS8;  // "warp-owns-rows"
        case 's': return force[1] == 'm' && force[2] == 'a' ? (one_warp_ok ? V_SMALL1 : V_SMEM8) : V_SMEM8;
        default: break;
        }
    }
    if (n <= 30 && m <= 30) return V_TINY1;
    // a batch that cannot even fill the SMs is a latency problem: all eight sweep warps on each LP
    // (a lone 50x50 LP: 2.7 us per pivot on smem8, 3.7 on small1)
    if (B <= 2 * di.sm_count) return V_SMEM8;
    if (one_warp_ok && nphys * npairs <= 1500) return V_SMALL1;
    if (lpb_batch_variant_smem_bytes(LPB_V_MID4, n, m) * 3 + 3 * 1024 <= (long long)di.max_smem_per_sm) return V_MID4;
    return V_SMEM8;
}

extern "C" int lpb200_simplex_batch_ex_device(int B, int n, int m, const void *A, int a_dtype, const double *b,
                                              const double *c, double tol, int max_pivots, int rule,
                                              int32_t *status, double *x, double *objective, int32_t *npivots,
                                              int32_t *pivot_log, int log_cap, void *stream_v)
{
    int rc = check_batch_args(B, n, m, tol, max_pivots, log_cap);
    if (rc) return rc;
    if (dtype_bytes(a_dtype) == 0) {
        set_error("lpb200_simplex_batch: a_dtype must be one of LPB200_DTYPE_*");
        return LPB200_E_INVALID;
    }
    if (rule != LPB200_RULE_REFERENCE && rule != LPB200_RULE_BLAND && rule != LPB200_RULE_DANTZIG &&
        rule != LPB200_RULE_DEVEX) {
        set_error("lpb200_simplex_batch: rule must be one of LPB200_RULE_*");
        return LPB200_E_INVALID;
    }
    if (B == 0) return 0;
    cudaStream_t stream = (cudaStream_t)stream_v;
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
#define LPB_LAUNCH(v) \
    lpb_batch_variant_launch(v, di, B, n, m, A, a_dtype, b, c, tol, max_pivots, status, x, objective, npivots, pivot_log, log_cap, stream)
    if (rule != LPB200_RULE_REFERENCE) {  // corrected modes: the two-warp CTA for diet-size problems, else 8+1 warps
        const char *force = getenv("LPB_BATCH_KERNEL");
        const bool tiny = (n <= 30 && m <= 30 && !(force && force[0] == 's' && force[1] == 'm' && force[2] == 'e'));
        if (rule == LPB200_RULE_DEVEX) return tiny ? LPB_LAUNCH(LPB_V_DEVEX_TINY1) : LPB_LAUNCH(LPB_V_DEVEX8);
        if (rule == LPB200_RULE_DANTZIG) return tiny ? LPB_LAUNCH(LPB_V_DANTZIG_TINY1) : LPB_LAUNCH(LPB_V_DANTZIG8);
        return tiny ? LPB_LAUNCH(LPB_V_BLAND_TINY1) : LPB_LAUNCH(LPB_V_BLAND8);
    }
    {
        const int v = pick_variant(B, n, m, di);
        if (v == V_ROWS8)
            return rows8::launch(di, B, n, m, A, a_dtype, b, c, tol, max_pivots, status, x, objective, npivots,
                                 pivot_log, log_cap, stream);
        if (v != V_SMEM8) return LPB_LAUNCH(v);
    }
#undef LPB_LAUNCH
    return smem8::launch(di, B, n, m, A, a_dtype, b, c, tol, max_pivots, status, x, objective, npivots, pivot_log,
                         log_cap, stream);
}

extern "C" int lpb200_simplex_batch_device(int B, int n, int m, const double *A, const double *b,
                                           const double *c, double tol, int max_pivots, int32_t *status,
                                           double *x, double *objective, int32_t *npivots,
                                           int32_t *pivot_log, int log_cap, void *stream_v)
{
    return lpb200_simplex_batch_ex_device(B, n, m, A, LPB200_DTYPE_F64, b, c, tol, max_pivots,
                                          LPB200_RULE_REFERENCE, status, x, objective, npivots, pivot_log, log_cap,
                                          stream_v);
}

// Device scratch of the host-pointer entry point: two buffer sets and two streams, owned by the
// calling thread and kept between calls (cudaMalloc/cudaFree of ~200 MB cost more than solving a
// few thousand LPs); grown on demand, released when the thread exits.
constexpr size_t kStageBytes = (size_t)4 << 20;  // calls smaller than this go through one pinned staging buffer

// sets 0-1: the copy-engine lane; sets 2..: the host-narrowing lane (see below).  The narrowing lane is deeper:
// its small copies queue behind the other lane's 96 MB transfers on the PCIe link, so a chunk takes a few
// milliseconds to leave the GPU although the cores narrow one in about a millisecond.
constexpr int kCopySets = 2, kNarrowSets = 4, kMaxSets = kCopySets + kNarrowSets;

struct HostArena {
    int device = -1;
    size_t cap = 0;
    char *mem[kMaxSets] = {};
    cudaStream_t stream[kMaxSets] = {};
    cudaEvent_t done[kMaxSets] = {};  // blocking-sync events: "the set's last chunk has left the GPU"
    char *stage = nullptr;  // pinned host mirror of a small call's buffer set (one copy in, one copy out)
    int8_t *narrow[kNarrowSets] = {};  // pinned int8 staging of the narrowing lane's sets
    size_t narrow_cap = 0;

    void release_memory()
    {
        for (int s = 0; s < kMaxSets; ++s) {
            if (mem[s]) cudaFree(mem[s]);
            mem[s] = nullptr;
        }
        cap = 0;
    }
    void release()
    {
        release_memory();
        for (int s = 0; s < kMaxSets; ++s) {
            if (stream[s]) cudaStreamDestroy(stream[s]);
            if (done[s]) cudaEventDestroy(done[s]);
            stream[s] = nullptr;
            done[s] = nullptr;
        }
        if (stage) cudaFreeHost(stage);
        stage = nullptr;
        for (int s = 0; s < kNarrowSets; ++s) {
            if (narrow[s]) cudaFreeHost(narrow[s]);
            narrow[s] = nullptr;
        }
        narrow_cap = 0;
        device = -1;
    }
    int ensure_narrow(size_t bytes)
    {
        if (bytes <= narrow_cap) return 0;
        for (int s = 0; s < kMaxSets; ++s)
            if (stream[s]) LPB_CUDA(cudaStreamSynchronize(stream[s]));
        for (int s = 0; s < kNarrowSets; ++s) {
            if (narrow[s]) cudaFreeHost(narrow[s]);
            narrow[s] = nullptr;
        }
        narrow_cap = 0;
        for (int s = 0; s < kNarrowSets; ++s) LPB_CUDA(cudaMallocHost((void **)&narrow[s], bytes));
        narrow_cap = bytes;
        return 0;
    }
    int ensure(int dev, int nset, size_t bytes)
    {
        if (dev != device) {
            release();
            device = dev;
        }
        if (bytes > cap) {  // grow geometrically (at least 8 MB): a run of slightly larger batches reallocates once
            size_t want = cap * 2 > bytes ? cap * 2 : bytes;
            if (want < ((size_t)8 << 20)) want = (size_t)8 << 20;
            for (int s = 0; s < kMaxSets; ++s)
                if (stream[s]) LPB_CUDA(cudaStreamSynchronize(stream[s]));
            release_memory();
            cap = want;
        }
        for (int s = 0; s < nset; ++s) {
            if (!stream[s]) LPB_CUDA(cudaStreamCreateWithFlags(&stream[s], cudaStreamNonBlocking));
            if (!done[s]) LPB_CUDA(cudaEventCreateWithFlags(&done[s], cudaEventBlockingSync | cudaEventDisableTiming));
            if (!mem[s]) LPB_CUDA(cudaMalloc((void **)&mem[s], cap));
        }
        if (!stage && bytes <= kStageBytes) LPB_CUDA(cudaMallocHost((void **)&stage, kStageBytes));
        return 0;
    }
    ~HostArena() { release(); }
};
static thread_local HostArena g_arena;
static thread_local long long g_last_lanes[4] = {0, 0, 0, 0};

// Whether the host-narrowing lane pays is MEASURED per calling thread, not guessed from the pool size: ranks that
// share one host compete for its memory bandwidth, and then the cores' reads only take bandwidth from the copy
// engines (round 1: 8 GPUs x 4 threads lost 9 %, 4 GPUs x 8 threads were never measured and regressed).  A probe call
// lets the copy-engine lane run alone for its first kProbeChunks chunks and only then starts the narrowing lane:
//   r1 = chunk rate of the copy engine alone (its first chunks, the pipeline filling up, left out),
//   r2 = aggregate chunk rate of both lanes afterwards;
// the second lane is kept while r2 > 1.10 * r1.  The verdict holds for the thread's next kGateEvery large calls, then
// one call probes again.  LPB_HOST_NARROW=0 / 1 forces the lane off / on (no probe); LPB_DEBUG_GATE=1 prints the rates.
constexpr int kProbeChunks = 7, kProbeSkip = 2, kGateEvery = 64;
struct LaneGate {
    int verdict = 0;  // 0: not measured yet, +1: both lanes, -1: copy engine only
    int age = 0;      // large calls since the last probe
    double r1 = 0.0, r2 = 0.0;
};
static thread_local LaneGate g_gate;

// Host-pointer entry point: the batch is cut into chunks that are copied in, solved and
// copied out on two alternating streams, so the PCIe copy of chunk k+1 overlaps the solve of
// chunk k and device memory stays bounded (two chunk-sized buffer sets).
//
// Large fp64 batches run TWO LANES that pull chunks off one counter (work stealing): the copy-engine
// lane ships a chunk's A as the caller holds it (fp64, PCIe-bound: ~55 GB/s), while the host cores
// narrow another chunk's A to int8 (host_narrow.cpp; only when every element survives the round trip
// bit for bit, else the chunk goes over as fp64) and ship 1/8 of the bytes.  The kernel widens int8
// exactly, so every answer is bit-identical to the single-lane path; the end-to-end rate becomes
// (PCIe rate + narrowing rate of the host cores) instead of the PCIe rate.  LPB_HOST_NARROW=0 turns
// the second lane off; LPB_HOST_THREADS sets the pool size (default: the cores of the process).
constexpr size_t kNarrowSlice = (size_t)1 << 18;  // elements per pool work item (2 MB of fp64)
struct NarrowJob {
    const double *src;
    int8_t *dst;
    size_t total;
    std::atomic<int> failed;
};
static void narrow_item(void *ctx, int i)
{
    NarrowJob *j = static_cast<NarrowJob *>(ctx);
    if (j->failed.load(std::memory_order_relaxed)) return;  // the chunk goes over as fp64 anyway
    const size_t lo = (size_t)i * kNarrowSlice;
    const size_t len = j->total - lo < kNarrowSlice ? j->total - lo : kNarrowSlice;
    if (!narrow_f64_to_i8(j->src + lo, j->dst + lo, len)) j->failed.store(1, std::memory_order_relaxed);
}

struct LaneResult {
    int rc = 0;
    std::string error;
    long long chunks = 0, narrowed = 0, a_bytes = 0;
};
extern "C" int lpb200_simplex_batch_ex(int B, int n, int m, const void *A_v, int a_dtype, const double *b,
                                       const double *c, double tol, int max_pivots, int rule, int32_t *status,
                                       double *x, double *objective, int32_t *npivots, int32_t *pivot_log,
                                       int log_cap)
{
    int rc = check_batch_args(B, n, m, tol, max_pivots, log_cap);
    if (rc) return rc;
    const size_t esz = (size_t)dtype_bytes(a_dtype);
    if (esz == 0) {
        set_error("lpb200_simplex_batch: a_dtype must be one of LPB200_DTYPE_*");
        return LPB200_E_INVALID;
    }
    if (B == 0) return 0;
    const char *A = static_cast<const char *>(A_v);
    DeviceInfo di;
    rc = get_device_info(&di);
    if (rc) return rc;
    const size_t lpA = (size_t)n * m, lpb = (size_t)n, lpc = (size_t)m;
    const size_t lplog = pivot_log ? (size_t)log_cap * 2 : 0;
    const size_t lp_bytes = lpA * esz + (lpb + 2 * lpc + 1) * 8 + 8 + lplog * 4;
    size_t chunk_bytes = (size_t)96 << 20;  // ~96 MB per chunk
    if (const char *e = getenv("LPB_HOST_CHUNK_BYTES")) {  // tests shrink it to drive both lanes with small batches
        const long long v = atoll(e);
        if (v > 0) chunk_bytes = (size_t)v;
    }
    long long chunk = (long long)(chunk_bytes / lp_bytes);
    if (chunk < 1) chunk = 1;
    if (chunk > B) chunk = B;
    const int nset = (chunk < B) ? 2 : 1;
    // second lane: fp64 A, enough chunks to share; whether it pays is measured (LaneGate above)
    bool two_lanes = (a_dtype == LPB200_DTYPE_F64) && (B + chunk - 1) / chunk >= 8 && HostPool::instance().threads() >= 2;
    bool probe = false;
    if (two_lanes) {
        const char *e = getenv("LPB_HOST_NARROW");
        if (e && e[0] == '0') {
            two_lanes = false;
        } else if (!(e && e[0] == '1')) {
            if (g_gate.verdict == 0 || g_gate.age >= kGateEvery) probe = true;
            else if (g_gate.verdict < 0) two_lanes = false;
            ++g_gate.age;
        }
    }
    const int nsets_used = two_lanes ? kMaxSets : nset;
    struct Set {
        char *A;
        double *b, *c, *x, *obj;
        int32_t *st, *np, *log;
        cudaStream_t stream;
    } set[kMaxSets] = {};
    g_last_lanes[0] = g_last_lanes[1] = g_last_lanes[2] = g_last_lanes[3] = 0;
    rc = 0;
#define LPB_TRY(call)                                                 \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess && rc == 0) rc = cuda_fail(_e, #call);  \
    } while (0)
    // one device arena per buffer set, kept by the calling thread between calls (see HostArena)
    size_t off[9];
    {
        const size_t sizes[8] = {(size_t)chunk * lpA * esz, (size_t)chunk * lpb * 8, (size_t)chunk * lpc * 8,
                                 (size_t)chunk * lpc * 8,   (size_t)chunk * 8,       (size_t)chunk * 4,
                                 (size_t)chunk * 4,         (size_t)chunk * lplog * 4};
        off[0] = 0;
        for (int i = 0; i < 8; ++i) off[i + 1] = off[i] + (sizes[i] + 255) / 256 * 256;
    }
    rc = g_arena.ensure(di.device, nsets_used, off[8]);
    if (rc == 0 && two_lanes) rc = g_arena.ensure_narrow((size_t)chunk * lpA);
    if (rc) return rc;
    for (int s = 0; s < nsets_used; ++s) {
        char *base = g_arena.mem[s];
        set[s].A = base + off[0];
        set[s].b = reinterpret_cast<double *>(base + off[1]);
        set[s].c = reinterpret_cast<double *>(base + off[2]);
        set[s].x = reinterpret_cast<double *>(base + off[3]);
        set[s].obj = reinterpret_cast<double *>(base + off[4]);
        set[s].st = reinterpret_cast<int32_t *>(base + off[5]);
        set[s].np = reinterpret_cast<int32_t *>(base + off[6]);
        set[s].log = lplog ? reinterpret_cast<int32_t *>(base + off[7]) : nullptr;
        set[s].stream = g_arena.stream[s];
    }
    if (nset == 1 && off[8] <= kStageBytes) {
        // small call: the inputs are gathered in the pinned mirror of the buffer set and go over in ONE
        // copy, the outputs come back in one -- 2 driver copies instead of 8 pageable ones
        Set &q = set[0];
        char *h = g_arena.stage;
        memcpy(h + off[0], A, (size_t)B * lpA * esz);
        memcpy(h + off[1], b, (size_t)B * lpb * 8);
        memcpy(h + off[2], c, (size_t)B * lpc * 8);
        LPB_TRY(cudaMemcpyAsync(g_arena.mem[0], h, off[3], cudaMemcpyHostToDevice, q.stream));
        if (lplog) LPB_TRY(cudaMemsetAsync(q.log, 0xff, (size_t)B * lplog * 4, q.stream));
        if (rc == 0)
            rc = lpb200_simplex_batch_ex_device(B, n, m, q.A, a_dtype, q.b, q.c, tol, max_pivots, rule, q.st, q.x,
                                                q.obj, q.np, q.log, log_cap, q.stream);
        LPB_TRY(cudaMemcpyAsync(h + off[3], g_arena.mem[0] + off[3], off[8] - off[3], cudaMemcpyDeviceToHost,
                                q.stream));
        LPB_TRY(cudaStreamSynchronize(q.stream));
        if (rc == 0) {
            memcpy(x, h + off[3], (size_t)B * lpc * 8);
            memcpy(objective, h + off[4], (size_t)B * 8);
            memcpy(status, h + off[5], (size_t)B * 4);
            memcpy(npivots, h + off[6], (size_t)B * 4);
            if (lplog) memcpy(pivot_log, h + off[7], (size_t)B * lplog * 4);
        }
#undef LPB_TRY
        return rc;
    }
    // One lane: claims chunks off the shared counter and runs them on its two buffer sets.
    HostArena *const arena = &g_arena;  // thread_local: the copier thread must use the CALLER's arena
    std::atomic<long long> next_chunk{0};
    const long long nchunks = (B + chunk - 1) / chunk;
    // probe bookkeeping (LaneGate): when the copy-engine lane saw each of its chunks leave the GPU
    std::atomic<int> copy_done{0};
    std::atomic<bool> copy_finished{false};
    std::vector<std::chrono::steady_clock::time_point> copy_times;
    if (probe) copy_times.reserve((size_t)nchunks);
    auto run_lane = [&](int lane, bool narrow, bool backpressure, LaneResult *out) {
        int rc = 0;
#define LPB_TRY(call)                                                 \
    do {                                                              \
        cudaError_t _e = (call);                                      \
        if (_e != cudaSuccess && rc == 0) rc = cuda_fail(_e, #call);  \
    } while (0)
        LPB_TRY(cudaSetDevice(di.device));
        int k = 0, misses = 0;
        while (rc == 0) {
            const long long ci = next_chunk.fetch_add(1, std::memory_order_relaxed);
            if (ci >= nchunks) break;
            const long long lo = ci * chunk;
            const long long cnt = (B - lo < chunk) ? (B - lo) : chunk;
            const int si = (lane == 0) ? (k++ % kCopySets) : kCopySets + (k++ % kNarrowSets);
            Set &q = set[si];
            // the set's previous chunk has left the GPU: bounds the lane to two chunks in flight (so the other
            // lane gets its share of the counter) and frees the set's pinned staging buffer
            if (backpressure) LPB_TRY(cudaEventSynchronize(arena->done[si]));
            if (probe && lane == 0 && k > kCopySets) {  // the chunk this set carried two rounds ago has left the GPU
                copy_times.push_back(std::chrono::steady_clock::now());
                copy_done.fetch_add(1, std::memory_order_release);
            }
            int dtype = a_dtype;
            const char *srcA = A + lo * lpA * esz;
            size_t bytesA = (size_t)cnt * lpA * esz;
            if (narrow) {
                NarrowJob job = {reinterpret_cast<const double *>(srcA), arena->narrow[si - kCopySets], (size_t)cnt * lpA, {0}};
                const int items = (int)((job.total + kNarrowSlice - 1) / kNarrowSlice);
                HostPool::instance().parallel_for(items, narrow_item, &job);
                if (job.failed.load() == 0) {
                    dtype = LPB200_DTYPE_I8;
                    srcA = reinterpret_cast<const char *>(job.dst);
                    bytesA = job.total;
                    ++out->narrowed;
                } else if (++misses >= 2 && out->narrowed == 0) {
                    narrow = false;  // not small-integer data: stop scanning, this lane copies fp64 like the other
                }
            }
            LPB_TRY(cudaMemcpyAsync(q.A, srcA, bytesA, cudaMemcpyHostToDevice, q.stream));
            LPB_TRY(cudaMemcpyAsync(q.b, b + lo * lpb, cnt * lpb * 8, cudaMemcpyHostToDevice, q.stream));
            LPB_TRY(cudaMemcpyAsync(q.c, c + lo * lpc, cnt * lpc * 8, cudaMemcpyHostToDevice, q.stream));
            if (lplog) LPB_TRY(cudaMemsetAsync(q.log, 0xff, cnt * lplog * 4, q.stream));
            if (rc == 0)
                rc = lpb200_simplex_batch_ex_device((int)cnt, n, m, q.A, dtype, q.b, q.c, tol, max_pivots, rule, q.st,
                                                    q.x, q.obj, q.np, q.log, log_cap, q.stream);
            LPB_TRY(cudaMemcpyAsync(status + lo, q.st, cnt * 4, cudaMemcpyDeviceToHost, q.stream));
            LPB_TRY(cudaMemcpyAsync(x + lo * lpc, q.x, cnt * lpc * 8, cudaMemcpyDeviceToHost, q.stream));
            LPB_TRY(cudaMemcpyAsync(objective + lo, q.obj, cnt * 8, cudaMemcpyDeviceToHost, q.stream));
            LPB_TRY(cudaMemcpyAsync(npivots + lo, q.np, cnt * 4, cudaMemcpyDeviceToHost, q.stream));
            if (lplog)
                LPB_TRY(cudaMemcpyAsync(pivot_log + lo * lplog, q.log, cnt * lplog * 4, cudaMemcpyDeviceToHost,
                                        q.stream));
            if (backpressure) LPB_TRY(cudaEventRecord(arena->done[si], q.stream));
            ++out->chunks;
            out->a_bytes += (long long)bytesA;
        }
        if (probe && lane == 0) {
            for (int s = 0; s < kCopySets && rc == 0; ++s) {  // the last chunks in flight
                if (k > s) {
                    LPB_TRY(cudaEventSynchronize(arena->done[(k - 1 - s) % kCopySets]));
                    copy_times.push_back(std::chrono::steady_clock::now());
                }
            }
            copy_finished.store(true, std::memory_order_release);
        }
#undef LPB_TRY
        out->rc = rc;
        if (rc) out->error = lpb200_last_error();
    };
    LaneResult lane_res[2];
    if (two_lanes) {
        using clk = std::chrono::steady_clock;
        std::thread copier(run_lane, 0, false, true, &lane_res[0]);  // the copy-engine lane only enqueues and waits
        clk::time_point t_narrow = clk::now();
        if (probe) {  // the copy engine alone for its first chunks: that is the rate the second lane has to beat
            while (copy_done.load(std::memory_order_acquire) < kProbeChunks && !copy_finished.load(std::memory_order_acquire))
                std::this_thread::sleep_for(std::chrono::microseconds(50));
            t_narrow = clk::now();
        }
        run_lane(1, true, true, &lane_res[1]);                       // the caller drives the host pool
        copier.join();
        if (probe && rc == 0) {
            for (int s = 0; s < nsets_used; ++s)
                if (set[s].stream) cudaStreamSynchronize(set[s].stream);
            const clk::time_point t_end = clk::now();
            // completions of the copy lane before the narrowing lane started; the first kProbeSkip of them are the
            // pipeline filling up (first copy in, first kernel, first results out) and would make the copy engine look
            // slower than it is -- at eight ranks that bias switched the lane on and cost 10 % (1.95 M against 2.2 M LPs/s)
            size_t k1 = 0;
            while (k1 < copy_times.size() && copy_times[k1] <= t_narrow) ++k1;
            if (k1 >= (size_t)kProbeSkip + 3 && nchunks > (long long)k1 + 4) {
                const double d1 = std::chrono::duration<double>(copy_times[k1 - 1] - copy_times[kProbeSkip]).count();
                const double d2 = std::chrono::duration<double>(t_end - copy_times[k1 - 1]).count();
                g_gate.r1 = d1 > 0 ? (double)(k1 - 1 - kProbeSkip) / d1 : 0.0;
                g_gate.r2 = d2 > 0 ? (double)(nchunks - (long long)k1) / d2 : 0.0;
                g_gate.verdict = (g_gate.r2 > 1.10 * g_gate.r1) ? 1 : -1;
                g_gate.age = 0;
                if (getenv("LPB_DEBUG_GATE"))
                    fprintf(stderr, "[lpb gate] copy engine alone %.1f chunks/s, both lanes %.1f chunks/s -> narrowing lane %s\n",
                            g_gate.r1, g_gate.r2, g_gate.verdict > 0 ? "on" : "off");
            }
        }
    } else {
        run_lane(0, false, false, &lane_res[0]);
    }
    for (int l = 0; l < 2 && rc == 0; ++l) {
        if (lane_res[l].rc) {
            rc = lane_res[l].rc;
            set_error(lane_res[l].error);
        }
    }
    for (int s = 0; s < nsets_used; ++s) {
        if (set[s].stream) {
            cudaError_t e = cudaStreamSynchronize(set[s].stream);
            if (e != cudaSuccess && rc == 0) rc = cuda_fail(e, "cudaStreamSynchronize");
        }
    }
    g_last_lanes[0] = lane_res[0].chunks;
    g_last_lanes[1] = lane_res[1].chunks;
    g_last_lanes[2] = lane_res[1].narrowed;
    g_last_lanes[3] = lane_res[0].a_bytes + lane_res[1].a_bytes;
    return rc;
}

// Chunks served by the copy-engine lane / the narrowing lane / narrowed to int8, and the bytes of A that
// crossed PCIe, in the calling thread's last chunked lpb200_simplex_batch[_ex] call (diagnostics for
// bench.py and the tests).
extern "C" int lpb200_simplex_batch_last_lanes(int64_t *out4)
{
    if (out4 == nullptr) return LPB200_E_INVALID;
    for (int i = 0; i < 4; ++i) out4[i] = g_last_lanes[i];
    return 0;
}

extern "C" int lpb200_simplex_batch(int B, int n, int m, const double *A, const double *b, const double *c,
                                    double tol, int max_pivots, int32_t *status, double *x,
                                    double *objective, int32_t *npivots, int32_t *pivot_log, int log_cap)
{
    return lpb200_simplex_batch_ex(B, n, m, A, LPB200_DTYPE_F64, b, c, tol, max_pivots, LPB200_RULE_REFERENCE,
                                   status, x, objective, npivots, pivot_log, log_cap);
}
