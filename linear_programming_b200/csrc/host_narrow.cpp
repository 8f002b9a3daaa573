// host_narrow.cpp -- host-core side of the end-to-end path of liblp_b200.so (no CUDA in this file).
//
// The reference parses its problems with int() (week-2/diet/simplex.py:712-717) and the course bounds the
// coefficients to |A[i][j]| <= 100, so a caller that holds A as fp64 (the reference's own storage,
// simplex.py:259-262 float()) is shipping 8 bytes per small integer over PCIe -- and the host->device copy
// of A, 98 % of a problem, is what bounds lpb200_simplex_batch from host memory (55 GB/s against a kernel that
// consumes 310 GB/s).  The host cores of the box sit idle during that copy.  This file lets them help:
// narrow_f64_to_i8() rewrites a slice of A as int8 when, and only when, EVERY element survives the round
// trip bit for bit ((double)(int8)v has the bits of v: integers in [-127, 127], and +0.0 but not -0.0); the
// kernel widens int8 exactly (LPB200_DTYPE_I8), so the answers are those of the fp64 copy, bit for bit.
// A slice that fails the test is shipped as fp64.  HostPool is the small persistent thread pool that runs it.
#include "host_narrow.h"

#include "../../include/lp_b200.h"

#include <immintrin.h>
#include <sched.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

namespace lpb {

// Scalar version; also the tail of the AVX2 one.  Returns true when all n elements are exact.
static bool narrow_scalar(const double *src, int8_t *dst, size_t n)
{
    uint64_t bad = 0;
    for (size_t i = 0; i < n; ++i) {
        const double v = src[i];
        const int q = (v >= -127.0 && v <= 127.0) ? (int)v : 0;  // NaN fails both compares; -128 is left to the fp64 path like the AVX2 loop does
        const double r = (double)(int8_t)q;
        uint64_t bv, br;
        memcpy(&bv, &v, 8);
        memcpy(&br, &r, 8);
        bad |= bv ^ br;
        dst[i] = (int8_t)q;
    }
    return bad == 0;
}

__attribute__((target("avx2"))) static bool narrow_avx2(const double *src, int8_t *dst, size_t n)
{
    __m256i bad = _mm256_setzero_si256();
    __m128i mx = _mm_setzero_si128();  // unsigned max of |q|: INT_MIN (out-of-range / NaN conversions) stays huge
    size_t i = 0;
    for (; i + 16 <= n; i += 16) {
        const __m256d a0 = _mm256_loadu_pd(src + i), a1 = _mm256_loadu_pd(src + i + 4);
        const __m256d a2 = _mm256_loadu_pd(src + i + 8), a3 = _mm256_loadu_pd(src + i + 12);
        const __m128i q0 = _mm256_cvttpd_epi32(a0), q1 = _mm256_cvttpd_epi32(a1);
        const __m128i q2 = _mm256_cvttpd_epi32(a2), q3 = _mm256_cvttpd_epi32(a3);
        bad = _mm256_or_si256(bad, _mm256_xor_si256(_mm256_castpd_si256(_mm256_cvtepi32_pd(q0)), _mm256_castpd_si256(a0)));
        bad = _mm256_or_si256(bad, _mm256_xor_si256(_mm256_castpd_si256(_mm256_cvtepi32_pd(q1)), _mm256_castpd_si256(a1)));
        bad = _mm256_or_si256(bad, _mm256_xor_si256(_mm256_castpd_si256(_mm256_cvtepi32_pd(q2)), _mm256_castpd_si256(a2)));
        bad = _mm256_or_si256(bad, _mm256_xor_si256(_mm256_castpd_si256(_mm256_cvtepi32_pd(q3)), _mm256_castpd_si256(a3)));
        mx = _mm_max_epu32(mx, _mm_max_epu32(_mm_max_epu32(_mm_abs_epi32(q0), _mm_abs_epi32(q1)),
                                             _mm_max_epu32(_mm_abs_epi32(q2), _mm_abs_epi32(q3))));
        const __m128i w01 = _mm_packs_epi32(q0, q1), w23 = _mm_packs_epi32(q2, q3);  // saturating, checked through mx
        _mm_storeu_si128(reinterpret_cast<__m128i *>(dst + i), _mm_packs_epi16(w01, w23));
    }
    uint64_t b[4];
    uint32_t mm[4];
    _mm256_storeu_si256(reinterpret_cast<__m256i *>(b), bad);
    _mm_storeu_si128(reinterpret_cast<__m128i *>(mm), mx);
    bool ok = (b[0] | b[1] | b[2] | b[3]) == 0;
    for (int k = 0; k < 4; ++k) ok = ok && mm[k] <= 127u;  // -128 is exact too, but is left to the fp64 path
    if (i < n) ok = narrow_scalar(src + i, dst + i, n - i) && ok;
    return ok;
}

bool narrow_f64_to_i8(const double *src, int8_t *dst, size_t n)
{
    static const bool have_avx2 = __builtin_cpu_supports("avx2");
    return have_avx2 ? narrow_avx2(src, dst, n) : narrow_scalar(src, dst, n);
}

// ---- HostPool ----
static int usable_cores()
{
    cpu_set_t set;
    CPU_ZERO(&set);
    if (sched_getaffinity(0, sizeof(set), &set) == 0) {
        const int c = CPU_COUNT(&set);
        if (c > 0) return c;
    }
    const unsigned h = std::thread::hardware_concurrency();
    return h ? (int)h : 1;
}

HostPool &HostPool::instance()
{
    static HostPool pool;
    return pool;
}

HostPool::HostPool()
{
    int want = usable_cores();
    if (const char *e = getenv("LPB_HOST_THREADS")) {
        const int v = atoi(e);
        if (v >= 1) want = v;
    }
    if (want > 32) want = 32;
    nthreads_ = want;
    owner_pid_ = (int)getpid();
    for (int t = 1; t < nthreads_; ++t) workers_.emplace_back([this] { worker(); });  // the caller is thread 0
}

HostPool::~HostPool()
{
    if ((int)getpid() != owner_pid_) {  // forked child: the std::thread objects refer to threads that do not exist here
        for (auto &w : workers_) w.detach();
        return;
    }
    {
        std::lock_guard<std::mutex> lock(mu_);
        stop_ = true;
    }
    cv_start_.notify_all();
    for (auto &w : workers_) w.join();
}

void HostPool::drain()
{
    for (;;) {
        const int i = next_.fetch_add(1, std::memory_order_relaxed);
        if (i >= nitems_) break;
        fn_(ctx_, i);
    }
}

void HostPool::worker()
{
    unsigned seen = 0;
    for (;;) {
        {
            std::unique_lock<std::mutex> lock(mu_);
            cv_start_.wait(lock, [&] { return stop_ || generation_ != seen; });
            if (stop_) return;
            seen = generation_;
        }
        drain();
        if (pending_.fetch_sub(1, std::memory_order_acq_rel) == 1) {
            std::lock_guard<std::mutex> lock(mu_);
            cv_done_.notify_one();
        }
    }
}

void HostPool::parallel_for(int nitems, void (*fn)(void *, int), void *ctx)
{
    std::lock_guard<std::mutex> serial(call_mu_);  // one parallel_for at a time (callers on several threads queue up)
    if (nitems <= 0) return;
    // a fork()ed child inherits the pool object but none of its worker threads: run the items inline there
    if (workers_.empty() || nitems == 1 || (int)getpid() != owner_pid_) {
        for (int i = 0; i < nitems; ++i) fn(ctx, i);
        return;
    }
    {
        std::lock_guard<std::mutex> lock(mu_);
        fn_ = fn;
        ctx_ = ctx;
        nitems_ = nitems;
        next_.store(0, std::memory_order_relaxed);
        pending_.store((int)workers_.size(), std::memory_order_release);
        ++generation_;
    }
    cv_start_.notify_all();
    drain();
    std::unique_lock<std::mutex> lock(mu_);
    cv_done_.wait(lock, [&] { return pending_.load(std::memory_order_acquire) == 0; });
}

}  // namespace lpb

// ---- C ABI: the same narrowing for callers that want to hold A as int8 themselves ----
namespace {
struct SliceJob {
    const double *src;
    int8_t *dst;
    size_t total, slice;
    std::atomic<int> failed;
};
void slice_item(void *ctx, int i)
{
    SliceJob *j = static_cast<SliceJob *>(ctx);
    if (j->failed.load(std::memory_order_relaxed)) return;
    const size_t lo = (size_t)i * j->slice;
    const size_t len = j->total - lo < j->slice ? j->total - lo : j->slice;
    if (!lpb::narrow_f64_to_i8(j->src + lo, j->dst + lo, len)) j->failed.store(1, std::memory_order_relaxed);
}
}  // namespace

extern "C" int lpb200_narrow_f64_to_i8(const double *src, int8_t *dst, int64_t count)
{
    if (count < 0 || (count > 0 && (src == nullptr || dst == nullptr))) return LPB200_E_INVALID;
    if (count == 0) return 1;
    SliceJob job = {src, dst, (size_t)count, (size_t)1 << 18, {0}};
    const int items = (int)((job.total + job.slice - 1) / job.slice);
    lpb::HostPool::instance().parallel_for(items, slice_item, &job);
    return job.failed.load() ? 0 : 1;
}

extern "C" int lpb200_host_threads(void) { return lpb::HostPool::instance().threads(); }
