// host_narrow.h -- host-core helpers of the end-to-end (host-pointer) path: exact fp64 -> int8 narrowing
// of the constraint matrix and the persistent thread pool that runs it (host_narrow.cpp).
#pragma once

#include <stddef.h>
#include <stdint.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

namespace lpb {

// dst[i] = (int8)src[i] for i < n; returns true when every element is exactly representable, i.e. the
// kernel's widening of dst[i] reproduces src[i] bit for bit.  On false, dst is unspecified.
bool narrow_f64_to_i8(const double *src, int8_t *dst, size_t n);

// Persistent pool: the calling thread plus (threads - 1) workers.  Thread count = the cores this process
// may run on (sched_getaffinity), at most 32, or LPB_HOST_THREADS.
class HostPool {
public:
    static HostPool &instance();
    int threads() const { return nthreads_; }
    // fn(ctx, i) for every i in [0, nitems), items handed out dynamically; returns when all are done
    void parallel_for(int nitems, void (*fn)(void *, int), void *ctx);

private:
    HostPool();
    ~HostPool();
    void worker();
    void drain();
    int nthreads_ = 1;
    int owner_pid_ = 0;
    std::vector<std::thread> workers_;
    std::mutex mu_, call_mu_;
    std::condition_variable cv_start_, cv_done_;
    unsigned generation_ = 0;
    bool stop_ = false;
    void (*fn_)(void *, int) = nullptr;
    void *ctx_ = nullptr;
    int nitems_ = 0;
    std::atomic<int> next_{0}, pending_{0};
};

}  // namespace lpb
