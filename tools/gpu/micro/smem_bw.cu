// Microbenchmark: shared-memory streaming bandwidth of the whole GPU with the access shape of the
// one-CTA-per-LP sweep (simplex_batched_impl.cuh): every thread reads a 16-byte word (LDS.128), touches
// it and writes it back (STS.128), conflict-free, over a buffer of the size of a 100x100 tableau, two
// 288-thread CTAs per SM.  Prints one JSON line; bench.py divides the batched kernel's algorithmic bytes
// per second by "smem_GBps" (profiles/r2_smem_peak.json).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o smem_bw smem_bw.cu && ./smem_bw
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int kThreads = 256;
constexpr int kWords = 102 * 102 / 2;  // double2 words of a 100x100 tableau (83 KB)

template <int kUnroll>
__global__ void __launch_bounds__(kThreads, 2) k_stream(int iters, double bump, double *sink)
{
    extern __shared__ __align__(16) unsigned char raw[];
    double2 *buf = reinterpret_cast<double2 *>(raw);
    for (int i = threadIdx.x; i < kWords; i += kThreads) buf[i] = make_double2(1.0 + i, 2.0 + i);
    __syncthreads();
    double acc = 0.0;
    const int span = kThreads * kUnroll;
    for (int it = 0; it < iters; ++it) {
        for (int base = 0; base + span <= kWords; base += span) {
            double2 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) v[u] = buf[base + u * kThreads + threadIdx.x];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                v[u].x += bump;
                v[u].y += bump;
                buf[base + u * kThreads + threadIdx.x] = v[u];
            }
            acc += v[0].x;
        }
        __syncthreads();
    }
    if (acc == -1.0) sink[0] = acc;
}

int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    const int smem = kWords * 16;
    double *sink;
    cudaMalloc(&sink, 8);
    cudaFuncSetAttribute(k_stream<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaFuncSetAttribute(k_stream<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int iters = 4000;
    double best = 0.0;
    int best_unroll = 0;
    for (int pass = 0; pass < 2; ++pass) {
        for (int rep = 0; rep < 4; ++rep) {
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            cudaEventRecord(e0);
            if (pass == 0) k_stream<4><<<2 * sms, kThreads, smem>>>(iters, 0.0, sink);
            else k_stream<8><<<2 * sms, kThreads, smem>>>(iters, 0.0, sink);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            const int unroll = pass == 0 ? 4 : 8;
            const int span = kThreads * unroll;
            const double words = (double)(kWords / span * span) * iters * 2 * sms;  // per launch
            const double gbs = words * 32.0 / (ms * 1e-3) / 1e9;                   // 16 B read + 16 B written
            if (rep > 0 && gbs > best) { best = gbs; best_unroll = unroll; }
        }
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "cuda error\n"); return 1; }
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"smem_GBps\": %.1f, \"sms\": %d, \"unroll\": %d, \"ctas_per_sm\": 2, \"threads\": %d, "
           "\"nominal_GBps_at_max_clock\": %.1f, \"max_clock_mhz\": %.0f, "
           "\"how\": \"LDS.128 + STS.128 streaming over an 83 KB buffer per CTA, conflict-free, best of 3 after warm-up, CUDA events\"}\n",
           best, sms, best_unroll, kThreads, sms * 128.0 * clk * 1e3 / 1e9, clk / 1e3);
    return 0;
}
