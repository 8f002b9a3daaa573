// Microbenchmark: latency of dependent fp64 operations and of shared-memory loads on sm_100a,
// measured with clock64() by one warp (what bounds the one-CTA-per-LP sweep).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 --fmad=false -o fp64_latency fp64_latency.cu && ./fp64_latency
#include <cstdio>
#include <cuda_runtime.h>

constexpr int N = 2048;

__global__ void k_lat(double seed, double *out, long long *cyc)
{
    __shared__ double sh[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sh[i] = (double)((i * 7 + 1) & 1023);
    __syncthreads();
    double a = seed + threadIdx.x, b = 1.0000001, c = 0.5;
    long long t0, t1;
    // dependent DADD chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) a = __dadd_rn(a, b);
    t1 = clock64();
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    // dependent DMUL chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) a = __dmul_rn(a, b);
    t1 = clock64();
    if (threadIdx.x == 0) cyc[1] = t1 - t0;
    // dependent DMUL -> DADD (the axpy of the sweep)
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) a = __dadd_rn(c, __dmul_rn(a, b));
    t1 = clock64();
    if (threadIdx.x == 0) cyc[2] = t1 - t0;
    // 8 independent DADD chains (throughput per warp)
    double v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = a + j;
    t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __dadd_rn(v[j], b);
    }
    t1 = clock64();
    if (threadIdx.x == 0) cyc[3] = t1 - t0;
#pragma unroll
    for (int j = 0; j < 8; ++j) a += v[j];
    // compare -> select chain (snap): DSETP + FSELx2 dependent
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) a = (fabs(a) <= 1e-9) ? b : a + 0.0 * c;
    t1 = clock64();
    if (threadIdx.x == 0) cyc[4] = t1 - t0;
    // dependent shared-memory loads (pointer chase, 64-bit)
    int idx = threadIdx.x;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) idx = (int)sh[idx & 1023];
    t1 = clock64();
    if (threadIdx.x == 0) cyc[5] = t1 - t0;
    out[threadIdx.x] = a + idx;
}

int main()
{
    double *out;
    long long *cyc, h[8];
    cudaMalloc(&out, 1024 * 8);
    cudaMalloc(&cyc, 8 * 8);
    const char *names[6] = {"dependent DADD", "dependent DMUL", "dependent DMUL+DADD pair", "8 independent DADD chains (per DADD)",
                            "dependent abs-compare+select(+DADD,DMUL)", "dependent LDS.64 (+cvt)"};
    for (int warps = 1; warps <= 8; warps *= 2) {
        k_lat<<<1, 32 * warps>>>(1.0, out, cyc);
        k_lat<<<1, 32 * warps>>>(1.0, out, cyc);
        cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        printf("warps per SM %d:\n", warps);
        for (int i = 0; i < 6; ++i) {
            double per = (double)h[i] / N;
            if (i == 3) per /= 8;
            printf("  %-44s %.1f cycles\n", names[i], per);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
