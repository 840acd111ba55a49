// ubench_chain.cu — dependent-issue latencies that bound the lexicographic sweep's per-row chain on sm_100a:
// DFMA, DADD, 64-bit SHFL, SHFL+DFMA (one Kogge-Stone level), predicated select, LDS.  One warp, clock64() around
// a fully dependent chain.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_chain ubench_chain.cu
#include <cstdio>
#include <cuda_runtime.h>

#define N 512

__global__ void k_lat(double* out, long long* cyc, double a, double b) {
    __shared__ double sm[64];
    const int lane = threadIdx.x & 31;
    sm[lane] = a * lane; sm[lane + 32] = b;
    __syncthreads();
    double x = a + lane;
    long long t0, t1;
    // 0: DFMA chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = fma(x, a, b);
    t1 = clock64(); if (lane == 0) cyc[0] = t1 - t0;
    // 1: DADD chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = x + b;
    t1 = clock64(); if (lane == 0) cyc[1] = t1 - t0;
    // 2: SHFL(double) chain
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = __shfl_up_sync(0xffffffffu, x, 1);
    t1 = clock64(); if (lane == 0) cyc[2] = t1 - t0;
    // 3: SHFL + DFMA (one scan level, multiplier form: no select)
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) { const double t = __shfl_up_sync(0xffffffffu, x, 1); x = fma(a, t, x); }
    t1 = clock64(); if (lane == 0) cyc[3] = t1 - t0;
    // 4: SHFL + predicated DFMA (the current form)
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) { const double t = __shfl_up_sync(0xffffffffu, x, 1); if (lane >= (i & 15) + 1) x = fma(a, t, x); }
    t1 = clock64(); if (lane == 0) cyc[4] = t1 - t0;
    // 5: LDS dependent (address from the loaded value)
    int idx = lane;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) { const double t = sm[idx & 63]; idx = (int)t & 31; x += t; }
    t1 = clock64(); if (lane == 0) cyc[5] = t1 - t0;
    // 6: FFMA chain (fp32)
    float y = (float)x;
    const float af = (float)a, bf = (float)b;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) y = fmaf(y, af, bf);
    t1 = clock64(); if (lane == 0) cyc[6] = t1 - t0;
    // 7: SHFL(float) + FFMA
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) { const float t = __shfl_up_sync(0xffffffffu, y, 1); y = fmaf(af, t, y); }
    t1 = clock64(); if (lane == 0) cyc[7] = t1 - t0;
    out[threadIdx.x] = x + y;
}

// DFMA throughput: W warps per SM, 8 independent chains per thread
__global__ void k_tput(double* out, double a, double b, int iters) {
    double x0 = a + threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 8 * 4); cudaMalloc(&cyc, 64 * 8);
    k_lat<<<1, 32>>>(out, cyc, 0.999, 1e-3);
    k_lat<<<1, 32>>>(out, cyc, 0.999, 1e-3);
    long long h[8];
    cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    const char* names[8] = {"DFMA", "DADD", "SHFL64", "SHFL64+DFMA", "SHFL64+pred DFMA", "LDS64 dependent(+I2F)", "FFMA", "SHFL32+FFMA"};
    for (int i = 0; i < 8; ++i) printf("%-24s %.2f cycles per dependent op\n", names[i], (double)h[i] / N);
    // fp64 FMA peak: 148 SMs x 4 blocks x 256 threads
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k_tput<<<148 * 4, 256>>>(out, 0.999, 1e-3, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double flop = 2.0 * 8 * iters * 148.0 * 4 * 256;
        printf("DFMA throughput: %.2f TFLOP/s fp64 (%.3f ms)\n", flop / ms / 1e9, ms);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
