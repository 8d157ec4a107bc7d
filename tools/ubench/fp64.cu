// Micro-benchmark: FP64 FMA / shared-memory FP64 dot throughput per SM.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(1024) kd(double *out, int iters, double z, long long *cyc) {
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3 + i;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = fma(a[i], z, 1e-9);
    }
    long long t1 = clock64();
    double s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void __launch_bounds__(1024) kf(float *out, int iters, float z, long long *cyc) {
    float a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3f + i;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], z, 1e-9f);
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    double *out; long long *cyc; cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cyc, 148 * 8);
    long long h[148];
    for (int threads : {128, 256, 1024}) {
        kd<<<148, threads>>>(out, 2048, 0.999, cyc); cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        printf("DFMA threads=%4d: %.2f fma/clk/SM\n", threads, (double)threads * 2048 * 8 / h[0]);
        kf<<<148, threads>>>((float *)out, 2048, 0.999f, cyc); cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        printf("FFMA threads=%4d: %.2f fma/clk/SM\n", threads, (double)threads * 2048 * 8 / h[0]);
    }
    return 0;
}
