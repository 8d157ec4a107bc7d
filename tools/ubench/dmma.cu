// Micro-benchmark: FP64 tensor-core (DMMA m8n8k4) issue rate per SM, 1 CTA per SM.
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(1024) k(double *out, int iters, long long *cyc) {
    double c[8][2];
    for (int i = 0; i < 8; ++i) { c[i][0] = 0; c[i][1] = 0; }
    double a = threadIdx.x * 1e-3, b = 1.0 - threadIdx.x * 1e-4;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma884(c[i][0], c[i][1], a, b);
    }
    long long t1 = clock64();
    double s = 0; for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
    double *out; long long *cyc; cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cyc, 148 * 8);
    long long h[148];
    for (int threads : {32, 128, 512, 1024}) {
        k<<<148, threads>>>(out, 1024, cyc); cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double n = (double)(threads / 32) * 1024 * 8;
        printf("DMMA warps=%2d: %.2f clk per DMMA per SM  (%.1f FMA/clk/SM)\n", threads / 32, h[0] / n, n * 256 / h[0]);
    }
    return 0;
}
