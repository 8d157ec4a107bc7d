// Micro-benchmark: the K1 generator loop in isolation (no MMA, no cross-warp barriers).
//   variants: bit0 = LDS.128 (zt,b) table reads, bit1 = tcgen05.st of the results, bit2 = tcgen05.wait::st per chunk
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../gpt_b200/csrc/tc_ptx.cuh"
using namespace tcx;

__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t *>(&h); }
__device__ __forceinline__ void split_pack(float a0, float a1, uint32_t &hi, uint32_t &lo) {
    __half2 h = __floats2half2_rn(a0, a1);
    hi = h2u(h);
    float2 hf = __half22float2(h);
    lo = h2u(__floats2half2_rn(a0 - hf.x, a1 - hf.y));
}
__device__ __forceinline__ void split_pack2(float2 a, uint32_t &hi, uint32_t &lo) {
    const __half2 h = __float22half2_rn(a);
    hi = h2u(h);
    const float2 l = __ffma2_rn(__half22float2(h), make_float2(-1.f, -1.f), a);
    lo = h2u(__float22half2_rn(l));
}

template <int V>
__global__ void __launch_bounds__(768, 1) k(float *out, int chunks, long long *cyc, int nwarps_active) {
    extern __shared__ __align__(16) unsigned char smem[];
    float2 *sZb = reinterpret_cast<float2 *>(smem);
    __shared__ uint32_t tmem_base;
    for (int t = threadIdx.x; t < 1024; t += blockDim.x) sZb[t] = make_float2(0.001f * t, 0.5f + t);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&tmem_base, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tmem_base;
    long long t0 = clock64();
    uint32_t accu = 0;
    if (warp < nwarps_active) {
        const int q = warp & 3, g = warp >> 2;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        float x = 0.01f * (threadIdx.x + blockIdx.x);
        for (int c = g; c < chunks; c += nwarps_active / 4) {
            const uint32_t st_addr = tmem + lane_addr + (c % 7) * 64;
            const float2 *zb64 = sZb + (c % 16) * 64;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int cc = 0; cc < 16; ++cc) {
                    float4 v;
                    if (V & 1) v = reinterpret_cast<const float4 *>(zb64 + half * 32)[cc];
                    else v = make_float4(0.37f + cc, 0.11f, 0.41f, 0.2f + cc);
                    if (V & 8) {
                        const float2 arg = __ffma2_rn(make_float2(x, x), make_float2(v.x, v.y), make_float2(v.z, v.w));
                        split_pack2(make_float2(__cosf(arg.x), __cosf(arg.y)), hi[cc], lo[cc]);
                    } else
                    split_pack(__cosf(fmaf(x, v.x, v.y)), __cosf(fmaf(x, v.z, v.w)), hi[cc], lo[cc]);
                }
                if (V & 2) {
                    tmem_st16(st_addr + half * 16, hi);
                    tmem_st16(st_addr + 32 + half * 16, lo);
                } else {
#pragma unroll
                    for (int cc = 0; cc < 16; ++cc) accu ^= hi[cc] + lo[cc];
                }
            }
            if (V & 4) { wait_st(); fence_before_sync(); }
            x += 0.001f;
        }
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)accu;
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int V>
void run(const char *name, int nw) {
    float *out; long long *cyc;
    cudaMalloc(&out, 148 * 768 * 4); cudaMalloc(&cyc, 148 * 8);
    const int chunks = 4096;
    cudaFuncSetAttribute(k<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 150 * 1024);
    k<V><<<148, 768, 150 * 1024>>>(out, 64, cyc, nw);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int rep = 0; rep < 20; ++rep) k<V><<<148, 768, 150 * 1024>>>(out, chunks, cyc, nw);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    printf("%-40s warps=%2d  %.1f clk/chunk  %.2f features/clk/SM  eff clock %.0f MHz %s\n", name, nw, avg / chunks,
           128.0 * 64 * chunks / avg, avg * 20 / (ms * 1e3), e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int nw : {16}) {
        run<0>("math only", nw);
        run<1>("math + LDS table", nw);
        run<3>("math + LDS + tcgen05.st", nw);
        run<7>("math + LDS + tcgen05.st + wait::st", nw);
        run<15>("packed math + LDS + st + wait::st", nw);
    }
    return 0;
}
