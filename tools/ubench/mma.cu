// Micro-benchmark: issue cost / throughput of tcgen05.mma kind::f16 with A in tensor memory and a small N,
// as the forward kernel issues them (M=128, K=16, B from shared memory with the no-swizzle layout).
//   mma <busy>   busy=1: sixteen other warps run a MUFU/FFMA loop meanwhile
#include <cstdio>
#include <cstdlib>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include "../../gpt_b200/csrc/tc_ptx.cuh"
using namespace tcx;

template <int N, int V>
__global__ void __launch_bounds__(768, 1) k(long long *out, float *sink, int iters, int per_commit, int busy) {
    constexpr int nacc = V;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar[8];
    __shared__ uint32_t tbase;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(&bar[i], 1); fence_barrier_init(); }
    for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
    if (warp == 4) tmem_alloc(&tbase, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = tbase;
    if (warp >= 4 && warp < 4 + (V == 5 ? 2 : V == 6 ? 4 : 1)) {
        const int wi = warp - 4;
        constexpr uint32_t idesc = make_idesc_f16(128, N, 0);
        constexpr uint32_t lbo = (2 * N / 8) * 128, sbo = 128;
        const uint64_t bd0 = make_desc(smem_u32(smem), lbo, sbo);
        const long long t0 = clock64();
        uint32_t ph = 0;
        int nb = 0;
        for (int it = 0; it < iters; ++it) {
            if (elect_one()) {
#pragma unroll
                for (int s = 0; s < 12; ++s) {
                    if constexpr (nacc == 1) {          // identical operands: no per-MMA address arithmetic at all
                        mma_f16_ts(tmem + 256, tmem, bd0, idesc, 1);
                    } else if constexpr (nacc == 2) {   // the forward kernel's pattern
                        const uint64_t bd = make_desc(smem_u32(smem) + (s & 3) * 2048, lbo, sbo);
                        mma_f16_ts(tmem + 256, tmem + (s & 7) * 8, bd, idesc, 1);
                    } else if constexpr (nacc == 3) {   // descriptor by 64-bit add on a precomputed base
                        mma_f16_ts(tmem + 256, tmem + (s & 7) * 8, bd0 + (uint64_t)((s & 3) * 128), idesc, 1);
                    } else if constexpr (nacc >= 5) {   // several issuing warps, each its own accumulator
                        mma_f16_ts(tmem + 256 + wi * N, tmem + (s & 7) * 8, bd0 + (uint64_t)((s & 3) * 128), idesc, 1);
                    } else {                  // two accumulators, descriptor by add
                        mma_f16_ts(tmem + 256 + (s & 1) * N, tmem + (s & 7) * 8, bd0 + (uint64_t)((s & 3) * 128), idesc, 1);
                    }
                }
                if ((it + 1) % per_commit == 0) mma_commit(&bar[wi]);
            }
            __syncwarp();
            if ((it + 1) % per_commit == 0) { ++nb; if (per_commit > 1 || true) { mbar_wait_warp(&bar[wi], ph); ph ^= 1; } }
        }
        const long long t1 = clock64();
        if (lane == 0 && wi == 0) out[blockIdx.x] = t1 - t0;
    } else if (warp >= 8 && busy) {
        float a = threadIdx.x * 1e-3f, acc = 0.f;
        for (int i = 0; i < busy; ++i) {
#pragma unroll
            for (int j = 0; j < 8; ++j) { a = fmaf(a, 1.0001f, 0.5f); acc += __cosf(a); }
        }
        if (acc == 12345.f) sink[threadIdx.x] = acc;
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 512);
}

template <int N, int V>
void run(int busy) {
    long long *d;
    float *sink;
    cudaMalloc(&d, 148 * 8);
    cudaMalloc(&sink, 4096);
    cudaFuncSetAttribute(k<N, V>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    {
        const int iters = 512, pc = 4, nacc = V;
        k<N, V><<<148, 768, 65536>>>(d, sink, iters, pc, busy);
        k<N, V><<<148, 768, 65536>>>(d, sink, iters, pc, busy);
        long long h[148];
        cudaError_t e = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
        double s = 0;
        for (int i = 0; i < 148; ++i) s += (double)h[i];
        printf("N=%3d busy=%d variant %d (5: two issuing warps, 6: four): %.1f clk per MMA of one warp (floor N/2 = %d)\n", N, busy, nacc,
               s / 148 / iters / 12, N / 2);
    }
}

int main(int argc, char **argv) {
    const int busy = argc > 1 ? atoi(argv[1]) : 0;
    run<32, 1>(busy); run<32, 3>(busy); run<32, 5>(busy); run<32, 6>(busy);
    run<64, 1>(busy); run<64, 5>(busy); run<64, 6>(busy);
    return 0;
}
