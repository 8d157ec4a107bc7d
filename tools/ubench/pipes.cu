// Micro-benchmark: per-SM throughput of the instruction mixes the feature generators use (B200, sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

template <int MODE>
__global__ void __launch_bounds__(1024) k(float *out, int iters, float z, float b, long long *cyc) {
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = threadIdx.x * 0.001f + i;
    uint32_t acc = 0;
    float facc = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
            float a0, a1;
            if (MODE == 0) {            // FFMA + FMUL + MUFU.COS
                a0 = __cosf(fmaf(x[i], z, b)); a1 = __cosf(fmaf(x[i + 1], z, b));
                facc += a0 + a1;
            } else if (MODE == 1) {     // + one F2FP pack (TC1 body)
                a0 = __cosf(fmaf(x[i], z, b)); a1 = __cosf(fmaf(x[i + 1], z, b));
                __half2 h = __floats2half2_rn(a0, a1);
                acc ^= *reinterpret_cast<uint32_t *>(&h);
            } else if (MODE == 2) {     // full fp16 hi/lo split (TC3 body)
                a0 = __cosf(fmaf(x[i], z, b)); a1 = __cosf(fmaf(x[i + 1], z, b));
                __half2 h = __floats2half2_rn(a0, a1);
                float2 hf = __half22float2(h);
                __half2 l = __floats2half2_rn(a0 - hf.x, a1 - hf.y);
                acc ^= *reinterpret_cast<uint32_t *>(&h) + *reinterpret_cast<uint32_t *>(&l);
            } else if (MODE == 3) {     // pack only (no MUFU)
                a0 = fmaf(x[i], z, b); a1 = fmaf(x[i + 1], z, b);
                __half2 h = __floats2half2_rn(a0, a1);
                acc ^= *reinterpret_cast<uint32_t *>(&h);
            } else if (MODE == 4) {     // tf32-style split: raw hi, lo = a - (a & mask)
                a0 = __cosf(fmaf(x[i], z, b)); a1 = __cosf(fmaf(x[i + 1], z, b));
                float l0 = a0 - __uint_as_float(__float_as_uint(a0) & 0xFFFFE000u);
                float l1 = a1 - __uint_as_float(__float_as_uint(a1) & 0xFFFFE000u);
                acc ^= __float_as_uint(a0) + __float_as_uint(a1) + __float_as_uint(l0) + __float_as_uint(l1);
            } else if (MODE == 5) {     // split without unpack: hi by magic rounding, 2 packs
                a0 = __cosf(fmaf(x[i], z, b)); a1 = __cosf(fmaf(x[i + 1], z, b));
                float h0 = (a0 + 6144.f) - 6144.f, h1 = (a1 + 6144.f) - 6144.f;
                __half2 h = __floats2half2_rn(h0, h1);
                __half2 l = __floats2half2_rn(a0 - h0, a1 - h1);
                acc ^= *reinterpret_cast<uint32_t *>(&h) + *reinterpret_cast<uint32_t *>(&l);
            } else if (MODE == 6) {     // polynomial cos on the FMA pipe (argument in revolutions), no MUFU
                float t0_ = fmaf(x[i], z, b), t1_ = fmaf(x[i + 1], z, b);
                float r0 = t0_ - rintf(t0_), r1 = t1_ - rintf(t1_);
                float s0 = r0 * r0, s1 = r1 * r1;
                float p0 = fmaf(s0, -1.7143907e+0f, 7.9035363e+0f), p1 = fmaf(s1, -1.7143907e+0f, 7.9035363e+0f);
                p0 = fmaf(p0, s0, -2.6426256e+1f); p1 = fmaf(p1, s1, -2.6426256e+1f);
                p0 = fmaf(p0, s0, 6.0244641e+1f); p1 = fmaf(p1, s1, 6.0244641e+1f);
                p0 = fmaf(p0, s0, -8.5456817e+1f); p1 = fmaf(p1, s1, -8.5456817e+1f);
                p0 = fmaf(p0, s0, 6.4939394e+1f); p1 = fmaf(p1, s1, 6.4939394e+1f);
                p0 = fmaf(p0, s0, -1.9739209e+1f); p1 = fmaf(p1, s1, -1.9739209e+1f);
                a0 = fmaf(p0, s0, 1.f); a1 = fmaf(p1, s1, 1.f);
                facc += a0 + a1;
            } else if (MODE == 7) {     // MUFU.COS alone on a pre-scaled argument chain (ex2 as a stand-in for raw MUFU)
                a0 = exp2f(x[i]); a1 = exp2f(x[i + 1]);
                facc += a0 + a1;
            }
            x[i] += 1.0f; x[i + 1] += 1.0f;
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = facc + (float)acc;
}

template <int MODE>
void run(const char *name, int threads) {
    float *out; long long *cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 4096;
    k<MODE><<<148, threads>>>(out, 16, 0.37f, 0.11f, cyc);
    k<MODE><<<148, threads>>>(out, iters, 0.37f, 0.11f, cyc);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    double feats = (double)threads * iters * 8;
    printf("%-44s threads=%4d  %.2f features/clk/SM  (%.1f clk per warp-feature)\n", name, threads, feats / avg, avg / (feats / 32) );
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int threads : {512, 1024}) {
        run<0>("ffma+fmul+mufu.cos", threads);
        run<1>("cos + pack(hi)            [TC1 body]", threads);
        run<2>("cos + pack+unpack+sub+pack [TC3 body]", threads);
        run<3>("ffma + pack only", threads);
        run<4>("cos + tf32 split (lop3+fadd)", threads);
        run<5>("cos + magic-round split, 2 packs", threads);
        run<6>("polynomial cos (FMA pipe)", threads);
        run<7>("ex2 only", threads);
    }
    return 0;
}
