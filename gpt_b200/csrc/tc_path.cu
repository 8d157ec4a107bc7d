// Tensor-core path (GPTB_PREC_TC3 / TC1): K1 feature x factor forward and K3 factor-gradient
// back-projection on tcgen05, with the random Fourier features generated on the fly.
//
//   K1  T[k][l][i]    = amp * sum_j cos(x_ik*Zt[j,k] + b[j,k]) * U[j,l,k]      feature  GPT_SGLD.jl:71-84
//                                                                            phidotU  GPT_SGLD.jl:193-205
//   K3  gradU[j,l,k] += amp * sum_i cos(x_ik*Zt[j,k] + b[j,k]) * E[k][l][i]    computePsi + grad-U :276-286,:405-408
//
// Neither the N x D x n feature tensor (phi) nor Psi ever exists in memory.  In both kernels the
// features are the A operand of the MMA and are written by the generator warps straight into
// TENSOR MEMORY (tcgen05.st), never through shared memory: with N = r <= 64 the MMA would
// otherwise be bound by the shared-memory reads of A (4 KB per 16-deep K step).  Shared memory
// only carries the small B operand (U or E image chunks, 1-D bulk copies, see tc_layout.cuh).
//
// Precision: fp16 x fp16 -> fp32 MMAs at twice the TF32 rate.  A single pass carries 2^-11 operand
// rounding (~1e-3 after the D-fold product), so the parity mode TC3 splits both operands,
// a = hi + lo (fp16 each), and issues hi*[hi|lo] (N = 2*rpad) and lo*hi (N = rpad):
// 22+ significant bits at 1.5x the cost of one TF32 pass.  cos is MUFU.COS on an FP32 argument.
//
// CTA = 512 threads, one CTA per SM (persistent):
//   warps 0-3   epilogue (TMEM -> registers -> global), one TMEM lane quarter each
//   warp  4     MMA issuer (one elected lane), also owns the TMEM allocation
//   warp  5     B-operand loader (cp.async.bulk + mbarrier)
//   warps 8-15  two generator warpgroups, alternating K chunks of 64
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_layout.cuh"
#include "tc_ptx.cuh"

namespace {

using namespace tcx;

constexpr int TC_THREADS = 512;
constexpr int KC = 64;        // K elements per chunk (4 MMA K-steps of 16)
constexpr int NSTA = 4;       // A stages in TMEM (64 columns each: 32 hi + 32 lo)
constexpr int NSTB = 4;       // B stages in shared memory
constexpr int A_COLS = NSTA * 64;

struct Bars {
    uint64_t a_full[NSTA], a_empty[NSTA], b_full[NSTB], b_empty[NSTB], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
};

__device__ __forceinline__ uint32_t h2_u32(__half2 h) { return *reinterpret_cast<uint32_t *>(&h); }

// hi/lo fp16 split of two fp32 values, packed (first value in the low half).
template <bool THREE>
__device__ __forceinline__ void split_pack(float a0, float a1, uint32_t &hi, uint32_t &lo) {
    __half2 h = __floats2half2_rn(a0, a1);
    hi = h2_u32(h);
    if (THREE) {
        float2 hf = __half22float2(h);
        lo = h2_u32(__floats2half2_rn(a0 - hf.x, a1 - hf.y));
    }
}

struct FwdParams {
    const float *XbF;      // [D][mpad]
    const __half *Uimg;    // see tc_layout.cuh
    const float2 *ZbF;     // [D][npad] (zt, b) pairs; for sin/cos features pair p = (zt_p, unused)
    const float *uscale;   // -> {uscale, 1/uscale}
    float *T;              // [D][r][mpad]
    float amp;
    int D, n, npad, r;
    int64_t bs, mpad, mext;   // live rows, row stride of XbF/T, rows processed (bs rounded up to 128)
    int sincos;
    int swap_lbo_sbo;      // probe switch
};

template <int RPAD, bool THREE>
__global__ void __launch_bounds__(TC_THREADS, 1) fwd_tc_kernel(const FwdParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr int NB = THREE ? 2 * RPAD : RPAD;          // MMA N of the wide pass / accumulator width
    constexpr int B_STAGE_BYTES = 2 * RPAD * KC * 2;     // image always carries hi|lo
    constexpr int ACC_COLS = NB;
    unsigned char *sB = smem_raw;
    float2 *sZb = reinterpret_cast<float2 *>(smem_raw + NSTB * B_STAGE_BYTES);
    Bars *bars = reinterpret_cast<Bars *>(smem_raw + NSTB * B_STAGE_BYTES + (size_t)p.D * p.npad * sizeof(float2));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntiles = (int)(p.mext / 128);
    const int nunits = ntiles * p.D;
    const int NCH = p.npad / KC;

    for (int t = tid; t < p.D * p.npad; t += TC_THREADS) sZb[t] = p.ZbF[t];
    if (tid == 0) {
        for (int s = 0; s < NSTA; ++s) { mbar_init(&bars->a_full[s], 128); mbar_init(&bars->a_empty[s], 1); }
        for (int s = 0; s < NSTB; ++s) { mbar_init(&bars->b_full[s], 1); mbar_init(&bars->b_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&bars->acc_full[s], 1); mbar_init(&bars->acc_empty[s], 128); }
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(&bars->tmem_base, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = bars->tmem_base;
    const uint32_t acc_base = tmem + A_COLS;

    if (warp >= 8) {
        // ---------------------------------------------------------------- feature generators
        const int g = (warp - 8) >> 2;
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        uint32_t cg = 0;
        for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
            const int tile = u / p.D, k = u % p.D;
            const float x = p.XbF[(size_t)k * p.mpad + (size_t)tile * 128 + row];
            for (int ch = 0; ch < NCH; ++ch, ++cg) {
                if ((int)(cg & 1) != g) continue;
                const int sa = cg % NSTA;
                mbar_wait(&bars->a_empty[sa], ((cg / NSTA) & 1) ^ 1);
                fence_after_sync();
                const uint32_t st_addr = tmem + lane_addr + sa * 64;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t hi[16], lo[16];
                    if (!p.sincos) {
                        const float4 *zb = reinterpret_cast<const float4 *>(sZb + (size_t)k * p.npad + ch * KC + half * 32);
#pragma unroll
                        for (int c = 0; c < 16; ++c) {
                            float4 v = zb[c];
                            float a0 = __cosf(fmaf(x, v.x, v.y));
                            float a1 = __cosf(fmaf(x, v.z, v.w));
                            split_pack<THREE>(a0, a1, hi[c], lo[c]);
                        }
                    } else {
                        const float2 *zb = sZb + (size_t)k * p.npad + (ch * KC + half * 32) / 2;
#pragma unroll
                        for (int c = 0; c < 16; ++c) {
                            float a = x * zb[c].x;
                            split_pack<THREE>(__sinf(a), __cosf(a), hi[c], lo[c]);
                        }
                    }
                    tmem_st16(st_addr + half * 16, hi);
                    if (THREE) tmem_st16(st_addr + 32 + half * 16, lo);
                }
                wait_st();
                fence_before_sync();
                mbar_arrive(&bars->a_full[sa]);
            }
        }
    } else if (warp == 4) {
        // ---------------------------------------------------------------- MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc_wide = make_idesc_f16(128, NB, 0);
            constexpr uint32_t idesc_narrow = make_idesc_f16(128, RPAD, 0);
            uint32_t lbo = (2 * RPAD / 8) * 128, sbo = 128;
            if (p.swap_lbo_sbo) { uint32_t t2 = lbo; lbo = sbo; sbo = t2; }
            uint32_t cg = 0, it = 0;
            for (int u = blockIdx.x; u < nunits; u += gridDim.x, ++it) {
                const uint32_t ab = it & 1;
                mbar_wait(&bars->acc_empty[ab], ((it >> 1) & 1) ^ 1);
                fence_after_sync();
                const uint32_t d = acc_base + ab * ACC_COLS;
                for (int ch = 0; ch < NCH; ++ch, ++cg) {
                    const int sa = cg % NSTA, sb = cg % NSTB;
                    mbar_wait(&bars->b_full[sb], (cg / NSTB) & 1);
                    mbar_wait(&bars->a_full[sa], (cg / NSTA) & 1);
                    fence_after_sync();
                    const uint32_t a_hi = tmem + sa * 64, a_lo = a_hi + 32;
                    const uint32_t b_addr = smem_u32(sB + sb * B_STAGE_BYTES);
#pragma unroll
                    for (int s4 = 0; s4 < KC / 16; ++s4) {
                        const uint64_t bd = make_desc(b_addr + s4 * 2 * ((2 * RPAD / 8) * 128), lbo, sbo);
                        mma_f16_ts(d, a_hi + s4 * 8, bd, idesc_wide, (ch | s4) != 0);
                        if (THREE) mma_f16_ts(d, a_lo + s4 * 8, bd, idesc_narrow, 1);
                    }
                    mma_commit(&bars->a_empty[sa]);
                    mma_commit(&bars->b_empty[sb]);
                }
                mma_commit(&bars->acc_full[ab]);
            }
        }
        __syncwarp();
    } else if (warp == 5) {
        // ---------------------------------------------------------------- B loader
        if (lane == 0) {
            uint32_t cg = 0;
            for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
                const int k = u % p.D;
                const __half *src = p.Uimg + (size_t)k * p.npad * 2 * RPAD;
                for (int ch = 0; ch < NCH; ++ch, ++cg) {
                    const int sb = cg % NSTB;
                    mbar_wait(&bars->b_empty[sb], ((cg / NSTB) & 1) ^ 1);
                    mbar_expect_tx(&bars->b_full[sb], B_STAGE_BYTES);
                    bulk_g2s(sB + sb * B_STAGE_BYTES, src + (size_t)ch * KC * 2 * RPAD, B_STAGE_BYTES, &bars->b_full[sb]);
                }
            }
        }
        __syncwarp();
    } else if (warp < 4) {
        // ---------------------------------------------------------------- epilogue
        const int row = warp * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
        const float scale = p.amp * p.uscale[1];
        uint32_t it = 0;
        for (int u = blockIdx.x; u < nunits; u += gridDim.x, ++it) {
            const int tile = u / p.D, k = u % p.D;
            const uint32_t ab = it & 1;
            mbar_wait(&bars->acc_full[ab], (it >> 1) & 1);
            fence_after_sync();
            const uint32_t d = acc_base + ab * ACC_COLS + lane_addr;
            const int64_t i = (int64_t)tile * 128 + row;
            const bool live = i < p.bs;
            float *Tk = p.T + (size_t)k * p.r * p.mpad + i;
#pragma unroll
            for (int c0 = 0; c0 < RPAD; c0 += 16) {
                uint32_t v[16], v2[16];
                tmem_ld16(d + c0, v);
                if (THREE) tmem_ld16(d + RPAD + c0, v2);
                wait_ld();
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const int l = c0 + c;
                    float val = __uint_as_float(v[c]);
                    if (THREE) val += __uint_as_float(v2[c]);
                    if (l < p.r) Tk[(size_t)l * p.mpad] = live ? val * scale : 0.f;
                }
            }
            fence_before_sync();
            mbar_arrive(&bars->acc_empty[ab]);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 512);
}

struct BwdParams {
    const float *XbF;      // [D][mpad]
    const __half *Eimg;    // see tc_layout.cuh
    const float *ZtF, *bF; // [D][npad]
    const float *escale;   // -> {escale, 1/escale}
    double *gradU;         // [D][r][n], accumulated with atomics
    float amp;
    int D, n, npad, r;
    int64_t mpad, mext;
    int sincos;
    int swap_lbo_sbo;
};

template <int RPAD, bool THREE>
__global__ void __launch_bounds__(TC_THREADS, 1) bwd_tc_kernel(const BwdParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr int NB = THREE ? 2 * RPAD : RPAD;
    constexpr int B_IMG_BYTES = 2 * RPAD * KC * 2;
    constexpr int X_BYTES = KC * 4;                         // the chunk's 64 input values ride in the same stage
    constexpr int B_STAGE_BYTES = B_IMG_BYTES + X_BYTES;
    constexpr int ACC_COLS = NB;
    unsigned char *sB = smem_raw;
    Bars *bars = reinterpret_cast<Bars *>(smem_raw + NSTB * B_STAGE_BYTES);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int JB = p.npad / 128;
    const int NCH = (int)(p.mext / KC);
    const int64_t total = (int64_t)p.D * JB * NCH;
    const int64_t it0 = total * blockIdx.x / gridDim.x, it1 = total * (blockIdx.x + 1) / gridDim.x;

    if (tid == 0) {
        for (int s = 0; s < NSTA; ++s) { mbar_init(&bars->a_full[s], 128); mbar_init(&bars->a_empty[s], 1); }
        for (int s = 0; s < NSTB; ++s) { mbar_init(&bars->b_full[s], 1); mbar_init(&bars->b_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&bars->acc_full[s], 1); mbar_init(&bars->acc_empty[s], 128); }
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(&bars->tmem_base, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = bars->tmem_base;
    const uint32_t acc_base = tmem + A_COLS;

    if (warp >= 8) {
        // ---------------------------------------------------------------- feature generators (lane = feature)
        const int g = (warp - 8) >> 2;
        const int q = warp & 3;
        const int jl = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        int cur_grp = -1;
        float zt = 0.f, bb = 0.f;
        int k = 0;
        uint32_t cnt = 0;
        for (int64_t it = it0; it < it1; ++it, ++cnt) {
            const int grp = (int)(it / NCH), ch = (int)(it % NCH);
            if (grp != cur_grp) {
                cur_grp = grp;
                k = grp / JB;
                const int j = (grp % JB) * 128 + jl;
                if (!p.sincos) {
                    zt = p.ZtF[(size_t)k * p.npad + j];
                    bb = p.bF[(size_t)k * p.npad + j];
                } else {  // rows 2p / 2p+1 are sin / cos of x*Zt[p]:  cos(a) = sin(a + pi/2)
                    zt = p.ZtF[(size_t)k * p.npad + (j >> 1)];
                    bb = (j & 1) ? 1.5707963267948966f : 0.f;
                }
            }
            if ((int)(cnt & 1) != g) continue;
            const int sa = cnt % NSTA;
            mbar_wait(&bars->a_empty[sa], ((cnt / NSTA) & 1) ^ 1);
            fence_after_sync();
            const uint32_t st_addr = tmem + lane_addr + sa * 64;
            const int sbx = cnt % NSTB;
            mbar_wait(&bars->b_full[sbx], (cnt / NSTB) & 1);
            const float4 *xr = reinterpret_cast<const float4 *>(sB + sbx * B_STAGE_BYTES + B_IMG_BYTES);
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int c = 0; c < 16; c += 2) {
                    float4 xx = xr[half * 8 + c / 2];
                    float a0, a1, a2, a3;
                    if (!p.sincos) {
                        a0 = __cosf(fmaf(xx.x, zt, bb)); a1 = __cosf(fmaf(xx.y, zt, bb));
                        a2 = __cosf(fmaf(xx.z, zt, bb)); a3 = __cosf(fmaf(xx.w, zt, bb));
                    } else {
                        a0 = __sinf(fmaf(xx.x, zt, bb)); a1 = __sinf(fmaf(xx.y, zt, bb));
                        a2 = __sinf(fmaf(xx.z, zt, bb)); a3 = __sinf(fmaf(xx.w, zt, bb));
                    }
                    split_pack<THREE>(a0, a1, hi[c], lo[c]);
                    split_pack<THREE>(a2, a3, hi[c + 1], lo[c + 1]);
                }
                tmem_st16(st_addr + half * 16, hi);
                if (THREE) tmem_st16(st_addr + 32 + half * 16, lo);
            }
            wait_st();
            fence_before_sync();
            mbar_arrive(&bars->a_full[sa]);
        }
    } else if (warp == 4) {
        // ---------------------------------------------------------------- MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc_wide = make_idesc_f16(128, NB, 1);
            constexpr uint32_t idesc_narrow = make_idesc_f16(128, RPAD, 1);
            uint32_t lbo = (2 * RPAD / 8) * 128, sbo = 128;
            if (p.swap_lbo_sbo) { uint32_t t2 = lbo; lbo = sbo; sbo = t2; }
            uint32_t cnt = 0, seg = 0;
            bool fresh = true;
            for (int64_t it = it0; it < it1; ++it, ++cnt) {
                const int ch = (int)(it % NCH);
                const uint32_t ab = seg & 1;
                if (fresh) {
                    mbar_wait(&bars->acc_empty[ab], ((seg >> 1) & 1) ^ 1);
                    fence_after_sync();
                }
                const uint32_t d = acc_base + ab * ACC_COLS;
                const int sa = cnt % NSTA, sb = cnt % NSTB;
                mbar_wait(&bars->b_full[sb], (cnt / NSTB) & 1);
                mbar_wait(&bars->a_full[sa], (cnt / NSTA) & 1);
                fence_after_sync();
                const uint32_t a_hi = tmem + sa * 64, a_lo = a_hi + 32;
                const uint32_t b_addr = smem_u32(sB + sb * B_STAGE_BYTES);
#pragma unroll
                for (int s4 = 0; s4 < KC / 16; ++s4) {
                    const uint64_t bd = make_desc(b_addr + s4 * 2 * ((2 * RPAD / 8) * 128), lbo, sbo);
                    mma_f16_ts(d, a_hi + s4 * 8, bd, idesc_wide, (!fresh || s4 != 0) ? 1u : 0u);
                    if (THREE) mma_f16_ts(d, a_lo + s4 * 8, bd, idesc_narrow, 1);
                }
                fresh = false;
                mma_commit(&bars->a_empty[sa]);
                mma_commit(&bars->b_empty[sb]);
                if (ch == NCH - 1 || it == it1 - 1) {
                    mma_commit(&bars->acc_full[ab]);
                    ++seg;
                    fresh = true;
                }
            }
        }
        __syncwarp();
    } else if (warp == 5) {
        // ---------------------------------------------------------------- B loader
        if (lane == 0) {
            uint32_t cnt = 0;
            for (int64_t it = it0; it < it1; ++it, ++cnt) {
                const int grp = (int)(it / NCH), ch = (int)(it % NCH);
                const int k = grp / JB;
                const __half *src = p.Eimg + (size_t)k * p.mpad * 2 * RPAD + (size_t)ch * KC * 2 * RPAD;
                const int sb = cnt % NSTB;
                mbar_wait(&bars->b_empty[sb], ((cnt / NSTB) & 1) ^ 1);
                mbar_expect_tx(&bars->b_full[sb], B_STAGE_BYTES);
                bulk_g2s(sB + sb * B_STAGE_BYTES + B_IMG_BYTES, p.XbF + (size_t)k * p.mpad + (size_t)ch * KC, X_BYTES, &bars->b_full[sb]);
                bulk_g2s(sB + sb * B_STAGE_BYTES, src, B_IMG_BYTES, &bars->b_full[sb]);
            }
        }
        __syncwarp();
    } else if (warp < 4) {
        // ---------------------------------------------------------------- epilogue (lane = feature)
        const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
        const float scale = p.amp * p.escale[1];
        uint32_t seg = 0;
        for (int64_t it = it0; it < it1; ++seg) {
            const int grp = (int)(it / NCH);
            const int64_t seg_end = min((long long)it1, (long long)(grp + 1) * NCH);
            it = seg_end;
            const int k = grp / JB;
            const int j = (grp % JB) * 128 + warp * 32 + lane;
            const uint32_t ab = seg & 1;
            mbar_wait(&bars->acc_full[ab], (seg >> 1) & 1);
            fence_after_sync();
            const uint32_t d = acc_base + ab * ACC_COLS + lane_addr;
            double *gk = p.gradU + (size_t)k * p.r * p.n + j;
#pragma unroll
            for (int c0 = 0; c0 < RPAD; c0 += 16) {
                uint32_t v[16], v2[16];
                tmem_ld16(d + c0, v);
                if (THREE) tmem_ld16(d + RPAD + c0, v2);
                wait_ld();
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const int l = c0 + c;
                    float val = __uint_as_float(v[c]);
                    if (THREE) val += __uint_as_float(v2[c]);
                    if (l < p.r && j < p.n) atomicAdd(gk + (size_t)l * p.n, (double)(val * scale));
                }
            }
            fence_before_sync();
            mbar_arrive(&bars->acc_empty[ab]);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------- packing
__global__ void absmax_kernel(const double *__restrict__ a, size_t n, float *__restrict__ out) {
    float m = 0.f;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        m = fmaxf(m, fabsf((float)a[i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(reinterpret_cast<int *>(out), __float_as_int(m));
}

// power-of-two scale that maps `mx` into [256, 512): fp16 keeps 11 bits for hi and the lo part stays normal.
__device__ __forceinline__ float pow2_scale_for(float mx) {
    if (!(mx > 0.f) || !isfinite(mx)) return 1.f;
    int e;
    frexpf(mx, &e);  // mx = f * 2^e, f in [0.5,1)
    return ldexpf(1.f, 9 - e);
}

__global__ void set_scale_kernel(float *state_max, float *state_scale, int reset_max) {
    float s = pow2_scale_for(*state_max);
    state_scale[0] = s;
    state_scale[1] = 1.f / s;
    if (reset_max) *state_max = 0.f;
}

__global__ void pack_U_kernel(const double *__restrict__ U /*[D][r][n]*/, int D, int n, int r, int npad, int rpad,
                              const float *__restrict__ scale, __half *__restrict__ Uimg) {
    const float sc = scale[0];
    const size_t total = (size_t)D * npad * rpad;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        int j = (int)(t % npad);
        int l = (int)((t / npad) % rpad);
        int k = (int)(t / ((size_t)npad * rpad));
        float v = (j < n && l < r) ? (float)(U[((size_t)k * r + l) * n + j] * (double)sc) : 0.f;
        __half hh = __float2half_rn(v);
        __half ll = __float2half_rn(v - __half2float(hh));
        Uimg[gptb_uimg_offset(k, j, l, npad, rpad)] = hh;
        Uimg[gptb_uimg_offset(k, j, rpad + l, npad, rpad)] = ll;
    }
}

__global__ void interleave_zb_kernel(const float *zt, const float *b, size_t n, float2 *out) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = make_float2(zt[i], b[i]);
}

int probe_swap() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("GPTB_TC_SWAP_LBO_SBO");
        v = (e && e[0] == '1') ? 1 : 0;
    }
    return v;
}

}  // namespace

size_t gptb_tc_uimg_bytes(const gptb_handle *h) { return (size_t)h->cfg.D * h->npad * 2 * h->rpad * sizeof(__half); }
size_t gptb_tc_eimg_bytes(const gptb_handle *h, int64_t mpad) {
    return (size_t)h->cfg.D * mpad * 2 * h->rpad * sizeof(__half);
}

// state slots: [0] max|E| running, [1..2] escale, 1/escale, [3] max|U|, [4..5] uscale, 1/uscale (sampler params),
//              [6] max|U| (prediction params), [7..8] its scales
int gptb_tc_pack_U_slot(gptb_handle *h, const double *U, void *Uimg, int max_slot, int scale_slot) {
    const gptb_config &c = h->cfg;
    const size_t cnt = (size_t)c.D * c.r * c.n;
    GPTB_CUDA_OK(h, cudaMemsetAsync(h->tc_state + max_slot, 0, sizeof(float), h->stream));
    absmax_kernel<<<std::min<int>(h->sm_count, (int)((cnt + 1023) / 1024)), 1024, 0, h->stream>>>(U, cnt, h->tc_state + max_slot);
    set_scale_kernel<<<1, 1, 0, h->stream>>>(h->tc_state + max_slot, h->tc_state + scale_slot, 0);
    const size_t total = (size_t)c.D * h->npad * h->rpad;
    pack_U_kernel<<<(int)std::min<size_t>((total + 255) / 256, (size_t)h->sm_count * 8), 256, 0, h->stream>>>(
        U, c.D, c.n, c.r, h->npad, h->rpad, h->tc_state + scale_slot, (__half *)Uimg);
    h->launches += 3;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_tc_pack_params(gptb_handle *h) { return gptb_tc_pack_U_slot(h, h->U, h->Uimg, 3, 4); }
int gptb_tc_pack_U(gptb_handle *h, const double *U, void *Uimg) { return gptb_tc_pack_U_slot(h, U, Uimg, 6, 7); }

// Turn the max|E| measured by the previous contraction into this step's power-of-two image scale.
int gptb_tc_update_escale(gptb_handle *h) {
    set_scale_kernel<<<1, 1, 0, h->stream>>>(h->tc_state + 0, h->tc_state + 1, 1);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_tc_prepare_features(gptb_handle *h) {
    const size_t cnt = (size_t)h->cfg.D * h->npad;
    if (!h->ZbF) GPTB_CUDA_OK(h, cudaMalloc((void **)&h->ZbF, cnt * sizeof(float2)));
    interleave_zb_kernel<<<64, 256, 0, h->stream>>>(h->ZtF, h->bF, cnt, (float2 *)h->ZbF);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

template <int RPAD>
static int launch_fwd_r(gptb_handle *h, const FwdParams &p, int grid, size_t smem) {
    if (h->cfg.precision == GPTB_PREC_TC3) {
        auto k = fwd_tc_kernel<RPAD, true>;
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<grid, TC_THREADS, smem, h->stream>>>(p);
    } else {
        auto k = fwd_tc_kernel<RPAD, false>;
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<grid, TC_THREADS, smem, h->stream>>>(p);
    }
    return GPTB_OK;
}

template <int RPAD>
static int launch_bwd_r(gptb_handle *h, const BwdParams &p, int grid, size_t smem) {
    if (h->cfg.precision == GPTB_PREC_TC3) {
        auto k = bwd_tc_kernel<RPAD, true>;
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<grid, TC_THREADS, smem, h->stream>>>(p);
    } else {
        auto k = bwd_tc_kernel<RPAD, false>;
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<grid, TC_THREADS, smem, h->stream>>>(p);
    }
    return GPTB_OK;
}

int gptb_launch_fwd_tc(gptb_handle *h, const float *XbF, const void *Uimg, int64_t bs, int64_t mpad, float *T) {
    const gptb_config &c = h->cfg;
    if (!h->ZbF_valid) {
        GPTB_TRY(gptb_tc_prepare_features(h));
        h->ZbF_valid = true;
    }
    FwdParams p;
    p.XbF = XbF;
    p.Uimg = (const __half *)Uimg;
    p.ZbF = (const float2 *)h->ZbF;
    p.uscale = h->tc_state + (Uimg == h->Uimg ? 4 : 7);
    p.T = T;
    p.amp = (float)h->amp;
    p.D = c.D; p.n = c.n; p.npad = h->npad; p.r = c.r;
    p.bs = bs;
    p.mpad = mpad;
    p.mext = (bs + 127) / 128 * 128;   // only the 128-row tiles that hold live rows are generated
    p.sincos = c.feature_kind == GPTB_FEAT_SINCOS;
    p.swap_lbo_sbo = probe_swap();
    const int nunits = (int)(p.mext / 128) * c.D;
    const FwdParams &q = p;
    const int grid = std::max(1, std::min(h->sm_count, nunits));
    // at least 120 KB so that two CTAs (each wants all 512 TMEM columns) never share an SM
    const size_t smem = std::max<size_t>((size_t)NSTB * (2 * h->rpad * KC * 2) + (size_t)c.D * h->npad * sizeof(float2) + sizeof(Bars) + 64, 120 * 1024);
    if (smem > 227 * 1024) {
        gptb_set_error(h, "tensor-core forward: D*n=%d too large for the shared-memory feature table", c.D * h->npad);
        return GPTB_ERR_UNSUPPORTED;
    }
    int st;
    switch (h->rpad) {
        case 16: st = launch_fwd_r<16>(h, q, grid, smem); break;
        case 32: st = launch_fwd_r<32>(h, q, grid, smem); break;
        case 48: st = launch_fwd_r<48>(h, q, grid, smem); break;
        default: st = launch_fwd_r<64>(h, q, grid, smem); break;
    }
    GPTB_TRY(st);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_bwd_tc(gptb_handle *h, const float *XbF, const void *Eimg, int64_t bs, int64_t mpad, double *gradU) {
    const gptb_config &c = h->cfg;
    BwdParams p;
    p.mext = (bs + 127) / 128 * 128;
    p.XbF = XbF;
    p.Eimg = (const __half *)Eimg;
    p.ZtF = h->ZtF;
    p.bF = h->bF;
    p.escale = h->tc_state + 1;
    p.gradU = gradU;
    p.amp = (float)h->amp;
    p.D = c.D; p.n = c.n; p.npad = h->npad; p.r = c.r;
    p.mpad = mpad;
    p.sincos = c.feature_kind == GPTB_FEAT_SINCOS;
    p.swap_lbo_sbo = probe_swap();
    const int64_t total = (int64_t)c.D * (h->npad / 128) * (p.mext / KC);
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(h->sm_count, total));
    const size_t smem = std::max<size_t>((size_t)NSTB * (2 * h->rpad * KC * 2) + sizeof(Bars) + 64, 120 * 1024);
    int st;
    switch (h->rpad) {
        case 16: st = launch_bwd_r<16>(h, p, grid, smem); break;
        case 32: st = launch_bwd_r<32>(h, p, grid, smem); break;
        case 48: st = launch_bwd_r<48>(h, p, grid, smem); break;
        default: st = launch_bwd_r<64>(h, p, grid, smem); break;
    }
    GPTB_TRY(st);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}
