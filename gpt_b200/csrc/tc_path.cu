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
// otherwise be bound by the shared-memory reads of A (4 KB per 16-deep K step).
//
// Both kernels are bound by MUFU.COS (16/clk/SM) when the B operand stays out of the way, so the
// B operand is made to move as little as possible (measured: streaming U chunks per 128-row tile
// made K1 L2-bound at 4.3 TB/s):
//   K1  a CTA owns a contiguous range of (k, row tile) units in k-major order and keeps the whole
//       operand image of U^(k) resident in shared memory (reloaded only when k changes);
//   K3  a CTA owns a contiguous range of (k, feature-block group, row chunk) and uses every E chunk
//       it loads for all feature blocks of the group (one accumulator per block in TMEM).
//
// Precision: fp16 x fp16 -> fp32 MMAs.  A single pass carries 2^-11 operand rounding (~1e-3 after
// the D-fold product), so the parity mode TC3 splits both operands, a = hi + lo (fp16 each), and
// accumulates hi*hi + hi*lo + lo*hi into the same fp32 accumulator: 22 significant bits.
// cos is MUFU.COS on an FP32 argument.
//
// CTA = 768 threads, one CTA per SM (persistent):
//   warps 0-3   epilogue (TMEM -> registers -> global), one TMEM lane quarter each
//   warp  4     MMA issuer (one elected lane), also owns the TMEM allocation
//   warp  5     B-operand loader (cp.async.bulk + mbarrier)
//   warps 8-23  four generator warpgroups, taking K chunks of 64 round-robin
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_layout.cuh"
#include "tc_ptx.cuh"

namespace {

using namespace tcx;

#ifndef GPTB_HSPLIT
#define GPTB_HSPLIT 1
#endif
constexpr int HSPLIT = GPTB_HSPLIT;     // warps per TMEM lane quarter that share one chunk (each takes 64/HSPLIT of its K columns)
constexpr int NGEN = 4 / HSPLIT;   // generator groups (4*HSPLIT warps each); chunks go to the groups round-robin.
                              // Fewer, wider groups keep fewer A stages "in generation" and leave more of the ring
                              // as slack between the generators and the MMA issuers.
constexpr int TC_THREADS = 256 + 512;
// K1 may run more generator groups than K3: the generator warps are bound one by one (dependent chains, barrier round
// trips), not by the SM sub-partition, and K1 has the A stages for a fifth group (K3's TMEM is half accumulators).
#ifndef GPTB_NGEN_FWD
#define GPTB_NGEN_FWD 4
#endif
constexpr int NGEN_F = GPTB_NGEN_FWD / HSPLIT;
constexpr int TC_THREADS_F = 256 + 128 * GPTB_NGEN_FWD;
constexpr int KC = 64;        // K elements per chunk (4 MMA K-steps of 16)
constexpr int SA_COLS = 64;   // TMEM columns of one A stage (32 hi + 32 lo, two fp16 per column)
constexpr int NSTA_MAX = 8;
constexpr int NSTB = 8;       // B stages in shared memory (streaming kernels)

__host__ __device__ constexpr int nsta_for(int acc_cols) {
    return (512 - acc_cols) / SA_COLS > NSTA_MAX ? NSTA_MAX : (512 - acc_cols) / SA_COLS;
}

struct Bars {
    uint64_t a_full[NSTA_MAX], a_empty[NSTA_MAX], b_full[NSTB], b_empty[NSTB], acc_full[2], acc_empty[2];
    uint64_t u_full, u_empty;
    uint32_t tmem_base;
    uint32_t pad;
};

__device__ __forceinline__ uint32_t h2_u32(__half2 h) { return *reinterpret_cast<uint32_t *>(&h); }

// hi/lo fp16 split of two fp32 values, packed (first value in the low half).  The generators are bound by
// instruction issue as much as by MUFU, so everything that can is done two features per instruction
// (FFMA2 / F2FP pack): per feature 0.5 FFMA2 + FMUL + MUFU + 0.5 F2FP + HADD2.F32 + 0.5 FFMA2 + 0.5 F2FP.
template <bool THREE>
__device__ __forceinline__ void split_pack(float2 a, uint32_t &hi, uint32_t &lo) {
    const __half2 h = __float22half2_rn(a);
    hi = h2_u32(h);
    if (THREE) {
        const float2 l = __ffma2_rn(__half22float2(h), make_float2(-1.f, -1.f), a);
        lo = h2_u32(__float22half2_rn(l));
    }
}

// The MMAs of one 16-deep K step, all into accumulator d (RPAD columns):
//   A_hi*B_hi (+ A_hi*B_lo + A_lo*B_hi in TC3).  b_addr = shared address of the K step's image rows.
template <int RPAD, bool THREE, int B_MN_MAJOR>
__device__ __forceinline__ void mma_kstep(uint32_t d, uint32_t a_hi, uint32_t a_lo, uint32_t b_addr, uint32_t acc) {
    constexpr uint32_t idesc = make_idesc_f16(128, RPAD, B_MN_MAJOR);
    constexpr uint32_t lbo = (2 * RPAD / 8) * 128, sbo = 128;
    const uint64_t bd_hi = make_desc(b_addr, lbo, sbo);
    mma_f16_ts(d, a_hi, bd_hi, idesc, acc);
    if (THREE) {
        const uint64_t bd_lo = make_desc(b_addr + (RPAD / 8) * 128, lbo, sbo);
        mma_f16_ts(d, a_hi, bd_lo, idesc, 1);
        mma_f16_ts(d, a_lo, bd_hi, idesc, 1);
    }
}

// ============================================================================================ K1
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
    int P, NCHp;           // K split: P parts of NCHp chunks each (P > 1 only when the image of U^(k) exceeds shared memory;
                           // T is then accumulated with atomics into a zeroed buffer)
};

// 64 features of one row -> the hi / lo words of one A stage.
// cos features: table entry p (float4) = (zt[2p], zt[2p+1], b[2p], b[2p+1]);  sin/cos: entry p (float2) = (zt[p], -).
template <bool THREE>
__device__ __forceinline__ void gen_fwd_chunk(const float2 *zb64, float x, bool sincos, uint32_t st_addr, int half0) {
    const float2 x2 = make_float2(x, x);
#pragma unroll
    for (int hh = 0; hh < 2 / HSPLIT; ++hh) {
        const int half = half0 + hh;
        uint32_t hi[16], lo[16];
        if (!sincos) {
            const float4 *zb = reinterpret_cast<const float4 *>(zb64 + half * 32);
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const float4 v = zb[c];
                const float2 arg = __ffma2_rn(x2, make_float2(v.x, v.y), make_float2(v.z, v.w));
                split_pack<THREE>(make_float2(__cosf(arg.x), __cosf(arg.y)), hi[c], lo[c]);
            }
        } else {
            const float2 *zb = zb64 + half * 16;   // pair p of this chunk -> features 2p (sin), 2p+1 (cos)
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                const float a = x * zb[c].x;
                split_pack<THREE>(make_float2(__sinf(a), __cosf(a)), hi[c], lo[c]);
            }
        }
        tmem_st16(st_addr + half * 16, hi);
        if (THREE) tmem_st16(st_addr + 32 + half * 16, lo);
    }
}

// Resident-U forward.  Units u = (k*P + part)*ntiles + tile, CTA c owns [nunits*c/G, nunits*(c+1)/G).
//
// One thread can issue a small-N tcgen05.mma only every ~44 clk (~65 clk while the generators keep its SM
// sub-partition busy; tools/ubench/mma.cu), three times the tensor pipe's N/2 clk floor at N = 32, and twelve of
// them per chunk made the single issuer the bottleneck of the kernel.  For RPAD <= 64 two warps issue (NI = 2):
// issuer w takes the chunks ch = w (mod 2) of every unit and accumulates into its own columns, the epilogue adds
// the two partial tiles.
template <int RPAD, bool THREE>
__global__ void __launch_bounds__(TC_THREADS_F, 1) fwd_tc_kernel(const FwdParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr int NI = RPAD <= 64 ? 2 : 1;
    constexpr int NSTA = nsta_for(2 * NI * RPAD);
    constexpr int ACC_BASE = NSTA * SA_COLS;
    constexpr int ROW_BYTES = 2 * RPAD * 2;                  // image bytes per feature j (hi|lo)
    const size_t u_cap = (size_t)p.NCHp * KC * ROW_BYTES;   // shared-memory bytes of one K part of the image
    unsigned char *sU = smem_raw;
    float2 *sZb = reinterpret_cast<float2 *>(smem_raw + u_cap);
    Bars *bars = reinterpret_cast<Bars *>(smem_raw + u_cap + (size_t)p.NCHp * KC * sizeof(float2));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntiles = (int)(p.mext / 128);
    const int nunits = ntiles * p.D * p.P;
    const int NCH = p.npad / KC;
    const int u0 = (int)((int64_t)nunits * blockIdx.x / gridDim.x);
    const int u1 = (int)((int64_t)nunits * (blockIdx.x + 1) / gridDim.x);

    if (tid == 0) {
        for (int s = 0; s < NSTA_MAX; ++s) { mbar_init(&bars->a_full[s], 4 * HSPLIT); mbar_init(&bars->a_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&bars->acc_full[s], NI); mbar_init(&bars->acc_empty[s], 128); }
        mbar_init(&bars->u_full, 1);
        mbar_init(&bars->u_empty, NI);
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(&bars->tmem_base, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = bars->tmem_base;
    const uint32_t acc_base = tmem + ACC_BASE;

    if (warp >= 8) {
        // ---------------------------------------------------------------- feature generators (lane = row)
        const int g = (warp - 8) / (4 * HSPLIT);
        const int q = warp & 3;
        const int half0 = (((warp - 8) >> 2) % HSPLIT) * (2 / HSPLIT);
        const int row = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        // this warpgroup takes chunks cg = g, g+NGEN, ... of the CTA's chunk sequence; (u, ch) follow incrementally
        uint32_t gen = 0;
        int u = u0, ch = g, cur_grp = -1, nch = 0, sa = g % NSTA;
        uint32_t pa = ((g / NSTA) & 1) ^ 1;
        float x_next = 0.f;
        // (grp, tile, k) of unit u and of unit u + 1, kept incrementally (the divisions cost ~100 issue slots per unit)
        int grp_c = u0 / ntiles;
        int grp_n = grp_c, tile_n = u0 - grp_c * ntiles, k_n = grp_c / p.P;
        if (u0 < u1) x_next = p.XbF[(size_t)k_n * p.mpad + (size_t)tile_n * 128 + row];   // x of the first unit
        auto advance_next = [&]() {
            if (++tile_n == ntiles) {
                tile_n = 0;
                ++grp_n;
                k_n = grp_n / p.P;
            }
        };
        advance_next();   // (grp_n, tile_n, k_n) now describe unit u0 + 1
        for (; u < u1; ++u) {
            const int grp = grp_c;
            if (grp != cur_grp) {   // the (zt, b) table of this K part arrives with the U image
                cur_grp = grp;
                mbar_wait(&bars->u_full, gen & 1);
                ++gen;
                const int k = grp / p.P;
                nch = min(p.NCHp, NCH - (grp - k * p.P) * p.NCHp);
            }
            const float x = x_next;
            if (u + 1 < u1) x_next = p.XbF[(size_t)k_n * p.mpad + (size_t)tile_n * 128 + row];   // one unit ahead
            grp_c = grp_n;
            advance_next();
            for (; ch < nch; ch += NGEN_F) {
                mbar_wait(&bars->a_empty[sa], pa);
                fence_after_sync();
                const uint32_t st_addr = tmem + lane_addr + sa * SA_COLS;
                gen_fwd_chunk<THREE>(sZb + (p.sincos ? ch * KC / 2 : ch * KC), x, p.sincos != 0, st_addr, half0);
                wait_st();
                fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars->a_full[sa]);   // one arrival per warp: 32 lanes on one mbarrier serialise
                sa += NGEN_F;
                if (sa >= NSTA) { sa -= NSTA; pa ^= 1; }
            }
            ch -= nch;
        }
    } else if (warp == 4 || (NI == 2 && warp == 6)) {
        // ---------------------------------------------------------------- MMA issuers (whole warp runs the loop)
        const int w = (warp - 4) >> 1;
        uint32_t it = 0, gen = 0, cbase = 0;   // cbase: chunks of the CTA before this unit (-> A stage and parity)
        int cur_k = -1;
        for (int u = u0; u < u1; ++u, ++it) {
            const int grp = u / ntiles;
            if (grp != cur_k) {
                cur_k = grp;
                mbar_wait(&bars->u_full, gen & 1);
                ++gen;
            }
            const int nch = min(p.NCHp, NCH - (grp % p.P) * p.NCHp);
            const bool last_of_k = (u + 1 == u1) || ((u + 1) / ntiles != grp);
            const uint32_t ab = it & 1;
            mbar_wait(&bars->acc_empty[ab], ((it >> 1) & 1) ^ 1);
            fence_after_sync();
            const uint32_t d = acc_base + (ab * NI + w) * RPAD;
            for (int ch = w; ch < nch; ch += NI) {
                const uint32_t c = cbase + ch, sa = c % NSTA, pa = (c / NSTA) & 1;
                mbar_wait(&bars->a_full[sa], pa);
                fence_after_sync();
                const uint32_t a_hi = tmem + sa * SA_COLS, a_lo = a_hi + 32;
                const uint32_t b_addr = smem_u32(sU + (size_t)ch * KC * ROW_BYTES);
                if (elect_one()) {
#pragma unroll
                    for (int s4 = 0; s4 < KC / 16; ++s4)
                        mma_kstep<RPAD, THREE, 0>(d, a_hi + s4 * 8, a_lo + s4 * 8, b_addr + s4 * 16 * ROW_BYTES,
                                                  (ch >= NI || s4 != 0) ? 1u : 0u);
                    mma_commit(&bars->a_empty[sa]);
                }
                __syncwarp();
            }
            if (elect_one()) {
                // an issuer without a chunk in this unit (nch < NI) has nothing to track: commit arrives at once
                mma_commit(&bars->acc_full[ab]);
                if (last_of_k) mma_commit(&bars->u_empty);
            }
            __syncwarp();
            cbase += nch;
        }
    } else if (warp == 5) {
        // ---------------------------------------------------------------- U / (zt,b) loader: once per k
        if (lane == 0) {
            uint32_t gen = 0;
            int cur_k = -1;
            for (int u = u0; u < u1; ++u) {
                const int grp = u / ntiles;
                if (grp == cur_k) continue;
                cur_k = grp;
                const int k = grp / p.P, ch0 = (grp - k * p.P) * p.NCHp, nch = min(p.NCHp, NCH - ch0);
                mbar_wait(&bars->u_empty, (gen & 1) ^ 1);
                const size_t u_bytes = (size_t)nch * KC * ROW_BYTES;
                const uint32_t zb_bytes = (uint32_t)(nch * KC * sizeof(float2)) / (p.sincos ? 2 : 1);
                mbar_expect_tx(&bars->u_full, (uint32_t)u_bytes + zb_bytes);
                const unsigned char *src = reinterpret_cast<const unsigned char *>(p.Uimg) +
                                           ((size_t)k * p.npad + (size_t)ch0 * KC) * ROW_BYTES;
                for (size_t off = 0; off < u_bytes; off += 32768) {
                    const uint32_t nb = (uint32_t)(u_bytes - off < 32768 ? u_bytes - off : 32768);
                    bulk_g2s(sU + off, src + off, nb, &bars->u_full);
                }
                bulk_g2s(sZb, p.ZbF + (size_t)k * p.npad + (size_t)ch0 * KC / (p.sincos ? 2 : 1), zb_bytes, &bars->u_full);
                ++gen;
            }
        }
        __syncwarp();
    } else if (warp < 4) {
        // ---------------------------------------------------------------- epilogue
        const int row = warp * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
        const float scale = p.amp * p.uscale[1];
        uint32_t it = 0;
        for (int u = u0; u < u1; ++u, ++it) {
            const int grp = u / ntiles, tile = u - grp * ntiles, k = grp / p.P;
            const uint32_t ab = it & 1;
            mbar_wait_relaxed(&bars->acc_full[ab], (it >> 1) & 1);
            fence_after_sync();
            const uint32_t d = acc_base + ab * NI * RPAD + lane_addr;
            const bool two = NI == 2 && min(p.NCHp, NCH - (grp % p.P) * p.NCHp) >= 2;
            const int64_t i = (int64_t)tile * 128 + row;
            const bool live = i < p.bs;
            float *Tk = p.T + (size_t)k * p.r * p.mpad + i;
#pragma unroll
            for (int c0 = 0; c0 < RPAD; c0 += 16) {
                uint32_t v[16], v1[16];
                tmem_ld16(d + c0, v);
                if (two) tmem_ld16(d + RPAD + c0, v1);
                wait_ld();
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const int l = c0 + c;
                    if (l < p.r) {
                        float acc = __uint_as_float(v[c]);
                        if (two) acc += __uint_as_float(v1[c]);
                        const float val = live ? acc * scale : 0.f;
                        if (p.P == 1) Tk[(size_t)l * p.mpad] = val;
                        else atomicAdd(Tk + (size_t)l * p.mpad, val);
                    }
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

// ============================================================================================ K3
struct BwdParams {
    const float *XbF;      // [D][mpad]
    const __half *Eimg;    // see tc_layout.cuh
    const float *ZtF, *bF; // [D][npad]
    const float *escale;   // -> {escale, 1/escale}
    float *gradF;          // [D][npad][RPAD] fp32 partial sums, reduced with vector RED; finalize_grad_kernel folds
                           // them into the FP64 gradient and clears them
    float amp;
    int D, n, npad, r;
    int64_t mpad, mext;
    int sincos;
    int JG, NJG;           // feature blocks (128 features) per group, groups per dimension
    int nsta;              // A stages that fit beside the JG accumulators
};

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Work items it = (k*NJG + jg)*NCH + ch, CTA c owns [total*c/G, total*(c+1)/G).  For every item the loader
// brings one 64-row chunk of the E image (and the 64 inputs x_ik) and the generators / issuer run JG
// "A steps" on it, one per 128-feature block of the group.
template <int RPAD, bool THREE>
__global__ void __launch_bounds__(TC_THREADS, 1) bwd_tc_kernel(const BwdParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr int B_IMG_BYTES = 2 * RPAD * KC * 2;
    constexpr int X_BYTES = KC * 4;                         // the chunk's 64 input values ride in the same stage
    constexpr int B_STAGE_BYTES = B_IMG_BYTES + X_BYTES;
    constexpr int ROW_BYTES = 2 * RPAD * 2;
    unsigned char *sB = smem_raw;
    Bars *bars = reinterpret_cast<Bars *>(smem_raw + NSTB * B_STAGE_BYTES);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int NIB = 2;   // MMA issuers (see fwd_tc_kernel): issuer w takes the feature blocks jbl = w (mod 2)
    const int JB = p.npad / 128, JG = p.JG, NJG = p.NJG, NSTA = p.nsta;
    const int NCH = (int)(p.mext / KC);
    const int64_t total = (int64_t)p.D * NJG * NCH;
    const int64_t it0 = total * blockIdx.x / gridDim.x, it1 = total * (blockIdx.x + 1) / gridDim.x;
    const int grp0 = (int)(it0 / NCH), ch0 = (int)(it0 - (int64_t)grp0 * NCH);
    const uint32_t n_it = (uint32_t)(it1 - it0);
    const uint32_t n_as = n_it * JG;

    if (tid == 0) {
        for (int s = 0; s < NSTA_MAX; ++s) { mbar_init(&bars->a_full[s], 4 * HSPLIT); mbar_init(&bars->a_empty[s], 1); }
        for (int s = 0; s < NSTB; ++s) { mbar_init(&bars->b_full[s], 1); mbar_init(&bars->b_empty[s], NIB); }
        mbar_init(&bars->acc_full[0], NIB);
        mbar_init(&bars->acc_empty[0], 128);
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(&bars->tmem_base, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = bars->tmem_base;
    const uint32_t acc_base = tmem + NSTA * SA_COLS;

    if (warp >= 8) {
        // ---------------------------------------------------------------- feature generators (lane = feature)
        const int g = (warp - 8) / (4 * HSPLIT);
        const int q = warp & 3;
        const int half0 = (((warp - 8) >> 2) % HSPLIT) * (2 / HSPLIT);
        const int jl = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        // position of A step `as`: item lit (-> grp, ch) and block jbl
        int jbl = g, ch = ch0;
        int kk = grp0 / NJG, jg = grp0 - kk * NJG;   // (k, feature-block group) of the current item, kept incrementally:
        uint32_t lit = 0;                            // two integer divisions per A step were 5% of the kernel's instructions
        auto next_item = [&]() {
            ++lit;
            if (++ch == NCH) {
                ch = 0;
                if (++jg == NJG) { jg = 0; ++kk; }
            }
        };
        while (jbl >= JG) { jbl -= JG; next_item(); }
        int sa = g % NSTA;
        uint32_t pa = ((g / NSTA) & 1) ^ 1;
        // (zt, b) of this lane's feature in A step (k, jg, jbl); the load latency overlaps the barrier waits
        auto load_zb = [&](int k, int jg_, int jbl_, float &zt, float &bb) {
            int j = (jg_ * JG + jbl_) * 128 + jl;
            if (j >= p.npad) j = p.npad - 1;    // padding block of a ragged last group: computed, never stored
            const size_t base = (size_t)k * p.npad;
            if (!p.sincos) {
                zt = p.ZtF[base + j];
                bb = p.bF[base + j];
            } else {  // rows 2p / 2p+1 are sin / cos of x*Zt[p]:  cos(a) = sin(a + pi/2)
                zt = p.ZtF[base + (j >> 1)];
                bb = (j & 1) ? 1.5707963267948966f : 0.f;
            }
        };
        // one A step: 64 rows x this lane's feature -> stage sa (hi | lo words), then hand the stage to the issuers
        auto a_step = [&](float zt, float bb, int sb, uint32_t lit_cur, bool wait_b) {
            const float2 zt2 = make_float2(zt, zt), bb2 = make_float2(bb, bb);
            mbar_wait(&bars->a_empty[sa], pa);
            fence_after_sync();
            if (wait_b) mbar_wait(&bars->b_full[sb], (lit_cur / NSTB) & 1);
            const uint32_t st_addr = tmem + lane_addr + sa * SA_COLS;
            const float4 *xr = reinterpret_cast<const float4 *>(sB + sb * B_STAGE_BYTES + B_IMG_BYTES);
#pragma unroll
            for (int hh = 0; hh < 2 / HSPLIT; ++hh) {
                const int half = half0 + hh;
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int c = 0; c < 16; c += 2) {
                    const float4 xx = xr[half * 8 + c / 2];
                    const float2 g0 = __ffma2_rn(make_float2(xx.x, xx.y), zt2, bb2);
                    const float2 g1 = __ffma2_rn(make_float2(xx.z, xx.w), zt2, bb2);
                    if (!p.sincos) {
                        split_pack<THREE>(make_float2(__cosf(g0.x), __cosf(g0.y)), hi[c], lo[c]);
                        split_pack<THREE>(make_float2(__cosf(g1.x), __cosf(g1.y)), hi[c + 1], lo[c + 1]);
                    } else {
                        split_pack<THREE>(make_float2(__sinf(g0.x), __sinf(g0.y)), hi[c], lo[c]);
                        split_pack<THREE>(make_float2(__sinf(g1.x), __sinf(g1.y)), hi[c + 1], lo[c + 1]);
                    }
                }
                tmem_st16(st_addr + half * 16, hi);
                if (THREE) tmem_st16(st_addr + 32 + half * 16, lo);
            }
            wait_st();
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->a_full[sa]);   // one arrival per warp: 32 lanes on one mbarrier serialise
            sa += NGEN;
            while (sa >= NSTA) { sa -= NSTA; pa ^= 1; }
        };
        // (Keeping the (zt, b) of the group's blocks in registers across items instead of re-reading them here was
        // measured 20% SLOWER for the whole kernel, twice, A/B in the same run — as was fetching them one step ahead.)
        for (uint32_t as = g; as < n_as; as += NGEN) {
            float zt, bb;
            load_zb(kk, jg, jbl, zt, bb);
            const int sb = lit % NSTB;
            const uint32_t lit_cur = lit;
            jbl += NGEN;
            while (jbl >= JG) { jbl -= JG; next_item(); }
            a_step(zt, bb, sb, lit_cur, true);
        }
    } else if (warp == 4 || warp == 6) {
        // ---------------------------------------------------------------- MMA issuers (whole warp runs the loop)
        const int w = (warp - 4) >> 1;
        uint32_t seg = 0;
        int ch = ch0;
        bool fresh = true;
        for (uint32_t lit = 0; lit < n_it; ++lit) {
            if (fresh) {
                mbar_wait(&bars->acc_empty[0], (seg & 1) ^ 1);
                fence_after_sync();
            }
            const int sb = lit % NSTB;
            mbar_wait(&bars->b_full[sb], (lit / NSTB) & 1);
            const uint32_t b_addr = smem_u32(sB + sb * B_STAGE_BYTES);
            const bool last = (ch == NCH - 1) || (lit == n_it - 1);
            for (int jbl = w; jbl < JG; jbl += NIB) {
                const uint32_t as = lit * JG + jbl, sa = as % (uint32_t)NSTA, pa = (as / (uint32_t)NSTA) & 1;
                mbar_wait(&bars->a_full[sa], pa);
                fence_after_sync();
                const uint32_t a_hi = tmem + sa * SA_COLS, a_lo = a_hi + 32;
                const uint32_t d = acc_base + jbl * RPAD;
                if (elect_one()) {
#pragma unroll
                    for (int s4 = 0; s4 < KC / 16; ++s4)
                        mma_kstep<RPAD, THREE, 1>(d, a_hi + s4 * 8, a_lo + s4 * 8, b_addr + s4 * 16 * ROW_BYTES,
                                                  (!fresh || s4 != 0) ? 1u : 0u);
                    mma_commit(&bars->a_empty[sa]);
                }
                __syncwarp();
            }
            if (elect_one()) {   // (an issuer without a block of its own, JG == 1, tracks nothing: arrives at once)
                mma_commit(&bars->b_empty[sb]);
                if (last) mma_commit(&bars->acc_full[0]);
            }
            __syncwarp();
            fresh = false;
            ++ch;
            if (last) { ch = 0; ++seg; fresh = true; }
        }
    } else if (warp == 5) {
        // ---------------------------------------------------------------- B loader
        if (lane == 0) {
            int grp = grp0, ch = ch0;
            for (uint32_t lit = 0; lit < n_it; ++lit) {
                const int k = grp / NJG;
                const int sb = lit % NSTB;
                const __half *src = p.Eimg + (size_t)k * p.mpad * 2 * RPAD + (size_t)ch * KC * 2 * RPAD;
                mbar_wait(&bars->b_empty[sb], ((lit / NSTB) & 1) ^ 1);
                mbar_expect_tx(&bars->b_full[sb], B_STAGE_BYTES);
                bulk_g2s(sB + sb * B_STAGE_BYTES + B_IMG_BYTES, p.XbF + (size_t)k * p.mpad + (size_t)ch * KC, X_BYTES,
                         &bars->b_full[sb]);
                bulk_g2s(sB + sb * B_STAGE_BYTES, src, B_IMG_BYTES, &bars->b_full[sb]);
                if (++ch == NCH) { ch = 0; ++grp; }
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
            it = min((long long)it1, (long long)(grp + 1) * NCH);
            const int k = grp / NJG, jg = grp - k * NJG;
            mbar_wait_relaxed(&bars->acc_full[0], seg & 1);
            fence_after_sync();
            for (int jbl = 0; jbl < JG; ++jbl) {
                const int jb = jg * JG + jbl;
                if (jb >= JB) break;
                const int j = jb * 128 + warp * 32 + lane;
                float *gk = p.gradF + ((size_t)k * p.npad + j) * RPAD;
                const uint32_t d = acc_base + jbl * RPAD + lane_addr;
#pragma unroll
                for (int c0 = 0; c0 < RPAD; c0 += 16) {
                    uint32_t v[16];
                    tmem_ld16(d + c0, v);
                    wait_ld();
#pragma unroll
                    for (int c = 0; c < 16; c += 4)
                        red_add_v4(gk + c0 + c, __uint_as_float(v[c]) * scale, __uint_as_float(v[c + 1]) * scale,
                                   __uint_as_float(v[c + 2]) * scale, __uint_as_float(v[c + 3]) * scale);
                }
            }
            fence_before_sync();
            mbar_arrive(&bars->acc_empty[0]);
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 512);
}

// gradU[k][l][j] (FP64, Julia order) += gradF[k][j][l]; gradF is cleared for the next step.
// One CTA per (k, 32-feature block): coalesced reads along l, transposed through shared memory, coalesced
// writes along j.
__global__ void __launch_bounds__(256) finalize_grad_kernel(float *__restrict__ gradF, double *__restrict__ gradU, int n,
                                                            int npad, int r, int rpad) {
    __shared__ float tile[32][65];
    const int k = blockIdx.y, j0 = blockIdx.x * 32;
    float *src = gradF + ((size_t)k * npad + j0) * rpad;
    for (int t = threadIdx.x; t < 32 * rpad; t += 256) {
        const int jj = t / rpad, l = t - jj * rpad;
        tile[jj][l] = src[t];
        src[t] = 0.f;
    }
    __syncthreads();
    for (int t = threadIdx.x; t < 32 * r; t += 256) {
        const int l = t >> 5, jj = t & 31;
        if (j0 + jj < n) gradU[((size_t)k * r + l) * n + j0 + jj] += (double)tile[jj][l];
    }
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

// E-image scale guard.  The image is written with the scale derived from the PREVIOUS step's max|E| (128x headroom to
// fp16 overflow, 128x down to where the lo parts leave fp16's normal range).  After K2 this kernel looks at the max|E|
// the step actually produced: outside [2, 32768) after scaling it installs the exact scale and raises the flag that
// makes the (otherwise empty) second K2 launch rewrite the image.
__global__ void check_escale_kernel(float *state /*[0] max|E|, [1] scale, [2] 1/scale*/, int *flag) {
    const float mx = state[0], M = mx * state[1];
    const int redo = (mx > 0.f && isfinite(mx) && !(M >= 2.f && M < 32768.f)) ? 1 : 0;
    if (redo) {
        const float s = pow2_scale_for(mx);
        state[1] = s;
        state[2] = 1.f / s;
    }
    *flag = redo;
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

// (zt, b) table of the forward generators: cos features -> float4 per feature pair (zt0, zt1, b0, b1);
// sin/cos features -> float2 (zt[p], b[p]) per frequency p.
__global__ void interleave_zb_kernel(const float *zt, const float *b, size_t n, int pair_layout, float2 *out) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        if (!pair_layout) {
            out[i] = make_float2(zt[i], b[i]);
        } else {
            const size_t pr = i >> 1;
            out[i] = (i & 1) ? make_float2(b[2 * pr], b[2 * pr + 1]) : make_float2(zt[2 * pr], zt[2 * pr + 1]);
        }
    }
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

int gptb_tc_check_escale(gptb_handle *h) {
    check_escale_kernel<<<1, 1, 0, h->stream>>>(h->tc_state + 0, h->nan_flag + 2);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_tc_prepare_features(gptb_handle *h) {
    const size_t cnt = (size_t)h->cfg.D * h->npad;
    if (!h->ZbF) GPTB_CUDA_OK(h, cudaMalloc((void **)&h->ZbF, cnt * sizeof(float2)));
    interleave_zb_kernel<<<64, 256, 0, h->stream>>>(h->ZtF, h->bF, cnt, h->cfg.feature_kind == GPTB_FEAT_COS ? 1 : 0,
                                                    (float2 *)h->ZbF);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

template <int RPAD>
static int launch_fwd_r(gptb_handle *h, const FwdParams &p, int grid, size_t smem) {
    if (h->cfg.precision == GPTB_PREC_TC3) {
        auto k = fwd_tc_kernel<RPAD, true>;
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<grid, TC_THREADS_F, smem, h->stream>>>(p);
    } else {
        auto k = fwd_tc_kernel<RPAD, false>;
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k<<<grid, TC_THREADS_F, smem, h->stream>>>(p);
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

// Forward with an explicit operand geometry: the image has NC columns per hi / lo half (NC a multiple of 16, at most
// 128), `ncols` of them live; T is [D][ncols][mpad].  The sampler uses (r, rpad); the posterior-predictive sweep packs
// several samples' factors side by side (ncols = Sg*r), so the features are generated once per sample group.
static int launch_fwd_tc_cols_range(gptb_handle *h, const float *XbF, const void *Uimg, const float *uscale, int64_t bs,
                                    int64_t mpad, float *T, int ncols, int NC, int k0, int kn);

int gptb_launch_fwd_tc_cols(gptb_handle *h, const float *XbF, const void *Uimg, const float *uscale, int64_t bs,
                            int64_t mpad, float *T, int ncols, int NC) {
    return launch_fwd_tc_cols_range(h, XbF, Uimg, uscale, bs, mpad, T, ncols, NC, 0, h->cfg.D);
}

// Dimensions [k0, k0 + kn) of the forward pass: every per-dimension array is entered at k0, the kernel sees D = kn.
static int launch_fwd_tc_cols_range(gptb_handle *h, const float *XbF, const void *Uimg, const float *uscale, int64_t bs,
                                    int64_t mpad, float *T, int ncols, int NC, int k0, int kn) {
    const gptb_config &c = h->cfg;
    if (!h->ZbF_valid) {
        GPTB_TRY(gptb_tc_prepare_features(h));
        h->ZbF_valid = true;
    }
    FwdParams p;
    p.XbF = XbF + (size_t)k0 * mpad;
    p.Uimg = (const __half *)Uimg + (size_t)k0 * h->npad * 2 * NC;
    p.ZbF = (const float2 *)h->ZbF + (size_t)k0 * h->npad;
    p.uscale = uscale;
    T += (size_t)k0 * ncols * mpad;
    p.T = T;
    p.amp = (float)h->amp;
    p.D = kn; p.n = c.n; p.npad = h->npad; p.r = ncols;
    p.bs = bs;
    p.mpad = mpad;
    p.mext = (bs + 127) / 128 * 128;   // only the 128-row tiles that hold live rows are generated
    p.sincos = c.feature_kind == GPTB_FEAT_SINCOS;
    // K parts: the image of one part (plus its (zt,b) table) must fit in shared memory
    const int NCH = h->npad / KC;
    const size_t chunk_bytes = (size_t)KC * (2 * NC * 2 + sizeof(float2));
    const size_t budget = 227 * 1024 - sizeof(Bars) - 1024;
    const int max_chunks = (int)(budget / chunk_bytes);
    p.P = (NCH + max_chunks - 1) / max_chunks;
    p.NCHp = (NCH + p.P - 1) / p.P;
    if (p.P > 1) GPTB_CUDA_OK(h, cudaMemsetAsync(T, 0, sizeof(float) * (size_t)kn * ncols * mpad, h->stream));
    const int nunits = (int)(p.mext / 128) * kn * p.P;
    const int grid = std::max(1, std::min(h->sm_count - h->sm_reserve, nunits));
    // at least 120 KB so that two CTAs (each wants all 512 TMEM columns) never share an SM
    const size_t smem = std::max<size_t>((size_t)p.NCHp * chunk_bytes + sizeof(Bars) + 64, 120 * 1024);
    int st;
    switch (NC) {
        case 16: st = launch_fwd_r<16>(h, p, grid, smem); break;
        case 32: st = launch_fwd_r<32>(h, p, grid, smem); break;
        case 48: st = launch_fwd_r<48>(h, p, grid, smem); break;
        case 64: st = launch_fwd_r<64>(h, p, grid, smem); break;
        case 80: st = launch_fwd_r<80>(h, p, grid, smem); break;
        case 96: st = launch_fwd_r<96>(h, p, grid, smem); break;
        case 112: st = launch_fwd_r<112>(h, p, grid, smem); break;
        case 128: st = launch_fwd_r<128>(h, p, grid, smem); break;
        default:
            gptb_set_error(h, "tensor-core forward: unsupported operand width %d", NC);
            return GPTB_ERR_UNSUPPORTED;
    }
    GPTB_TRY(st);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_fwd_tc(gptb_handle *h, const float *XbF, const void *Uimg, int64_t bs, int64_t mpad, float *T) {
    return gptb_launch_fwd_tc_cols(h, XbF, Uimg, h->tc_state + (Uimg == h->Uimg ? 4 : 7), bs, mpad, T, h->cfg.r, h->rpad);
}
int gptb_launch_fwd_tc_range(gptb_handle *h, const float *XbF, const void *Uimg, int64_t bs, int64_t mpad, float *T, int k0,
                             int kn) {
    return launch_fwd_tc_cols_range(h, XbF, Uimg, h->tc_state + (Uimg == h->Uimg ? 4 : 7), bs, mpad, T, h->cfg.r, h->rpad, k0,
                                    kn);
}

// Operand image of a group of Sg posterior samples side by side: column s*r + l of dimension k is U_s[:, l, k].
__global__ void pack_Ugroup_kernel(const double *__restrict__ Ug /*[Sg][D][r][n]*/, int Sg, int D, int n, int r, int npad,
                                   int NC, const float *__restrict__ scale, __half *__restrict__ img) {
    const float sc = scale[0];
    const size_t total = (size_t)D * npad * NC;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int j = (int)(t % npad);
        const int nn = (int)((t / npad) % NC);
        const int k = (int)(t / ((size_t)npad * NC));
        const int s = nn / r, l = nn - s * r;
        float v = 0.f;
        if (j < n && s < Sg) v = (float)(Ug[(((size_t)s * D + k) * r + l) * n + j] * (double)sc);
        const __half hh = __float2half_rn(v);
        img[gptb_uimg_offset(k, j, nn, npad, NC)] = hh;
        img[gptb_uimg_offset(k, j, NC + nn, npad, NC)] = __float2half_rn(v - __half2float(hh));
    }
}

// max|U| over `cnt` doubles -> power-of-two scale in tc_state[7..8] (prediction slot)
int gptb_tc_pred_scale(gptb_handle *h, const double *U, size_t cnt, bool first) {
    if (first) GPTB_CUDA_OK(h, cudaMemsetAsync(h->tc_state + 6, 0, sizeof(float), h->stream));
    absmax_kernel<<<std::min<int>(h->sm_count, (int)((cnt + 1023) / 1024)), 1024, 0, h->stream>>>(U, cnt, h->tc_state + 6);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}
int gptb_tc_pred_scale_commit(gptb_handle *h) {
    set_scale_kernel<<<1, 1, 0, h->stream>>>(h->tc_state + 6, h->tc_state + 7, 0);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}
int gptb_tc_pack_group(gptb_handle *h, const double *Ug, int Sg, int NC, void *img) {
    const gptb_config &c = h->cfg;
    const size_t total = (size_t)c.D * h->npad * NC;
    pack_Ugroup_kernel<<<(int)std::min<size_t>((total + 255) / 256, (size_t)h->sm_count * 8), 256, 0, h->stream>>>(
        Ug, Sg, c.D, c.n, c.r, h->npad, NC, h->tc_state + 7, (__half *)img);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// Scale and pack one sample group on its own: state -> {max|U| of the group, 2^k, 2^-k}; the forward kernel is given
// state + 1 as its scale.  (The node-grid sweep uploads U_store group by group, so no global maximum exists up front.)
int gptb_tc_pack_group_scaled(gptb_handle *h, const double *Ug, int Sg, int NC, void *img, float *state, cudaStream_t s) {
    const gptb_config &c = h->cfg;
    const size_t cnt = (size_t)Sg * c.D * c.r * c.n, total = (size_t)c.D * h->npad * NC;
    GPTB_CUDA_OK(h, cudaMemsetAsync(state, 0, sizeof(float), s));
    absmax_kernel<<<std::min<int>(h->sm_count, (int)((cnt + 1023) / 1024)), 1024, 0, s>>>(Ug, cnt, state);
    set_scale_kernel<<<1, 1, 0, s>>>(state, state + 1, 0);
    pack_Ugroup_kernel<<<(int)std::min<size_t>((total + 255) / 256, (size_t)h->sm_count * 8), 256, 0, s>>>(
        Ug, Sg, c.D, c.n, c.r, h->npad, NC, state + 1, (__half *)img);
    h->launches += 3;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_tc_ensure_gradF(gptb_handle *h) {
    if (h->gradFbase) return GPTB_OK;
    const size_t gf = (size_t)h->cfg.D * h->npad * h->rpad;
    h->gradF_qstride = (h->cfg.Q + 31) & ~31;
    h->gradF_qpad = GPTB_GW_SLOTS * h->gradF_qstride;
    GPTB_CUDA_OK(h, cudaMalloc((void **)&h->gradFbase, (gf + h->gradF_qpad) * sizeof(float)));
    GPTB_CUDA_OK(h, cudaMemsetAsync(h->gradFbase, 0, (gf + h->gradF_qpad) * sizeof(float), h->stream));
    h->gradF = h->gradFbase + h->gradF_qpad;
    return GPTB_OK;
}

int gptb_launch_bwd_tc(gptb_handle *h, const float *XbF, const void *Eimg, int64_t bs, int64_t mpad, double *gradU) {
    return gptb_launch_bwd_tc_range(h, XbF, Eimg, bs, mpad, gradU, 0, h->cfg.D);
}

int gptb_launch_bwd_tc_range(gptb_handle *h, const float *XbF, const void *Eimg, int64_t bs, int64_t mpad, double *gradU,
                             int k0, int kn) {
    const gptb_config &c = h->cfg;
    GPTB_TRY(gptb_tc_ensure_gradF(h));
    BwdParams p;
    p.mext = (bs + 127) / 128 * 128;
    float *gradF_g = h->gradF + (size_t)k0 * h->npad * h->rpad;
    p.XbF = XbF + (size_t)k0 * mpad;
    p.Eimg = (const __half *)Eimg + (size_t)k0 * mpad * 2 * h->rpad;
    p.ZtF = h->ZtF + (size_t)k0 * h->npad;
    p.bF = h->bF + (size_t)k0 * h->npad;
    p.escale = h->tc_state + 1;
    p.gradF = gradF_g;
    p.amp = (float)h->amp;
    p.D = kn; p.n = c.n; p.npad = h->npad; p.r = c.r;
    p.mpad = mpad;
    p.sincos = c.feature_kind == GPTB_FEAT_SINCOS;
    const int JB = h->npad / 128;
#ifndef GPTB_K3_ACC_COLS
#define GPTB_K3_ACC_COLS 256
#endif
    const int jg_max = GPTB_K3_ACC_COLS / h->rpad;    // accumulators take at most half of TMEM
    p.NJG = (JB + jg_max - 1) / jg_max;
    p.JG = (JB + p.NJG - 1) / p.NJG;
    p.nsta = nsta_for(p.JG * h->rpad);
    const int64_t total = (int64_t)kn * p.NJG * (p.mext / KC);
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(h->sm_count - h->sm_reserve, total));
    const size_t smem = std::max<size_t>((size_t)NSTB * (2 * h->rpad * KC * 2 + KC * 4) + sizeof(Bars) + 64, 120 * 1024);
    int st;
    switch (h->rpad) {
        case 16: st = launch_bwd_r<16>(h, p, grid, smem); break;
        case 32: st = launch_bwd_r<32>(h, p, grid, smem); break;
        case 48: st = launch_bwd_r<48>(h, p, grid, smem); break;
        default: st = launch_bwd_r<64>(h, p, grid, smem); break;
    }
    GPTB_TRY(st);
    if (!h->f32_grad) {   // (in the FP32 gradient flow the update reads and clears gradF itself)
        finalize_grad_kernel<<<dim3(h->npad / 32, kn), 256, 0, h->stream>>>(gradF_g, gradU + (size_t)k0 * c.r * c.n, c.n,
                                                                           h->npad, c.r, h->rpad);
        h->launches++;
    }
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}
