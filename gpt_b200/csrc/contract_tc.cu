// K2 (tensor-core path) — core contraction, residual and leave-one-out back-projection in FP32,
// two minibatch rows per thread.
//
// Replaces (all /root/reference/GPT_SGLD.jl):
//   computeV     :208-220   V[q,i]   = prod_k T[k, I[q,k], i]
//   computefhat  :223-230   fhat[i]  = sum_q w[q] V[q,i]
//   grad-w       :393       gradw   += sum_i e_i V[:,i],      e_i = (N/bs)(y_i - fhat_i)/signal_var
//   computeU_phi :246-258   U_phi[q,i,k] = V[q,i] / T[k, I[q,k], i]
//   computeA     :261-273   A[l,k,i] = sum_{q: I[q,k]=l} w[q] U_phi[q,i,k]
// and emits the fp16 hi|lo operand image of E[k][l][i] = e_i * A[l,k,i] for K3 (tc_layout.cuh).
//
// The kernel is a gather machine: per row 2*Q*D shared-memory reads whose index (I[q,k]) is the same for
// every row, so rows sit on the lanes (conflict-free, unit stride) and the bound is shared-memory bandwidth
// (8*Q*D bytes per row) as long as the instruction count per gather stays near one.  Hence:
//   * a CTA owns 64 rows and a lane two adjacent rows: every gather is one LDS.64 feeding packed FP32 math
//     (FMUL2 / FADD2), and the pre-multiplied offset tables are read four at a time (one 16-byte uniform load);
//   * pass 1: warp s forms V[q,:] for q = s, s+8, ... and its share of fhat; V goes to shared memory;
//   * gradw: a warp reduces e_i V[q,i] for up to 32 of its q at once with a 31-shuffle transpose-reduction;
//   * pass 2: every (k,l) owns the list of q with I[q,k] = l (built in gptb_set_index), so
//     A[l,k,i] = (sum_{q in list} w_q V[q,i]) / T[k,l,i] — the reference's own division (:253) taken once per
//     (k,l).  If T[k,l,i] is exactly 0 the quotient is replaced by the explicit leave-one-out product (the
//     reference would produce Inf/NaN there).  A warp handles 8 consecutive l at a time, so it writes whole
//     32-byte pieces (two rows x 8 fp16) of the MMA operand image.
#include <cuda_fp16.h>

#include <vector>

#include "common.cuh"
#include "tc_layout.cuh"
#include "tc_ptx.cuh"

namespace {

constexpr int CRB = 64;          // rows per CTA
#ifndef GPTB_CNW
#define GPTB_CNW 20
#endif
constexpr int CNW = GPTB_CNW;    // warps per CTA.  The tiles (128+ KB at C4) allow one CTA per SM, so the CTA itself
                                 // brings the warps that hide the LDS -> FMUL2 / FADD2 chains; 20 divides the work of
                                 // the headline shape evenly (Q = 200 -> 10 q per warp, D*rpad/8 = 40 units -> 2 per warp)
constexpr int CT_THREADS = 32 * CNW;
#ifndef GPTB_K2_TLOAD
#define GPTB_K2_TLOAD 1
#endif
#ifndef GPTB_K2_CHAINS
#define GPTB_K2_CHAINS 2
#endif
constexpr int NCH = GPTB_K2_CHAINS;   // independent V products per warp and iteration of pass 1

struct ContractTcParams {
    const float *T;          // [D][r][mpad]
    const double *w;
    const double *wA;        // weights of computeA when they differ from w (RMSprop sampler: A sees the updated w, :1194), or null
    const int32_t *off1;     // [Q][DP]  (k*r + I[q,k]) * 32  (float2 index into Ts), DP = D rounded up to 4
    const int32_t *lptr;     // [D*rpad + 1] start of each (k,l) list in `lists` (lists padded to multiples of 4)
    const int32_t *lists;    // q * 32 (float2 index into Vs) for the q of each list; padding entries -> zero row q = Q
    const double *yb;
    double *fhat, *gradw;
    float *gradwF;           // FP32 gradient flow: gradw is summed here instead (one all-reduce buffer with gradF),
    int gw_stride;           //   into copy blockIdx.x % GPTB_GW_SLOTS (stride gw_stride floats)
    __half *Eimg;
    const float *escale_img;
    float *emax_out;
    int64_t bs, mpad;
    double escale;
    int DP, r, rpad, Q, want_grad;
    int n_lists;             // entries of `lists` (multiple of 4)
    const int *redo_flag;    // redo != 0: the launch only rewrites the E image with a corrected fp16 scale, and only if
    int redo;                //   *redo_flag says the first pass left the safe range (tc_path.cu: check_escale_kernel)
    int t_rowmajor;          // T is [D][mpad][rpad] (row-major per sample, written by the node-grid interpolation)
    float *Erow;             // optional plain FP32 copy of E, [D][mpad][rpad] (row-major per sample: the node-grid
                             //   spreading kernel reads one 4*rpad-byte line per row), instead of the operand image
};

__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
// shared-memory gather of this lane's two rows: `base` already carries the lane offset, `off` is a byte offset
__device__ __forceinline__ float2 lds2(const char *base, int off) { return *reinterpret_cast<const float2 *>(base + off); }

template <int DT>
__global__ void __launch_bounds__(CT_THREADS, 1) contract_tc_kernel(const ContractTcParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int D = DT;
    constexpr int DP4 = (DT + 3) / 4;
    const int r = p.r, rpad = p.rpad, Q = p.Q, DR = D * r;
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    if (p.redo == 1 && *p.redo_flag == 0) return;                // (uniform across the grid; redo == 2: unconditional)
    const int64_t i0 = (int64_t)blockIdx.x * CRB + 2 * lane;     // this lane's rows: i0, i0 + 1
    const bool live0 = i0 < p.bs, live1 = i0 + 1 < p.bs;

    float2 *Ts = reinterpret_cast<float2 *>(smem_raw);           // [DR][32]
    float2 *Vs = Ts + (size_t)DR * 32;                           // [Q + 1][32], row Q is zero
    float *ws = reinterpret_cast<float *>(Vs + (size_t)(Q + 1) * 32);   // [Q rounded up to 4]
    float2 *Fs = reinterpret_cast<float2 *>(ws + ((Q + 3) & ~3)); // [CNW][32]
    // the (uniform) gather offsets, staged once per CTA as BYTE offsets: every use is an LDS broadcast
    int4 *off1s = reinterpret_cast<int4 *>(Fs + CNW * 32);       // [Q][DP4]
    int4 *lists_s = off1s + (size_t)Q * DP4;                     // [n_lists / 4]
    int *lptr_s = reinterpret_cast<int *>(lists_s + p.n_lists / 4);   // [D*rpad + 1]

    // ---- prologue: the 64-row tile of T (one 256-byte row piece per (k,l)) arrives by 16-byte cp.async, all pieces in
    // flight at once and no registers held; the tables are staged meanwhile
    if (!p.t_rowmajor) {
        const float *src = p.T + (size_t)blockIdx.x * CRB;
        const uint32_t dst = tcx::smem_u32(Ts);
        for (int t = tid; t < DR * 16; t += CT_THREADS) {
            const int c = t >> 4, part = t & 15;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (uint32_t)t * 16u),
                         "l"(src + (size_t)c * p.mpad + part * 4)
                         : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    } else {
        // [k][row][rpad] -> Ts[k*r + l][row]: a thread takes one row and four columns at a time (16-byte loads); lanes
        // run along the rows, so the shared-memory stores are conflict-free and the eight loads that share a 128-byte
        // line hit L1 after the first
        float *Tf = reinterpret_cast<float *>(Ts);
        const int q4 = rpad >> 2, total = CRB * D * q4;
        // (four loads in flight per thread before the first store: the loop is latency-, not bandwidth-bound)
        constexpr int TB = 4;
#if GPTB_K2_TLOAD != 0
        const int qh = q4 / (GPTB_K2_TLOAD == 1 ? 4 : 2);                  // column groups per row and warp pass
        const int qh_sh = (qh & (qh - 1)) == 0 ? 31 - __clz(qh) : -1;      // a power of two (rpad = 16, 32, 64): a shift, no division
#endif
        for (int t0 = tid; t0 < total; t0 += TB * CT_THREADS) {
            float4 v[TB];
            float *d[TB];
            int l0[TB];
#pragma unroll
            for (int j = 0; j < TB; ++j) {
                const int t = t0 + j * CT_THREADS;
                const int tt = t < total ? t : 0;                 // (past the end: a valid address, nothing stored)
#if GPTB_K2_TLOAD == 0
                const int rowt = tt & (CRB - 1), rest = tt >> 6;   // a warp: 32 rows x one 16-byte column group
                const int k = rest / q4, q = rest - k * q4;
#else
                // a warp: RS rows x (32 / RS) adjacent column groups — fewer lines per load instruction, at the price of
                // a (32 / RS)-way bank conflict in the transposing stores
                constexpr int RS = GPTB_K2_TLOAD == 1 ? 8 : 16, QS = 32 / RS;
                const int ln = tt & 31, rowgrp = (tt >> 5) % (CRB / RS), kq = tt / (32 * (CRB / RS));
                const int k = qh_sh >= 0 ? (kq >> qh_sh) : kq / qh, q = (kq - k * qh) * QS + (ln % QS);
                const int rowt = rowgrp * RS + ln / QS;
#endif
                v[j] = __ldg(reinterpret_cast<const float4 *>(p.T + ((size_t)k * p.mpad + (size_t)blockIdx.x * CRB + rowt) * rpad) + q);
                l0[j] = t < total ? 4 * q : r;                    // past the end: nothing is stored
                d[j] = Tf + (size_t)(k * r + 4 * q) * CRB + rowt;
            }
#pragma unroll
            for (int j = 0; j < TB; ++j) {
                if (l0[j] < r) d[j][0] = v[j].x;
                if (l0[j] + 1 < r) d[j][CRB] = v[j].y;
                if (l0[j] + 2 < r) d[j][2 * CRB] = v[j].z;
                if (l0[j] + 3 < r) d[j][3 * CRB] = v[j].w;
            }
        }
    }
    for (int t = tid; t < Q * DP4; t += CT_THREADS) {
        int4 v = __ldg(reinterpret_cast<const int4 *>(p.off1) + t);
        v.x <<= 3; v.y <<= 3; v.z <<= 3; v.w <<= 3;
        off1s[t] = v;
    }
    for (int t = tid; t < p.n_lists / 4; t += CT_THREADS) {
        int4 v = __ldg(reinterpret_cast<const int4 *>(p.lists) + t);
        v.x <<= 3; v.y <<= 3; v.z <<= 3; v.w <<= 3;
        lists_s[t] = v;
    }
    for (int t = tid; t <= D * rpad; t += CT_THREADS) lptr_s[t] = __ldg(p.lptr + t);
    for (int t = tid; t < Q; t += CT_THREADS) ws[t] = (float)p.w[t];
    if (wp == 0) Vs[(size_t)Q * 32 + lane] = make_float2(0.f, 0.f);
    // the residual's inputs, fetched before the wait so that their latency is hidden too
    double y0 = 0.0, y1 = 0.0;
    if (p.want_grad) {
        if (live0) y0 = p.yb[i0];
        if (live1) y1 = p.yb[i0 + 1];
    }
    const float sc = *p.escale_img;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- pass 1: V and fhat
    const char *Tb = reinterpret_cast<const char *>(Ts + lane);
    float2 f = make_float2(0.f, 0.f);
    for (int q0 = wp; q0 < Q; q0 += NCH * CNW) {     // NCH independent product chains per iteration
        const int4 *o[NCH];
        float2 v[NCH];
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
            const int qj = q0 + j * CNW;
            o[j] = off1s + (size_t)(qj < Q ? qj : q0) * DP4;
            v[j] = make_float2(1.f, 1.f);
        }
#pragma unroll
        for (int g4 = 0; g4 < DP4; ++g4) {
            int4 a[NCH];
#pragma unroll
            for (int j = 0; j < NCH; ++j) a[j] = o[j][g4];
            if (g4 * 4 + 0 < D) {
#pragma unroll
                for (int j = 0; j < NCH; ++j) v[j] = mul2(v[j], lds2(Tb, a[j].x));
            }
            if (g4 * 4 + 1 < D) {
#pragma unroll
                for (int j = 0; j < NCH; ++j) v[j] = mul2(v[j], lds2(Tb, a[j].y));
            }
            if (g4 * 4 + 2 < D) {
#pragma unroll
                for (int j = 0; j < NCH; ++j) v[j] = mul2(v[j], lds2(Tb, a[j].z));
            }
            if (g4 * 4 + 3 < D) {
#pragma unroll
                for (int j = 0; j < NCH; ++j) v[j] = mul2(v[j], lds2(Tb, a[j].w));
            }
        }
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
            const int qj = q0 + j * CNW;
            if (qj < Q) {
                Vs[(size_t)qj * 32 + lane] = v[j];
                const float wj = ws[qj];
                f = __ffma2_rn(make_float2(wj, wj), v[j], f);
            }
        }
    }
    Fs[wp * 32 + lane] = f;
    __syncthreads();
    f = Fs[lane];
#pragma unroll
    for (int t = 1; t < CNW; ++t) f = add2(f, Fs[t * 32 + lane]);
    if (wp == 0 && !p.redo) {
        if (live0) p.fhat[i0] = (double)f.x;
        if (live1) p.fhat[i0 + 1] = (double)f.y;
    }
    if (!p.want_grad) return;
    float2 e;
    e.x = live0 ? (float)(p.escale * (y0 - (double)f.x)) : 0.f;
    e.y = live1 ? (float)(p.escale * (y1 - (double)f.y)) : 0.f;

    // ---- gradw: sum over the CTA's rows of e_i V[q,i]; V[q,i] <- w_q V[q,i] in place for pass 2
    constexpr int NB = 16;   // q per warp and batch
    for (int qb = wp; qb < Q; qb += CNW * NB) {
        float v[NB];
#pragma unroll
        for (int j = 0; j < NB; ++j) {
            const int q = qb + CNW * j;
            float val = 0.f;
            if (q < Q) {
                const float2 vv = Vs[(size_t)q * 32 + lane];
                val = fmaf(e.x, vv.x, e.y * vv.y);
                const float wq = p.wA ? (float)p.wA[q] : ws[q];
                Vs[(size_t)q * 32 + lane] = mul2(vv, make_float2(wq, wq));
            }
            v[j] = val;
        }
        // transpose-reduction: after the steps lane L holds the sum over all lanes of v[L % NB]
#pragma unroll
        for (int off = NB / 2; off >= 1; off >>= 1) {
            const bool up = (lane & off) != 0;
#pragma unroll
            for (int j = 0; j < off; ++j) {
                const float keep = up ? v[j + off] : v[j];
                const float send = up ? v[j] : v[j + off];
                v[j] = keep + __shfl_xor_sync(0xffffffffu, send, off);
            }
        }
#pragma unroll
        for (int off = NB; off < 32; off <<= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
        const int q = qb + CNW * (lane % NB);
        if (lane < NB && q < Q && !p.redo) {
            if (p.gradwF) atomicAdd(&p.gradwF[(size_t)(blockIdx.x % GPTB_GW_SLOTS) * p.gw_stride + q], v[0]);
            else atomicAdd(&p.gradw[q], (double)v[0]);
        }
    }
    __syncthreads();

    // ---- pass 2: E[k][l][i] = e_i * (sum_{q: I[q,k]=l} w_q V[q,i]) / T[k][l][i]
    // A warp takes 4 consecutive l at a time (8-byte pieces of the operand image's 16-byte rows).  The four lists
    // advance together, each at its own length (padded to 4 entries that address the zero row), so up to 16
    // independent gathers are in flight per warp.
    constexpr int LU = 4;
    const int nlu = rpad / LU, nlb = rpad / 8;
    const char *Vb = reinterpret_cast<const char *>(Vs + lane);
    float mx = 0.f;
    for (int u = wp; u < D * nlu; u += CNW) {
        const int k = u / nlu, lu = u - k * nlu;
        float2 out[LU];
        int pb[LU + 1];
#pragma unroll
        for (int j = 0; j <= LU; ++j) pb[j] = lptr_s[k * rpad + lu * LU + j];
        int maxlen = 0;
#pragma unroll
        for (int j = 0; j < LU; ++j) maxlen = max(maxlen, pb[j + 1] - pb[j]);
        float2 s0[LU], s1[LU];
#pragma unroll
        for (int j = 0; j < LU; ++j) s0[j] = s1[j] = make_float2(0.f, 0.f);
        for (int g = 0; g < maxlen; g += 4) {
#pragma unroll
            for (int j = 0; j < LU; ++j) {
                const int pp = pb[j] + g;
                if (pp < pb[j + 1]) {   // warp-uniform
                    const int4 oo = lists_s[pp >> 2];
                    s0[j] = add2(s0[j], add2(lds2(Vb, oo.x), lds2(Vb, oo.y)));
                    s1[j] = add2(s1[j], add2(lds2(Vb, oo.z), lds2(Vb, oo.w)));
                }
            }
        }
#pragma unroll
        for (int j = 0; j < LU; ++j) {
            const int l = lu * LU + j;
            const int p0 = pb[j], p1 = pb[j + 1];
            const float2 sum = add2(s0[j], s1[j]);
            float2 val = make_float2(0.f, 0.f);
            if (p1 > p0) {  // uniform across the CTA
                const float2 t = Ts[(size_t)(k * r + l) * 32 + lane];
                // (dead rows of a ragged last tile hold T = 0: they must not send the warp down the slow path)
                if (!((live0 && t.x == 0.f) || (live1 && t.y == 0.f))) {
                    val.x = live0 ? e.x * __fdividef(sum.x, t.x) : 0.f;
                    val.y = live1 ? e.y * __fdividef(sum.y, t.y) : 0.f;
                } else {  // exact zero factor: explicit leave-one-out products
                    float2 acc = make_float2(0.f, 0.f);
                    for (int pp = p0; pp < p1; ++pp) {
                        const int q = __ldg(p.lists + pp) >> 5;
                        if (q >= Q) continue;
                        const float wq0 = p.wA ? (float)p.wA[q] : ws[q];
                        float2 v = make_float2(wq0, wq0);
                        for (int kk = 0; kk < D; ++kk)
                            if (kk != k) v = mul2(v, Ts[__ldg(p.off1 + (size_t)q * p.DP + kk) + lane]);
                        acc = add2(acc, v);
                    }
                    val.x = live0 ? (t.x != 0.f ? e.x * __fdividef(sum.x, t.x) : e.x * acc.x) : 0.f;
                    val.y = live1 ? (t.y != 0.f ? e.y * __fdividef(sum.y, t.y) : e.y * acc.y) : 0.f;
                }
            }
            out[j] = val;
        }
        // fp16 hi|lo operand image for the backward MMA (tc_layout.cuh); the power-of-two scale comes from the
        // previous step's max|E| and keeps the entries in fp16's normal range.  Rows i0 and i0+1 are adjacent
        // 16-byte rows of the same core matrix; this unit owns 8 bytes (4 values of l) of each.
        uint32_t hi[LU], lo[LU];   // [0..1]: row i0, l pairs; [2..3]: row i0 + 1
#pragma unroll
        for (int j = 0; j < LU; j += 2) {
            mx = fmaxf(mx, fmaxf(fmaxf(fabsf(out[j].x), fabsf(out[j].y)), fmaxf(fabsf(out[j + 1].x), fabsf(out[j + 1].y))));
            const float2 a0 = make_float2(out[j].x * sc, out[j + 1].x * sc);   // row i0, l = j, j+1
            const float2 a1 = make_float2(out[j].y * sc, out[j + 1].y * sc);   // row i0 + 1
            const __half2 h0 = __float22half2_rn(a0), h1 = __float22half2_rn(a1);
            const float2 g0 = __half22float2(h0), g1 = __half22float2(h1);
            const __half2 l0 = __float22half2_rn(make_float2(a0.x - g0.x, a0.y - g0.y));
            const __half2 l1 = __float22half2_rn(make_float2(a1.x - g1.x, a1.y - g1.y));
            hi[j / 2] = *reinterpret_cast<const uint32_t *>(&h0);
            hi[LU / 2 + j / 2] = *reinterpret_cast<const uint32_t *>(&h1);
            lo[j / 2] = *reinterpret_cast<const uint32_t *>(&l0);
            lo[LU / 2 + j / 2] = *reinterpret_cast<const uint32_t *>(&l1);
        }
        if (p.Erow) {   // node-grid path: E itself, [k][i][rpad]
            mx = fmaxf(mx, fmaxf(fmaxf(fabsf(out[0].x), fabsf(out[0].y)), fmaxf(fabsf(out[1].x), fabsf(out[1].y))));
            mx = fmaxf(mx, fmaxf(fmaxf(fabsf(out[2].x), fabsf(out[2].y)), fmaxf(fabsf(out[3].x), fabsf(out[3].y))));
            float *er = p.Erow + ((size_t)k * p.mpad + i0) * rpad + lu * LU;
            *reinterpret_cast<float4 *>(er) = make_float4(out[0].x, out[1].x, out[2].x, out[3].x);
            *reinterpret_cast<float4 *>(er + rpad) = make_float4(out[0].y, out[1].y, out[2].y, out[3].y);
            continue;
        }
        const int lb = (lu * LU) >> 3, lin = (lu * LU) & 7;
        __half *base = p.Eimg + gptb_eimg_offset(k, i0, p.mpad, rpad) + lin;
        uint2 *dh = reinterpret_cast<uint2 *>(base + (size_t)lb * 64);
        uint2 *dl = reinterpret_cast<uint2 *>(base + (size_t)(nlb + lb) * 64);
        dh[0] = make_uint2(hi[0], hi[1]);
        dh[2] = make_uint2(hi[2], hi[3]);    // row i0 + 1: 16 bytes on
        dl[0] = make_uint2(lo[0], lo[1]);
        dl[2] = make_uint2(lo[2], lo[3]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0 && mx > 0.f && !p.redo) atomicMax(reinterpret_cast<int *>(p.emax_out), __float_as_int(mx));
}

}  // namespace

// Offset tables of the kernel above, from the 0-based core index table Iz[D][Q] (host).
int gptb_contract_tc_build_tables(gptb_handle *h, const int32_t *Iz) {
    const gptb_config &c = h->cfg;
    const int DP = (c.D + 3) & ~3;
    std::vector<int32_t> off1((size_t)c.Q * DP, 0);
    for (int q = 0; q < c.Q; ++q)
        for (int k = 0; k < c.D; ++k) off1[(size_t)q * DP + k] = (k * c.r + Iz[(size_t)k * c.Q + q]) * 32;
    // CSR over (k,l): the q with I[q,k] = l in increasing q (the order computeA's findin visits them, :267),
    // every list padded to a multiple of four entries that address the all-zero row q = Q
    const size_t ngrp = (size_t)c.D * h->rpad;
    std::vector<int32_t> cnt(ngrp, 0), ptr(ngrp + 1, 0);
    for (int k = 0; k < c.D; ++k)
        for (int q = 0; q < c.Q; ++q) cnt[(size_t)k * h->rpad + Iz[(size_t)k * c.Q + q]]++;
    for (size_t g = 0; g < ngrp; ++g) ptr[g + 1] = ptr[g] + ((cnt[g] + 3) & ~3);
    std::vector<int32_t> lists((size_t)ptr[ngrp] + 4, c.Q * 32);
    h->ct_nlists = (int)lists.size();
    {
        std::vector<int32_t> fill(ptr.begin(), ptr.end() - 1);
        for (int k = 0; k < c.D; ++k)
            for (int q = 0; q < c.Q; ++q) lists[fill[(size_t)k * h->rpad + Iz[(size_t)k * c.Q + q]]++] = q * 32;
    }
    if (h->ct_off1) cudaFree(h->ct_off1);
    if (h->ct_lptr) cudaFree(h->ct_lptr);
    if (h->ct_lists) cudaFree(h->ct_lists);
    h->ct_off1 = h->ct_lptr = h->ct_lists = nullptr;
    GPTB_CUDA_OK(h, cudaMalloc((void **)&h->ct_off1, off1.size() * sizeof(int32_t)));
    GPTB_CUDA_OK(h, cudaMalloc((void **)&h->ct_lptr, ptr.size() * sizeof(int32_t)));
    GPTB_CUDA_OK(h, cudaMalloc((void **)&h->ct_lists, lists.size() * sizeof(int32_t)));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->ct_off1, off1.data(), off1.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->ct_lptr, ptr.data(), ptr.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->ct_lists, lists.data(), lists.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    return GPTB_OK;
}

size_t gptb_contract_tc_smem(const gptb_handle *h);
bool gptb_contract_tc_usable(const gptb_handle *h) {
    return h->ct_off1 && h->cfg.D <= 16 && gptb_contract_tc_smem(h) <= 227 * 1024;
}

size_t gptb_contract_tc_smem(const gptb_handle *h) {
    const gptb_config &c = h->cfg;
    const size_t DP4 = (c.D + 3) / 4;
    return ((size_t)c.D * c.r * 32 + (size_t)(c.Q + 1) * 32 + (size_t)CNW * 32) * sizeof(float2) +
           (size_t)((c.Q + 3) & ~3) * sizeof(float) + (size_t)c.Q * DP4 * 16 + (size_t)h->ct_nlists * 4 +
           ((size_t)c.D * h->rpad + 4) * 4 + 64;
}

// Returns GPTB_ERR_UNSUPPORTED (without an error message) when the shape does not fit; the caller then
// uses the generic kernel of contract.cu.
int gptb_launch_contract_tc(gptb_handle *h, const float *T, const double *w, const double *yb, int64_t bs, int64_t mpad,
                            double escale, bool want_grad, double *fhat, double *gradw, void *Eimg) {
    const gptb_config &c = h->cfg;
    const size_t smem = gptb_contract_tc_smem(h);
    if (smem > 227 * 1024 || c.D > 16 || !h->ct_off1) return GPTB_ERR_UNSUPPORTED;
    ContractTcParams p;
    p.T = T;
    p.w = w;
    p.wA = nullptr;
    p.off1 = h->ct_off1;
    p.lptr = h->ct_lptr;
    p.lists = h->ct_lists;
    p.yb = yb;
    p.fhat = fhat;
    p.gradw = gradw;
    p.gradwF = (h->f32_grad && want_grad) ? h->gradFbase : nullptr;
    p.gw_stride = h->gradF_qstride;
    p.Eimg = (__half *)Eimg;
    p.escale_img = h->tc_state + 1;
    p.emax_out = h->tc_state + 0;
    p.bs = bs;
    p.mpad = mpad;
    p.escale = escale;
    p.DP = (c.D + 3) & ~3; p.r = c.r; p.rpad = h->rpad; p.Q = c.Q;
    p.want_grad = want_grad ? 1 : 0;
    p.n_lists = h->ct_nlists;
    p.redo = h->k2_redo;
    p.wA = h->k2_wA;
    p.redo_flag = h->nan_flag + 2;
    p.Erow = want_grad ? h->k2_Erow : nullptr;
    p.t_rowmajor = h->k2_T_rowmajor ? 1 : 0;
    const int64_t mext = (bs + 127) / 128 * 128;  // rows beyond bs inside the last 128-tile are dead (zeros)
    dim3 grid((unsigned)(mext / CRB));
#define GPTB_CONTRACT_CASE(DT)                                                                              \
    case DT: {                                                                                              \
        auto kfn = contract_tc_kernel<DT>;                                                                  \
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        kfn<<<grid, CT_THREADS, smem, h->stream>>>(p);                                                      \
    } break;
    switch (c.D) {
        GPTB_CONTRACT_CASE(1) GPTB_CONTRACT_CASE(2) GPTB_CONTRACT_CASE(3) GPTB_CONTRACT_CASE(4)
        GPTB_CONTRACT_CASE(5) GPTB_CONTRACT_CASE(6) GPTB_CONTRACT_CASE(7) GPTB_CONTRACT_CASE(8)
        GPTB_CONTRACT_CASE(9) GPTB_CONTRACT_CASE(10) GPTB_CONTRACT_CASE(11) GPTB_CONTRACT_CASE(12)
        GPTB_CONTRACT_CASE(13) GPTB_CONTRACT_CASE(14) GPTB_CONTRACT_CASE(15) GPTB_CONTRACT_CASE(16)
        default: return GPTB_ERR_UNSUPPORTED;
    }
#undef GPTB_CONTRACT_CASE
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// ------------------------------------------------------------------------------------------------
// Forward-only contraction of the posterior-predictive sweep: fhat_s[i] = sum_q w_s[q] prod_k T[k, s*r + I[q,k], i]
// for the Sg samples whose factor products sit side by side in T ([D][ncols_total][mpad], K1 with a grouped image).
// grid = (row blocks of 64, Sg).  Only the T tile lives in shared memory (no V), so several CTAs share an SM and
// hide the gather latency.  Optionally accumulates mean_acc[i] += mean_scale * fhat_s[i] (pred_samples' average).
// ------------------------------------------------------------------------------------------------
namespace {

struct ContractFwdParams {
    const float *T;
    const double *w;         // [Sg][Q]
    const int32_t *off1;     // [Q][DP]
    double *fh;              // [Sg][mpad] or null
    double *mean_acc;        // [mpad] or null
    double mean_scale;
    int64_t bs, mpad;
    int DP, r, Q, ncols_total;
};

template <int DT>
__global__ void __launch_bounds__(256) contract_fwd_kernel(const ContractFwdParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int D = DT;
    constexpr int DP4 = (DT + 3) / 4;
    constexpr int NW = 8;
    const int r = p.r, Q = p.Q, DR = D * r;
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5, s = blockIdx.y;
    const int64_t i0 = (int64_t)blockIdx.x * CRB + 2 * lane;
    float2 *Ts = reinterpret_cast<float2 *>(smem_raw);           // [DR][32]
    float *ws = reinterpret_cast<float *>(Ts + (size_t)DR * 32); // [Q rounded up to 4]
    float2 *Fs = reinterpret_cast<float2 *>(ws + ((Q + 3) & ~3)); // [NW][32]
    for (int t = tid; t < Q; t += 256) ws[t] = (float)p.w[(size_t)s * Q + t];
    for (int c = wp; c < DR; c += NW) {
        const int k = c / r, l = c - k * r;
        Ts[(size_t)c * 32 + lane] =
            *reinterpret_cast<const float2 *>(p.T + ((size_t)k * p.ncols_total + (size_t)s * r + l) * p.mpad + i0);
    }
    __syncthreads();
    const float2 *Trow = Ts + lane;
    float2 f = make_float2(0.f, 0.f);
#pragma unroll 2
    for (int q = wp; q < Q; q += NW) {
        const int4 *o = reinterpret_cast<const int4 *>(p.off1) + (size_t)q * DP4;
        float2 v = make_float2(1.f, 1.f);
#pragma unroll
        for (int g4 = 0; g4 < DP4; ++g4) {
            const int4 oo = __ldg(o + g4);
            if (g4 * 4 + 0 < D) v = mul2(v, Trow[oo.x]);
            if (g4 * 4 + 1 < D) v = mul2(v, Trow[oo.y]);
            if (g4 * 4 + 2 < D) v = mul2(v, Trow[oo.z]);
            if (g4 * 4 + 3 < D) v = mul2(v, Trow[oo.w]);
        }
        const float wq = ws[q];
        f = __ffma2_rn(make_float2(wq, wq), v, f);
    }
    Fs[wp * 32 + lane] = f;
    __syncthreads();
    if (wp == 0) {
        f = Fs[lane];
#pragma unroll
        for (int t = 1; t < NW; ++t) f = add2(f, Fs[t * 32 + lane]);
        const bool live0 = i0 < p.bs, live1 = i0 + 1 < p.bs;
        if (p.fh) {
            if (live0) p.fh[(size_t)s * p.mpad + i0] = (double)f.x;
            if (live1) p.fh[(size_t)s * p.mpad + i0 + 1] = (double)f.y;
        }
        if (p.mean_acc) {
            if (live0) atomicAdd(&p.mean_acc[i0], p.mean_scale * (double)f.x);
            if (live1) atomicAdd(&p.mean_acc[i0 + 1], p.mean_scale * (double)f.y);
        }
    }
}

}  // namespace

int gptb_launch_contract_fwd(gptb_handle *h, const float *T, const double *w_group, int Sg, int64_t bs, int64_t mpad,
                             int ncols_total, double *fh_group, double *mean_acc, double mean_scale) {
    const gptb_config &c = h->cfg;
    const size_t smem = (size_t)c.D * c.r * 32 * sizeof(float2) + (size_t)((c.Q + 3) & ~3) * sizeof(float) +
                        8 * 32 * sizeof(float2) + 64;
    if (smem > 227 * 1024 || !h->ct_off1) return GPTB_ERR_UNSUPPORTED;
    ContractFwdParams p;
    p.T = T; p.w = w_group; p.off1 = h->ct_off1; p.fh = fh_group; p.mean_acc = mean_acc; p.mean_scale = mean_scale;
    p.bs = bs; p.mpad = mpad;
    p.DP = (c.D + 3) & ~3; p.r = c.r; p.Q = c.Q; p.ncols_total = ncols_total;
    const int64_t mext = (bs + 127) / 128 * 128;
    dim3 grid((unsigned)(mext / CRB), (unsigned)Sg);
#define GPTB_CONTRACT_CASE(DT)                                                                              \
    case DT: {                                                                                              \
        auto kfn = contract_fwd_kernel<DT>;                                                                 \
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        kfn<<<grid, 256, smem, h->stream>>>(p);                                                             \
    } break;
    switch (c.D) {
        GPTB_CONTRACT_CASE(1) GPTB_CONTRACT_CASE(2) GPTB_CONTRACT_CASE(3) GPTB_CONTRACT_CASE(4)
        GPTB_CONTRACT_CASE(5) GPTB_CONTRACT_CASE(6) GPTB_CONTRACT_CASE(7) GPTB_CONTRACT_CASE(8)
        GPTB_CONTRACT_CASE(9) GPTB_CONTRACT_CASE(10) GPTB_CONTRACT_CASE(11) GPTB_CONTRACT_CASE(12)
        GPTB_CONTRACT_CASE(13) GPTB_CONTRACT_CASE(14) GPTB_CONTRACT_CASE(15) GPTB_CONTRACT_CASE(16)
        default: return GPTB_ERR_UNSUPPORTED;
    }
#undef GPTB_CONTRACT_CASE
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// ------------------------------------------------------------------------------------------------
// Forward-only contraction of the node-grid posterior sweep (pred_nodes.cu): the same sum as contract_fwd_kernel,
//     fhat_s[i] = sum_q w_s[q] prod_k T[k, s*r + I[q,k], i]                       (computeV, computefhat :208-230)
// but T comes row-major per sample from the interpolation kernel, T2[k][row][ncols] with the columns of sample s at
// col0 + s*r, and is transposed into the [k*r + l][row] tile while it is staged.  grid = (64-row tiles, samples).
// ------------------------------------------------------------------------------------------------
namespace {

struct ContractFwd2Params {
    const float *T2;
    const double *w;         // [ns][Q]
    const int32_t *off1;     // [Q][DP]
    double *fh;              // [ns][fh_stride] or null
    double *mean_acc;        // [nrows] or null
    double mean_scale;
    int64_t nrows, t2_stride, fh_stride;
    int DP, r, Q, ncols, col0;
};

template <int DT>
__global__ void __launch_bounds__(256) contract_fwd2_kernel(const ContractFwd2Params p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int D = DT;
    constexpr int DP4 = (DT + 3) / 4;
    constexpr int NW = 8;
    const int r = p.r, Q = p.Q, DR = D * r;
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5, s = blockIdx.y;
    const int64_t tile0 = (int64_t)blockIdx.x * CRB;
    const int64_t i0 = tile0 + 2 * lane;
    float2 *Ts = reinterpret_cast<float2 *>(smem_raw);           // [DR][32]
    float *Tf = reinterpret_cast<float *>(Ts);                   // the same as [DR][64 rows]
    float *ws = reinterpret_cast<float *>(Ts + (size_t)DR * 32); // [Q rounded up to 4]
    float2 *Fs = reinterpret_cast<float2 *>(ws + ((Q + 3) & ~3)); // [NW][32]
    // the (uniform) gather offsets, staged once per CTA: with ~200 KB of the SM's 228 KB in shared memory the L1 is too
    // small to keep them beside the streaming tile loads (ncu: long_scoreboard led the stalls when they were __ldg'd)
    int4 *off1s = reinterpret_cast<int4 *>(Fs + NW * 32);        // [Q][DP4]
    for (int t = tid; t < Q * DP4; t += 256) off1s[t] = __ldg(reinterpret_cast<const int4 *>(p.off1) + t);
    for (int t = tid; t < Q; t += 256) ws[t] = (float)p.w[(size_t)s * Q + t];
    const int cbase = p.col0 + s * r;
    if (((r | p.ncols | cbase) & 3) == 0) {
        // 16-byte pieces: thread = (row, k, piece); lanes along the rows (conflict-free stores)
        const int q4 = r >> 2, total = CRB * D * q4;
        for (int t = tid; t < total; t += 256) {
            const int rowt = t & (CRB - 1), rest = t >> 6;
            const int k = rest / q4, q = rest - k * q4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (tile0 + rowt < p.nrows)
                v = __ldg(reinterpret_cast<const float4 *>(p.T2 + ((size_t)k * p.t2_stride + tile0 + rowt) * p.ncols + cbase) + q);
            float *d = Tf + (size_t)(k * r + 4 * q) * CRB + rowt;
            d[0] = v.x; d[CRB] = v.y; d[2 * CRB] = v.z; d[3 * CRB] = v.w;
        }
    } else {
        for (int t = tid; t < CRB * DR; t += 256) {
            const int rowt = t & (CRB - 1), c = t >> 6;
            const int k = c / r, l = c - k * r;
            float v = 0.f;
            if (tile0 + rowt < p.nrows) v = __ldg(p.T2 + ((size_t)k * p.t2_stride + tile0 + rowt) * p.ncols + cbase + l);
            Tf[(size_t)c * CRB + rowt] = v;
        }
    }
    __syncthreads();
    const float2 *Trow = Ts + lane;
    float2 f = make_float2(0.f, 0.f);
#pragma unroll 2
    for (int q = wp; q < Q; q += NW) {
        const int4 *o = off1s + (size_t)q * DP4;
        float2 v = make_float2(1.f, 1.f);
#pragma unroll
        for (int g4 = 0; g4 < DP4; ++g4) {
            const int4 oo = o[g4];
            if (g4 * 4 + 0 < D) v = mul2(v, Trow[oo.x]);
            if (g4 * 4 + 1 < D) v = mul2(v, Trow[oo.y]);
            if (g4 * 4 + 2 < D) v = mul2(v, Trow[oo.z]);
            if (g4 * 4 + 3 < D) v = mul2(v, Trow[oo.w]);
        }
        const float wq = ws[q];
        f = __ffma2_rn(make_float2(wq, wq), v, f);
    }
    Fs[wp * 32 + lane] = f;
    __syncthreads();
    if (wp == 0) {
        f = Fs[lane];
#pragma unroll
        for (int t = 1; t < NW; ++t) f = add2(f, Fs[t * 32 + lane]);
        const bool live0 = i0 < p.nrows, live1 = i0 + 1 < p.nrows;
        if (p.fh) {
            if (live0) p.fh[(size_t)s * p.fh_stride + i0] = (double)f.x;
            if (live1) p.fh[(size_t)s * p.fh_stride + i0 + 1] = (double)f.y;
        }
        if (p.mean_acc) {
            if (live0) atomicAdd(&p.mean_acc[i0], p.mean_scale * (double)f.x);
            if (live1) atomicAdd(&p.mean_acc[i0 + 1], p.mean_scale * (double)f.y);
        }
    }
}


// ------------------------------------------------------------------------------------------------------------
// Posterior sweep with the samples on the lanes: T2[k][row][l][SL] (SL = 8, 16 or 32 samples side by side, written
// by the interpolation as whole 4*SL-byte pieces).  A warp owns 32/SL rows at a time, lane = (row, sample), and
// brings their D*r lines (20 KB at BASELINE config 5) into its own shared-memory tile [k*r + l][32 lanes] with 16-byte
// cp.async — a straight copy, no transposition; a gather T[k][I[q,k]] is then one conflict-free LDS.32 of every
// lane's own sample.  Warps are independent (load / compute phases of different warps overlap); one CTA per SM, as
// many warps as tiles fit.  (Gathering straight from global memory through L1 was measured 4x slower: the lines of a
// row are re-read ~Q/r times each and did not stay in L1 — 1.4 TB of L2 traffic at C5.)
struct ContractFwd3Params {
    const float *T2;
    const double *w;         // [ns][Q]
    const int32_t *off1;     // [Q][DP]: (k*r + I[q,k]) * 32 = float index of the (k,l) line in a warp's tile
    double *fh;              // [ns][fh_stride] or null
    double *mean_acc;        // [nrows] or null (plain read-modify-write: one warp per row, groups run in stream order)
    double mean_scale;
    int64_t nrows, t2_stride, fh_stride;
    int DP, r, Q, ns, SL, sl_shift;
    int wpt;                 // warps per tile (1..4)
};

template <int DT>
__global__ void __launch_bounds__(1024) contract_fwd3_kernel(const ContractFwd3Params p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int D = DT;
    constexpr int DP4 = (DT + 3) / 4;
    const int r = p.r, Q = p.Q, SL = p.SL, DR = D * r;
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    // wpt warps share a tile: each copies its part of it and takes its part of the q, the partial sums meet in shared
    // memory.  (Shared memory holds 9 tiles at BASELINE config 5; one warp per tile left the LSU at 63 % with 13 %
    // occupancy: 116 ms; two warps per tile 106 ms.)
    const int wpt = p.wpt, pr = wp / wpt, half = wp - pr * wpt, npair = (blockDim.x >> 5) / wpt;
    int4 *offs = reinterpret_cast<int4 *>(smem_raw);                      // [Q][DP4] byte offsets of the (k,l) lines
    float *ws = reinterpret_cast<float *>(offs + (size_t)Q * DP4);        // [Q][SL]
    float *fx = ws + (((size_t)Q * SL + 31) & ~(size_t)31);               // [npair][3][32] partial sums of the other warps
    float *tiles = fx + (size_t)npair * 96;                               // [npair][DR][32]
    for (int t = tid; t < Q * DP4; t += blockDim.x) {
        int4 v = __ldg(reinterpret_cast<const int4 *>(p.off1) + t);
        v.x <<= 2; v.y <<= 2; v.z <<= 2; v.w <<= 2;
        offs[t] = v;
    }
    for (int t = tid; t < Q * SL; t += blockDim.x) {
        const int q = t >> p.sl_shift, sidx = t & (SL - 1);
        ws[t] = sidx < p.ns ? (float)p.w[(size_t)sidx * Q + q] : 0.f;
    }
    __syncthreads();
    const int rpw = 32 >> p.sl_shift;                     // rows per tile
    const int sidx = lane & (SL - 1), rsub = lane >> p.sl_shift;
    const int64_t ngrp = (p.nrows + rpw - 1) / rpw;
    const float *wl = ws + sidx;
    float *tile = tiles + (size_t)pr * DR * 32;
    const char *tb = reinterpret_cast<const char *>(tile + lane);
    const uint32_t tile_s = tcx::smem_u32(tile);
    // this lane's 16-byte pieces of the copy: piece (kl, within) with kl = it*8 + 4*half + lane/8, within = lane%8 ->
    // lanes 4*within .. 4*within+3 of line kl, i.e. row (4*within) >> sl_shift, samples (4*within) & (SL-1) ..
    const int within = lane & 7, c_rsub = (4 * within) >> p.sl_shift, c_s = (4 * within) & (SL - 1);
    const int qh = ((Q + wpt - 1) / wpt + 1) & ~1;        // warp j of the tile: q in [j*qh, (j+1)*qh)
    const int q_lo = min(half * qh, Q), q_hi = min(q_lo + qh, Q);
    const int bar_id = 1 + pr, bar_n = 32 * wpt;
    for (int64_t gi = (int64_t)blockIdx.x * npair + pr; gi < ngrp; gi += (int64_t)gridDim.x * npair) {
        const int64_t row0 = gi * rpw;
        {
            const int64_t crow = min(row0 + c_rsub, p.nrows - 1);         // rows past the end: a valid row, never stored
            for (int kl = (lane >> 3) + 4 * half; kl < DR; kl += 4 * wpt) {
                const int k = kl / r, l = kl - k * r;
                const float *src = p.T2 + (((size_t)k * p.t2_stride + crow) * r + l) * SL + c_s;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tile_s + (uint32_t)(kl * 32 + 4 * within) * 4u), "l"(src)
                             : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(bar_n) : "memory");   // every part of the tile has landed
        }
        float f0 = 0.f, f1 = 0.f;
        int q = q_lo;
        for (; q + 1 < q_hi; q += 2) {                    // two independent product chains
            const int4 *o0 = offs + (size_t)q * DP4, *o1 = o0 + DP4;
            float v0 = 1.f, v1 = 1.f;
#pragma unroll
            for (int g4 = 0; g4 < DP4; ++g4) {
                const int4 a = o0[g4], b = o1[g4];
                if (g4 * 4 + 0 < D) { v0 *= *reinterpret_cast<const float *>(tb + a.x); v1 *= *reinterpret_cast<const float *>(tb + b.x); }
                if (g4 * 4 + 1 < D) { v0 *= *reinterpret_cast<const float *>(tb + a.y); v1 *= *reinterpret_cast<const float *>(tb + b.y); }
                if (g4 * 4 + 2 < D) { v0 *= *reinterpret_cast<const float *>(tb + a.z); v1 *= *reinterpret_cast<const float *>(tb + b.z); }
                if (g4 * 4 + 3 < D) { v0 *= *reinterpret_cast<const float *>(tb + a.w); v1 *= *reinterpret_cast<const float *>(tb + b.w); }
            }
            f0 = fmaf(wl[(size_t)q * SL], v0, f0);
            f1 = fmaf(wl[(size_t)(q + 1) * SL], v1, f1);
        }
        if (q < q_hi) {
            const int4 *o0 = offs + (size_t)q * DP4;
            float v0 = 1.f;
#pragma unroll
            for (int g4 = 0; g4 < DP4; ++g4) {
                const int4 a = o0[g4];
                if (g4 * 4 + 0 < D) v0 *= *reinterpret_cast<const float *>(tb + a.x);
                if (g4 * 4 + 1 < D) v0 *= *reinterpret_cast<const float *>(tb + a.y);
                if (g4 * 4 + 2 < D) v0 *= *reinterpret_cast<const float *>(tb + a.z);
                if (g4 * 4 + 3 < D) v0 *= *reinterpret_cast<const float *>(tb + a.w);
            }
            f0 = fmaf(wl[(size_t)q * SL], v0, f0);
        }
        float f = f0 + f1;
        if (half) fx[(pr * 3 + half - 1) * 32 + lane] = f;
        asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(bar_n) : "memory");   // partial sums written; all warps are done with the tile
        if (half) continue;
        for (int j = 1; j < wpt; ++j) f += fx[(pr * 3 + j - 1) * 32 + lane];
        const int64_t row = row0 + rsub;
        const bool out = row < p.nrows && sidx < p.ns;
        if (p.fh && out) p.fh[(size_t)sidx * p.fh_stride + row] = (double)f;
        if (p.mean_acc) {
            double sum = out ? (double)f : 0.0;
            for (int o = SL >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (row < p.nrows && sidx == 0) p.mean_acc[row] += p.mean_scale * sum;
        }
    }
}

}  // namespace

int gptb_launch_contract_fwd2(gptb_handle *h, const float *T2, const double *w_group, int ns, int col0, int ncols, int64_t nrows,
                              int64_t t2_stride, double *fh, int64_t fh_stride, double *mean_acc, double mean_scale,
                              cudaStream_t s) {
    const gptb_config &c = h->cfg;
    const size_t smem = (size_t)c.D * c.r * 32 * sizeof(float2) + (size_t)((c.Q + 3) & ~3) * sizeof(float) +
                        8 * 32 * sizeof(float2) + (size_t)c.Q * ((c.D + 3) / 4) * 16 + 64;
    if (smem > 227 * 1024 || !h->ct_off1) return GPTB_ERR_UNSUPPORTED;
    ContractFwd2Params p;
    p.T2 = T2; p.w = w_group; p.off1 = h->ct_off1; p.fh = fh; p.mean_acc = mean_acc; p.mean_scale = mean_scale;
    p.nrows = nrows; p.t2_stride = t2_stride; p.fh_stride = fh_stride;
    p.DP = (c.D + 3) & ~3; p.r = c.r; p.Q = c.Q; p.ncols = ncols; p.col0 = col0;
    dim3 grid((unsigned)((nrows + CRB - 1) / CRB), (unsigned)ns);
#define GPTB_CONTRACT_CASE(DT)                                                                              \
    case DT: {                                                                                              \
        auto kfn = contract_fwd2_kernel<DT>;                                                                \
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        kfn<<<grid, 256, smem, s>>>(p);                                                                     \
    } break;
    switch (c.D) {
        GPTB_CONTRACT_CASE(1) GPTB_CONTRACT_CASE(2) GPTB_CONTRACT_CASE(3) GPTB_CONTRACT_CASE(4)
        GPTB_CONTRACT_CASE(5) GPTB_CONTRACT_CASE(6) GPTB_CONTRACT_CASE(7) GPTB_CONTRACT_CASE(8)
        GPTB_CONTRACT_CASE(9) GPTB_CONTRACT_CASE(10) GPTB_CONTRACT_CASE(11) GPTB_CONTRACT_CASE(12)
        GPTB_CONTRACT_CASE(13) GPTB_CONTRACT_CASE(14) GPTB_CONTRACT_CASE(15) GPTB_CONTRACT_CASE(16)
        default: return GPTB_ERR_UNSUPPORTED;
    }
#undef GPTB_CONTRACT_CASE
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_contract_fwd3(gptb_handle *h, const float *T2, const double *w_group, int ns, int SL, int64_t nrows,
                              int64_t t2_stride, double *fh, int64_t fh_stride, double *mean_acc, double mean_scale,
                              cudaStream_t s) {
    const gptb_config &c = h->cfg;
    if (!h->ct_off1 || (SL != 8 && SL != 16 && SL != 32) || ns > SL) return GPTB_ERR_UNSUPPORTED;
    const size_t tables = (size_t)c.Q * ((c.D + 3) / 4) * 16 + (((size_t)c.Q * SL + 31) & ~(size_t)31) * sizeof(float) +
                          15 * 96 * sizeof(float);
    const size_t tile = (size_t)c.D * c.r * 32 * sizeof(float);
    if (tables + tile > 227 * 1024) return GPTB_ERR_UNSUPPORTED;
    int npair = (int)std::min<size_t>(15, (227 * 1024 - tables) / tile);     // tiles (15 named barriers beside barrier 0)
    int wpt = h->pred_warps > 0 ? std::min(h->pred_warps, 4) : 2;             // warps per tile (GPTB_OPT_PRED_WARPS; measured 1: 116 ms, 2: 106, 3: 109, 4: 112 at C5)
    while (wpt > 1 && npair * wpt > 32) --wpt;
    const int nwarps = wpt * npair;
    const size_t smem = tables + (size_t)npair * tile;
    ContractFwd3Params p;
    p.T2 = T2; p.w = w_group; p.off1 = h->ct_off1; p.fh = fh; p.mean_acc = mean_acc; p.mean_scale = mean_scale;
    p.nrows = nrows; p.t2_stride = t2_stride; p.fh_stride = fh_stride;
    p.DP = (c.D + 3) & ~3; p.r = c.r; p.Q = c.Q; p.ns = ns; p.SL = SL; p.sl_shift = SL == 8 ? 3 : (SL == 16 ? 4 : 5);
    p.wpt = wpt;
    const dim3 grid((unsigned)h->sm_count);
#define GPTB_CONTRACT_CASE(DT)                                                                              \
    case DT: {                                                                                              \
        auto kfn = contract_fwd3_kernel<DT>;                                                                \
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        kfn<<<grid, nwarps * 32, smem, s>>>(p);                                                             \
    } break;
    switch (c.D) {
        GPTB_CONTRACT_CASE(1) GPTB_CONTRACT_CASE(2) GPTB_CONTRACT_CASE(3) GPTB_CONTRACT_CASE(4)
        GPTB_CONTRACT_CASE(5) GPTB_CONTRACT_CASE(6) GPTB_CONTRACT_CASE(7) GPTB_CONTRACT_CASE(8)
        GPTB_CONTRACT_CASE(9) GPTB_CONTRACT_CASE(10) GPTB_CONTRACT_CASE(11) GPTB_CONTRACT_CASE(12)
        GPTB_CONTRACT_CASE(13) GPTB_CONTRACT_CASE(14) GPTB_CONTRACT_CASE(15) GPTB_CONTRACT_CASE(16)
        default: return GPTB_ERR_UNSUPPORTED;
    }
#undef GPTB_CONTRACT_CASE
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}
