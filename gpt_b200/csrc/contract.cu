// K2 — core contraction, residual and leave-one-out back-projection.
//
// Replaces (all /root/reference/GPT_SGLD.jl):
//   computeV     :208-220   V[q,i]   = prod_k T[k, I[q,k], i]
//   computefhat  :223-230   fhat[i]  = sum_q w[q] V[q,i]
//   grad-w       :393       gradw   += sum_i e_i V[:,i],      e_i = (N/bs)(y_i - fhat_i)/signal_var
//   computeU_phi :246-258   U_phi[q,i,k] = V[q,i] / T[k, I[q,k], i]
//   computeA     :261-273   A[l,k,i] = sum_{q: I[q,k]=l} w[q] U_phi[q,i,k]
// and emits E[k][l][i] = e_i * A[l,k,i], the right-hand operand of the factor-gradient contraction
// (GPT_SGLD.jl:405-408).
//
// Work decomposition: a CTA owns RB = 32 minibatch rows (one warp of rows) and NS = 8 "slices" of threads.
//   pass 1: slice s forms V[q,i] for q = s, s+NS, ...  and its share of fhat_i; V goes to shared memory.
//   pass 2: every (k,l) pair owns the list of q with I[q,k] = l (built once in gptb_set_index), so
//           A[l,k,i] = (sum_{q in list} w_q V[q,i]) / T[k,l,i]  — the reference's own division (:253), taken
//           once per (k,l) instead of once per (q,k).  A slice handles blocks of 8 consecutive l, so it writes
//           whole 16-byte rows of the MMA operand image.  If T[k,l,i] is exactly 0 the quotient is replaced by
//           the explicit leave-one-out product (the reference would produce Inf/NaN there).
// T[k][l][i] / E[k][l][i] have the row index fastest and the gather index is uniform across a warp, so all
// shared-memory traffic is conflict-free unit-stride data plus broadcasts.
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_layout.cuh"

namespace {

constexpr int RB = 32;
constexpr int NS = 8;
constexpr int CT_THREADS = RB * NS;
constexpr int VS = RB + 1;  // padded row stride of Vs (the gradw reduction walks it by q)

struct ContractParams {
    const void *T;
    const double *w;
    const int32_t *I;        // [D][Q] 0-based
    const int32_t *grp_ptr;  // [D*rpad + 1]
    const uint16_t *grp_q;   // [Q*D] q sorted by (k, l)
    const double *yb;
    double *fhat, *gradw;
    void *E;
    __half *Eimg;
    const float *escale_img;
    float *emax_out;
    int64_t bs, mpad;
    double escale;
    int D, r, rpad, Q, want_grad;
    const int *redo_flag;    // see contract_tc.cu: a conditional (1) or unconditional (2) second launch that only rewrites E
    int redo;
    const double *wA;        // weights of computeA when they differ from w (RMSprop sampler), or null
};

template <typename real, int DT>
__global__ void __launch_bounds__(CT_THREADS) contract_kernel(const ContractParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int D = DT > 0 ? DT : p.D;
    constexpr int DU = DT > 0 ? DT : 16;
    const int r = p.r, rpad = p.rpad, Q = p.Q, DR = D * r;
    const int tid = threadIdx.x, row = tid & 31, s = tid >> 5;
    if (p.redo == 1 && *p.redo_flag == 0) return;
    const int64_t i = (int64_t)blockIdx.x * RB + row;
    const bool live = i < p.bs;
    const real *T = reinterpret_cast<const real *>(p.T);

    real *Ts = reinterpret_cast<real *>(smem_raw);             // [DR][RB]
    real *Vs = Ts + (size_t)DR * RB;                           // [Q][VS]
    real *ws = Vs + (size_t)Q * VS;                            // [Q]
    real *Fs = ws + ((Q + 3) & ~3);                            // [NS][RB]
    real *Es = Fs + NS * RB;                                   // [RB]
    int32_t *gp = reinterpret_cast<int32_t *>(Es + RB);        // [D*rpad+1]
    uint16_t *gq = reinterpret_cast<uint16_t *>(gp + ((D * rpad + 4) & ~3));  // [Q*D]
    unsigned char *I8 = reinterpret_cast<unsigned char *>(gq + (((size_t)Q * D + 7) & ~(size_t)7));  // [Q][D]

    real *wsA = ws;            // weights of pass 2
    for (int t = tid; t < Q; t += CT_THREADS) ws[t] = (real)p.w[t];
    for (int t = tid; t < Q * D; t += CT_THREADS) {
        I8[t] = (unsigned char)p.I[(size_t)(t % D) * Q + t / D];
        gq[t] = p.grp_q[t];
    }
    for (int t = tid; t <= D * rpad; t += CT_THREADS) gp[t] = p.grp_ptr[t];
    for (int c = s; c < DR; c += NS) Ts[(size_t)c * RB + row] = T[(size_t)c * p.mpad + i];
    __syncthreads();

    // ---- pass 1: V and fhat
    real f = 0;
    for (int q = s; q < Q; q += NS) {
        const unsigned char *iq = I8 + q * D;
        real v = 1;
#pragma unroll
        for (int k = 0; k < DU; ++k)
            if (k < D) v *= Ts[(size_t)(k * r + iq[k]) * RB + row];
        Vs[(size_t)q * VS + row] = v;
        f += ws[q] * v;
    }
    Fs[s * RB + row] = f;
    __syncthreads();
    f = 0;
#pragma unroll
    for (int t = 0; t < NS; ++t) f += Fs[t * RB + row];
    if (s == 0 && live && !p.redo) p.fhat[i] = (double)f;
    if (!p.want_grad) return;
    const real e = live ? (real)(p.escale * (p.yb[i] - (double)f)) : (real)0;
    if (s == 0) Es[row] = e;
    __syncthreads();

    // ---- gradw: sum over the CTA's rows of e_i V[q,i]
    for (int q = tid; q < Q && !p.redo; q += CT_THREADS) {
        double g = 0.0;
#pragma unroll 8
        for (int ii = 0; ii < RB; ++ii) g += (double)(Es[ii] * Vs[(size_t)q * VS + ii]);
        atomicAdd(&p.gradw[q], g);
    }

    // ---- pass 2: E[k][l][i] = e_i * (sum_{q: I[q,k]=l} w_q V[q,i]) / T[k][l][i]
    const int nlb = rpad / 8;
    const float sc = p.Eimg ? *p.escale_img : 1.f;
    float mx = 0.f;
    real *E = reinterpret_cast<real *>(p.E);
    for (int u = s; u < D * nlb; u += NS) {
        const int k = u / nlb, lb = u % nlb;
        real out[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int l = lb * 8 + j;
            const int p0 = gp[k * rpad + l], p1 = gp[k * rpad + l + 1];
            real val = 0;
            if (p1 > p0) {  // uniform across the CTA
                real sum = 0;
                for (int pp = p0; pp < p1; ++pp) {
                    const int q = gq[pp];
                    sum += (p.wA ? (real)p.wA[q] : wsA[q]) * Vs[(size_t)q * VS + row];
                }
                const real t = Ts[(size_t)(k * r + l) * RB + row];
                if (t != (real)0) {
                    val = e * (sum / t);
                } else if (live) {  // exact zero factor: explicit leave-one-out products
                    real acc = 0;
                    for (int pp = p0; pp < p1; ++pp) {
                        const int q = gq[pp];
                        const unsigned char *iq = I8 + q * D;
                        real v = p.wA ? (real)p.wA[q] : wsA[q];
                        for (int kk = 0; kk < D; ++kk)
                            if (kk != k) v *= Ts[(size_t)(kk * r + iq[kk]) * RB + row];
                        acc += v;
                    }
                    val = e * acc;
                }
            }
            out[j] = val;
        }
        if (E != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int l = lb * 8 + j;
                if (l < r) E[(size_t)(k * r + l) * p.mpad + i] = out[j];
            }
        }
        if (p.Eimg != nullptr) {
            // fp16 hi|lo operand image for the backward MMA (tc_layout.cuh); the power-of-two scale comes from
            // the previous step's max|E| and keeps the entries in fp16's normal range.
            __align__(16) __half hi[8], lo[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float v = (float)out[j];
                mx = fmaxf(mx, fabsf(v));
                v *= sc;
                __half hh = __float2half_rn(v);
                hi[j] = hh;
                lo[j] = __float2half_rn(v - __half2float(hh));
            }
            __half *base = p.Eimg + gptb_eimg_offset(k, i, p.mpad, rpad);
            *reinterpret_cast<uint4 *>(base + (size_t)lb * 64) = *reinterpret_cast<const uint4 *>(hi);
            *reinterpret_cast<uint4 *>(base + (size_t)(nlb + lb) * 64) = *reinterpret_cast<const uint4 *>(lo);
        }
    }
    if (p.Eimg != nullptr) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (row == 0 && mx > 0.f && !p.redo) atomicMax(reinterpret_cast<int *>(p.emax_out), __float_as_int(mx));
    }
}

template <typename real>
int launch_typed(gptb_handle *h, ContractParams &p) {
    const gptb_config &c = h->cfg;
    const int DR = c.D * c.r;
    size_t smem = (size_t)DR * RB * sizeof(real) + (size_t)c.Q * VS * sizeof(real) + (size_t)((c.Q + 3) & ~3) * sizeof(real) +
                  (size_t)(NS * RB + RB) * sizeof(real) + (size_t)((c.D * h->rpad + 4) & ~3) * sizeof(int32_t) +
                  (((size_t)c.Q * c.D + 7) & ~(size_t)7) * sizeof(uint16_t) + (size_t)c.Q * c.D + 64;
    if (smem > 227 * 1024) {
        gptb_set_error(h, "contract: D*r=%d, Q=%d need %zu bytes of shared memory (limit 227 KB)", DR, c.Q, smem);
        return GPTB_ERR_UNSUPPORTED;
    }
    const int64_t mext = (p.bs + 127) / 128 * 128;  // rows beyond bs inside the last 128-tile are dead (zeros)
    dim3 grid((unsigned)(mext / RB));
#define GPTB_CONTRACT_CASE(DT)                                                                              \
    case DT: {                                                                                              \
        auto kfn = contract_kernel<real, DT>;                                                               \
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        kfn<<<grid, CT_THREADS, smem, h->stream>>>(p);                                                      \
    } break;
    switch (c.D) {
        GPTB_CONTRACT_CASE(1) GPTB_CONTRACT_CASE(2) GPTB_CONTRACT_CASE(3) GPTB_CONTRACT_CASE(4)
        GPTB_CONTRACT_CASE(5) GPTB_CONTRACT_CASE(6) GPTB_CONTRACT_CASE(7) GPTB_CONTRACT_CASE(8)
        GPTB_CONTRACT_CASE(9) GPTB_CONTRACT_CASE(10) GPTB_CONTRACT_CASE(11) GPTB_CONTRACT_CASE(12)
        default:
            GPTB_CONTRACT_CASE(0)
    }
#undef GPTB_CONTRACT_CASE
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

}  // namespace

int gptb_launch_contract(gptb_handle *h, const void *T, const void *w, const double *yb, int64_t bs, int64_t mpad,
                         double escale, bool want_grad, double *fhat, double *gradw, void *E, void *Eimg,
                         bool real_is_double) {
    if (bs <= 0) return GPTB_OK;
    if (!real_is_double) {
        const int st = gptb_launch_contract_tc(h, (const float *)T, (const double *)w, yb, bs, mpad, escale, want_grad, fhat,
                                               gradw, Eimg);
        if (st != GPTB_ERR_UNSUPPORTED) return st;   // shapes that do not fit its shared-memory tile fall through
    }
    ContractParams p;
    p.T = T;
    p.w = (const double *)w;
    p.I = h->I;
    p.grp_ptr = h->grp_ptr;
    p.grp_q = h->grp_q;
    p.yb = yb;
    p.fhat = fhat;
    p.gradw = gradw;
    p.E = E;
    p.Eimg = real_is_double ? nullptr : (__half *)Eimg;
    p.escale_img = h->tc_state + 1;
    p.emax_out = h->tc_state + 0;
    p.bs = bs;
    p.mpad = mpad;
    p.escale = escale;
    p.D = h->cfg.D; p.r = h->cfg.r; p.rpad = h->rpad; p.Q = h->cfg.Q;
    p.want_grad = want_grad ? 1 : 0;
    p.redo = (!real_is_double && h->k2_redo == 1) ? 1 : (h->k2_redo == 2 ? 2 : 0);
    p.wA = h->k2_wA;
    p.redo_flag = h->nan_flag + 2;
    return real_is_double ? launch_typed<double>(h, p) : launch_typed<float>(h, p);
}
