// K2 — core contraction, residual and leave-one-out back-projection, one thread per minibatch row.
//
// Replaces (all /root/reference/GPT_SGLD.jl):
//   computeV     :208-220   V[q,i]   = prod_k T[k, I[q,k], i]
//   computefhat  :223-230   fhat[i]  = sum_q w[q] V[q,i]
//   grad-w       :393       gradw   += sum_i e_i V[:,i],      e_i = (N/bs)(y_i - fhat_i)/signal_var
//   computeU_phi :246-258   leave-one-out products  prod_{j!=k} T[j, I[q,j], i]
//   computeA     :261-273   A[l,k,i] = sum_{q: I[q,k]=l} w[q] * leave-one-out
// and emits E[k][l][i] = e_i * A[l,k,i], the right-hand operand of the factor-gradient contraction
// (GPT_SGLD.jl:405-408).  The leave-one-out product is formed division-free from prefix/suffix
// products (the reference divides V by the left-out factor, :253, which is Inf/NaN when a factor
// is 0; the two agree to O(1e-16) relative otherwise).
//
// Layout: T[k][l][i] and E[k][l][i] have the row index fastest, so a warp's 32 rows read/write
// 128 contiguous bytes for every (k,l); the gather index I[q,k] is uniform across the warp, so the
// shared-memory reads are conflict-free broadcasts of the index plus unit-stride data.
#include <cuda_fp16.h>

#include "common.cuh"
#include "tc_layout.cuh"

namespace {

template <typename real>
struct Smem {
    __device__ static real *get() {
        extern __shared__ __align__(16) unsigned char gptb_contract_smem[];
        return reinterpret_cast<real *>(gptb_contract_smem);
    }
};

// One block handles RB = blockDim.x rows.  Shared memory: Ts[D*r][RB], As[D*r][RB] (if want_grad),
// I8[Q][D] bytes, ws[Q] real, gw[warps][Q] double partials.
template <typename real, int DT>
__global__ void contract_kernel(const real *__restrict__ T, const double *__restrict__ w,
                                const int32_t *__restrict__ I /*[D][Q] 0-based*/, const double *__restrict__ yb,
                                int64_t bs, int64_t mpad, int Dr, int r, int Q, double escale, int want_grad,
                                double *__restrict__ fhat, double *__restrict__ gradw, real *__restrict__ E,
                                __half *__restrict__ Eimg, int rpad, const float *__restrict__ escale_img_ptr,
                                float *__restrict__ emax_out) {
    const int D = DT > 0 ? DT : Dr;
    const int RB = blockDim.x;
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.x * RB;
    const int64_t i = i0 + tid;
    const bool live = i < bs;
    const int DR = D * r;

    real *Ts = Smem<real>::get();
    real *As = Ts + (size_t)DR * RB;
    real *ws = As + (want_grad ? (size_t)DR * RB : 0);
    double *gw = reinterpret_cast<double *>(ws + ((Q + 1) & ~1));
    unsigned char *I8 = reinterpret_cast<unsigned char *>(gw + (want_grad ? (size_t)(RB / 32) * Q : 0));

    for (int t = tid; t < Q; t += RB) ws[t] = (real)w[t];
    for (int t = tid; t < Q * D; t += RB) {
        int q = t / D, k = t % D;
        I8[t] = (unsigned char)I[(size_t)k * Q + q];
    }
    for (int c = 0; c < DR; ++c) {
        Ts[(size_t)c * RB + tid] = (i < mpad) ? T[(size_t)c * mpad + i] : (real)0;
        if (want_grad) As[(size_t)c * RB + tid] = (real)0;
    }
    __syncthreads();

    // pass 1: fhat
    real f = 0;
    for (int q = 0; q < Q; ++q) {
        real p = ws[q];
        const unsigned char *iq = I8 + q * D;
#pragma unroll
        for (int k = 0; k < (DT > 0 ? DT : 16); ++k)
            if (k < D) p *= Ts[(size_t)(k * r + iq[k]) * RB + tid];
        f += p;
    }
    if (live) fhat[i] = (double)f;
    if (!want_grad) return;

    // pass 2: residual-weighted leave-one-out scatter + gradw
    const real e = live ? (real)(escale * (yb[i] - (double)f)) : (real)0;
    const int warp = tid >> 5, lane = tid & 31;
    for (int q = 0; q < Q; ++q) {
        const unsigned char *iq = I8 + q * D;
        real t[DT > 0 ? DT : 16], pre[(DT > 0 ? DT : 16) + 1];
        pre[0] = 1;
#pragma unroll
        for (int k = 0; k < (DT > 0 ? DT : 16); ++k)
            if (k < D) {
                t[k] = Ts[(size_t)(k * r + iq[k]) * RB + tid];
                pre[k + 1] = pre[k] * t[k];
            }
        real V = (DT > 0) ? pre[DT] : pre[D];
        real suf = ws[q] * e;
#pragma unroll
        for (int k = (DT > 0 ? DT : 16) - 1; k >= 0; --k)
            if (k < D) {
                As[(size_t)(k * r + iq[k]) * RB + tid] += pre[k] * suf;
                suf *= t[k];
            }
        double g = (double)(V * e);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) g += __shfl_xor_sync(0xffffffffu, g, o);
        if (lane == 0) gw[(size_t)warp * Q + q] = g;
    }
    __syncthreads();
    for (int q = tid; q < Q; q += RB) {
        double s = 0.0;
        for (int wv = 0; wv < RB / 32; ++wv) s += gw[(size_t)wv * Q + q];
        atomicAdd(&gradw[q], s);
    }
    if (E != nullptr && i < mpad) {
        for (int c = 0; c < DR; ++c) E[(size_t)c * mpad + i] = As[(size_t)c * RB + tid];
    }
    if (Eimg != nullptr && i < mpad) {
        // fp16 hi|lo operand image for the backward MMA (see tc_layout.cuh).  The power-of-two scale
        // keeps the entries in fp16's normal range; it comes from the previous step's max|E| (device side).
        const float sc = *escale_img_ptr;
        float mx = 0.f;
        const int nb = rpad / 8;
        for (int k = 0; k < D; ++k) {
            __half *base = Eimg + gptb_eimg_offset(k, i, mpad, rpad);
            for (int bq = 0; bq < nb; ++bq) {
                __align__(16) __half hi[8], lo[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    int l = bq * 8 + u;
                    float v = (l < r) ? (float)As[(size_t)(k * r + l) * RB + tid] : 0.f;
                    mx = fmaxf(mx, fabsf(v));
                    v *= sc;
                    __half hh = __float2half_rn(v);
                    hi[u] = hh;
                    lo[u] = __float2half_rn(v - __half2float(hh));
                }
                *reinterpret_cast<uint4 *>(base + (size_t)bq * 64) = *reinterpret_cast<const uint4 *>(hi);
                *reinterpret_cast<uint4 *>(base + (size_t)(nb + bq) * 64) = *reinterpret_cast<const uint4 *>(lo);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0 && mx > 0.f) atomicMax(reinterpret_cast<int *>(emax_out), __float_as_int(mx));
    }
}

template <typename real>
int launch_typed(gptb_handle *h, const void *T, const double *w, const double *yb, int64_t bs, int64_t mpad,
                 double escale, bool want_grad, double *fhat, double *gradw, void *E, void *Eimg,
                 const float *escale_img, float *emax_out) {
    const gptb_config &c = h->cfg;
    const int DR = c.D * c.r;
    auto smem_for = [&](int RB) -> size_t {
        size_t s = (size_t)DR * RB * sizeof(real) * (want_grad ? 2 : 1);
        s += (size_t)((c.Q + 1) & ~1) * sizeof(real);
        if (want_grad) s += (size_t)(RB / 32) * c.Q * sizeof(double);
        s += (size_t)c.Q * c.D;
        return s + 16;
    };
    int RB = 128;
    while (RB > 32 && smem_for(RB) > 200 * 1024) RB >>= 1;
    size_t smem = smem_for(RB);
    if (smem > 227 * 1024) {
        gptb_set_error(h, "contract: D*r=%d too large for shared memory", DR);
        return GPTB_ERR_UNSUPPORTED;
    }
    const int64_t mext = (bs + 127) / 128 * 128;  // rows beyond bs inside the last 128-tile are dead (zeros)
    dim3 grid((unsigned)(mext / RB));
#define GPTB_CONTRACT_CASE(DT)                                                                                     \
    case DT: {                                                                                                     \
        auto kfn = contract_kernel<real, DT>;                                                                      \
        GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
        kfn<<<grid, RB, smem, h->stream>>>((const real *)T, w, h->I, yb, bs, mpad, c.D, c.r, c.Q, escale,          \
                                           want_grad ? 1 : 0, fhat, gradw, (real *)E, (__half *)Eimg, h->rpad,     \
                                           escale_img, emax_out);                                                  \
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
    if (real_is_double)
        return launch_typed<double>(h, T, (const double *)w, yb, bs, mpad, escale, want_grad, fhat, gradw, E, nullptr,
                                    nullptr, nullptr);
    return launch_typed<float>(h, T, (const double *)w, yb, bs, mpad, escale, want_grad, fhat, gradw, E, Eimg,
                               h->tc_state + 1, h->tc_state + 0);
}
