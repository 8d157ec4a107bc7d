// Node-grid evaluation of the per-dimension feature sums (tensor-core path).
//
// For one input dimension k the forward quantity of phidotU (GPT_SGLD.jl:193-205),
//     T_k(x)[l] = amp * sum_j cos(x*Zt[j,k] + b[j,k]) * U[j,l,k],
// is a function of ONE scalar x and band-limited by zmax_k = max_j |Zt[j,k]|.  The n features of dimension k therefore
// only have to be generated at the points of a node grid that covers the input domain [lo_k, hi_k] — G cells of
// half-width h_k <= theta/zmax_k with P Chebyshev nodes each — instead of at every minibatch row:
//
//   forward   K1 (tc_path.cu, unchanged) runs on the G*P nodes  ->  Tn[k][l][node]
//             node_coef_kernel : node values -> Chebyshev coefficients per cell (a P x P cosine transform)
//             interp_kernel    : T[k][l][i] = sum_p T_p(u_ik) * coef[k][cell_ik][p][l]          (u in [-1,1] inside the cell)
//   backward  spread_kernel    : M[k][cell][p][l] += T_p(u_ik) * E[k][i][l]     (the adjoint of the interpolation, rows
//                                visited in cell order so that a warp keeps the sums of a cell in registers)
//             node_adj_kernel  : moments -> node weights (adjoint cosine transform), max|.| for the fp16 image scale
//             node_pack_kernel : fp16 hi|lo operand image of the node weights
//             K3 (tc_path.cu, unchanged) runs on the G*P nodes  ->  gradU  (computePsi + grad-U, :276-286, :405-408)
//
// With theta = 2 and P = 12 the interpolation error is ~5e-11 relative (tools/proto_nodes.py; bound
// theta^P / (2^(P-1) P!)), far below the fp16-split MMA's own 1e-6; the cosines evaluated per step drop from
// m*n*D (6.6e8 at BASELINE config 4) to G*P*n*D (~1e6), and every feature is generated once for the forward AND the
// backward pass in the sense that matters: per node, not per row.  A thread-block-cluster fusion of K1 -> K2 -> K3 that
// keeps the rows' own features on chip cannot work on this part: a row carries D*n = 10^4 features (40 KB as fp16
// hi|lo), K2 needs all D dimensions of a row before K3 may start, so at most ~12 rows per SM could be in flight while
// every SM re-streams the 2.6 MB operand image of U — and DSMEM moves 17-21 B/clk/SM (B300_MICROARCH.md).
#include <cuda_fp16.h>
#include <math.h>

#include <algorithm>
#include <vector>

#include "common.cuh"
#include "tc_layout.cuh"

namespace {

constexpr int PMAX = 16;

// Chebyshev nodes of the first kind t_j = cos(pi (j + 1/2) / P) and the cosine-transform matrix
//   C[p][j] = (2 - [p == 0]) / P * cos(pi p (j + 1/2) / P):   coef_p = sum_j C[p][j] f(t_j),  f(u) = sum_p coef_p T_p(u).
__constant__ float c_cheb[PMAX * PMAX];

// T_0..T_{P-1} at u by the three-term recurrence (uniform across the warp in both kernels below).
template <int P>
__device__ __forceinline__ void cheb_basis(float u, float (&t)[P]) {
    t[0] = 1.f;
    if (P > 1) t[1] = u;
    const float u2 = 2.f * u;
#pragma unroll
    for (int p = 2; p < P; ++p) t[p] = fmaf(u2, t[p - 1], -t[p - 2]);
}

// ------------------------------------------------------------------------------------------------ coefficients
// grid (ceil(G*rpad/256), kn).  Tn[k][l][c*P + j] -> Cf[k][c][l][PP]  (PP = P rounded up to 4: a lane reads the
// coefficients of its column with 16-byte loads; columns r..rpad-1 and entries P..PP-1 are zero).
__global__ void __launch_bounds__(256) node_coef_kernel(const float *__restrict__ Tn, float *__restrict__ Cf, int G, int P,
                                                        int PP, int r, int rpad, int mnpad) {
    const int k = blockIdx.y;
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= G * rpad) return;
    const int l = t % rpad, c = t / rpad;
    float f[PMAX];
#pragma unroll
    for (int j = 0; j < PMAX; ++j) f[j] = (j < P && l < r) ? Tn[((size_t)k * r + l) * mnpad + c * P + j] : 0.f;
    float *out = Cf + (((size_t)k * G + c) * rpad + l) * PP;
    for (int p = 0; p < PP; ++p) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < PMAX; ++j) a = fmaf(c_cheb[p * PMAX + j], f[j], a);   // (rows / entries >= P of the matrix are zero)
        out[p] = a;
    }
}

// ------------------------------------------------------------------------------------------------ interpolation
// grid (ceil(mext / IROWS), kn), 256 threads.  A CTA holds the coefficient table of its dimension in shared memory;
// a warp takes 32 consecutive rows at a time, lane = column l: per row P conflict-free LDS + P FFMA, and the 32 x 32
// block of results goes through a padded shared-memory tile so that T[k][l][i] is written with unit stride over rows.
constexpr int IROWS = 512;
constexpr int ITHREADS = 256;

template <int P, int NL /* 32-column passes: rpad / 32 rounded up */>
__global__ void __launch_bounds__(ITHREADS) interp_kernel(const float *__restrict__ Cf, const int32_t *__restrict__ cell,
                                                          const float *__restrict__ uu, float *__restrict__ T, int r, int rpad,
                                                          int G, int64_t bs, int64_t mpad, int64_t mext) {
    extern __shared__ __align__(16) float sm[];
    constexpr int PP = (P + 3) & ~3;
    const int k = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float *tab = sm;                                                   // [G][rpad][PP]
    float *tile = sm + (size_t)G * rpad * PP + warp * (32 * 33);       // per-warp [32][33]
    {
        const float4 *src = reinterpret_cast<const float4 *>(Cf + (size_t)k * G * rpad * PP);
        float4 *dst = reinterpret_cast<float4 *>(tab);
        for (int t = tid; t < G * rpad * PP / 4; t += ITHREADS) dst[t] = src[t];
    }
    __syncthreads();
    const int64_t row0 = (int64_t)blockIdx.x * IROWS;
    const int32_t *ck = cell + (size_t)k * mpad;
    const float *uk = uu + (size_t)k * mpad;
    float *Tk = T + (size_t)k * r * mpad;
    for (int64_t i0 = row0 + warp * 32; i0 < min(row0 + IROWS, mext); i0 += (ITHREADS / 32) * 32) {
        const int c_l = ck[i0 + lane];
        const float u_l = uk[i0 + lane];
#pragma unroll
        for (int pass = 0; pass < NL; ++pass) {
            const int l = pass * 32 + lane;
            const float4 *tl = reinterpret_cast<const float4 *>(tab) + (size_t)(l < rpad ? l : 0) * (PP / 4);
#pragma unroll 4
            for (int j = 0; j < 32; ++j) {
                const int c = __shfl_sync(0xffffffffu, c_l, j);
                const float u = __shfl_sync(0xffffffffu, u_l, j);
                float tb[PP];
                cheb_basis<PP>(u, tb);                                // (entries P..PP-1 meet zero coefficients)
                const float4 *a = tl + (size_t)c * rpad * (PP / 4);
                float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
                for (int q = 0; q < PP / 4; ++q) {
                    const float4 v = a[q];
                    acc0 = fmaf(tb[4 * q], v.x, acc0);
                    acc1 = fmaf(tb[4 * q + 1], v.y, acc1);
                    acc0 = fmaf(tb[4 * q + 2], v.z, acc0);
                    acc1 = fmaf(tb[4 * q + 3], v.w, acc1);
                }
                tile[j * 33 + lane] = acc0 + acc1;
            }
            __syncwarp();
            // rows beyond bs (dead rows of the last tile) hold T = 0, as K1 leaves them
            const bool live = i0 + lane < bs;
#pragma unroll 8
            for (int ll = 0; ll < 32; ++ll) {
                const int lo = pass * 32 + ll;
                if (lo < r) Tk[(size_t)lo * mpad + i0 + lane] = live ? tile[lane * 33 + ll] : 0.f;
            }
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------------------------------------ cell sort
// grid (kn), 1024 threads: a stable counting sort of the batch rows of one dimension by cell.  Warp w owns the
// contiguous rows [w*chunk, (w+1)*chunk) and a private histogram, so there are no shared-memory atomics between warps;
// equal cells inside a warp are aggregated with match_any.
__global__ void __launch_bounds__(1024) cell_sort_kernel(const int32_t *__restrict__ cell, int32_t *__restrict__ sorted, int G,
                                                         int64_t bs, int64_t mpad) {
    extern __shared__ int hist[];          // [32][G]
    const int k = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int32_t *ck = cell + (size_t)k * mpad;
    int32_t *sk = sorted + (size_t)k * mpad;
    for (int t = tid; t < 32 * G; t += 1024) hist[t] = 0;
    __syncthreads();
    const int64_t chunk = ((bs + 31) / 32 + 31) / 32 * 32;   // rows per warp, a multiple of 32
    const int64_t lo = (int64_t)warp * chunk, hi = min(bs, lo + chunk);
    int *hw = hist + warp * G;
    for (int64_t i0 = lo; i0 < hi; i0 += 32) {
        const int64_t i = i0 + lane;
        const bool v = i < hi;
        const int c = v ? ck[i] : -1;
        const unsigned m = __match_any_sync(0xffffffffu, c);
        if (v && lane == __ffs(m) - 1) hw[c] += __popc(m);
        __syncwarp();
    }
    __syncthreads();
    // exclusive prefix over (cell, warp) in cell-major order
    for (int c = tid; c < G; c += 1024) {
        int run = 0;
        for (int w = 0; w < 32; ++w) {
            const int v = hist[w * G + c];
            hist[w * G + c] = run;
            run += v;
        }
        hist[32 * G + c] = run;            // total of the cell
    }
    __syncthreads();
    if (warp == 0) {                        // scan of the cell totals, 32 at a time
        int base = 0;
        for (int c0 = 0; c0 < G; c0 += 32) {
            const int c = c0 + lane;
            int v = c < G ? hist[32 * G + c] : 0, inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int nb = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += nb;
            }
            if (c < G) hist[32 * G + c] = base + inc - v;
            base += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    __syncthreads();
    for (int64_t i0 = lo; i0 < hi; i0 += 32) {
        const int64_t i = i0 + lane;
        const bool v = i < hi;
        const int c = v ? ck[i] : -1;
        const unsigned m = __match_any_sync(0xffffffffu, c);
        if (v) {
            const int pos = hist[32 * G + c] + hw[c] + __popc(m & ((1u << lane) - 1u));
            sk[pos] = (int32_t)i;
        }
        __syncwarp();
        if (v && lane == __ffs(m) - 1) hw[c] += __popc(m);
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------ spreading
// grid (ceil(bs / SROWS), kn), 256 threads; lane = column l.  A warp walks 32 * SITER consecutive rows of the
// dimension's cell-sorted order; the P moment sums of the current cell stay in registers and go to the global
// table (FP32 RED, one 4*rpad-byte line per p) only when the cell changes.
constexpr int SROWS = 512;
constexpr int STHREADS = 256;

template <int P, int NL>
__global__ void __launch_bounds__(STHREADS) spread_kernel(const float *__restrict__ Erow, const int32_t *__restrict__ cell,
                                                          const float *__restrict__ uu, const int32_t *__restrict__ sorted,
                                                          float *__restrict__ Mn, int rpad, int mnpad, int64_t bs,
                                                          int64_t mpad) {
    extern __shared__ __align__(16) float sm[];
    float (*stage)[NL][32][32] = reinterpret_cast<float (*)[NL][32][32]>(sm);   // per warp: the E rows of 32 batch rows
    const int k = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int32_t *ck = cell + (size_t)k * mpad, *sk = sorted + (size_t)k * mpad;
    const float *uk = uu + (size_t)k * mpad;
    const float *Ek = Erow + (size_t)k * mpad * rpad;
    float *Mk = Mn + (size_t)k * mnpad * rpad;
    constexpr int per_warp = SROWS / (STHREADS / 32);
    const int64_t s0 = (int64_t)blockIdx.x * SROWS + warp * per_warp, s1 = min(bs, s0 + per_warp);
    float acc[NL][P];
#pragma unroll
    for (int a = 0; a < NL; ++a)
#pragma unroll
        for (int p = 0; p < P; ++p) acc[a][p] = 0.f;
    int cur = -1;
    auto flush = [&]() {
        if (cur < 0) return;
#pragma unroll
        for (int a = 0; a < NL; ++a) {
            const int l = a * 32 + lane;
            if (l < rpad) {
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    atomicAdd(Mk + ((size_t)cur * P + p) * rpad + l, acc[a][p]);
                    acc[a][p] = 0.f;
                }
            }
        }
    };
    for (int64_t sb = s0; sb < s1; sb += 32) {
        const int64_t s = sb + lane;
        int row_l = 0, c_l = -1;
        float u_l = 0.f;
        if (s < s1) {
            row_l = sk[s];
            c_l = ck[row_l];
            u_l = uk[row_l];
        }
        // all 32 row loads of the block are issued before any is used (one row = one 4*rpad-byte line; the rows of a
        // cell are scattered over the batch, so every load is a separate trip to L2)
#pragma unroll
        for (int a = 0; a < NL; ++a) {
            const int l = a * 32 + lane;
            float ev[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int row = __shfl_sync(0xffffffffu, row_l, j);      // (row 0 for the lanes past the end: a valid address)
                ev[j] = l < rpad ? Ek[(size_t)row * rpad + l] : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) stage[warp][a][j][lane] = ev[j];
        }
        __syncwarp();
        const int cnt = (int)min((int64_t)32, s1 - sb);
        for (int j = 0; j < cnt; ++j) {
            const int c = __shfl_sync(0xffffffffu, c_l, j);
            const float u = __shfl_sync(0xffffffffu, u_l, j);
            if (c != cur) {   // warp-uniform
                flush();
                cur = c;
            }
            float tb[P];
            cheb_basis<P>(u, tb);
#pragma unroll
            for (int a = 0; a < NL; ++a) {
                const float e = stage[warp][a][j][lane];
#pragma unroll
                for (int p = 0; p < P; ++p) acc[a][p] = fmaf(tb[p], e, acc[a][p]);
            }
        }
        __syncwarp();
    }
    flush();
}

// ------------------------------------------------------------------------------------------------ adjoint transform
// grid (ceil(G*rpad/256), kn): moments M[c*P + p][l] -> node weights En[c*P + j][l] = sum_p C[p][j] M[p], in place,
// and the maximum magnitude (for the power-of-two scale of the fp16 image).
__global__ void __launch_bounds__(256) node_adj_kernel(float *__restrict__ Mn, int G, int P, int rpad, int mnpad,
                                                       float *__restrict__ max_out) {
    const int k = blockIdx.y;
    const int t = blockIdx.x * 256 + threadIdx.x;
    float mx = 0.f;
    if (t < G * rpad) {
        const int l = t % rpad, c = t / rpad;
        float *mp = Mn + ((size_t)k * mnpad + (size_t)c * P) * rpad + l;
        float m[PMAX];
#pragma unroll
        for (int p = 0; p < PMAX; ++p) m[p] = p < P ? mp[(size_t)p * rpad] : 0.f;
        for (int j = 0; j < P; ++j) {
            float e = 0.f;
#pragma unroll
            for (int p = 0; p < PMAX; ++p) e = fmaf(c_cheb[p * PMAX + j], m[p], e);
            mp[(size_t)j * rpad] = e;
            mx = fmaxf(mx, fabsf(e));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0 && mx > 0.f) atomicMax(reinterpret_cast<int *>(max_out), __float_as_int(mx));
}

__device__ __forceinline__ float pow2_scale(float mx) {
    if (!(mx > 0.f) || !isfinite(mx)) return 1.f;
    int e;
    frexpf(mx, &e);
    return ldexpf(1.f, 9 - e);   // maps mx into [256, 512), as tc_path.cu does for the row image
}

// grid (ceil(mnpad * rpad/8 / 256), kn): fp16 hi|lo operand image of the node weights (tc_layout.cuh, the layout K3
// reads), scaled by the exact power of two for THIS step's maximum; the weight table is cleared for the next step.
__global__ void __launch_bounds__(256) node_pack_kernel(float *__restrict__ Mn, __half *__restrict__ img, int mnode, int mnpad,
                                                        int rpad, const float *__restrict__ max_in, float *__restrict__ scale_out,
                                                        float *__restrict__ max_next) {
    const int k = blockIdx.y;
    const int nlb = rpad / 8;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const float sc = pow2_scale(*max_in);
    if (t == 0 && k == 0) {
        scale_out[0] = sc;
        scale_out[1] = 1.f / sc;
        *max_next = 0.f;
    }
    if (t >= mnpad * nlb) return;
    const int lb = t % nlb, i = t / nlb;
    float *src = Mn + ((size_t)k * mnpad + i) * rpad + lb * 8;
    __align__(16) __half hi[8], lo[8];
    if (i < mnode) {
        const float4 a = *reinterpret_cast<const float4 *>(src), b = *reinterpret_cast<const float4 *>(src + 4);
        const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float s = v[e] * sc;
            hi[e] = __float2half_rn(s);
            lo[e] = __float2half_rn(s - __half2float(hi[e]));
        }
        *reinterpret_cast<float4 *>(src) = make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(src + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) hi[e] = lo[e] = __float2half_rn(0.f);
    }
    __half *base = img + gptb_eimg_offset(k, i, mnpad, rpad);
    *reinterpret_cast<uint4 *>(base + (size_t)lb * 64) = *reinterpret_cast<const uint4 *>(hi);
    *reinterpret_cast<uint4 *>(base + (size_t)(nlb + lb) * 64) = *reinterpret_cast<const uint4 *>(lo);
}

template <typename T>
int nd_alloc(gptb_handle *h, T **p, size_t count) {
    if (*p) cudaFree(*p);
    *p = nullptr;
    cudaError_t e = cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) {
        gptb_set_error(h, "node grid: cudaMalloc(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
        *p = nullptr;
        return GPTB_ERR_NOMEM;
    }
    return GPTB_OK;
}

}  // namespace

void gptb_nodes_free(gptb_handle *h) {
    gptb_nodes &g = h->nodes;
    cudaFree(g.lo); cudaFree(g.invh); cudaFree(g.XnF); cudaFree(g.Tn); cudaFree(g.Cf); cudaFree(g.Mn); cudaFree(g.Enimg);
    cudaFree(g.cell); cudaFree(g.uu); cudaFree(g.sorted); cudaFree(g.Erow);
    g = gptb_nodes();
}

// Builds the grid for the current domain / features / options.  Leaves g.m = 0 (and returns GPTB_OK) when the node grid
// is not applicable: no domain, coefficient table larger than shared memory, ...
int gptb_nodes_setup(gptb_handle *h) {
    const gptb_config &c = h->cfg;
    gptb_nodes &g = h->nodes;
    h->grid_valid = true;
    g.m = 0;
    if (!h->have_domain || !h->have_features || !gptb_is_tc(h)) return GPTB_OK;
    const int P = h->node_P;
    int G = 1;
    for (int k = 0; k < c.D; ++k) {
        const double range = h->dom_hi[k] - h->dom_lo[k];
        if (!(range >= 0.0) || !std::isfinite(range)) return GPTB_OK;
        G = std::max(G, (int)std::ceil(range * h->zmax[k] / (2.0 * h->node_theta)));
    }
    const int64_t m = (int64_t)G * P;
    // the interpolation kernel keeps one dimension's table (m x rpad floats) plus its transpose tiles in shared memory;
    // the sort keeps 33 x G counters
    const int PP = (P + 3) & ~3;
    if ((size_t)G * PP * h->rpad * 4 + (ITHREADS / 32) * 32 * 33 * 4 > 200 * 1024 || (size_t)33 * G * 4 > 200 * 1024) return GPTB_OK;
    const int mnpad = (int)((m + 127) / 128 * 128);
    std::vector<double> lo(c.D), invh(c.D);
    std::vector<float> xn((size_t)c.D * mnpad, 0.f);
    for (int k = 0; k < c.D; ++k) {
        double range = h->dom_hi[k] - h->dom_lo[k];
        double hk = range / (2.0 * G);
        if (!(hk > 0.0)) hk = 0.5 / std::max(h->zmax[k], 1e-30);   // constant column: any cell width will do
        lo[k] = h->dom_lo[k];
        invh[k] = 1.0 / hk;
        for (int cc = 0; cc < G; ++cc)
            for (int j = 0; j < P; ++j)
                xn[(size_t)k * mnpad + cc * P + j] = (float)(lo[k] + (2.0 * cc + 1.0 + std::cos(M_PI * (j + 0.5) / P)) * hk);
        for (int i = (int)m; i < mnpad; ++i) xn[(size_t)k * mnpad + i] = (float)lo[k];
    }
    std::vector<float> cm((size_t)PMAX * PMAX, 0.f);
    for (int p = 0; p < P; ++p)
        for (int j = 0; j < P; ++j) cm[(size_t)p * PMAX + j] = (float)((p == 0 ? 1.0 : 2.0) / P * std::cos(M_PI * p * (j + 0.5) / P));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyToSymbol(c_cheb, cm.data(), sizeof(float) * PMAX * PMAX));
    GPTB_TRY(nd_alloc(h, &g.lo, c.D));
    GPTB_TRY(nd_alloc(h, &g.invh, c.D));
    GPTB_TRY(nd_alloc(h, &g.XnF, (size_t)c.D * mnpad));
    GPTB_TRY(nd_alloc(h, &g.Tn, (size_t)c.D * c.r * mnpad));
    GPTB_TRY(nd_alloc(h, &g.Cf, (size_t)c.D * G * PP * h->rpad));
    GPTB_TRY(nd_alloc(h, &g.Mn, (size_t)c.D * mnpad * h->rpad));
    GPTB_TRY(nd_alloc(h, (unsigned char **)&g.Enimg, gptb_tc_eimg_bytes(h, mnpad)));
    GPTB_CUDA_OK(h, cudaMemcpy(g.lo, lo.data(), sizeof(double) * c.D, cudaMemcpyHostToDevice));
    GPTB_CUDA_OK(h, cudaMemcpy(g.invh, invh.data(), sizeof(double) * c.D, cudaMemcpyHostToDevice));
    GPTB_CUDA_OK(h, cudaMemcpy(g.XnF, xn.data(), sizeof(float) * xn.size(), cudaMemcpyHostToDevice));
    GPTB_CUDA_OK(h, cudaMemset(g.Mn, 0, sizeof(float) * (size_t)c.D * mnpad * h->rpad));
    GPTB_CUDA_OK(h, cudaMemset(h->tc_state + 10, 0, 2 * sizeof(float)));
    g.G = G; g.P = P; g.m = (int)m; g.mnpad = mnpad;
    g.batch_mpad = 0;
    return GPTB_OK;
}

// Per-batch scratch (cell, u, cell-sorted order, FP32 E rows), sized with the handle's batch capacity.
int gptb_nodes_ensure_batch(gptb_handle *h) {
    gptb_nodes &g = h->nodes;
    if (g.batch_mpad == h->mpad) return GPTB_OK;
    const gptb_config &c = h->cfg;
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    GPTB_TRY(nd_alloc(h, &g.cell, (size_t)c.D * h->mpad));
    GPTB_TRY(nd_alloc(h, &g.uu, (size_t)c.D * h->mpad));
    GPTB_TRY(nd_alloc(h, &g.sorted, (size_t)c.D * h->mpad));
    GPTB_TRY(nd_alloc(h, &g.Erow, (size_t)c.D * h->mpad * h->rpad));
    g.batch_mpad = h->mpad;
    return GPTB_OK;
}

#define GPTB_P_SWITCH(P, NL, CALL)                \
    switch ((P) * 4 + (NL)) {                     \
        case 4 * 4 + 1: { CALL(4, 1); } break;    \
        case 6 * 4 + 1: { CALL(6, 1); } break;    \
        case 8 * 4 + 1: { CALL(8, 1); } break;    \
        case 10 * 4 + 1: { CALL(10, 1); } break;  \
        case 12 * 4 + 1: { CALL(12, 1); } break;  \
        case 14 * 4 + 1: { CALL(14, 1); } break;  \
        case 16 * 4 + 1: { CALL(16, 1); } break;  \
        case 4 * 4 + 2: { CALL(4, 2); } break;    \
        case 6 * 4 + 2: { CALL(6, 2); } break;    \
        case 8 * 4 + 2: { CALL(8, 2); } break;    \
        case 10 * 4 + 2: { CALL(10, 2); } break;  \
        case 12 * 4 + 2: { CALL(12, 2); } break;  \
        case 14 * 4 + 2: { CALL(14, 2); } break;  \
        case 16 * 4 + 2: { CALL(16, 2); } break;  \
        default:                                  \
            gptb_set_error(h, "node grid: unsupported order P=%d", (P)); \
            return GPTB_ERR_UNSUPPORTED;          \
    }

// Forward for the dimensions [k0, k0+kn): K1 on the nodes, coefficients.  (The interpolation needs every dimension's
// cells of the batch and runs once, see gptb_nodes_interp.)
int gptb_nodes_forward_range(gptb_handle *h, const void *Uimg, int k0, int kn, cudaStream_t s) {
    gptb_nodes &g = h->nodes;
    const gptb_config &c = h->cfg;
    (void)s;
    GPTB_TRY(gptb_launch_fwd_tc_range(h, g.XnF, Uimg, g.m, g.mnpad, g.Tn, k0, kn));
    const int PP = (g.P + 3) & ~3;
    node_coef_kernel<<<dim3((g.G * h->rpad + 255) / 256, kn), 256, 0, h->stream>>>(
        g.Tn + (size_t)k0 * c.r * g.mnpad, g.Cf + (size_t)k0 * g.G * PP * h->rpad, g.G, g.P, PP, c.r, h->rpad, g.mnpad);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_nodes_interp(gptb_handle *h, int64_t bs, float *T) {
    gptb_nodes &g = h->nodes;
    const gptb_config &c = h->cfg;
    const int64_t mext = (bs + 127) / 128 * 128;
    const int NL = (h->rpad + 31) / 32;
    const int PP = (g.P + 3) & ~3;
    const size_t smem = ((size_t)g.G * PP * h->rpad + (ITHREADS / 32) * 32 * 33) * sizeof(float);
    const dim3 grid((unsigned)((mext + IROWS - 1) / IROWS), c.D);
#define GPTB_CALL(PP, NN)                                                                                         \
    auto kfn = interp_kernel<PP, NN>;                                                                             \
    GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));           \
    kfn<<<grid, ITHREADS, smem, h->stream>>>(g.Cf, g.cell, g.uu, T, c.r, h->rpad, g.G, bs, h->mpad, mext);
    GPTB_P_SWITCH(g.P, NL, GPTB_CALL)
#undef GPTB_CALL
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// Cell-sorted row order of the staged batch (depends on the batch only, not on the parameters).
int gptb_nodes_sort(gptb_handle *h, int64_t bs, cudaStream_t s) {
    gptb_nodes &g = h->nodes;
    const size_t smem = (size_t)33 * g.G * sizeof(int);
    static bool optin = false;
    if (!optin) {
        cudaFuncSetAttribute(cell_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        optin = true;
    }
    cell_sort_kernel<<<h->cfg.D, 1024, smem, s>>>(g.cell, g.sorted, g.G, bs, h->mpad);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// Backward: spread the FP32 E rows K2 wrote (g.Erow) onto the nodes, build the node image, run K3 on the nodes.
int gptb_nodes_spread(gptb_handle *h, int64_t bs) {
    gptb_nodes &g = h->nodes;
    const gptb_config &c = h->cfg;
    const int NL = (h->rpad + 31) / 32;
    const dim3 grid((unsigned)((bs + SROWS - 1) / SROWS), c.D);
    const size_t smem = (size_t)(STHREADS / 32) * NL * 32 * 32 * sizeof(float);
#define GPTB_CALL(PP, NN)                                                                                       \
    auto kfn = spread_kernel<PP, NN>;                                                                           \
    GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));         \
    kfn<<<grid, STHREADS, smem, h->stream>>>(g.Erow, g.cell, g.uu, g.sorted, g.Mn, h->rpad, g.mnpad, bs, h->mpad);
    GPTB_P_SWITCH(g.P, NL, GPTB_CALL)
#undef GPTB_CALL
    const int slot = (int)(g.steps & 1);
    node_adj_kernel<<<dim3((g.G * h->rpad + 255) / 256, c.D), 256, 0, h->stream>>>(g.Mn, g.G, g.P, h->rpad, g.mnpad,
                                                                                  h->tc_state + 10 + slot);
    node_pack_kernel<<<dim3((g.mnpad * (h->rpad / 8) + 255) / 256, c.D), 256, 0, h->stream>>>(
        g.Mn, (__half *)g.Enimg, g.m, g.mnpad, h->rpad, h->tc_state + 10 + slot, h->tc_state + 1, h->tc_state + 10 + (slot ^ 1));
    g.steps++;
    h->launches += 3;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_nodes_backward_range(gptb_handle *h, double *gradU, int k0, int kn) {
    gptb_nodes &g = h->nodes;
    return gptb_launch_bwd_tc_range(h, g.XnF, g.Enimg, g.m, g.mnpad, gradU, k0, kn);
}
