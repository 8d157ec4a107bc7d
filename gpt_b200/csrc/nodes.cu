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
//             cell_sort_kernel : records (row, cell, u) of the batch in cell order, per dimension
//             interp_kernel    : T2[k][i][l] = sum_p T_p(u_ik) * coef[k][cell_ik][l][p]         (u in [-1,1] inside the cell)
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

// The same for two records at once: packed FP32 (FFMA2 takes the negated third operand and a scalar multiplier for free).
template <int P>
__device__ __forceinline__ void cheb_basis2(float2 u, float2 (&t)[P]) {
    t[0] = make_float2(1.f, 1.f);
    if (P > 1) t[1] = u;
    const float2 u2 = make_float2(2.f * u.x, 2.f * u.y);
#pragma unroll
    for (int p = 2; p < P; ++p) t[p] = __ffma2_rn(u2, t[p - 1], make_float2(-t[p - 2].x, -t[p - 2].y));
}
__device__ __forceinline__ float2 splat(float c) { return make_float2(c, c); }

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

// The same for `ncols` columns of a wider table (posterior sweep: the columns of several samples side by side):
// Tn[k][col][c*P + j] (tn_cols columns per dimension) -> Cf[k][c][col0 + col][PP] of a table with ctot columns.
// grid (ceil(G*ncols/256), D).
__global__ void __launch_bounds__(256) node_coef_cols_kernel(const float *__restrict__ Tn, float *__restrict__ Cf, int G, int P,
                                                             int PP, int ncols, int tn_cols, int ctot, int col0, int mnpad) {
    const int k = blockIdx.y;
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= G * ncols) return;
    const int col = t % ncols, c = t / ncols;
    float f[PMAX];
#pragma unroll
    for (int j = 0; j < PMAX; ++j) f[j] = j < P ? Tn[((size_t)k * tn_cols + col) * mnpad + c * P + j] : 0.f;
    float *out = Cf + (((size_t)k * G + c) * ctot + col0 + col) * PP;
    for (int p = 0; p < PP; ++p) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < PMAX; ++j) a = fmaf(c_cheb[p * PMAX + j], f[j], a);
        out[p] = a;
    }
}

// The same with the samples on the fastest axis (posterior sweep, pred_nodes.cu): Tn holds ns samples x r columns
// (column s*r + l); coefficient column l*SL + s0 + s of a table with r*SL columns.
__global__ void __launch_bounds__(256) node_coef_lanes_kernel(const float *__restrict__ Tn, float *__restrict__ Cf, int G, int P,
                                                              int PP, int r, int ns, int tn_cols, int SL, int s0, int mnpad) {
    const int k = blockIdx.y;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const int ncols = ns * r;
    if (t >= G * ncols) return;
    const int col = t % ncols, c = t / ncols;
    const int sl = col / r, l = col - sl * r;
    float f[PMAX];
#pragma unroll
    for (int j = 0; j < PMAX; ++j) f[j] = j < P ? Tn[((size_t)k * tn_cols + col) * mnpad + c * P + j] : 0.f;
    float *out = Cf + (((size_t)k * G + c) * (r * SL) + l * SL + s0 + sl) * PP;
    for (int p = 0; p < PP; ++p) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < PMAX; ++j) a = fmaf(c_cheb[p * PMAX + j], f[j], a);
        out[p] = a;
    }
}

// ------------------------------------------------------------------------------------------------ cell sort
// grid (ceil(bs / SCHUNK), D), 256 threads: the batch rows of one dimension are sorted by cell inside chunks of SCHUNK
// rows (a stable counting sort: warp w owns 256 contiguous rows and a private histogram, equal cells inside a warp are
// aggregated with match_any) and written out as records (row, cell, u) in sorted order.  Both row kernels below walk
// the records: consecutive records share their cell, so a warp keeps that cell's coefficients / sums in registers.
constexpr int SCHUNK = 2048;
constexpr int SORT_T = 256;

__global__ void __launch_bounds__(SORT_T) cell_sort_kernel(const int32_t *__restrict__ cell, const float *__restrict__ uu,
                                                           int32_t *__restrict__ srow, int32_t *__restrict__ sc,
                                                           float *__restrict__ su, int G, int64_t bs, int64_t mpad) {
    extern __shared__ int hist[];          // [8][G] per-warp counts, then [G] cell bases
    constexpr int NW = SORT_T / 32;
    const int k = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int32_t *ck = cell + (size_t)k * mpad;
    const float *uk = uu + (size_t)k * mpad;
    const int64_t c0 = (int64_t)blockIdx.x * SCHUNK;
    for (int t = tid; t < (NW + 1) * G; t += SORT_T) hist[t] = 0;
    __syncthreads();
    const int64_t lo = c0 + warp * (SCHUNK / NW), hi = min(bs, lo + SCHUNK / NW);
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
    for (int c = tid; c < G; c += SORT_T) {    // exclusive prefix over the warps of each cell, total into hist[NW][c]
        int run = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            const int v = hist[w * G + c];
            hist[w * G + c] = run;
            run += v;
        }
        hist[NW * G + c] = run;
    }
    __syncthreads();
    if (warp == 0) {                            // exclusive scan of the cell totals, 32 at a time
        int base = 0;
        for (int cb = 0; cb < G; cb += 32) {
            const int c = cb + lane;
            const int v = c < G ? hist[NW * G + c] : 0;
            int inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int nb = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += nb;
            }
            if (c < G) hist[NW * G + c] = base + inc - v;
            base += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    __syncthreads();
    const size_t ob = (size_t)k * mpad + c0;
    for (int64_t i0 = lo; i0 < hi; i0 += 32) {
        const int64_t i = i0 + lane;
        const bool v = i < hi;
        const int c = v ? ck[i] : -1;
        const unsigned m = __match_any_sync(0xffffffffu, c);
        if (v) {
            const int pos = hist[NW * G + c] + hw[c] + __popc(m & ((1u << lane) - 1u));
            srow[ob + pos] = (int32_t)i;
            sc[ob + pos] = c;
            su[ob + pos] = uk[i];
        }
        __syncwarp();
        if (v && lane == __ffs(m) - 1) hw[c] += __popc(m);
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------ interpolation
// grid (ceil(nrows / (8 * IWROWS)), D), 256 threads.  A warp walks IWROWS (8..64) consecutive records of one dimension; lane
// = column (NL blocks of 32 columns: the sampler has ncols = rpad, the posterior sweep packs the columns of several
// samples side by side).  The PP Chebyshev coefficients of (cell, column) sit in registers and are reloaded (16-byte
// loads from L2) only when the cell changes; per row there is one basis recurrence shared by all columns, PP FFMA per
// column and one 128-byte store per column block: T2[k][row][ncols] is row-major per sample, so the scattered rows
// of a cell are still written as whole lines (the contraction kernels transpose while staging their tile).
//   Cf: [k][cell][ctot][PP], this launch uses the columns [col0, col0 + ncols);  records / T2 rows start at row0.
constexpr int ITHREADS = 256;

template <int P, int NL>
__global__ void __launch_bounds__(ITHREADS) interp_kernel(const float *__restrict__ Cf, const int32_t *__restrict__ srow,
                                                          const int32_t *__restrict__ sc, const float *__restrict__ su,
                                                          float *__restrict__ T2, int ctot, int col0, int ncols, int G,
                                                          int64_t row0, int64_t nrows, int64_t rec_stride,
                                                          int64_t t2_stride, int IWROWS) {
    // blockIdx.z: block of NL*32 columns of the ncols this launch produces (the T2 rows are ncols wide)
    // IWROWS records per warp, chosen by the launcher: few enough that the grid fills the chip with ~32 warps per SM
    // (a chunk of the posterior sweep has only 65536 (row, dimension) items), many enough that a cell's coefficients
    // are reused
    constexpr int PP = (P + 3) & ~3;
    const int k = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t kb = (size_t)k * rec_stride + row0;
    const float4 *Ck = reinterpret_cast<const float4 *>(Cf + (size_t)k * G * ctot * PP);
    const int zoff = blockIdx.z * NL * 32;
    float *Tk = T2 + (size_t)k * t2_stride * ncols + zoff;
    // stores: this lane's column of row `row` is at Tlane + row * rowbytes (+ 128 bytes per further column block) — one
    // 32 x 32 -> 64-bit multiply-add per store instead of a 64-bit index computation
    char *const Tlane = reinterpret_cast<char *>(Tk + lane);
    const uint32_t rowbytes = (uint32_t)ncols * 4u;
    const int zcols = min(NL * 32, ncols - zoff);       // columns of this block
    col0 += zoff;
    const int64_t s0 = ((int64_t)blockIdx.x * (ITHREADS / 32) + warp) * IWROWS, s1 = min(nrows, s0 + IWROWS);
    float4 cf[NL][PP / 4];
    int cur = -1;
    for (int64_t sb = s0; sb < s1; sb += 32) {
        const int64_t s = sb + lane;
        int row_l = 0, c_l = -1;
        float u_l = 0.f;
        if (s < s1) {
            row_l = srow[kb + s] - (int32_t)row0;   // records hold batch-global rows; T2 rows are relative to row0
            c_l = sc[kb + s];
            u_l = su[kb + s];
        }
        const int cnt = (int)min((int64_t)32, s1 - sb);
        // two records per iteration: when both lie in the same cell (the records are cell-sorted, so nearly always)
        // one packed recurrence and PP packed FMAs per column block serve both rows
        auto load_cell = [&](int c) {
            cur = c;
#pragma unroll
            for (int a = 0; a < NL; ++a) {
                const int col = col0 + min(a * 32 + lane, zcols - 1);
#pragma unroll
                for (int q = 0; q < PP / 4; ++q) cf[a][q] = __ldg(Ck + ((size_t)c * ctot + col) * (PP / 4) + q);
            }
        };
        auto single = [&](int row, float u) {
            float tb[PP];
            cheb_basis<PP>(u, tb);                                // (entries P..PP-1 meet zero coefficients)
#pragma unroll
            for (int a = 0; a < NL; ++a) {
                float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
                for (int q = 0; q < PP / 4; ++q) {
                    acc0 = fmaf(tb[4 * q], cf[a][q].x, acc0);
                    acc1 = fmaf(tb[4 * q + 1], cf[a][q].y, acc1);
                    acc0 = fmaf(tb[4 * q + 2], cf[a][q].z, acc0);
                    acc1 = fmaf(tb[4 * q + 3], cf[a][q].w, acc1);
                }
                const int l = a * 32 + lane;
                if (l < zcols) *reinterpret_cast<float *>(Tlane + (uint64_t)(uint32_t)row * rowbytes + a * 128) = acc0 + acc1;
            }
        };
        for (int j = 0; j < cnt; j += 2) {
            const int jb = min(j + 1, cnt - 1);
            const int c_a = __shfl_sync(0xffffffffu, c_l, j), c_b = __shfl_sync(0xffffffffu, c_l, jb);
            const int row_a = __shfl_sync(0xffffffffu, row_l, j), row_b = __shfl_sync(0xffffffffu, row_l, jb);
            const float u_a = __shfl_sync(0xffffffffu, u_l, j), u_b = __shfl_sync(0xffffffffu, u_l, jb);
            if (c_a != cur) load_cell(c_a);   // warp-uniform
            if (c_b != c_a) {
                single(row_a, u_a);
                load_cell(c_b);
                single(row_b, u_b);
                continue;
            }
            float2 tb[PP];
            cheb_basis2<PP>(make_float2(u_a, u_b), tb);
#pragma unroll
            for (int a = 0; a < NL; ++a) {
                float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
#pragma unroll
                for (int q = 0; q < PP / 4; ++q) {
                    acc0 = __ffma2_rn(tb[4 * q], splat(cf[a][q].x), acc0);
                    acc1 = __ffma2_rn(tb[4 * q + 1], splat(cf[a][q].y), acc1);
                    acc0 = __ffma2_rn(tb[4 * q + 2], splat(cf[a][q].z), acc0);
                    acc1 = __ffma2_rn(tb[4 * q + 3], splat(cf[a][q].w), acc1);
                }
                const int l = a * 32 + lane;
                if (l < zcols) {
                    *reinterpret_cast<float *>(Tlane + (uint64_t)(uint32_t)row_a * rowbytes + a * 128) = acc0.x + acc1.x;
                    if (jb != j) *reinterpret_cast<float *>(Tlane + (uint64_t)(uint32_t)row_b * rowbytes + a * 128) = acc0.y + acc1.y;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ spreading
// grid (ceil(bs / SROWS), D), 128 threads; lane = column l.  A warp walks 128 consecutive records in blocks of 32; the
// E rows of a block (one 4*rpad-byte line per row, scattered over the batch) are brought into shared memory with
// 16-byte cp.async, the next block in flight while the current one is used.  The P moment sums of the current cell
// stay in registers and go to the global table (FP32 RED, one line per p) only when the cell changes.
constexpr int SROWS = 512;
constexpr int STHREADS = 128;

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src)
                 : "memory");
}

template <int P, int NL>
__global__ void __launch_bounds__(STHREADS) spread_kernel(const float *__restrict__ Erow, const int32_t *__restrict__ srow,
                                                          const int32_t *__restrict__ sc, const float *__restrict__ su,
                                                          float *__restrict__ Mn, int rpad, int mnpad, int64_t bs,
                                                          int64_t mpad) {
    extern __shared__ __align__(16) float sm[];
    const int k = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *stage = sm + (size_t)warp * 2 * 32 * rpad;     // [2][32][rpad] per warp
    const size_t kb = (size_t)k * mpad;
    const float *Ek = Erow + kb * rpad;
    float *Mk = Mn + (size_t)k * mnpad * rpad;
    constexpr int per_warp = SROWS / (STHREADS / 32);
    const int64_t s0 = (int64_t)blockIdx.x * SROWS + warp * per_warp, s1 = min(bs, s0 + per_warp);
    float2 acc[NL][P];     // two partial sums per moment (records are taken in pairs), added at the flush
#pragma unroll
    for (int a = 0; a < NL; ++a)
#pragma unroll
        for (int p = 0; p < P; ++p) acc[a][p] = make_float2(0.f, 0.f);
    int cur = -1;
    auto flush = [&]() {
        if (cur < 0) return;
#pragma unroll
        for (int a = 0; a < NL; ++a) {
            const int l = a * 32 + lane;
            if (l < rpad) {
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    atomicAdd(Mk + ((size_t)cur * P + p) * rpad + l, acc[a][p].x + acc[a][p].y);
                    acc[a][p] = make_float2(0.f, 0.f);
                }
            }
        }
    };
    int row_n = 0, c_n = -1;
    float u_n = 0.f;
    // records of block `sb` into (row_n, c_n, u_n), its E rows into stage buffer `buf` (asynchronously)
    auto fetch = [&](int64_t sb, int buf) {
        const int64_t s = sb + lane;
        row_n = 0; c_n = -1; u_n = 0.f;
        if (s < s1) {
            row_n = srow[kb + s];
            c_n = sc[kb + s];
            u_n = su[kb + s];
        }
        const int ppr = rpad / 4;                               // 16-byte pieces per row
        const int psh = (ppr & (ppr - 1)) == 0 ? 31 - __clz(ppr) : -1;   // a power of two (rpad = 16, 32, 64): shifts, no division
        float *dst = stage + (size_t)buf * 32 * rpad;
        const char *Eb = reinterpret_cast<const char *>(Ek);
        const uint32_t rowbytes = (uint32_t)rpad * 4u;
        for (int t = lane; t < 32 * ppr; t += 32) {             // (32 * ppr is a multiple of 32: the shuffle is convergent)
            const int j = psh >= 0 ? (t >> psh) : t / ppr, part = t - j * ppr;
            const int row = __shfl_sync(0xffffffffu, row_n, j);   // (row 0 for records past the end: a valid address)
            cp_async16(dst + j * rpad + part * 4, Eb + (uint64_t)(uint32_t)row * rowbytes + part * 16);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (s0 < s1) fetch(s0, 0);
    int buf = 0;
    for (int64_t sb = s0; sb < s1; sb += 32, buf ^= 1) {
        const int c_l = c_n;
        const float u_l = u_n;
        if (sb + 32 < s1) {
            fetch(sb + 32, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();
        const float *src = stage + (size_t)buf * 32 * rpad;
        const int cnt = (int)min((int64_t)32, s1 - sb);
        auto single = [&](int j, float u) {
            float tb[P];
            cheb_basis<P>(u, tb);
#pragma unroll
            for (int a = 0; a < NL; ++a) {
                const int l = a * 32 + lane;
                const float e = l < rpad ? src[j * rpad + l] : 0.f;
#pragma unroll
                for (int p = 0; p < P; ++p) acc[a][p].x = fmaf(tb[p], e, acc[a][p].x);
            }
        };
        for (int j = 0; j < cnt; j += 2) {   // two records per iteration (packed FP32) when they share a cell
            const int jb = min(j + 1, cnt - 1);
            const int c_a = __shfl_sync(0xffffffffu, c_l, j), c_b = __shfl_sync(0xffffffffu, c_l, jb);
            const float u_a = __shfl_sync(0xffffffffu, u_l, j), u_b = __shfl_sync(0xffffffffu, u_l, jb);
            if (c_a != cur) {   // warp-uniform
                flush();
                cur = c_a;
            }
            if (c_b != c_a) {
                single(j, u_a);
                flush();
                cur = c_b;
                single(jb, u_b);
                continue;
            }
            float2 tb[P];
            cheb_basis2<P>(make_float2(u_a, u_b), tb);
#pragma unroll
            for (int a = 0; a < NL; ++a) {
                const int l = a * 32 + lane;
                float2 e = make_float2(0.f, 0.f);
                if (l < rpad) {
                    e.x = src[j * rpad + l];
                    if (jb != j) e.y = src[jb * rpad + l];
                }
#pragma unroll
                for (int p = 0; p < P; ++p) acc[a][p] = __ffma2_rn(tb[p], e, acc[a][p]);
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
    cudaFree(g.cell); cudaFree(g.uu); cudaFree(g.srow); cudaFree(g.sc); cudaFree(g.su); cudaFree(g.Erow); cudaFree(g.T2);
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
    // bounds of convenience: the sort keeps 9 x G counters in shared memory, and a grid with more nodes than that is
    // no shortcut any more
    if ((size_t)(SORT_T / 32 + 1) * G * 4 > 64 * 1024 || (size_t)G * PP * h->rpad * 4 > (size_t)8 << 20) return GPTB_OK;
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
    GPTB_TRY(nd_alloc(h, &g.srow, (size_t)c.D * h->mpad));
    GPTB_TRY(nd_alloc(h, &g.sc, (size_t)c.D * h->mpad));
    GPTB_TRY(nd_alloc(h, &g.su, (size_t)c.D * h->mpad));
    GPTB_TRY(nd_alloc(h, &g.Erow, (size_t)c.D * h->mpad * h->rpad));
    GPTB_TRY(nd_alloc(h, &g.T2, (size_t)c.D * h->mpad * h->rpad));
    GPTB_CUDA_OK(h, cudaMemset(g.T2, 0, sizeof(float) * (size_t)c.D * h->mpad * h->rpad));   // dead rows stay finite
    g.batch_mpad = h->mpad;
    return GPTB_OK;
}

#define GPTB_P_SWITCH(P, NL, CALL) \
    switch ((P) * 8 + (NL)) { \
        case 4 * 8 + 1: { CALL(4, 1); } break; \
        case 4 * 8 + 2: { CALL(4, 2); } break; \
        case 6 * 8 + 1: { CALL(6, 1); } break; \
        case 6 * 8 + 2: { CALL(6, 2); } break; \
        case 8 * 8 + 1: { CALL(8, 1); } break; \
        case 8 * 8 + 2: { CALL(8, 2); } break; \
        case 10 * 8 + 1: { CALL(10, 1); } break; \
        case 10 * 8 + 2: { CALL(10, 2); } break; \
        case 12 * 8 + 1: { CALL(12, 1); } break; \
        case 12 * 8 + 2: { CALL(12, 2); } break; \
        case 14 * 8 + 1: { CALL(14, 1); } break; \
        case 14 * 8 + 2: { CALL(14, 2); } break; \
        case 16 * 8 + 1: { CALL(16, 1); } break; \
        case 16 * 8 + 2: { CALL(16, 2); } break; \
        case 12 * 8 + 3: { CALL(12, 3); } break; \
        case 12 * 8 + 4: { CALL(12, 4); } break; \
        case 12 * 8 + 5: { CALL(12, 5); } break; \
        default: \
            gptb_set_error(h, "node grid: unsupported order / column count (P=%d, %d column blocks)", (P), (NL)); \
            return GPTB_ERR_UNSUPPORTED; \
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

int gptb_nodes_launch_coef_cols(gptb_handle *h, const float *Tn, float *Cf, int G, int P, int ncols, int tn_cols, int ctot,
                                int col0, int mnpad, cudaStream_t s) {
    node_coef_cols_kernel<<<dim3((G * ncols + 255) / 256, h->cfg.D), 256, 0, s>>>(Tn, Cf, G, P, (P + 3) & ~3, ncols, tn_cols, ctot,
                                                                                  col0, mnpad);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_nodes_launch_coef_lanes(gptb_handle *h, const float *Tn, float *Cf, int G, int P, int ns, int tn_cols, int SL, int s0,
                                 int mnpad, cudaStream_t s) {
    node_coef_lanes_kernel<<<dim3((G * ns * h->cfg.r + 255) / 256, h->cfg.D), 256, 0, s>>>(Tn, Cf, G, P, (P + 3) & ~3, h->cfg.r, ns,
                                                                                           tn_cols, SL, s0, mnpad);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// The cosine-transform matrix of order P (shared by every grid of the library: the sampler's and the sweep's).
int gptb_nodes_upload_cheb(gptb_handle *h, int P) {
    std::vector<float> cm((size_t)PMAX * PMAX, 0.f);
    for (int p = 0; p < P; ++p)
        for (int j = 0; j < P; ++j) cm[(size_t)p * PMAX + j] = (float)((p == 0 ? 1.0 : 2.0) / P * std::cos(M_PI * p * (j + 0.5) / P));
    GPTB_CUDA_OK(h, cudaMemcpyToSymbol(c_cheb, cm.data(), sizeof(float) * PMAX * PMAX));
    return GPTB_OK;
}

// Launches the interpolation over the records [row0, row0 + nrows) of every dimension, columns [col0, col0 + ncols)
// of a coefficient table with ctot columns, into T2[k][row - row0][ncols] (t2_stride rows per dimension).
int gptb_nodes_launch_interp(gptb_handle *h, int P, const float *Cf, const int32_t *srow, const int32_t *sc, const float *su,
                             float *T2, int ctot, int col0, int ncols, int G, int64_t row0, int64_t nrows, int64_t rec_stride,
                             int64_t t2_stride, cudaStream_t s) {
    // column blocks per warp: up to 5 (160 columns) at order 12, 2 otherwise (what interp_kernel is instantiated for);
    // wider tables take several blocks along grid.z
    const int nl_all = (ncols + 31) / 32, nl_max = P == 12 ? 5 : 2;
    const int NL = std::min(nl_all, nl_max), nz = (nl_all + NL - 1) / NL;
    int IWROWS = 64;
    while (IWROWS > 8 && nrows * h->cfg.D * nz / IWROWS < (int64_t)h->sm_count * 32) IWROWS >>= 1;
    const dim3 grid((unsigned)((nrows + (ITHREADS / 32) * IWROWS - 1) / ((ITHREADS / 32) * IWROWS)), h->cfg.D, nz);
#define GPTB_CALL(PP, NN)                                                                                                   \
    interp_kernel<PP, NN><<<grid, ITHREADS, 0, s>>>(Cf, srow, sc, su, T2, ctot, col0, ncols, G, row0, nrows, rec_stride, t2_stride, \
                                                    IWROWS);
    GPTB_P_SWITCH(P, NL, GPTB_CALL)
#undef GPTB_CALL
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_nodes_interp(gptb_handle *h, int64_t bs) {
    gptb_nodes &g = h->nodes;
    return gptb_nodes_launch_interp(h, g.P, g.Cf, g.srow, g.sc, g.su, g.T2, h->rpad, 0, h->rpad, g.G, 0, bs, h->mpad, h->mpad,
                                    h->stream);
}

// Cell-sorted records of `bs` rows (they depend on the rows only, not on the parameters).
int gptb_nodes_launch_sort(gptb_handle *h, const int32_t *cell, const float *uu, int32_t *srow, int32_t *sc, float *su, int G,
                           int64_t bs, int64_t stride, cudaStream_t s) {
    const size_t smem = (size_t)(SORT_T / 32 + 1) * G * sizeof(int);
    // (per launch: function attributes are per device, and a process may drive several)
    cudaFuncSetAttribute(cell_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cell_sort_kernel<<<dim3((unsigned)((bs + SCHUNK - 1) / SCHUNK), h->cfg.D), SORT_T, smem, s>>>(cell, uu, srow, sc, su, G, bs,
                                                                                                  stride);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_nodes_sort(gptb_handle *h, int64_t bs, cudaStream_t s) {
    gptb_nodes &g = h->nodes;
    return gptb_nodes_launch_sort(h, g.cell, g.uu, g.srow, g.sc, g.su, g.G, bs, h->mpad, s);
}

// Backward: spread the FP32 E rows K2 wrote (g.Erow) onto the nodes, build the node image, run K3 on the nodes.
int gptb_nodes_spread(gptb_handle *h, int64_t bs) {
    gptb_nodes &g = h->nodes;
    const gptb_config &c = h->cfg;
    const int NL = (h->rpad + 31) / 32;
    const dim3 grid((unsigned)((bs + SROWS - 1) / SROWS), c.D);
    const size_t smem = (size_t)(STHREADS / 32) * 2 * 32 * h->rpad * sizeof(float);
#define GPTB_CALL(PP, NN)                                                                                       \
    auto kfn = spread_kernel<PP, NN>;                                                                           \
    GPTB_CUDA_OK(h, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));         \
    kfn<<<grid, STHREADS, smem, h->stream>>>(g.Erow, g.srow, g.sc, g.su, g.Mn, h->rpad, g.mnpad, bs, h->mpad);
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
