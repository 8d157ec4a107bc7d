// FP64 validation path (GPTB_PREC_FP64): features, factor contraction and factor gradient on the
// FP64 CUDA cores.  Same data flow as the tensor-core path, used for <=1e-10 parity with the
// reference's Float64 arithmetic and for the dense-phi compatibility entries.
//
//   features : feature  GPT_SGLD.jl:71-84, feature2 :88-105
//   fwd      : phidotU  GPT_SGLD.jl:193-205   T[k][l][i] = sum_j phi_k(x_i)[j] U[j,l,k]
//   bwd      : computePsi + grad-U  GPT_SGLD.jl:276-286,:405-408
//              gradU[j,l,k] = sum_i phi_k(x_i)[j] * (e_i A[l,k,i])   (Psi is never formed)
#include "common.cuh"

namespace {

constexpr int FB_ROWS = 32;   // rows per block tile
constexpr int FB_J = 32;      // features per chunk
constexpr int FB_THREADS = 256;
constexpr int FB_LG = FB_THREADS / 32;  // 8 column groups
constexpr int FB_MAXT = 8;    // supports r <= 64

// phi_k(x)[j] without the amplitude.  COS: cos(x*Zt[j] + b[j]).  SINCOS (feature2): rows 2p / 2p+1 (0-based)
// are sin / cos of x*Zt[p].
__device__ __forceinline__ double feat_value(int kind, double x, const double *Ztk, const double *bk, int j) {
    if (kind == GPTB_FEAT_COS) return cos(x * Ztk[j] + bk[j]);
    double a = x * Ztk[j >> 1];
    return (j & 1) ? cos(a) : sin(a);
}

__global__ void features_kernel(const double *__restrict__ X /*[N][D]*/, int64_t N, int D, int n, int kind,
                                double amp, const double *__restrict__ Zt, const double *__restrict__ b,
                                double *__restrict__ phi /*[N][D][n]*/) {
    int64_t total = N * (int64_t)D * n;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int j = (int)(t % n);
        int64_t ik = t / n;
        int k = (int)(ik % D);
        int64_t i = ik / D;
        phi[t] = amp * feat_value(kind, X[i * D + k], Zt + (size_t)k * n, b + (size_t)k * n, j);
    }
}

// grid (ceil(bs/32), D).  phib != null: dense features phi[row][k][j] (amplitude already inside).
__global__ void __launch_bounds__(FB_THREADS)
fwd_fp64_kernel(const double *__restrict__ Xb, const double *__restrict__ phib, const int32_t *__restrict__ rows,
                const double *__restrict__ U, const double *__restrict__ Zt, const double *__restrict__ b, int kind,
                double amp, int n, int r, int D, int64_t bs, int64_t mpad, double *__restrict__ T) {
    __shared__ double Ps[FB_J][FB_ROWS + 1];
    __shared__ double Us[FB_J][64];
    const int k = blockIdx.y;
    const int64_t i0 = (int64_t)blockIdx.x * FB_ROWS;
    const int tid = threadIdx.x, row = tid % 32, lg = tid / 32;
    const int64_t i = i0 + row;
    const bool live = i < bs;
    const double x = (phib == nullptr && live) ? Xb[(size_t)k * mpad + i] : 0.0;
    const int64_t prow = (phib != nullptr && live) ? (rows ? (int64_t)rows[i] : i) : 0;
    const double *Ztk = Zt + (size_t)k * n, *bk = b + (size_t)k * n;
    double acc[FB_MAXT];
#pragma unroll
    for (int t = 0; t < FB_MAXT; ++t) acc[t] = 0.0;

    for (int j0 = 0; j0 < n; j0 += FB_J) {
#pragma unroll
        for (int t = 0; t < FB_J / FB_LG; ++t) {
            int jj = lg + FB_LG * t, j = j0 + jj;
            double v = 0.0;
            if (live && j < n) v = phib ? phib[((size_t)prow * D + k) * n + j] : feat_value(kind, x, Ztk, bk, j);
            Ps[jj][row] = v;
        }
        for (int t = tid; t < FB_J * r; t += FB_THREADS) {
            int jj = t % FB_J, l = t / FB_J, j = j0 + jj;
            Us[jj][l] = (j < n) ? U[((size_t)k * r + l) * n + j] : 0.0;
        }
        __syncthreads();
#pragma unroll 4
        for (int jj = 0; jj < FB_J; ++jj) {
            double p = Ps[jj][row];
#pragma unroll
            for (int t = 0; t < FB_MAXT; ++t) {
                int l = lg + FB_LG * t;
                if (l < r) acc[t] = fma(p, Us[jj][l], acc[t]);
            }
        }
        __syncthreads();
    }
    const double s = phib ? 1.0 : amp;
    if (i < mpad) {
#pragma unroll
        for (int t = 0; t < FB_MAXT; ++t) {
            int l = lg + FB_LG * t;
            if (l < r) T[((size_t)k * r + l) * mpad + i] = live ? s * acc[t] : 0.0;
        }
    }
}

// grid (ceil(n/32), D, row splits).  gradU[k][l][j] += amp * sum_i phi[i][j] E[k][l][i]
__global__ void __launch_bounds__(FB_THREADS)
bwd_fp64_kernel(const double *__restrict__ Xb, const double *__restrict__ phib, const int32_t *__restrict__ rows,
                const double *__restrict__ E, const double *__restrict__ Zt, const double *__restrict__ b, int kind,
                double amp, int n, int r, int D, int64_t bs, int64_t mpad, double *__restrict__ gradU) {
    __shared__ double Ps[FB_ROWS][FB_J + 1];
    __shared__ double Es[FB_ROWS][64];
    const int k = blockIdx.y;
    const int j0 = blockIdx.x * FB_J;
    const int tid = threadIdx.x, jj = tid % 32, lg = tid / 32;
    const int j = j0 + jj;
    const double *Ztk = Zt + (size_t)k * n, *bk = b + (size_t)k * n;
    const int64_t chunk = (bs + gridDim.z - 1) / gridDim.z;
    const int64_t ibeg = (int64_t)blockIdx.z * chunk, iend = min((long long)bs, (long long)(ibeg + chunk));
    double acc[FB_MAXT];
#pragma unroll
    for (int t = 0; t < FB_MAXT; ++t) acc[t] = 0.0;

    for (int64_t i0 = ibeg; i0 < iend; i0 += FB_ROWS) {
#pragma unroll
        for (int t = 0; t < FB_ROWS / FB_LG; ++t) {
            int ii = lg + FB_LG * t;
            int64_t i = i0 + ii;
            double v = 0.0;
            if (i < iend && j < n) {
                if (phib) {
                    int64_t prow = rows ? (int64_t)rows[i] : i;
                    v = phib[((size_t)prow * D + k) * n + j];
                } else {
                    v = feat_value(kind, Xb[(size_t)k * mpad + i], Ztk, bk, j);
                }
            }
            Ps[ii][jj] = v;
        }
        for (int t = tid; t < FB_ROWS * r; t += FB_THREADS) {
            int ii = t % FB_ROWS, l = t / FB_ROWS;
            int64_t i = i0 + ii;
            Es[ii][l] = (i < iend) ? E[((size_t)k * r + l) * mpad + i] : 0.0;
        }
        __syncthreads();
#pragma unroll 4
        for (int ii = 0; ii < FB_ROWS; ++ii) {
            double p = Ps[ii][jj];
#pragma unroll
            for (int t = 0; t < FB_MAXT; ++t) {
                int l = lg + FB_LG * t;
                if (l < r) acc[t] = fma(p, Es[ii][l], acc[t]);
            }
        }
        __syncthreads();
    }
    const double s = phib ? 1.0 : amp;
    if (j < n) {
#pragma unroll
        for (int t = 0; t < FB_MAXT; ++t) {
            int l = lg + FB_LG * t;
            if (l < r) atomicAdd(&gradU[((size_t)k * r + l) * n + j], s * acc[t]);
        }
    }
}

}  // namespace

int gptb_launch_features_fp64(gptb_handle *h, const double *Xrm, int64_t N, double *phi) {
    const gptb_config &c = h->cfg;
    int64_t total = N * (int64_t)c.D * c.n;
    if (total == 0) return GPTB_OK;
    int blocks = (int)std::min<int64_t>((total + 255) / 256, (int64_t)h->sm_count * 16);
    features_kernel<<<blocks, 256, 0, h->stream>>>(Xrm, N, c.D, c.n, c.feature_kind, h->amp, h->Zt, h->b, phi);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_fwd_fp64(gptb_handle *h, const double *Xb, const double *phib, const int32_t *rows, const double *U,
                         int64_t bs, int64_t mpad, double *T) {
    const gptb_config &c = h->cfg;
    if (bs <= 0) return GPTB_OK;
    const int64_t mext = (bs + 127) / 128 * 128;
    dim3 grid((unsigned)(mext / FB_ROWS), c.D);
    fwd_fp64_kernel<<<grid, FB_THREADS, 0, h->stream>>>(Xb, phib, rows, U, h->Zt, h->b, c.feature_kind, h->amp, c.n,
                                                        c.r, c.D, bs, mpad, T);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_bwd_fp64(gptb_handle *h, const double *Xb, const double *phib, const int32_t *rows, const double *E,
                         int64_t bs, int64_t mpad, double *gradU) {
    const gptb_config &c = h->cfg;
    if (bs <= 0) return GPTB_OK;
    int jblocks = (c.n + FB_J - 1) / FB_J;
    int64_t want = ((int64_t)h->sm_count * 4 + (int64_t)jblocks * c.D - 1) / ((int64_t)jblocks * c.D);
    int64_t maxsplit = (bs + 4 * FB_ROWS - 1) / (4 * FB_ROWS);
    int splits = (int)std::max<int64_t>(1, std::min<int64_t>(want, maxsplit));
    dim3 grid(jblocks, c.D, splits);
    bwd_fp64_kernel<<<grid, FB_THREADS, 0, h->stream>>>(Xb, phib, rows, E, h->Zt, h->b, c.feature_kind, h->amp, c.n,
                                                        c.r, c.D, bs, mpad, gradU);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}
