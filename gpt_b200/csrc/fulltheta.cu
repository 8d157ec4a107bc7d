// Full-theta ("no tensor") random-Fourier-feature GP regression sampled by SGLD — the comparator of the tensor model
// (BASELINE config 3 "tensor vs full-theta RFF SGLD").
//
// Replaces (all /root/reference/GPT_SGLD.jl):
//   featureNotensor :109-120   phi[j,i] = sqrt(2/n)*sigma_RBF*cos(sum_d X[i,d]*Z[j,d]/length_scale[d] + b[j])
//   GPNT_SGLD       :809-847   theta += eps_t/2*(-theta/sigma_theta^2 + (N/bs)*phi_b*(y_b - phi_b'*theta)/signal_var)
//                                       + sqrt(eps_t)*randn(n),   eps_t = eps_theta*t^(-decay_rate)
//
// The n x N feature matrix (n up to 16000 in Plots/tensorsynth32000_fulltheta_SGLD.png) is never formed: both passes
// of a step generate the features of their tile on the fly.  The argument of the cosine is a D-term dot product
// (D <= 16), so there is no GEMM here worth a tensor core; the step is bound by the cosine — MUFU.COS in the FP32
// mode, the FP64 pipe in the validation mode — and the two passes are oriented so that their reductions stay in a
// register:
//   forward   thread = minibatch row i,  loops over the tile's features:  f_i  += phi[j,i]*theta_j   (Zt, b, theta in smem)
//   backward  thread = feature j,        loops over the tile's rows:      g_j  += phi[j,i]*e_i       (x, e in smem)
// each followed by one atomicAdd per thread and tile.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <random>
#include <string>
#include <vector>

#include "common.cuh"

struct gptb_nt_handle {
    int D = 0, n = 0, precision = GPTB_PREC_FP64, device = 0;
    uint64_t seed = 0;
    cudaStream_t stream = nullptr;
    mutable std::string err;
    bool have_features = false;
    double amp = 0.0;
    double *Zt = nullptr, *b = nullptr;      // [D][n], [n]
    int64_t N = 0;
    double *X = nullptr, *y = nullptr;       // [N][D] row-major, [N]
    int32_t *order = nullptr;                // [N] cumulative permutation
    double *theta = nullptr, *grad = nullptr, *fhat = nullptr, *noise = nullptr, *store = nullptr;
    int64_t fhat_cap = 0;
    int32_t *nan_flag = nullptr;
    double last_ms = 0.0;
    int64_t last_launches = 0, last_steps = 0;
};

namespace {

constexpr int NT_T = 128;      // threads per CTA = rows (forward) / features (backward) per tile
constexpr int NT_DMAX = 16;
constexpr uint64_t NT_STREAM_NOISE = 9ull << 56, NT_STREAM_INIT = 10ull << 56;

void nt_error(const gptb_nt_handle *h, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    gptb_set_error(nullptr, "%s", buf);
}

#define NT_CUDA_OK(h, expr)                                                                     \
    do {                                                                                        \
        cudaError_t _e = (expr);                                                                \
        if (_e != cudaSuccess) {                                                                \
            nt_error((h), "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return GPTB_ERR_CUDA;                                                               \
        }                                                                                       \
    } while (0)

template <typename T>
__device__ __forceinline__ T nt_cos(T a);
template <>
__device__ __forceinline__ double nt_cos<double>(double a) { return cos(a); }
template <>
__device__ __forceinline__ float nt_cos<float>(float a) { return __cosf(a); }

// Both passes are written for a compile-time D (the dot product unrolls without predicates) and two rows (forward) or
// two features (backward) per thread, with the tile's other operand in shared memory as [entry][DP] so that one
// entry costs ceil(D/4) 16-byte broadcast loads shared by both of the thread's rows / features.  (The first version
// — runtime D, one row per thread, [d][entry] tiles — issued 18 LDS per cosine and ran at 0.5 T cos/s: 32 us per pass
// at n = 16000, minibatch 1000.)
// rows (forward) / features (backward) per thread: two in FP32; one in the FP64 validation mode, whose software
// cosine wants the parallelism of more CTAs instead
template <typename T>
__host__ __device__ constexpr int nt_r() { return sizeof(T) == 8 ? 1 : 2; }

// forward: fhat[i] += amp * sum_{j in tile} cos(x_i . zt_j + b_j) * theta_j     grid (row tiles of NT_R*NT_T, feature tiles)
template <typename T, int D>
__global__ void __launch_bounds__(NT_T) nt_fwd_kernel(const double *__restrict__ X, const int32_t *__restrict__ rows,
                                                      int64_t bs, int n, const double *__restrict__ Zt,
                                                      const double *__restrict__ b, const double *__restrict__ theta,
                                                      double amp, double *__restrict__ fhat) {
    constexpr int DP = (D + 3) & ~3;
    constexpr int NT_R = nt_r<T>();
    __shared__ __align__(16) T zs[NT_T][DP];
    __shared__ T bs_s[NT_T], th_s[NT_T];
    const int tid = threadIdx.x;
    const int j0 = blockIdx.y * NT_T, nj = min(NT_T, n - j0);
    if (tid < nj) {
#pragma unroll
        for (int d = 0; d < DP; ++d) zs[tid][d] = d < D ? (T)Zt[(size_t)d * n + j0 + tid] : (T)0;
        bs_s[tid] = (T)b[j0 + tid];
        th_s[tid] = (T)theta[j0 + tid];
    }
    __syncthreads();
    const int64_t ibase = (int64_t)blockIdx.x * (NT_R * NT_T) + tid;     // rows ibase, ibase + NT_T
    T x[NT_R][D];
    bool live[NT_R];
#pragma unroll
    for (int q = 0; q < NT_R; ++q) {
        const int64_t i = ibase + (int64_t)q * NT_T;
        live[q] = i < bs;
        const int64_t row = live[q] ? (rows ? (int64_t)rows[i] : i) : 0;
#pragma unroll
        for (int d = 0; d < D; ++d) x[q][d] = live[q] ? (T)X[(size_t)row * D + d] : (T)0;
    }
    T f[NT_R];
#pragma unroll
    for (int q = 0; q < NT_R; ++q) f[q] = 0;
#pragma unroll 2
    for (int j = 0; j < nj; ++j) {
        T z[DP];
#pragma unroll
        for (int d4 = 0; d4 < DP; d4 += 4) {   // 16-byte (float) / 2 x 16-byte (double) broadcast loads
            z[d4] = zs[j][d4]; z[d4 + 1] = zs[j][d4 + 1]; z[d4 + 2] = zs[j][d4 + 2]; z[d4 + 3] = zs[j][d4 + 3];
        }
        const T bj = bs_s[j], tj = th_s[j];
#pragma unroll
        for (int q = 0; q < NT_R; ++q) {
            T a = bj;
#pragma unroll
            for (int d = 0; d < D; ++d) a = fma(x[q][d], z[d], a);
            f[q] = fma(nt_cos<T>(a), tj, f[q]);
        }
    }
#pragma unroll
    for (int q = 0; q < NT_R; ++q)
        if (live[q]) atomicAdd(&fhat[ibase + (int64_t)q * NT_T], amp * (double)f[q]);
}

// backward: grad[j] += amp * sum_{i in tile} cos(x_i . zt_j + b_j) * e_i,  e_i = escale*(y_i - fhat_i)
// grid (feature tiles of NT_R*NT_T, row tiles)
template <typename T, int D>
__global__ void __launch_bounds__(NT_T) nt_bwd_kernel(const double *__restrict__ X, const int32_t *__restrict__ rows,
                                                      const double *__restrict__ y, int64_t bs, int n,
                                                      const double *__restrict__ Zt, const double *__restrict__ b,
                                                      const double *__restrict__ fhat, double escale, double amp,
                                                      double *__restrict__ grad) {
    constexpr int DP = (D + 3) & ~3;
    constexpr int NT_R = nt_r<T>();
    __shared__ __align__(16) T xs[NT_T][DP];
    __shared__ T es[NT_T];
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.y * NT_T;
    const int ni = (int)min((int64_t)NT_T, bs - i0);
    if (tid < ni) {
        const int64_t row = rows ? rows[i0 + tid] : i0 + tid;
#pragma unroll
        for (int d = 0; d < DP; ++d) xs[tid][d] = d < D ? (T)X[(size_t)row * D + d] : (T)0;
        es[tid] = (T)(escale * (y[row] - fhat[i0 + tid]));
    }
    __syncthreads();
    const int jbase = blockIdx.x * (NT_R * NT_T) + tid;      // features jbase, jbase + NT_T
    T z[NT_R][D], bj[NT_R], g[NT_R];
    bool live[NT_R];
#pragma unroll
    for (int q = 0; q < NT_R; ++q) {
        const int j = jbase + q * NT_T;
        live[q] = j < n;
#pragma unroll
        for (int d = 0; d < D; ++d) z[q][d] = live[q] ? (T)Zt[(size_t)d * n + j] : (T)0;
        bj[q] = live[q] ? (T)b[j] : (T)0;
        g[q] = 0;
    }
#pragma unroll 2
    for (int i = 0; i < ni; ++i) {
        T x[DP];
#pragma unroll
        for (int d4 = 0; d4 < DP; d4 += 4) {
            x[d4] = xs[i][d4]; x[d4 + 1] = xs[i][d4 + 1]; x[d4 + 2] = xs[i][d4 + 2]; x[d4 + 3] = xs[i][d4 + 3];
        }
        const T ei = es[i];
#pragma unroll
        for (int q = 0; q < NT_R; ++q) {
            T a = bj[q];
#pragma unroll
            for (int d = 0; d < D; ++d) a = fma(z[q][d], x[d], a);
            g[q] = fma(nt_cos<T>(a), ei, g[q]);
        }
    }
#pragma unroll
    for (int q = 0; q < NT_R; ++q)
        if (live[q]) atomicAdd(&grad[jbase + q * NT_T], amp * (double)g[q]);
}

// dispatch on the runtime D (1..NT_DMAX) and the arithmetic type
#define NT_D_CASES(M) M(1) M(2) M(3) M(4) M(5) M(6) M(7) M(8) M(9) M(10) M(11) M(12) M(13) M(14) M(15) M(16)

// theta += eps/2*(-theta/sigma_theta^2 + grad) + sqrt(eps)*z  (:833-837); stores the sample, raises the NaN flag
// (:839), and clears grad / fhat for the next step.
__global__ void nt_update_kernel(double *__restrict__ theta, double *__restrict__ grad, double *__restrict__ fhat,
                                 int64_t fhat_n, int n, double eps, double sigma_theta, const double *__restrict__ noise,
                                 uint64_t key, uint64_t step, double *__restrict__ store_col, int32_t *__restrict__ nan_flag) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t < fhat_n) fhat[t] = 0.0;
    if (t >= n) return;
    double z;
    if (noise) {
        z = noise[t];
    } else {
        double a, bb;
        philox_normal2(key, NT_STREAM_NOISE | (uint64_t)(t >> 1), step, &a, &bb);
        z = (t & 1) ? bb : a;
    }
    const double th = theta[t];
    const double g = -th / (sigma_theta * sigma_theta) + grad[t];
    const double v = th + eps * g / 2 + sqrt(eps) * z;
    theta[t] = v;
    grad[t] = 0.0;
    if (store_col) store_col[t] = v;
    if (v != v) atomicExch(nan_flag, 1);
}

__global__ void nt_features_kernel(const double *__restrict__ X /*[N][D] row-major*/, int64_t N, int D, int n,
                                   const double *__restrict__ Zt, const double *__restrict__ b, double amp,
                                   double *__restrict__ phi /*[n,N] column-major*/) {
    const int64_t i = blockIdx.y;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        double a = b[j];
        for (int d = 0; d < D; ++d) a = fma(X[(size_t)i * D + d], Zt[(size_t)d * n + j], a);
        phi[(size_t)i * n + j] = amp * cos(a);
    }
}

__global__ void nt_transpose_kernel(const double *__restrict__ Xcm, int64_t N, int D, double *__restrict__ Xrm) {
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < N * D; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / D;
        const int d = (int)(t - i * D);
        Xrm[t] = Xcm[(size_t)d * N + i];
    }
}

template <typename T>
int nt_free(T **p) {
    if (*p) cudaFree(*p);
    *p = nullptr;
    return 0;
}

template <typename T>
int nt_alloc(gptb_nt_handle *h, T **p, size_t count) {
    nt_free(p);
    cudaError_t e = cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) {
        nt_error(h, "cudaMalloc(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
        return GPTB_ERR_NOMEM;
    }
    return GPTB_OK;
}

// column-major host X[N,D] -> row-major device buffer
int nt_upload_rows(gptb_nt_handle *h, const double *X, int64_t N, double **Xrm_out) {
    double *tmp = nullptr;
    int st = nt_alloc(h, &tmp, (size_t)N * h->D);
    if (st != GPTB_OK) return st;
    st = nt_alloc(h, Xrm_out, (size_t)N * h->D);
    if (st != GPTB_OK) {
        cudaFree(tmp);
        return st;
    }
    cudaError_t e = cudaMemcpyAsync(tmp, X, sizeof(double) * N * h->D, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) {
        nt_transpose_kernel<<<(int)std::min<int64_t>((N * h->D + 255) / 256, 4096), 256, 0, h->stream>>>(tmp, N, h->D, *Xrm_out);
        e = cudaStreamSynchronize(h->stream);
    }
    cudaFree(tmp);
    if (e != cudaSuccess) {
        nt_error(h, "upload of X: %s", cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    return GPTB_OK;
}

int nt_ensure_fhat(gptb_nt_handle *h, int64_t bs) {
    if (bs <= h->fhat_cap) return GPTB_OK;
    int st = nt_alloc(h, &h->fhat, (size_t)bs);
    if (st != GPTB_OK) return st;
    h->fhat_cap = bs;
    NT_CUDA_OK(h, cudaMemsetAsync(h->fhat, 0, sizeof(double) * bs, h->stream));
    return GPTB_OK;
}

// fhat += Phi(rows)' theta  (fhat must be zero on entry)
int nt_launch_fwd(gptb_nt_handle *h, const double *X, const int32_t *rows, int64_t bs, const double *theta, double *fhat) {
    const int NT_R = h->precision == GPTB_PREC_FP64 ? nt_r<double>() : nt_r<float>();
    dim3 grid((unsigned)((bs + NT_R * NT_T - 1) / (NT_R * NT_T)), (unsigned)((h->n + NT_T - 1) / NT_T));
#define NT_CASE(DD)                                                                                                  \
    case DD:                                                                                                         \
        if (h->precision == GPTB_PREC_FP64)                                                                          \
            nt_fwd_kernel<double, DD><<<grid, NT_T, 0, h->stream>>>(X, rows, bs, h->n, h->Zt, h->b, theta, h->amp, fhat); \
        else                                                                                                         \
            nt_fwd_kernel<float, DD><<<grid, NT_T, 0, h->stream>>>(X, rows, bs, h->n, h->Zt, h->b, theta, h->amp, fhat);  \
        break;
    switch (h->D) { NT_D_CASES(NT_CASE) default: return GPTB_ERR_UNSUPPORTED; }
#undef NT_CASE
    NT_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// grad += Phi(rows) e,  e = escale (y - fhat)
int nt_launch_bwd(gptb_nt_handle *h, const int32_t *rows, int64_t bs, double escale) {
    const int NT_R = h->precision == GPTB_PREC_FP64 ? nt_r<double>() : nt_r<float>();
    dim3 grid((unsigned)((h->n + NT_R * NT_T - 1) / (NT_R * NT_T)), (unsigned)((bs + NT_T - 1) / NT_T));
#define NT_CASE(DD)                                                                                                   \
    case DD:                                                                                                          \
        if (h->precision == GPTB_PREC_FP64)                                                                           \
            nt_bwd_kernel<double, DD><<<grid, NT_T, 0, h->stream>>>(h->X, rows, h->y, bs, h->n, h->Zt, h->b, h->fhat, escale, \
                                                                   h->amp, h->grad);                                  \
        else                                                                                                          \
            nt_bwd_kernel<float, DD><<<grid, NT_T, 0, h->stream>>>(h->X, rows, h->y, bs, h->n, h->Zt, h->b, h->fhat, escale,  \
                                                                  h->amp, h->grad);                                   \
        break;
    switch (h->D) { NT_D_CASES(NT_CASE) default: return GPTB_ERR_UNSUPPORTED; }
#undef NT_CASE
    NT_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

}  // namespace

extern "C" {

int gptb_nt_create(int32_t D, int32_t n, int32_t precision, int32_t device, uint64_t seed, gptb_nt_handle **out) {
    if (!out || D < 1 || D > NT_DMAX || n < 1 || (precision != GPTB_PREC_FP64 && precision != GPTB_PREC_TC3)) {
        gptb_set_error(nullptr, "gptb_nt_create: need 1 <= D <= %d, n >= 1, precision FP64 or TC3 (= FP32 + MUFU here)", NT_DMAX);
        return GPTB_ERR_ARG;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device >= ndev) {
        gptb_set_error(nullptr, "gptb_nt_create: no CUDA device %d (this library has no CPU path)", device);
        return GPTB_ERR_CUDA;
    }
    if (device < 0) cudaGetDevice(&device);   // negative = the current device, as in gptb_create
    gptb_nt_handle *h = new gptb_nt_handle;
    h->D = D; h->n = n; h->precision = precision; h->device = device; h->seed = seed;
    cudaSetDevice(device);
    if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc((void **)&h->nan_flag, sizeof(int32_t)) != cudaSuccess) {
        gptb_set_error(nullptr, "gptb_nt_create: %s", cudaGetErrorString(cudaGetLastError()));
        delete h;
        return GPTB_ERR_CUDA;
    }
    cudaMemsetAsync(h->nan_flag, 0, sizeof(int32_t), h->stream);
    *out = h;
    return GPTB_OK;
}

int gptb_nt_destroy(gptb_nt_handle *h) {
    if (!h) return GPTB_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    nt_free(&h->Zt); nt_free(&h->b); nt_free(&h->X); nt_free(&h->y); nt_free(&h->order); nt_free(&h->theta);
    nt_free(&h->grad); nt_free(&h->fhat); nt_free(&h->noise); nt_free(&h->store); nt_free(&h->nan_flag);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return GPTB_OK;
}

const char *gptb_nt_last_error(const gptb_nt_handle *h) { return h ? h->err.c_str() : ""; }

int gptb_nt_set_features(gptb_nt_handle *h, const double *Z, const double *b, const double *length_scale, int32_t ls_len,
                         double sigma_RBF) {
    if (!h || !Z || !b || !length_scale) return GPTB_ERR_ARG;
    if (ls_len != 1 && ls_len != h->D) {
        nt_error(h, "dimensions of X and length_scale do not match");
        return GPTB_ERR_SHAPE;
    }
    NT_CUDA_OK(h, cudaSetDevice(h->device));
    const int n = h->n, D = h->D;
    std::vector<double> zt((size_t)n * D);
    for (int d = 0; d < D; ++d) {
        const double ls = length_scale[ls_len == 1 ? 0 : d];
        for (int j = 0; j < n; ++j) zt[(size_t)d * n + j] = Z[(size_t)d * n + j] / ls;   // scale(Z, 1./length_scale), :113
    }
    int st = nt_alloc(h, &h->Zt, zt.size());
    if (st == GPTB_OK) st = nt_alloc(h, &h->b, (size_t)n);
    if (st != GPTB_OK) return st;
    NT_CUDA_OK(h, cudaMemcpyAsync(h->Zt, zt.data(), zt.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    NT_CUDA_OK(h, cudaMemcpyAsync(h->b, b, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    NT_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    h->amp = sqrt(2.0 / n) * sigma_RBF;   // :119
    h->have_features = true;
    return GPTB_OK;
}

int gptb_nt_features(gptb_nt_handle *h, const double *X, int64_t N, double *phi_out) {
    if (!h || !X || !phi_out || N < 0) return GPTB_ERR_ARG;
    if (!h->have_features) {
        nt_error(h, "features not set (gptb_nt_set_features)");
        return GPTB_ERR_STATE;
    }
    if (N == 0) return GPTB_OK;
    NT_CUDA_OK(h, cudaSetDevice(h->device));
    double *xr = nullptr, *ph = nullptr;
    int st = nt_upload_rows(h, X, N, &xr);
    if (st == GPTB_OK) st = nt_alloc(h, &ph, (size_t)N * h->n);
    if (st == GPTB_OK) {
        nt_features_kernel<<<dim3((unsigned)std::min((h->n + 255) / 256, 64), (unsigned)N), 256, 0, h->stream>>>(
            xr, N, h->D, h->n, h->Zt, h->b, h->amp, ph);
        cudaError_t e = cudaMemcpyAsync(phi_out, ph, sizeof(double) * N * h->n, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) {
            nt_error(h, "gptb_nt_features: %s", cudaGetErrorString(e));
            st = GPTB_ERR_CUDA;
        }
    }
    nt_free(&xr);
    nt_free(&ph);
    return st;
}

int gptb_nt_set_train(gptb_nt_handle *h, const double *X, const double *y, int64_t N) {
    if (!h || !X || !y || N < 1 || N > 0x7fffffff) return GPTB_ERR_ARG;
    NT_CUDA_OK(h, cudaSetDevice(h->device));
    int st = nt_upload_rows(h, X, N, &h->X);
    if (st == GPTB_OK) st = nt_alloc(h, &h->y, (size_t)N);
    if (st == GPTB_OK) st = nt_alloc(h, &h->order, (size_t)N);
    if (st != GPTB_OK) return st;
    NT_CUDA_OK(h, cudaMemcpyAsync(h->y, y, sizeof(double) * N, cudaMemcpyHostToDevice, h->stream));
    NT_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    h->N = N;
    return GPTB_OK;
}

int gptb_nt_pred(gptb_nt_handle *h, const double *theta, const double *Xtest, int64_t Nt, double *fhat_out) {
    if (!h || !theta || !Xtest || !fhat_out || Nt < 0) return GPTB_ERR_ARG;
    if (!h->have_features) {
        nt_error(h, "features not set (gptb_nt_set_features)");
        return GPTB_ERR_STATE;
    }
    if (Nt == 0) return GPTB_OK;
    NT_CUDA_OK(h, cudaSetDevice(h->device));
    double *xr = nullptr, *th = nullptr, *fh = nullptr;
    int st = nt_upload_rows(h, Xtest, Nt, &xr);
    if (st == GPTB_OK) st = nt_alloc(h, &th, (size_t)h->n);
    if (st == GPTB_OK) st = nt_alloc(h, &fh, (size_t)Nt);
    if (st == GPTB_OK) {
        cudaMemcpyAsync(th, theta, sizeof(double) * h->n, cudaMemcpyHostToDevice, h->stream);
        cudaMemsetAsync(fh, 0, sizeof(double) * Nt, h->stream);
        st = nt_launch_fwd(h, xr, nullptr, Nt, th, fh);
        if (st == GPTB_OK) {
            cudaError_t e = cudaMemcpyAsync(fhat_out, fh, sizeof(double) * Nt, cudaMemcpyDeviceToHost, h->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
            if (e != cudaSuccess) {
                nt_error(h, "gptb_nt_pred: %s", cudaGetErrorString(e));
                st = GPTB_ERR_CUDA;
            }
        }
    }
    nt_free(&xr); nt_free(&th); nt_free(&fh);
    return st;
}

int gptb_nt_run(gptb_nt_handle *h, double signal_var, double sigma_theta, int64_t m, double eps_theta, double decay_rate,
                int32_t burnin, int32_t maxepoch, uint64_t param_seed, const double *theta0, const int64_t *perms,
                const double *noise, double *theta_store_out) {
    if (!h || !theta_store_out || m < 1 || burnin < 0 || maxepoch < 0 || !(signal_var > 0) || !(sigma_theta > 0))
        return GPTB_ERR_ARG;
    if (!h->have_features || !h->X) {
        nt_error(h, "features / training set must be set before gptb_nt_run");
        return GPTB_ERR_STATE;
    }
    NT_CUDA_OK(h, cudaSetDevice(h->device));
    const int n = h->n;
    const int64_t N = h->N, numbatches = (N + m - 1) / m;              // :811
    const int64_t T = (int64_t)(maxepoch + burnin) * numbatches;       // :818
    if (T == 0) return GPTB_OK;
    int st = nt_alloc(h, &h->theta, (size_t)n);
    if (st == GPTB_OK) st = nt_alloc(h, &h->grad, (size_t)n);
    if (st == GPTB_OK) st = nt_alloc(h, &h->store, (size_t)n * T);
    if (st == GPTB_OK) st = nt_ensure_fhat(h, std::min(m, N));
    if (st == GPTB_OK && noise) st = nt_alloc(h, &h->noise, (size_t)n);
    if (st != GPTB_OK) return st;
    std::vector<double> th0(n);
    if (theta0) {
        memcpy(th0.data(), theta0, sizeof(double) * n);
    } else {                                                           // theta = sigma_theta*randn(n), :816
        for (int j = 0; j < n; j += 2) {
            double a, bb;
            philox_normal2(param_seed, NT_STREAM_INIT | (uint64_t)(j / 2), 0, &a, &bb);
            th0[j] = sigma_theta * a;
            if (j + 1 < n) th0[j + 1] = sigma_theta * bb;
        }
    }
    NT_CUDA_OK(h, cudaMemcpyAsync(h->theta, th0.data(), sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    NT_CUDA_OK(h, cudaMemsetAsync(h->grad, 0, sizeof(double) * n, h->stream));
    NT_CUDA_OK(h, cudaMemsetAsync(h->fhat, 0, sizeof(double) * h->fhat_cap, h->stream));
    NT_CUDA_OK(h, cudaMemsetAsync(h->nan_flag, 0, sizeof(int32_t), h->stream));
    std::vector<int32_t> order(N), next(N);
    for (int64_t i = 0; i < N; ++i) order[i] = (int32_t)i;
    std::mt19937_64 prng(param_seed ^ 0x9E3779B97F4A7C15ull);
    cudaEvent_t ev0, ev1;
    NT_CUDA_OK(h, cudaEventCreate(&ev0));
    NT_CUDA_OK(h, cudaEventCreate(&ev1));
    int64_t launches = 0, t = 0;
    bool first = true;
    for (int epoch = 0; epoch < maxepoch + burnin; ++epoch) {
        // phi=phi[:,perm]; y=y[perm] (:824-825): position i of the new order holds position perm[i] of the old one
        if (perms) {
            const int64_t *pe = perms + (size_t)epoch * N;
            for (int64_t i = 0; i < N; ++i) {
                if (pe[i] < 0 || pe[i] >= N) {
                    nt_error(h, "perms[%d][%lld] outside 0..N-1", epoch, (long long)i);
                    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
                    return GPTB_ERR_ARG;
                }
                next[i] = order[pe[i]];
            }
            order.swap(next);
        } else {
            std::shuffle(order.begin(), order.end(), prng);
        }
        NT_CUDA_OK(h, cudaStreamSynchronize(h->stream));   // the previous epoch's kernels read `order`
        NT_CUDA_OK(h, cudaMemcpyAsync(h->order, order.data(), sizeof(int32_t) * N, cudaMemcpyHostToDevice, h->stream));
        if (first) {
            NT_CUDA_OK(h, cudaEventRecord(ev0, h->stream));
            first = false;
        }
        for (int64_t batch = 0; batch < numbatches; ++batch) {
            ++t;
            const int64_t lo = m * batch, hi = std::min(m * (batch + 1), N), bs = hi - lo;   // :830-831
            const double eps = eps_theta * pow((double)t, -decay_rate);                      // :832
            const double escale = ((double)N / (double)bs) / signal_var;
            if (noise)
                NT_CUDA_OK(h, cudaMemcpyAsync(h->noise, noise + (size_t)(t - 1) * n, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
            st = nt_launch_fwd(h, h->X, h->order + lo, bs, h->theta, h->fhat);
            if (st != GPTB_OK) break;
            st = nt_launch_bwd(h, h->order + lo, bs, escale);
            if (st != GPTB_OK) break;
            const int64_t un = std::max<int64_t>(n, h->fhat_cap);
            nt_update_kernel<<<(unsigned)((un + 255) / 256), 256, 0, h->stream>>>(
                h->theta, h->grad, h->fhat, h->fhat_cap, n, eps, sigma_theta, noise ? h->noise : nullptr, h->seed ^ param_seed,
                (uint64_t)t, h->store + (size_t)(t - 1) * n, h->nan_flag);
            launches += 3;
        }
        if (st != GPTB_OK) break;
    }
    int32_t bad = 0;
    cudaError_t e = cudaEventRecord(ev1, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, h->nan_flag, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(theta_store_out, h->store, sizeof(double) * n * T, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    float ms = 0.f;
    if (e == cudaSuccess) cudaEventElapsedTime(&ms, ev0, ev1);
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    nt_free(&h->store);
    if (e != cudaSuccess) {
        nt_error(h, "gptb_nt_run: %s", cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    if (st != GPTB_OK) return st;
    h->last_ms = ms; h->last_launches = launches; h->last_steps = t;
    if (bad) {   // :839-842
        nt_error(h, "Get NaN in theta. Try smaller epsilon");
        memset(theta_store_out, 0, sizeof(double) * n * T);
        return GPTB_GEODESIC_NAN;
    }
    return GPTB_OK;
}

int gptb_nt_last_run_stats(gptb_nt_handle *h, double *ms_out, int64_t *launches_out, int64_t *steps_out) {
    if (!h) return GPTB_ERR_ARG;
    if (ms_out) *ms_out = h->last_ms;
    if (launches_out) *launches_out = h->last_launches;
    if (steps_out) *steps_out = h->last_steps;
    return GPTB_OK;
}

}  // extern "C"
