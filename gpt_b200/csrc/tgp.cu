// GPT_inf — the Gibbs sampler of TGP.jl:37-86 for the tensor model with Gaussian (not Stiefel) factors.
//
// Every iteration draws the core w from its Gaussian conditional (a q x q system, :62-66) and then, dimension by
// dimension, the whole factor U^(k) from ITS Gaussian conditional (an n*r x n*r system, :73-84).  The work that grows
// with the number of training rows N runs on the device, full batch, in FP64:
//     T[k][l][i]   = U_k[:,l]' b[:,k,i]                                   tgp_T_kernel        (phidotU)
//     V[q][i]      = prod_k T[k][I[q,k]][i]                               tgp_V_kernel        (:62)
//     V V', V y                                                           gram / gemv kernels (:64-65)
//     V_k = V ./ T[k][I[:,k]], C[l][i] = sum_{q: I[q,k]=l} w_q V_k[q][i]  tgp_C_kernel        (:74-79)
//     Ck[l*n+j][i] = C[l][i] b[j,k,i], Ck Ck', Ck y                       tgp_Ck_kernel, gram / gemv  (:80-82)
//     V = V_k .* T_new[k][I[:,k]]                                         tgp_Vup_kernel      (:84)
// So do the two small dense solves per conditional (Cholesky of a q x q and an n*r x n*r matrix: one CTA each, the matrix
// in L2): the host only launches, and reads the stored draws back.
//
// Reference quirks kept: the factor draw is invSigma_U \ randn(n*r) (covariance invSigma_U^-2, :83) while the w draw
// uses the Cholesky factor (:66); W and the factors are stored before the factors move (:68-71); rows of C for values
// of l that no I[:,k] takes are uninitialised in the reference (:76) and zero here.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace {

// T[l][i] for one dimension k: thread = (row i, column l)
__global__ void tgp_T_kernel(const double *__restrict__ phi /*[N][D][n]*/, const double *__restrict__ Uk /*[r][n]*/, int64_t N,
                             int D, int n, int r, int k, int64_t ld, double *__restrict__ T /*[r][ld]*/) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int l = blockIdx.y;
    if (i >= N) return;
    const double *p = phi + ((size_t)i * D + k) * n, *u = Uk + (size_t)l * n;
    double s = 0.0;
    for (int j = 0; j < n; ++j) s = fma(u[j], p[j], s);
    T[(size_t)l * ld + i] = s;
}

__global__ void tgp_V_kernel(const double *__restrict__ T /*[D][r][ld]*/, const int32_t *__restrict__ I /*[D][Q]*/, int64_t N, int D,
                             int r, int Q, int64_t ld, double *__restrict__ V /*[Q][ld]*/) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int q = blockIdx.y;
    if (i >= N) return;
    double v = 1.0;
    for (int k = 0; k < D; ++k) v *= T[((size_t)k * r + I[(size_t)k * Q + q]) * ld + i];
    V[(size_t)q * ld + i] = v;
}

// G[a][b] = sum_i A[a][i] A[b][i]: 16 x 16 outputs per CTA of 256 threads, the row index i in chunks of 64
__global__ void __launch_bounds__(256) tgp_gram_kernel(const double *__restrict__ A, int M, int64_t N, int64_t ld, double *__restrict__ G) {
    __shared__ double sa[16][65], sb[16][65];
    const int a0 = blockIdx.x * 16, b0 = blockIdx.y * 16;
    if (b0 > a0) return;                                   // lower triangle of tiles; mirrored below
    const int ta = threadIdx.x & 15, tb = threadIdx.x >> 4;
    double acc = 0.0;
    for (int64_t i0 = 0; i0 < N; i0 += 64) {
        for (int t = threadIdx.x; t < 16 * 64; t += 256) {
            const int rr = t >> 6, cc = t & 63;
            const bool in = i0 + cc < N;
            sa[rr][cc] = (in && a0 + rr < M) ? A[(size_t)(a0 + rr) * ld + i0 + cc] : 0.0;
            sb[rr][cc] = (in && b0 + rr < M) ? A[(size_t)(b0 + rr) * ld + i0 + cc] : 0.0;
        }
        __syncthreads();
#pragma unroll 16
        for (int cc = 0; cc < 64; ++cc) acc = fma(sa[ta][cc], sb[tb][cc], acc);
        __syncthreads();
    }
    if (a0 + ta < M && b0 + tb < M) {
        G[(size_t)(a0 + ta) * M + b0 + tb] = acc;
        G[(size_t)(b0 + tb) * M + a0 + ta] = acc;
    }
}

// v[a] = sum_i A[a][i] y[i]: one warp per row
__global__ void tgp_gemv_kernel(const double *__restrict__ A, const double *__restrict__ y, int M, int64_t N, int64_t ld,
                                double *__restrict__ v) {
    const int a = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (a >= M) return;
    double s = 0.0;
    for (int64_t i = lane; i < N; i += 32) s = fma(A[(size_t)a * ld + i], y[i], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) v[a] = s;
}

// V_k = V ./ T_k[I[:,k]] (in place) and C[l][i] = sum_{q: I[q,k]=l} w_q V_k[q][i]; thread = row i owns column i of both
__global__ void tgp_C_kernel(double *__restrict__ V /*[Q][ld]*/, const double *__restrict__ Tk /*[r][ld]*/,
                             const int32_t *__restrict__ Ik /*[Q]*/, const double *__restrict__ w, int64_t N, int r, int Q, int64_t ld,
                             double *__restrict__ C /*[r][ld]*/) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= N) return;
    for (int l = 0; l < r; ++l) C[(size_t)l * ld + i] = 0.0;
    for (int q = 0; q < Q; ++q) {
        const int l = Ik[q];
        const double vk = V[(size_t)q * ld + i] / Tk[(size_t)l * ld + i];        // :74
        V[(size_t)q * ld + i] = vk;
        C[(size_t)l * ld + i] += w[q] * vk;                                      // :75-79
    }
}

// Ck[l*n + j][i] = C[l][i] * b[j,k,i]   (:80)
__global__ void tgp_Ck_kernel(const double *__restrict__ C, const double *__restrict__ phi, int64_t N, int D, int n, int r, int k,
                              int64_t ld, double *__restrict__ Ck /*[n r][ld]*/) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int row = blockIdx.y;                            // l*n + j
    if (i >= N) return;
    const int l = row / n, j = row - l * n;
    Ck[(size_t)row * ld + i] = C[(size_t)l * ld + i] * phi[((size_t)i * D + k) * n + j];
}

// V = V_k .* T_new[k][I[:,k]]   (:84)
__global__ void tgp_Vup_kernel(double *__restrict__ V, const double *__restrict__ Tk, const int32_t *__restrict__ Ik, int64_t N, int Q,
                               int64_t ld) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int q = blockIdx.y;
    if (i >= N) return;
    V[(size_t)q * ld + i] *= Tk[(size_t)Ik[q] * ld + i];
}

// ---- the small dense systems, on the device: one CTA of 1024 threads per system, the matrix in global memory (L2).
// S (M x M, row-major, symmetric positive definite) <- S * scale + diag * I, then its upper Cholesky factor R in place
// (S = R'R, as chol(.,:U)); *bad is raised when a pivot is not positive (NaN / Inf in the Gram matrix).
__global__ void __launch_bounds__(1024) tgp_chol_kernel(double *__restrict__ S, int M, double scale, double diag, int32_t *bad) {
    __shared__ double piv;
    const int tid = threadIdx.x, nt = blockDim.x;
    for (size_t t = tid; t < (size_t)M * M; t += nt) {
        const int a = (int)(t / M), c = (int)(t - (size_t)a * M);
        S[t] = S[t] * scale + (a == c ? diag : 0.0);
    }
    __syncthreads();
    for (int j = 0; j < M; ++j) {          // right-looking: row j of R, then the rank-one update of the trailing block
        if (tid == 0) {
            const double d = S[(size_t)j * M + j];
            if (!(d > 0.0)) *bad = 1;
            piv = sqrt(d);
            S[(size_t)j * M + j] = piv;
        }
        __syncthreads();
        const double p = piv;
        for (int c = j + 1 + tid; c < M; c += nt) S[(size_t)j * M + c] /= p;
        __syncthreads();
        for (int a = j + 1 + (tid >> 5); a < M; a += nt >> 5) {      // a warp per row a >= j+1, lanes along c >= a
            const double ra = S[(size_t)j * M + a];
            for (int c = a + (tid & 31); c < M; c += 32) S[(size_t)a * M + c] -= ra * S[(size_t)j * M + c];
        }
        __syncthreads();
    }
}

// x <- pre * x, then x <- R' \ x (fwd != 0), then x <- R \ x (bwd != 0): column sweeps, one CTA
__global__ void __launch_bounds__(1024) tgp_trsv_kernel(const double *__restrict__ R, int M, double *__restrict__ x, double pre, int fwd,
                                                        int bwd) {
    __shared__ double xi;
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < M; i += nt) x[i] *= pre;
    __syncthreads();
    if (fwd)
        for (int i = 0; i < M; ++i) {      // R' is lower triangular with column i of R' = row i of R
            if (tid == 0) { xi = x[i] / R[(size_t)i * M + i]; x[i] = xi; }
            __syncthreads();
            const double v = xi;
            for (int c = i + 1 + tid; c < M; c += nt) x[c] -= R[(size_t)i * M + c] * v;
            __syncthreads();
        }
    if (bwd)
        for (int i = M - 1; i >= 0; --i) {
            if (tid == 0) { xi = x[i] / R[(size_t)i * M + i]; x[i] = xi; }
            __syncthreads();
            const double v = xi;
            for (int a = tid; a < i; a += nt) x[a] -= R[(size_t)a * M + i] * v;
            __syncthreads();
        }
}

// Philox draws of one conditional (production mode): out[0..M), two per block, counter (stream | i/2, step)
__global__ void tgp_noise_kernel(double *__restrict__ out, int M, uint64_t key, uint64_t stream, uint64_t step) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (2 * p >= M) return;
    double a, b;
    philox_normal2(key, stream | (uint64_t)p, step, &a, &b);
    out[2 * p] = a;
    if (2 * p + 1 < M) out[2 * p + 1] = b;
}

// out = a + s * b
__global__ void tgp_axpby_kernel(const double *__restrict__ a, const double *__restrict__ b, double s, int M, double *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < M) out[i] = a[i] + s * b[i];
}

template <typename T>
struct DevBuf {
    T *p = nullptr;
    ~DevBuf() { cudaFree(p); }
    int alloc(gptb_handle *h, size_t count) {
        cudaError_t e = cudaMalloc((void **)&p, std::max<size_t>(count, 1) * sizeof(T));
        if (e != cudaSuccess) {
            gptb_set_error(h, "gptb_tgp_gibbs: cudaMalloc(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
            return GPTB_ERR_NOMEM;
        }
        return GPTB_OK;
    }
};

}  // namespace

extern "C" int gptb_tgp_gibbs(gptb_handle *h, double sigma, int32_t num_iterations, int32_t burnin, const double *noise_w,
                              const double *noise_U, double *W_array, double *V_array) {
    if (!h || !(sigma > 0) || num_iterations < 1 || burnin < 0 || burnin > num_iterations) return GPTB_ERR_ARG;
    if (!h->have_params || !(h->have_train || h->have_phi) || !h->have_index) {
        gptb_set_error(h, "gptb_tgp_gibbs: features, index, training set and initial factors must be set");
        return GPTB_ERR_STATE;
    }
    if (h->nranks != 1) {
        gptb_set_error(h, "gptb_tgp_gibbs: one rank only");
        return GPTB_ERR_UNSUPPORTED;
    }
    const gptb_config &c = h->cfg;
    const int n = c.n, r = c.r, D = c.D, Q = c.Q, nr = n * r;
    if (nr > 4096 || Q > 4096) {   // the factor's conditional is an n*r x n*r dense system (TGP.jl:81): a sampler for small n*r
        gptb_set_error(h, "gptb_tgp_gibbs: n*r = %d (or q = %d) above 4096: the dense conditionals of TGP.jl:64,81 are out of reach", nr, Q);
        return GPTB_ERR_UNSUPPORTED;
    }
    const int64_t N = h->N, ld = (N + 31) / 32 * 32;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(gptb_synchronize(h));
    cudaStream_t st = h->stream;
    const double sigma_u = sqrt(1.0 / r), sigma_w = sqrt(pow((double)r, (double)D) / Q);     // :54-55
    const double s2 = sigma * sigma;

    DevBuf<double> phi_own, T, V, C, Ck, Gq, Gu, vq, vu, wd;
    const double *phi = h->phi;
    if (!h->have_phi) {       // b[:,:,i] = feature(X[i,:], ...) for every row (:49-50)
        GPTB_TRY(phi_own.alloc(h, (size_t)N * D * n));
        GPTB_TRY(gptb_launch_features_fp64(h, h->X, N, phi_own.p));
        phi = phi_own.p;
    }
    GPTB_TRY(T.alloc(h, (size_t)D * r * ld));
    GPTB_TRY(V.alloc(h, (size_t)Q * ld));
    GPTB_TRY(C.alloc(h, (size_t)r * ld));
    GPTB_TRY(Ck.alloc(h, (size_t)nr * ld));
    GPTB_TRY(Gq.alloc(h, (size_t)Q * Q));
    GPTB_TRY(Gu.alloc(h, (size_t)nr * nr));
    GPTB_TRY(vq.alloc(h, (size_t)Q));
    GPTB_TRY(vu.alloc(h, (size_t)nr));
    GPTB_TRY(wd.alloc(h, (size_t)Q));
    const unsigned nb = (unsigned)((N + 127) / 128);
    auto launch_T = [&](int k) {
        tgp_T_kernel<<<dim3(nb, r), 128, 0, st>>>(phi, h->U + (size_t)k * r * n, N, D, n, r, k, ld, T.p + (size_t)k * r * ld);
        h->launches++;
    };
    DevBuf<double> zq, zu, mu;
    DevBuf<int32_t> bad;
    GPTB_TRY(zq.alloc(h, (size_t)Q));
    GPTB_TRY(zu.alloc(h, (size_t)nr));
    GPTB_TRY(mu.alloc(h, (size_t)Q));
    GPTB_TRY(bad.alloc(h, 1));
    GPTB_CUDA_OK(h, cudaMemsetAsync(bad.p, 0, sizeof(int32_t), st));
    int32_t bad_host = 0;
    for (int m = 1; m <= num_iterations; ++m) {
        for (int k = 0; k < D; ++k) launch_T(k);
        tgp_V_kernel<<<dim3(nb, Q), 128, 0, st>>>(T.p, h->I, N, D, r, Q, ld, V.p);                    // :62
        tgp_gram_kernel<<<dim3((Q + 15) / 16, (Q + 15) / 16), 256, 0, st>>>(V.p, Q, N, ld, Gq.p);
        tgp_gemv_kernel<<<(Q + 7) / 8, 256, 0, st>>>(V.p, h->y, Q, N, ld, vq.p);
        // invSigma_w = V V'/sigma^2 + I/sigma_w^2 = R'R (:64);  Mu_w = invSigma_w \ (V y/sigma^2) (:65)
        tgp_chol_kernel<<<1, 1024, 0, st>>>(Gq.p, Q, 1.0 / s2, 1.0 / (sigma_w * sigma_w), bad.p);
        tgp_trsv_kernel<<<1, 1024, 0, st>>>(Gq.p, Q, vq.p, 1.0 / s2, 1, 1);
        // W = R \ randn(q) + Mu_w (:66)
        if (noise_w) GPTB_CUDA_OK(h, cudaMemcpyAsync(zq.p, noise_w + (size_t)(m - 1) * Q, sizeof(double) * Q, cudaMemcpyHostToDevice, st));
        else tgp_noise_kernel<<<(Q / 2 + 128) / 128, 128, 0, st>>>(zq.p, Q, c.seed, GPTB_STREAM_NOISE_W, (uint64_t)m | (2ull << 40));
        tgp_trsv_kernel<<<1, 1024, 0, st>>>(Gq.p, Q, zq.p, 1.0, 0, 1);
        tgp_axpby_kernel<<<(Q + 127) / 128, 128, 0, st>>>(zq.p, vq.p, 1.0, Q, wd.p);
        h->launches += 8;
        if (m > burnin) {                                                                              // :68-71
            if (W_array)
                GPTB_CUDA_OK(h, cudaMemcpyAsync(W_array + (size_t)(m - burnin - 1) * Q, wd.p, sizeof(double) * Q, cudaMemcpyDeviceToHost, st));
            if (V_array)
                GPTB_CUDA_OK(h, cudaMemcpyAsync(V_array + (size_t)(m - burnin - 1) * nr * D, h->U, sizeof(double) * nr * D,
                                                cudaMemcpyDeviceToHost, st));
        }
        for (int k = 0; k < D; ++k) {
            const double *Tk = T.p + (size_t)k * r * ld;
            tgp_C_kernel<<<nb, 128, 0, st>>>(V.p, Tk, h->I + (size_t)k * Q, wd.p, N, r, Q, ld, C.p);
            tgp_Ck_kernel<<<dim3(nb, nr), 128, 0, st>>>(C.p, phi, N, D, n, r, k, ld, Ck.p);
            tgp_gram_kernel<<<dim3((nr + 15) / 16, (nr + 15) / 16), 256, 0, st>>>(Ck.p, nr, N, ld, Gu.p);
            tgp_gemv_kernel<<<(nr + 7) / 8, 256, 0, st>>>(Ck.p, h->y, nr, N, ld, vu.p);
            // invSigma_U = Ck Ck'/sigma^2 + I/sigma_u^2 (:81);  U_k = invSigma_U \ randn + invSigma_U \ (Ck y/sigma^2) (:82-83):
            // one solve of the summed right-hand side
            tgp_chol_kernel<<<1, 1024, 0, st>>>(Gu.p, nr, 1.0 / s2, 1.0 / (sigma_u * sigma_u), bad.p);
            if (noise_U)
                GPTB_CUDA_OK(h, cudaMemcpyAsync(zu.p, noise_U + ((size_t)(m - 1) * D + k) * nr, sizeof(double) * nr, cudaMemcpyHostToDevice, st));
            else
                tgp_noise_kernel<<<(nr / 2 + 128) / 128, 128, 0, st>>>(zu.p, nr, c.seed, GPTB_STREAM_NOISE_U | ((uint64_t)k << 40),
                                                                       (uint64_t)m | (2ull << 40));
            // reshape(., n, r): entry l*n + j is U[j,l] — the handle's own layout of U^(k)
            tgp_axpby_kernel<<<(nr + 127) / 128, 128, 0, st>>>(zu.p, vu.p, 1.0 / s2, nr, h->U + (size_t)k * nr);
            tgp_trsv_kernel<<<1, 1024, 0, st>>>(Gu.p, nr, h->U + (size_t)k * nr, 1.0, 1, 1);
            launch_T(k);
            tgp_Vup_kernel<<<dim3(nb, Q), 128, 0, st>>>(V.p, Tk, h->I + (size_t)k * Q, N, Q, ld);      // :84
            h->launches += 9;
        }
        // a pivot that is not positive (NaN / Inf in a Gram matrix): looked at once per iteration
        GPTB_CUDA_OK(h, cudaMemcpyAsync(&bad_host, bad.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        GPTB_CUDA_OK(h, cudaStreamSynchronize(st));
        if (bad_host) {
            gptb_set_error(h, "gptb_tgp_gibbs: a conditional precision matrix of iteration %d is not positive definite (NaN / Inf in V?)", m);
            return GPTB_ERR_ARG;
        }
    }
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->w, wd.p, sizeof(double) * Q, cudaMemcpyDeviceToDevice, st));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(st));
    h->shadows_valid = false;
    return GPTB_OK;
}
