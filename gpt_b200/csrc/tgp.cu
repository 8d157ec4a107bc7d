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
// The two small dense solves per conditional (Cholesky of a q x q and an n*r x n*r matrix) run on the host, as the
// reference's `\` and `chol` do; only those matrices, their right-hand sides and the drawn parameters cross the bus.
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

// ---- host: S = R'R (upper R, as chol(.,:U)), solves with it.  S is M x M, symmetric positive definite, row-major.
bool chol_upper(std::vector<double> &S, int M) {
    for (int j = 0; j < M; ++j) {
        double d = S[(size_t)j * M + j];
        for (int t = 0; t < j; ++t) d -= S[(size_t)t * M + j] * S[(size_t)t * M + j];
        if (!(d > 0.0)) return false;
        d = sqrt(d);
        S[(size_t)j * M + j] = d;
        for (int c = j + 1; c < M; ++c) {
            double v = S[(size_t)j * M + c];
            for (int t = 0; t < j; ++t) v -= S[(size_t)t * M + j] * S[(size_t)t * M + c];
            S[(size_t)j * M + c] = v / d;
        }
    }
    return true;
}
void solve_Rt(const std::vector<double> &R, int M, std::vector<double> &x) {   // R' x = b (forward)
    for (int i = 0; i < M; ++i) {
        double v = x[i];
        for (int t = 0; t < i; ++t) v -= R[(size_t)t * M + i] * x[t];
        x[i] = v / R[(size_t)i * M + i];
    }
}
void solve_R(const std::vector<double> &R, int M, std::vector<double> &x) {    // R x = b (backward)
    for (int i = M - 1; i >= 0; --i) {
        double v = x[i];
        for (int t = i + 1; t < M; ++t) v -= R[(size_t)i * M + t] * x[t];
        x[i] = v / R[(size_t)i * M + i];
    }
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
    std::vector<double> Sq((size_t)Q * Q), bq(Q), W(Q), zq(Q), Su((size_t)nr * nr), bu(nr), zu(nr), Uhost((size_t)nr * D);
    const int Tn = num_iterations - burnin;
    for (int m = 1; m <= num_iterations; ++m) {
        for (int k = 0; k < D; ++k) launch_T(k);
        tgp_V_kernel<<<dim3(nb, Q), 128, 0, st>>>(T.p, h->I, N, D, r, Q, ld, V.p);                    // :62
        tgp_gram_kernel<<<dim3((Q + 15) / 16, (Q + 15) / 16), 256, 0, st>>>(V.p, Q, N, ld, Gq.p);
        tgp_gemv_kernel<<<(Q + 7) / 8, 256, 0, st>>>(V.p, h->y, Q, N, ld, vq.p);
        h->launches += 3;
        GPTB_CUDA_OK(h, cudaMemcpyAsync(Sq.data(), Gq.p, sizeof(double) * Q * Q, cudaMemcpyDeviceToHost, st));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(bq.data(), vq.p, sizeof(double) * Q, cudaMemcpyDeviceToHost, st));
        GPTB_CUDA_OK(h, cudaStreamSynchronize(st));
        for (int a = 0; a < Q; ++a) {
            for (int b = 0; b < Q; ++b) Sq[(size_t)a * Q + b] /= s2;
            Sq[(size_t)a * Q + a] += 1.0 / (sigma_w * sigma_w);                                        // :64
            bq[a] /= s2;
        }
        if (!chol_upper(Sq, Q)) {
            gptb_set_error(h, "gptb_tgp_gibbs: invSigma_w of iteration %d is not positive definite (NaN / Inf in V?)", m);
            return GPTB_ERR_ARG;
        }
        // Mu_w = invSigma_w \ b (:65);  W = R \ randn(q) + Mu_w (:66)
        solve_Rt(Sq, Q, bq);
        solve_R(Sq, Q, bq);
        for (int a = 0; a < Q; a += 2) {
            double z0, z1;
            if (noise_w) { z0 = noise_w[(size_t)(m - 1) * Q + a]; z1 = a + 1 < Q ? noise_w[(size_t)(m - 1) * Q + a + 1] : 0.0; }
            else philox_normal2(c.seed, GPTB_STREAM_NOISE_W | (uint64_t)(a / 2), (uint64_t)m | (2ull << 40), &z0, &z1);
            zq[a] = z0;
            if (a + 1 < Q) zq[a + 1] = z1;
        }
        solve_R(Sq, Q, zq);
        for (int a = 0; a < Q; ++a) W[a] = zq[a] + bq[a];
        GPTB_CUDA_OK(h, cudaMemcpyAsync(wd.p, W.data(), sizeof(double) * Q, cudaMemcpyHostToDevice, st));
        if (m > burnin) {                                                                              // :68-71
            if (W_array) memcpy(W_array + (size_t)(m - burnin - 1) * Q, W.data(), sizeof(double) * Q);
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
            h->launches += 4;
            GPTB_CUDA_OK(h, cudaMemcpyAsync(Su.data(), Gu.p, sizeof(double) * nr * nr, cudaMemcpyDeviceToHost, st));
            GPTB_CUDA_OK(h, cudaMemcpyAsync(bu.data(), vu.p, sizeof(double) * nr, cudaMemcpyDeviceToHost, st));
            GPTB_CUDA_OK(h, cudaStreamSynchronize(st));
            for (int a = 0; a < nr; ++a) {
                for (int b = 0; b < nr; ++b) Su[(size_t)a * nr + b] /= s2;
                Su[(size_t)a * nr + a] += 1.0 / (sigma_u * sigma_u);                                   // :81
            }
            for (int a = 0; a < nr; a += 2) {
                double z0, z1;
                if (noise_U) {
                    const double *zz = noise_U + ((size_t)(m - 1) * D + k) * nr;
                    z0 = zz[a]; z1 = a + 1 < nr ? zz[a + 1] : 0.0;
                } else {
                    philox_normal2(c.seed, GPTB_STREAM_NOISE_U | ((uint64_t)k << 40) | (uint64_t)(a / 2), (uint64_t)m | (2ull << 40), &z0, &z1);
                }
                zu[a] = z0;
                if (a + 1 < nr) zu[a + 1] = z1;
            }
            // U_k = invSigma_U \ randn + invSigma_U \ (Ck y / sigma^2)   (:82-83): one solve of the summed right-hand side
            for (int a = 0; a < nr; ++a) zu[a] += bu[a] / s2;
            if (!chol_upper(Su, nr)) {
                gptb_set_error(h, "gptb_tgp_gibbs: invSigma_U of iteration %d, dimension %d is not positive definite", m, k + 1);
                return GPTB_ERR_ARG;
            }
            solve_Rt(Su, nr, zu);
            solve_R(Su, nr, zu);
            // reshape(., n, r): entry l*n + j is U[j,l] — the handle's own layout of U^(k)
            GPTB_CUDA_OK(h, cudaMemcpyAsync(h->U + (size_t)k * nr, zu.data(), sizeof(double) * nr, cudaMemcpyHostToDevice, st));
            GPTB_CUDA_OK(h, cudaStreamSynchronize(st));       // (zu is reused by the next dimension)
            launch_T(k);
            tgp_Vup_kernel<<<dim3(nb, Q), 128, 0, st>>>(V.p, Tk, h->I + (size_t)k * Q, N, Q, ld);      // :84
            h->launches++;
        }
    }
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->w, W.data(), sizeof(double) * Q, cudaMemcpyHostToDevice, st));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(st));
    h->shadows_valid = false;
    (void)Tn;
    return GPTB_OK;
}
