// GPTclassification — GPT_SGLD.jl:452-680: C independent (w, U) sets under a softmax likelihood, SGLD on w and
// SGLDERM on the factors.
//
// The class-c gradients are the regression gradients (:393, :405-408) with the residual
//     e[i,c] = (N/bs) (1[y_i = c] - exp(fhat[i,c] - logsumexp(fhat[i,:])))                      (:506-524, :546-564)
// so every class is one handle of this library (its own parameters, operand images, T and gradient buffers; the same
// features, core index table and training inputs), and a minibatch runs
//     stage 1, per class:  gather, forward pass (K1 or the node grid), K2 up to fhat          (api.cu: compute_grad)
//     softmax_targets_kernel: yb_c[i] = fhat_c[i] + 1[y_i = c] - p[i,c]   — with signal_var = 1 the residual K2 forms
//                             from these targets, (N/bs)(yb_c - fhat_c), is e[i,c]
//     stage 2, per class:  K2 (V, fhat again, gradw, E), back-projection (K3 or the node grid)
//     update, per class:   the reference moves (w, U) TWICE per minibatch with the same gradient: once unconditionally
//                          with noise on the Stiefel manifold (:624-636), then under the langevin / stiefel switches
//                          (:639-671).  GPTB_FLAG_CLS_SINGLE_UPDATE keeps only the second move.
// Classes run on their handles' own streams, joined by events around the softmax kernel.
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

int gptb_i_grad_stage(gptb_handle *h, const int32_t *rows_dev, int64_t bs, int64_t N_total, int stage);   // api.cu
int gptb_i_stage_noise(gptb_handle *h, const double *noise_w, const double *noise_U, const double **nw_dev, const double **nU_dev);
int gptb_i_check_nan(gptb_handle *h);
int gptb_i_check_ready(gptb_handle *h);
int gptb_reset_order(gptb_handle *h);                        // batch.cu
int gptb_apply_perm(gptb_handle *h, const int64_t *perm_dev);

namespace {

constexpr int CLS_MAX = 64;

struct ClsPtrs {
    const double *fhat[CLS_MAX];
    double *yb[CLS_MAX];
};

// gradw -= w / sigma_w^2 (:524, sigma_w = 1): both moves use the prior of the w the gradient was taken at
__global__ void fold_prior_kernel(double *__restrict__ gradw, const double *__restrict__ w, int Q) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < Q) gradw[q] -= w[q];
}

// one thread per minibatch row: logsumexp over the classes (:508-511), then the targets of every class
__global__ void softmax_targets_kernel(ClsPtrs p, int C, const int32_t *__restrict__ labels, const int32_t *__restrict__ rows,
                                       int64_t bs) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < bs; i += (int64_t)gridDim.x * blockDim.x) {
        double mx = -INFINITY;
        for (int c = 0; c < C; ++c) mx = fmax(mx, p.fhat[c][i]);
        double s = 0.0;
        for (int c = 0; c < C; ++c) s += exp(p.fhat[c][i] - mx);
        const double lse = mx + log(s);
        const int yi = labels[rows[i]];
        for (int c = 0; c < C; ++c) {
            const double f = p.fhat[c][i];
            p.yb[c][i] = f + ((c == yi ? 1.0 : 0.0) - exp(f - lse));
        }
    }
}

}  // namespace

extern "C" int gptb_classify_run(gptb_handle *const *hc, int32_t C, const int32_t *labels, int64_t m, double epsw, double epsU,
                                 int32_t burnin, int32_t maxepoch, int32_t flags, const int64_t *perms, const double *noise_w,
                                 const double *noise_U, double *w_store, double *U_store) {
    if (!hc || C < 2 || C > CLS_MAX || !labels || m < 1 || burnin < 0 || maxepoch < 0) return GPTB_ERR_ARG;
    for (int c = 0; c < C; ++c)
        if (!hc[c]) return GPTB_ERR_ARG;
    gptb_handle *h0 = hc[0];
    const gptb_config &cf = h0->cfg;
    const int64_t N = h0->N;
    for (int c = 0; c < C; ++c) {
        gptb_handle *h = hc[c];
        GPTB_TRY(gptb_i_check_ready(h));
        if (h->cfg.n != cf.n || h->cfg.r != cf.r || h->cfg.D != cf.D || h->cfg.Q != cf.Q || h->N != N || h->device != h0->device) {
            gptb_set_error(h0, "gptb_classify_run: class %d's handle differs from class 0's in shape, training-set size or device", c);
            return GPTB_ERR_ARG;
        }
        if (h->nranks != 1) {
            gptb_set_error(h0, "gptb_classify_run: one rank only (the softmax needs every class's fhat of the same rows)");
            return GPTB_ERR_UNSUPPORTED;
        }
        for (int d = 0; d < c; ++d)
            if (hc[d] == h) return GPTB_ERR_ARG;
    }
    for (int64_t i = 0; i < N; ++i)
        if (labels[i] < 0 || labels[i] >= C) {
            gptb_set_error(h0, "gptb_classify_run: label %d of row %lld outside 0..C-1", labels[i], (long long)i);
            return GPTB_ERR_ARG;
        }
    GPTB_CUDA_OK(h0, cudaSetDevice(h0->device));
    const int Q = cf.Q;
    const size_t ucnt = (size_t)cf.n * cf.r * cf.D;
    const int64_t numbatches = (N + m - 1) / m;                       // :458
    const int64_t T = (int64_t)maxepoch * numbatches;
    const bool single = (flags & GPTB_FLAG_CLS_SINGLE_UPDATE) != 0;
    const int uflags = flags & (GPTB_FLAG_LANGEVIN | GPTB_FLAG_STIEFEL);
    const int per_step = single ? 1 : 2;                               // noise sets per minibatch

    int32_t *labels_dev = nullptr;
    int64_t *perm_dev = nullptr;
    double *store_w = nullptr, *store_U = nullptr;
    std::vector<cudaEvent_t> ev(C + 1, nullptr);
    int st = GPTB_OK;
    auto cleanup = [&]() {
        for (cudaEvent_t e : ev)
            if (e) cudaEventDestroy(e);
        cudaFree(labels_dev);
        cudaFree(perm_dev);
        cudaFree(store_w);
        cudaFree(store_U);
    };
#define CLS_CUDA(call)                                                   \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) {                                         \
            gptb_set_error(h0, "%s: %s", #call, cudaGetErrorString(e_)); \
            cleanup();                                                   \
            return GPTB_ERR_CUDA;                                        \
        }                                                                \
    } while (0)
#define CLS_TRY(call)             \
    do {                          \
        st = (call);              \
        if (st != GPTB_OK) {      \
            cleanup();            \
            return st;            \
        }                         \
    } while (0)
    for (auto &e : ev) CLS_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CLS_CUDA(cudaMalloc((void **)&labels_dev, sizeof(int32_t) * N));
    CLS_CUDA(cudaMemcpyAsync(labels_dev, labels, sizeof(int32_t) * N, cudaMemcpyHostToDevice, h0->stream));
    if (perms) CLS_CUDA(cudaMalloc((void **)&perm_dev, sizeof(int64_t) * N));
    if (T > 0 && w_store) CLS_CUDA(cudaMalloc((void **)&store_w, sizeof(double) * Q * C * T));
    if (T > 0 && U_store) CLS_CUDA(cudaMalloc((void **)&store_U, sizeof(double) * ucnt * C * T));
    // every class walks class 0's row order (the reference permutes the one phi / y pair, :480-481)
    CLS_TRY(gptb_reset_order(h0));
    CLS_CUDA(cudaStreamSynchronize(h0->stream));

    uint64_t step = 0;
    bool nan_abort = false;
    for (int epoch = 1; epoch <= burnin + maxepoch && !nan_abort; ++epoch) {
        if (perms) {
            CLS_CUDA(cudaMemcpyAsync(perm_dev, perms + (size_t)(epoch - 1) * N, sizeof(int64_t) * N, cudaMemcpyHostToDevice, h0->stream));
            CLS_TRY(gptb_apply_perm(h0, perm_dev));
        } else {
            CLS_TRY(gptb_shuffle_order(h0, h0->epoch_counter));
        }
        h0->epoch_counter++;
        CLS_CUDA(cudaEventRecord(ev[C], h0->stream));
        for (int c = 1; c < C; ++c) CLS_CUDA(cudaStreamWaitEvent(hc[c]->stream, ev[C], 0));
        for (int64_t batch = 1; batch <= numbatches; ++batch, ++step) {
            const int64_t lo = m * (batch - 1), hi = std::min(m * batch, N), bs = hi - lo;   // :486
            const int32_t *rows = h0->order + lo;
            ClsPtrs ptrs;
            for (int c = 0; c < C; ++c) {
                CLS_TRY(gptb_i_grad_stage(hc[c], rows, bs, N, 1));
                ptrs.fhat[c] = hc[c]->fhat;
                ptrs.yb[c] = hc[c]->yb;
                if (c > 0) {
                    CLS_CUDA(cudaEventRecord(ev[c], hc[c]->stream));
                    CLS_CUDA(cudaStreamWaitEvent(h0->stream, ev[c], 0));
                }
            }
            softmax_targets_kernel<<<(unsigned)std::min<int64_t>((bs + 127) / 128, 1024), 128, 0, h0->stream>>>(ptrs, C, labels_dev, rows, bs);
            CLS_CUDA(cudaGetLastError());
            CLS_CUDA(cudaEventRecord(ev[0], h0->stream));
            for (int c = 0; c < C; ++c) {
                gptb_handle *h = hc[c];
                if (c > 0) CLS_CUDA(cudaStreamWaitEvent(h->stream, ev[0], 0));
                CLS_TRY(gptb_i_grad_stage(h, rows, bs, N, 2));
                fold_prior_kernel<<<(Q + 127) / 128, 128, 0, h->stream>>>(h->grad, h->w, Q);
                for (int u = single ? 1 : 0; u < 2; ++u) {
                    // draws of move u: noise_w[step][u][Q][C], noise_U[step][u][n][r][D][C]
                    const size_t set = (size_t)step * per_step + (single ? 0 : u);
                    const double *nw = nullptr, *nU = nullptr;
                    const int fl = (u == 0 ? (GPTB_FLAG_LANGEVIN | GPTB_FLAG_STIEFEL) : uflags) | GPTB_IFLAG_PRIOR_IN_GRAD;
                    if (fl & GPTB_FLAG_LANGEVIN) {
                        // (the staging copy is ordered on the class's stream after move 0, which reads the same buffers)
                        CLS_TRY(gptb_i_stage_noise(h, noise_w ? noise_w + (set * C + c) * Q : nullptr,
                                                   noise_U ? noise_U + (set * C + c) * ucnt : nullptr, &nw, &nU));
                    }
                    CLS_TRY(gptb_launch_update_range(h, epsw, epsU, fl, nw, nU, h->step_counter, 0, cf.D, true, h->stream));
                    h->step_counter++;
                }
                if (epoch > burnin) {
                    const int64_t col = (int64_t)(epoch - burnin - 1) * numbatches + batch - 1;
                    if (store_w) CLS_CUDA(cudaMemcpyAsync(store_w + ((size_t)col * C + c) * Q, h->w, sizeof(double) * Q, cudaMemcpyDeviceToDevice, h->stream));
                    if (store_U) CLS_CUDA(cudaMemcpyAsync(store_U + ((size_t)col * C + c) * ucnt, h->U, sizeof(double) * ucnt, cudaMemcpyDeviceToDevice, h->stream));
                }
            }
        }
        // the reference tests for the NaN abort after every geodesic (:631, :650); the device flag is sticky and freezes
        // the parameters, so one look per epoch returns the same all-zero result
        for (int c = 0; c < C; ++c) {
            st = gptb_i_check_nan(hc[c]);
            if (st == GPTB_GEODESIC_NAN) {
                nan_abort = true;
                break;
            }
            if (st != GPTB_OK) {
                cleanup();
                return st;
            }
        }
    }
    for (int c = 0; c < C; ++c) CLS_CUDA(cudaStreamSynchronize(hc[c]->stream));
    if (nan_abort) {
        if (w_store) memset(w_store, 0, sizeof(double) * Q * C * T);                 // :632, :651
        if (U_store) memset(U_store, 0, sizeof(double) * ucnt * C * T);
        cleanup();
        return GPTB_GEODESIC_NAN;
    }
    if (store_w) CLS_CUDA(cudaMemcpy(w_store, store_w, sizeof(double) * Q * C * T, cudaMemcpyDeviceToHost));
    if (store_U) CLS_CUDA(cudaMemcpy(U_store, store_U, sizeof(double) * ucnt * C * T, cudaMemcpyDeviceToHost));
    cleanup();
#undef CLS_CUDA
#undef CLS_TRY
    return GPTB_OK;
}
