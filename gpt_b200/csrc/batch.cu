// Minibatch selection: the reference re-permutes the whole feature tensor every epoch
// (phi=phi[:,:,perm]; y=y[perm], GPT_SGLD.jl:373-374) and then slices contiguous ranges (:379-381).
// Here the data never move: a device-resident `order` array carries the cumulative permutation,
// and a gather kernel builds the k-major batch Xb[k][i] (unit stride over rows for the feature
// generators) for the rows order[start .. start+bs).
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace {

__global__ void gather_kernel(const double *__restrict__ X /*[N][D]*/, const double *__restrict__ y,
                              const int32_t *__restrict__ rows /*may be null: identity from row_base*/,
                              int64_t row_base, int64_t bs, int64_t mpad, int D, double *__restrict__ Xb,
                              float *__restrict__ XbF, double *__restrict__ yb, const double *__restrict__ glo,
                              const double *__restrict__ ginvh, int G, int32_t *__restrict__ cell, float *__restrict__ uu,
                              int32_t *__restrict__ oob) {
    // cell != null: the node-grid position of every entry, computed from the FP64 input — cell c = floor((x-lo)/(2h)),
    // u = (x-lo)/h - (2c+1) in [-1,1] (nodes.cu); a row outside the domain raises *oob
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < mpad; i += (int64_t)gridDim.x * blockDim.x) {
        const bool live = i < bs;
        int64_t src = 0;
        if (live) src = rows ? (int64_t)rows[i] : row_base + i;
        for (int k = 0; k < D; ++k) {
            double v = live ? X[src * D + k] : 0.0;
            if (Xb) Xb[(size_t)k * mpad + i] = v;
            if (XbF) XbF[(size_t)k * mpad + i] = (float)v;
            if (cell) {
                int c = 0;
                double u = 0.0;
                if (live) {
                    const double t = (v - glo[k]) * ginvh[k];
                    c = min(max((int)floor(0.5 * t), 0), G - 1);
                    u = t - (2.0 * c + 1.0);
                    if (!(u >= -1.000001 && u <= 1.000001)) *oob = 1;
                    u = fmin(fmax(u, -1.0), 1.0);
                }
                cell[(size_t)k * mpad + i] = c;
                uu[(size_t)k * mpad + i] = (float)u;
            }
        }
        if (y && yb) yb[i] = live ? y[src] : 0.0;
    }
}

// column-major host layout X[N,D] -> row-major rows [N][D]
__global__ void transpose_cm_to_rm(const double *__restrict__ Xcm, int64_t N, int D, double *__restrict__ Xrm) {
    int64_t total = N * D;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t i = t / D;
        int k = (int)(t % D);
        Xrm[t] = Xcm[(size_t)k * N + i];
    }
}

__global__ void iota_kernel(int32_t *o, int64_t N) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x)
        o[i] = (int32_t)i;
}

__global__ void perm_keys_kernel(uint64_t *keys, int64_t N, uint64_t key, uint64_t epoch) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
        philox4 p = philox4x32_10(key, GPTB_STREAM_PERM | (uint64_t)i, epoch);
        keys[i] = ((uint64_t)p.x << 32) | p.y;
    }
}

// order_out[i] = order_in[perm[i]]  — the cumulative application of :374
// (entries outside 0..N-1 raise *bad and read row 0: a bad host permutation must not become an out-of-bounds read)
__global__ void apply_perm_kernel(const int32_t *order_in, const int64_t *perm, int64_t N, int32_t *order_out, int32_t *bad) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t v = perm[i];
        if (v < 0 || v >= N) {
            *bad = 1;
            v = 0;
        }
        order_out[i] = order_in[v];
    }
}

}  // namespace

static int grid_for(const gptb_handle *h, int64_t n, int threads) {
    int64_t b = (n + threads - 1) / threads;
    return (int)std::max<int64_t>(1, std::min<int64_t>(b, (int64_t)h->sm_count * 8));
}

int gptb_launch_gather(gptb_handle *h, const int32_t *rows, int64_t start, int64_t bs, int64_t row_off) {
    // rows (device) already points at the first row of this rank's share, or is null for identity.
    (void)row_off;
    const bool tc = gptb_is_tc(h);
    const gptb_nodes &g = h->nodes;
    const bool gc = h->gather_cells;
    gather_kernel<<<grid_for(h, h->mpad, 128), 128, 0, h->stream>>>(h->X, h->y, rows, start, bs, h->mpad, h->cfg.D,
                                                                    tc ? nullptr : h->Xb, tc ? h->XbF : nullptr, h->yb, g.lo,
                                                                    g.invh, g.G, gc ? g.cell : nullptr, g.uu, h->nan_flag + 4);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_gather_from(gptb_handle *h, const double *Xrm, const double *y, int64_t bs, int64_t mpad, double *Xb,
                            float *XbF, double *yb) {
    const gptb_nodes &g = h->nodes;
    const bool gc = h->gather_cells;
    gather_kernel<<<grid_for(h, mpad, 128), 128, 0, h->stream>>>(Xrm, y, nullptr, 0, bs, mpad, h->cfg.D, Xb, XbF, yb, g.lo, g.invh,
                                                                 g.G, gc ? g.cell : nullptr, g.uu, h->nan_flag + 4);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_transpose(gptb_handle *h, const double *Xcm, int64_t N, double *Xrm) {
    if (N == 0) return GPTB_OK;
    transpose_cm_to_rm<<<grid_for(h, N * h->cfg.D, 256), 256, 0, h->stream>>>(Xcm, N, h->cfg.D, Xrm);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_reset_order(gptb_handle *h) {
    if (h->N == 0) return GPTB_OK;
    iota_kernel<<<grid_for(h, h->N, 256), 256, 0, h->stream>>>(h->order, h->N);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// Philox shuffle: sort (random key, order) pairs; equivalent to order = order[randperm(N)].
int gptb_shuffle_order(gptb_handle *h, uint64_t epoch) {
    const int64_t N = h->N;
    if (N <= 1) return GPTB_OK;
    if (!h->sort_keys) {
        GPTB_CUDA_OK(h, cudaMalloc(&h->sort_keys, sizeof(uint64_t) * N));
        GPTB_CUDA_OK(h, cudaMalloc(&h->sort_keys_tmp, sizeof(uint64_t) * N));
        GPTB_CUDA_OK(h, cudaMalloc(&h->order_tmp, sizeof(int32_t) * N));
        size_t bytes = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, bytes, h->sort_keys, h->sort_keys_tmp, h->order, h->order_tmp, (int)N);
        h->sort_tmp_bytes = bytes;
        GPTB_CUDA_OK(h, cudaMalloc(&h->sort_tmp, bytes));
    }
    perm_keys_kernel<<<grid_for(h, N, 256), 256, 0, h->stream>>>(h->sort_keys, N, h->cfg.seed, epoch);
    GPTB_CUDA_OK(h, cub::DeviceRadixSort::SortPairs(h->sort_tmp, h->sort_tmp_bytes, h->sort_keys, h->sort_keys_tmp,
                                                    h->order, h->order_tmp, (int)N, 0, 64, h->stream));
    std::swap(h->order, h->order_tmp);
    h->launches += 8;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// Host-supplied permutation (0-based, N entries, already on the device as int64).
int gptb_apply_perm(gptb_handle *h, const int64_t *perm_dev) {
    if (!h->order_tmp) GPTB_CUDA_OK(h, cudaMalloc(&h->order_tmp, sizeof(int32_t) * h->N));
    apply_perm_kernel<<<grid_for(h, h->N, 256), 256, 0, h->stream>>>(h->order, perm_dev, h->N, h->order_tmp, h->nan_flag + 3);
    std::swap(h->order, h->order_tmp);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}
