// K6 on the node grid — the posterior-predictive sweep (kin40kExperiment.jl:75-87, GPT_SGLD_p.jl:124-132):
//     per_sample[i, s] = pred(w_s, U_s, I, phitest)[i]   (GPT_SGLD.jl:233-243),     mean[i] = (1/S) sum_s per_sample[i, s].
//
// The direct tensor-core path (api.cu: pred_tc_grouped) generates the features of every test row once per group of four
// samples: 50 x the algorithmic cosine count at BASELINE config 5.  Here the features are generated at the G*P nodes of
// a grid over the test inputs' domain only (nodes.cu), once per sample group, and every (row, sample) pair costs
//   - D*r*P FMAs of interpolation  (interp_kernel on the cell-sorted records of the test rows: one basis recurrence
//     serves the 160 columns a warp owns), 4*D*r bytes of T2 written and read back (2 x 128 GB at config 5), and
//   - the Q*D shared-memory gathers of the contraction (contract_fwd3_kernel; contract_fwd2_kernel for the row-tile
//     layout): 4*Q*D bytes of shared-memory bandwidth per pair, 35 ms for 2e8 pairs on 148 SMs.
// Loop order: sample groups outside (up to 32 samples side by side on the fastest axis of T2: column l*SL + s, so
// that the contraction gathers one coalesced line per (k, l) for all samples at once), row chunks inside.  The T2 budget
// is GPTB_OPT_PRED_CHUNK_MB (4096 MB by default, two half-buffers): larger chunks are faster — the contraction is a
// persistent one-CTA-per-SM kernel whose launch stages 32 KB of tables per CTA and ends in a partial wave, and an
// L2-sized chunk (40 MB) buys nothing because the contraction is bound by its shared-memory gathers.
#include <math.h>

#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"

struct gptb_pred_cache {
    // one capacity slot per buffer (elements); buffers only ever grow and are kept between calls
    enum { B_X, B_ACC, B_CELL, B_SROW, B_SC, B_UU, B_SU, B_XNF, B_TN, B_CF, B_T2, B_STATE, B_LOHI, B_LO, B_INVH, B_UD, B_WD, B_FH,
           B_IMG, B_COUNT };
    size_t cap[B_COUNT] = {0};
    double *X = nullptr, *acc = nullptr;                 // [D][ntpad] inputs (column-major as passed), [ntpad] mean
    int32_t *cell = nullptr, *srow = nullptr, *sc = nullptr;
    float *uu = nullptr, *su = nullptr;
    float *XnF = nullptr, *Tn = nullptr, *Cf = nullptr, *T2 = nullptr, *state = nullptr;
    double *lohi = nullptr, *lo = nullptr, *invh = nullptr, *Ud = nullptr, *wd = nullptr, *fh = nullptr;
    unsigned char *img = nullptr;
    double *lohi_host = nullptr;                          // pinned
    cudaStream_t st2 = nullptr;                           // contraction stream (two T2 half-buffers alternate)
    cudaEvent_t ev_i[2] = {nullptr, nullptr}, ev_c[2] = {nullptr, nullptr}, ev_g = nullptr;
};

namespace {

constexpr int64_t PRED_CHUNK_ALIGN = 2048;   // = the sort's chunk: a row chunk starts on a sorted-chunk boundary

// lo / hi of every input column: one CTA per dimension.
__global__ void __launch_bounds__(1024) col_minmax_kernel(const double *__restrict__ X, int64_t Nt, int64_t ld,
                                                          double *__restrict__ lohi /*[2][D]*/, int D) {
    __shared__ double slo[32], shi[32];
    const int k = blockIdx.x;
    const double *col = X + (size_t)k * ld;
    double lo = INFINITY, hi = -INFINITY;
    for (int64_t i = threadIdx.x; i < Nt; i += 1024) {
        const double v = col[i];
        lo = fmin(lo, v);
        hi = fmax(hi, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) { slo[threadIdx.x >> 5] = lo; shi[threadIdx.x >> 5] = hi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 32; ++w) { lo = fmin(lo, slo[w]); hi = fmax(hi, shi[w]); }
        lohi[k] = lo;
        lohi[D + k] = hi;
    }
}

// cell c = floor((x - lo) / (2h)) and offset u = (x - lo)/h - (2c + 1) in [-1, 1] of every test input (as the
// sampler's gather does for a minibatch, batch.cu)
__global__ void pred_cells_kernel(const double *__restrict__ X, int64_t Nt, int64_t ld, const double *__restrict__ lo,
                                  const double *__restrict__ invh, int G, int32_t *__restrict__ cell, float *__restrict__ uu) {
    const int k = blockIdx.y;
    const double l = lo[k], ih = invh[k];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < Nt; i += (int64_t)gridDim.x * blockDim.x) {
        const double t = (X[(size_t)k * ld + i] - l) * ih;
        const int c = min(max((int)floor(0.5 * t), 0), G - 1);
        cell[(size_t)k * ld + i] = c;
        uu[(size_t)k * ld + i] = (float)fmin(fmax(t - (2.0 * c + 1.0), -1.0), 1.0);
    }
}

template <typename T>
int pc_grow(gptb_handle *h, T **p, size_t *cap, size_t count) {
    if (*p && *cap >= count) return GPTB_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    cudaError_t e = cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) {
        gptb_set_error(h, "posterior sweep: cudaMalloc(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
        *cap = 0;
        return GPTB_ERR_NOMEM;
    }
    *cap = count;
    return GPTB_OK;
}

}  // namespace

void gptb_pred_nodes_free(gptb_handle *h) {
    gptb_pred_cache *pc = h->pred_cache;
    if (!pc) return;
    cudaFree(pc->X); cudaFree(pc->acc); cudaFree(pc->cell); cudaFree(pc->srow); cudaFree(pc->sc); cudaFree(pc->uu); cudaFree(pc->su);
    cudaFree(pc->XnF); cudaFree(pc->Tn); cudaFree(pc->Cf); cudaFree(pc->T2); cudaFree(pc->state); cudaFree(pc->lohi);
    cudaFree(pc->lo); cudaFree(pc->invh); cudaFree(pc->Ud); cudaFree(pc->wd); cudaFree(pc->fh); cudaFree(pc->img);
    if (pc->lohi_host) cudaFreeHost(pc->lohi_host);
    if (pc->st2) {
        cudaStreamDestroy(pc->st2);
        for (int i = 0; i < 2; ++i) { cudaEventDestroy(pc->ev_i[i]); cudaEventDestroy(pc->ev_c[i]); }
        cudaEventDestroy(pc->ev_g);
    }
    delete pc;
    h->pred_cache = nullptr;
}

// Returns GPTB_ERR_UNSUPPORTED (no message) when the node grid does not apply; the caller then takes the direct path.
int gptb_pred_nodes(gptb_handle *h, const double *w_store, const double *U_store, int32_t S, const double *Xtest, int64_t Nt,
                    double *mean_out, double *per_sample_out) {
    const gptb_config &c = h->cfg;
    if (!gptb_is_tc(h) || !h->ct_off1 || h->algorithm == GPTB_ALG_DIRECT) return GPTB_ERR_UNSUPPORTED;
    if (h->algorithm == GPTB_ALG_AUTO && (Nt < 8192 || Nt * (int64_t)S < (1 << 20))) return GPTB_ERR_UNSUPPORTED;
    const int P = h->node_P, PP = (P + 3) & ~3;
    const int max_cols = P == 12 ? 160 : 64;              // widest column group of the tiles layout (interp_kernel instances)
    if (c.r > max_cols || Nt > 2147483647LL - PRED_CHUNK_ALIGN) return GPTB_ERR_UNSUPPORTED;
    if (!h->pred_cache) h->pred_cache = new gptb_pred_cache();
    gptb_pred_cache &pc = *h->pred_cache;
    typedef gptb_pred_cache PC;
    cudaStream_t st = h->stream;
    const size_t ucnt = (size_t)c.n * c.r * c.D;
    const int64_t ntpad = (Nt + PRED_CHUNK_ALIGN - 1) / PRED_CHUNK_ALIGN * PRED_CHUNK_ALIGN;

    // ---- test inputs, their domain, the grid
    GPTB_TRY(pc_grow(h, &pc.X, &pc.cap[PC::B_X], (size_t)c.D * ntpad));
    GPTB_TRY(pc_grow(h, &pc.lohi, &pc.cap[PC::B_LOHI], (size_t)2 * c.D));
    GPTB_TRY(pc_grow(h, &pc.lo, &pc.cap[PC::B_LO], (size_t)c.D));
    GPTB_TRY(pc_grow(h, &pc.invh, &pc.cap[PC::B_INVH], (size_t)c.D));
    GPTB_TRY(pc_grow(h, &pc.state, &pc.cap[PC::B_STATE], (size_t)4));
    if (!pc.lohi_host) GPTB_CUDA_OK(h, cudaMallocHost((void **)&pc.lohi_host, sizeof(double) * 64));
    GPTB_CUDA_OK(h, cudaMemcpy2DAsync(pc.X, ntpad * sizeof(double), Xtest, Nt * sizeof(double), Nt * sizeof(double), c.D,
                                      cudaMemcpyHostToDevice, st));
    col_minmax_kernel<<<c.D, 1024, 0, st>>>(pc.X, Nt, ntpad, pc.lohi, c.D);
    GPTB_CUDA_OK(h, cudaMemcpyAsync(pc.lohi_host, pc.lohi, sizeof(double) * 2 * c.D, cudaMemcpyDeviceToHost, st));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(st));
    h->launches++;
    int G = 1;
    for (int k = 0; k < c.D; ++k) {
        const double range = pc.lohi_host[c.D + k] - pc.lohi_host[k];
        if (!(range >= 0.0) || !std::isfinite(range)) return GPTB_ERR_UNSUPPORTED;   // NaN / Inf inputs: the direct path reports them
        G = std::max(G, (int)std::ceil(range * h->zmax[k] / (2.0 * h->node_theta)));
    }
    const int64_t m = (int64_t)G * P;
    if ((size_t)9 * G * 4 > 64 * 1024 || m > 65536) return GPTB_ERR_UNSUPPORTED;
    if (h->algorithm == GPTB_ALG_AUTO && Nt < 4 * m) return GPTB_ERR_UNSUPPORTED;
    const int mnpad = (int)((m + 127) / 128 * 128);
    {
        std::vector<double> lo(c.D), invh(c.D);
        std::vector<float> xn((size_t)c.D * mnpad);
        for (int k = 0; k < c.D; ++k) {
            double hk = (pc.lohi_host[c.D + k] - pc.lohi_host[k]) / (2.0 * G);
            if (!(hk > 0.0)) hk = 0.5 / std::max(h->zmax[k], 1e-30);
            lo[k] = pc.lohi_host[k];
            invh[k] = 1.0 / hk;
            for (int cc = 0; cc < G; ++cc)
                for (int j = 0; j < P; ++j)
                    xn[(size_t)k * mnpad + cc * P + j] = (float)(lo[k] + (2.0 * cc + 1.0 + std::cos(M_PI * (j + 0.5) / P)) * hk);
            for (int i = (int)m; i < mnpad; ++i) xn[(size_t)k * mnpad + i] = (float)lo[k];
        }
        GPTB_TRY(pc_grow(h, &pc.XnF, &pc.cap[PC::B_XNF], xn.size()));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(pc.XnF, xn.data(), sizeof(float) * xn.size(), cudaMemcpyHostToDevice, st));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(pc.lo, lo.data(), sizeof(double) * c.D, cudaMemcpyHostToDevice, st));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(pc.invh, invh.data(), sizeof(double) * c.D, cudaMemcpyHostToDevice, st));
        GPTB_CUDA_OK(h, cudaStreamSynchronize(st));   // (the vectors go out of scope)
    }
    GPTB_TRY(gptb_nodes_upload_cheb(h, P));

    // ---- cells and cell-sorted records of the test rows
    GPTB_TRY(pc_grow(h, &pc.cell, &pc.cap[PC::B_CELL], (size_t)c.D * ntpad));
    GPTB_TRY(pc_grow(h, &pc.uu, &pc.cap[PC::B_UU], (size_t)c.D * ntpad));
    GPTB_TRY(pc_grow(h, &pc.srow, &pc.cap[PC::B_SROW], (size_t)c.D * ntpad));
    GPTB_TRY(pc_grow(h, &pc.sc, &pc.cap[PC::B_SC], (size_t)c.D * ntpad));
    GPTB_TRY(pc_grow(h, &pc.su, &pc.cap[PC::B_SU], (size_t)c.D * ntpad));
    pred_cells_kernel<<<dim3((unsigned)std::min<int64_t>((Nt + 255) / 256, 8 * h->sm_count), c.D), 256, 0, st>>>(
        pc.X, Nt, ntpad, pc.lo, pc.invh, G, pc.cell, pc.uu);
    h->launches++;
    GPTB_TRY(gptb_nodes_launch_sort(h, pc.cell, pc.uu, pc.srow, pc.sc, pc.su, G, Nt, ntpad, st));

    // ---- geometry of the sample groups
    const size_t budget = 227 * 1024 - 2048;              // widest operand image that stays resident in K1's shared memory
    int nc_max = (int)(((budget / h->npad) - sizeof(float2)) / 4) / 16 * 16;
    nc_max = std::min(nc_max, 128);
    if (nc_max < h->rpad) return GPTB_ERR_UNSUPPORTED;
    const int Sg1 = std::max(1, std::min<int>(nc_max / c.r, S));          // samples per K1 launch
    const int NC1 = std::max(h->rpad, (Sg1 * c.r + 15) / 16 * 16);
    // Column layout of a sample group.  lanes (default): up to 32 samples side by side on the fastest axis, column
    // l*SL + s — the contraction then needs no transposed tile (contract_fwd3).  tiles: the samples' r columns one after
    // the other, 160/r samples per group, one 64-row tile per (tile, sample) CTA (contract_fwd2).
    // (the lanes contraction keeps one D*r*128-byte tile per warp beside its tables: very wide models use the tiles)
    const bool lanes = h->pred_layout == 0 && (size_t)c.Q * ((c.D + 3) / 4) * 16 + (size_t)c.Q * 128 + (size_t)c.D * c.r * 128 <= 220 * 1024;
    const int CGs = lanes ? std::min<int>(32, S) : std::max(1, std::min<int>(max_cols / c.r, S));   // samples per group
    auto sl_for = [](int gs) { return gs <= 8 ? 8 : (gs <= 16 ? 16 : 32); };
    const int ctot_max = lanes ? c.r * sl_for(CGs) : ((CGs * c.r + 3) & ~3);
    // two half-buffers of GPTB_OPT_PRED_CHUNK_MB / 2 each: the interpolation of chunk p+1 (HBM-write bound) runs on the
    // main stream beside the contraction of chunk p (shared-memory bound) on a second one
    // (measured at C5, 1M x 200: 150 ms with 256 MB, 130 ms with 1 GB, 123 ms with 2 GB, 117 ms with 4 GB of T2; without
    // the second stream 130 ms at 1 GB, 124 ms at 4 GB.  The budget is capped at an eighth of the free device memory.)
    const bool overlap = lanes;
    const int nbuf = overlap ? 2 : 1;
    size_t chunk_bytes = (size_t)h->pred_chunk_mb << 20;
    {
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) chunk_bytes = std::min(chunk_bytes, std::max<size_t>(free_b / 8, (size_t)64 << 20));
    }
    int64_t RC = (int64_t)(chunk_bytes / nbuf / ((size_t)c.D * ctot_max * 4)) / PRED_CHUNK_ALIGN * PRED_CHUNK_ALIGN;
    RC = std::max<int64_t>(PRED_CHUNK_ALIGN, std::min<int64_t>(RC, ntpad));
    if (overlap && !pc.st2) {
        GPTB_CUDA_OK(h, cudaStreamCreateWithFlags(&pc.st2, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&pc.ev_i[i], cudaEventDisableTiming));
            GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&pc.ev_c[i], cudaEventDisableTiming));
        }
        GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&pc.ev_g, cudaEventDisableTiming));
    }
    GPTB_TRY(pc_grow(h, &pc.Tn, &pc.cap[PC::B_TN], (size_t)c.D * Sg1 * c.r * mnpad));
    GPTB_TRY(pc_grow(h, &pc.Cf, &pc.cap[PC::B_CF], (size_t)c.D * G * ctot_max * PP));
    GPTB_TRY(pc_grow(h, &pc.T2, &pc.cap[PC::B_T2], (size_t)nbuf * c.D * RC * ctot_max));
    GPTB_TRY(pc_grow(h, &pc.img, &pc.cap[PC::B_IMG], (size_t)c.D * h->npad * 2 * NC1 * 2));
    GPTB_TRY(pc_grow(h, &pc.Ud, &pc.cap[PC::B_UD], (size_t)CGs * ucnt));
    GPTB_TRY(pc_grow(h, &pc.wd, &pc.cap[PC::B_WD], (size_t)S * c.Q));
    if (mean_out) {
        GPTB_TRY(pc_grow(h, &pc.acc, &pc.cap[PC::B_ACC], (size_t)ntpad));
        GPTB_CUDA_OK(h, cudaMemsetAsync(pc.acc, 0, sizeof(double) * ntpad, st));
    }
    if (per_sample_out) GPTB_TRY(pc_grow(h, &pc.fh, &pc.cap[PC::B_FH], (size_t)CGs * ntpad));
    GPTB_CUDA_OK(h, cudaMemsetAsync(pc.Cf, 0, sizeof(float) * (size_t)c.D * G * ctot_max * PP, st));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(pc.wd, w_store, sizeof(double) * S * c.Q, cudaMemcpyHostToDevice, st));

    // ---- sample groups outside, row chunks inside, everything on the handle's stream.  (Running the interpolation of
    // chunk p+1 on a second stream beside the contraction of chunk p was measured SLOWER, 215 vs 198 ms at C5: both
    // kernels queue on the LSU / L1 path — ncu: mio_throttle and long_scoreboard lead the contraction's stalls.)
    int64_t chunk_no = 0;
    for (int g0 = 0; g0 < S; g0 += CGs) {
        const int gs = std::min(CGs, S - g0);
        const int SL = sl_for(gs);
        const int ctot = lanes ? c.r * SL : ((gs * c.r + 3) & ~3);
        GPTB_CUDA_OK(h, cudaMemcpyAsync(pc.Ud, U_store + (size_t)g0 * ucnt, sizeof(double) * gs * ucnt, cudaMemcpyHostToDevice, st));
        if (lanes && gs < SL)   // unused sample slots of a ragged group: zero coefficients, so their T2 columns are zero
            GPTB_CUDA_OK(h, cudaMemsetAsync(pc.Cf, 0, sizeof(float) * (size_t)c.D * G * ctot * PP, st));
        for (int s0 = 0; s0 < gs; s0 += Sg1) {
            const int sg = std::min(Sg1, gs - s0);
            GPTB_TRY(gptb_tc_pack_group_scaled(h, pc.Ud + (size_t)s0 * ucnt, sg, NC1, pc.img, pc.state, st));
            GPTB_TRY(gptb_launch_fwd_tc_cols(h, pc.XnF, pc.img, pc.state + 1, m, mnpad, pc.Tn, sg * c.r, NC1));
            if (lanes) GPTB_TRY(gptb_nodes_launch_coef_lanes(h, pc.Tn, pc.Cf, G, P, sg, sg * c.r, SL, s0, mnpad, st));
            else GPTB_TRY(gptb_nodes_launch_coef_cols(h, pc.Tn, pc.Cf, G, P, sg * c.r, sg * c.r, ctot, s0 * c.r, mnpad, st));
        }
        cudaStream_t cs = overlap ? pc.st2 : st;           // the contractions' stream
        for (int64_t row0 = 0; row0 < Nt; row0 += RC, ++chunk_no) {
            const int64_t nr = std::min(RC, Nt - row0);
            const int b = overlap ? (int)(chunk_no & 1) : 0;
            float *T2b = pc.T2 + (size_t)b * c.D * RC * ctot_max;
            if (overlap && chunk_no >= 2) GPTB_CUDA_OK(h, cudaStreamWaitEvent(st, pc.ev_c[b], 0));   // half-buffer free again
            GPTB_TRY(gptb_nodes_launch_interp(h, P, pc.Cf, pc.srow, pc.sc, pc.su, T2b, ctot, 0, ctot, G, row0, nr, ntpad, RC, st));
            if (overlap) {
                GPTB_CUDA_OK(h, cudaEventRecord(pc.ev_i[b], st));
                GPTB_CUDA_OK(h, cudaStreamWaitEvent(cs, pc.ev_i[b], 0));
            }
            if (lanes)
                GPTB_TRY(gptb_launch_contract_fwd3(h, T2b, pc.wd + (size_t)g0 * c.Q, gs, SL, nr, RC,
                                                   per_sample_out ? pc.fh + row0 : nullptr, ntpad, mean_out ? pc.acc + row0 : nullptr,
                                                   1.0 / S, cs));
            else
                GPTB_TRY(gptb_launch_contract_fwd2(h, T2b, pc.wd + (size_t)g0 * c.Q, gs, 0, ctot, nr, RC,
                                                   per_sample_out ? pc.fh + row0 : nullptr, ntpad, mean_out ? pc.acc + row0 : nullptr,
                                                   1.0 / S, cs));
            if (overlap) GPTB_CUDA_OK(h, cudaEventRecord(pc.ev_c[b], cs));
        }
        if (per_sample_out)   // (the next group's contractions overwrite fh: they follow this copy on the same stream)
            GPTB_CUDA_OK(h, cudaMemcpy2DAsync(per_sample_out + (size_t)g0 * Nt, Nt * sizeof(double), pc.fh, ntpad * sizeof(double),
                                              Nt * sizeof(double), gs, cudaMemcpyDeviceToHost, cs));
    }
    if (overlap) {   // the main stream continues after the last contraction
        GPTB_CUDA_OK(h, cudaEventRecord(pc.ev_g, pc.st2));
        GPTB_CUDA_OK(h, cudaStreamWaitEvent(st, pc.ev_g, 0));
    }
    if (mean_out) GPTB_CUDA_OK(h, cudaMemcpyAsync(mean_out, pc.acc, sizeof(double) * Nt, cudaMemcpyDeviceToHost, st));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(st));
    return GPTB_OK;
}
