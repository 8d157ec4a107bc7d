// C-ABI implementation (include/gptb200.h): handle lifetime, host<->device staging and the
// orchestration of one SGLD step  gather -> K1 fwd -> K2 contract -> K3 bwd -> all-reduce -> K5 update.
#include <math.h>
#include <stdarg.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"
#include "tc_layout.cuh"

// defined in batch.cu / update.cu
int gptb_launch_gather_from(gptb_handle *h, const double *Xrm, const double *y, int64_t bs, int64_t mpad, double *Xb,
                            float *XbF, double *yb);
int gptb_launch_transpose(gptb_handle *h, const double *Xcm, int64_t N, double *Xrm);
int gptb_reset_order(gptb_handle *h);
int gptb_apply_perm(gptb_handle *h, const int64_t *perm_dev);
size_t gptb_update_small_stride(int r);

static thread_local std::string g_last_error;

void gptb_set_error(const gptb_handle *h, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    g_last_error = buf;
}

namespace {

template <typename T>
int dev_alloc(gptb_handle *h, T **p, size_t count) {
    if (*p) {
        cudaFree(*p);
        *p = nullptr;
    }
    if (count == 0) count = 1;
    cudaError_t e = cudaMalloc((void **)p, count * sizeof(T));
    if (e != cudaSuccess) {
        gptb_set_error(h, "cudaMalloc(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
        *p = nullptr;
        return e == cudaErrorMemoryAllocation ? GPTB_ERR_NOMEM : GPTB_ERR_CUDA;
    }
    return GPTB_OK;
}

template <typename T>
void dev_free(T **p) {
    if (*p) cudaFree(*p);
    *p = nullptr;
}

__global__ void d2f_kernel(const double *a, float *b, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        b[i] = (float)a[i];
}
__global__ void axpy_kernel(const double *x, double *y, double a, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        y[i] += a * x[i];
}
__global__ void gather_y_kernel(const double *y, const int32_t *rows, int64_t row_base, int64_t bs, int64_t mpad,
                                double *yb) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < mpad; i += (int64_t)gridDim.x * blockDim.x)
        yb[i] = i < bs ? y[rows ? (int64_t)rows[i] : row_base + i] : 0.0;
}
__global__ void idx64_to_32_kernel(const int64_t *a, int32_t *b, int64_t n, int64_t N, int32_t *bad) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t v = a[i];
        if (v < 0 || v >= N) {
            *bad = 1;
            v = 0;
        }
        b[i] = (int32_t)v;
    }
}

int check_ready(gptb_handle *h, bool need_train, bool need_params) {
    if (!h->have_features && !h->have_phi) {
        gptb_set_error(h, "features not set (gptb_set_features or gptb_set_train_phi)");
        return GPTB_ERR_STATE;
    }
    if (!h->have_index) {
        gptb_set_error(h, "core index table not set (gptb_set_index)");
        return GPTB_ERR_STATE;
    }
    if (need_train && !h->have_train && !h->have_phi) {
        gptb_set_error(h, "training set not set (gptb_set_train / gptb_set_train_phi)");
        return GPTB_ERR_STATE;
    }
    if (need_params && !h->have_params) {
        gptb_set_error(h, "parameters not set (gptb_set_params / gptb_init_params)");
        return GPTB_ERR_STATE;
    }
    return GPTB_OK;
}

// The dense-phi compatibility path always runs the FP64 kernels (it is HBM-bound on 8*n*D bytes/row).
inline bool use_fp64(const gptb_handle *h, bool phi_mode) { return phi_mode || h->cfg.precision == GPTB_PREC_FP64; }

int ensure_batch(gptb_handle *h, int64_t bs) {
    if (bs <= h->cap && h->cap > 0) return GPTB_OK;
    const gptb_config &c = h->cfg;
    int64_t cap = std::max<int64_t>(bs, 128);
    int64_t mpad = (cap + 127) / 128 * 128;
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    GPTB_TRY(dev_alloc(h, &h->Xb, (size_t)c.D * mpad));
    GPTB_TRY(dev_alloc(h, &h->XbF, (size_t)c.D * mpad));
    GPTB_TRY(dev_alloc(h, &h->yb, (size_t)mpad));
    GPTB_TRY(dev_alloc(h, &h->fhat, (size_t)mpad));
    GPTB_TRY(dev_alloc(h, (double **)&h->T, (size_t)c.D * c.r * mpad));
    GPTB_TRY(dev_alloc(h, (double **)&h->E, (size_t)c.D * c.r * mpad));
    GPTB_TRY(dev_alloc(h, &h->idx_dev, (size_t)mpad));
    if (gptb_is_tc(h)) {
        size_t eb = gptb_tc_eimg_bytes(h, mpad);
        GPTB_TRY(dev_alloc(h, (unsigned char **)&h->Eimg, eb));
        GPTB_CUDA_OK(h, cudaMemsetAsync(h->Eimg, 0, eb, h->stream));
    }
    h->cap = cap;
    h->mpad = mpad;
    return GPTB_OK;
}

int ensure_pin(gptb_handle *h, size_t bytes) {
    if (bytes <= h->pin_bytes) return GPTB_OK;
    if (h->pin) cudaFreeHost(h->pin);
    h->pin = nullptr;
    h->pin_bytes = 0;
    GPTB_CUDA_OK(h, cudaMallocHost((void **)&h->pin, bytes));
    h->pin_bytes = bytes;
    return GPTB_OK;
}

// Per-phase timing without host synchronisation: events go into a pool as the step is enqueued and
// are read back (after one stream sync) by drain_timers().
static void record_ev(gptb_handle *h, int tag) {
    if (h->ev_used == h->evpool.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        h->evpool.push_back(e);
        h->evtag.push_back(tag);
    }
    h->evtag[h->ev_used] = tag;
    cudaEventRecord(h->evpool[h->ev_used], h->stream);
    h->ev_used++;
}

static void drain_timers(gptb_handle *h) {
    if (h->ev_used == 0) return;
    cudaStreamSynchronize(h->stream);
    for (size_t i = 1; i < h->ev_used; ++i) {
        if (h->evtag[i] < 0) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, h->evpool[i - 1], h->evpool[i]) == cudaSuccess) h->phase_ms[h->evtag[i]] += ms;
    }
    h->ev_used = 0;
}

struct PhaseTimer {
    gptb_handle *h;
    explicit PhaseTimer(gptb_handle *hh) : h(hh) {
        if (h->profiling) {
            if (h->ev_used > 60000) drain_timers(h);
            record_ev(h, -1);
        }
    }
    void done(int phase) {
        if (h->profiling) record_ev(h, phase);
    }
    void finish() {}
};

// Pack / refresh the tensor-core shadows of the sampler's current parameters.
int refresh_shadows(gptb_handle *h) {
    if (!gptb_is_tc(h) || h->shadows_valid) return GPTB_OK;
    GPTB_TRY(gptb_tc_pack_params(h));
    h->shadows_valid = true;
    return GPTB_OK;
}

// Forward for the rows staged in Xb/XbF (or phi rows): T, then fhat (and E / gradw when want_grad).
int forward_contract(gptb_handle *h, bool phi_mode, const int32_t *rows, const double *w_dev, const double *U_dev,
                     const void *Uimg, int64_t bs, double escale, bool want_grad, PhaseTimer *pt, bool fwd_done = false) {
    const bool f64 = use_fp64(h, phi_mode);
    if (fwd_done) {
        // the pipelined step has launched the forward pass group by group
    } else if (f64) {
        GPTB_TRY(gptb_launch_fwd_fp64(h, h->Xb, phi_mode ? h->phi : nullptr, rows, U_dev, bs, h->mpad, (double *)h->T));
    } else {
        GPTB_TRY(gptb_launch_fwd_tc(h, h->XbF, Uimg, bs, h->mpad, (float *)h->T));
    }
    if (pt) pt->done(GPTB_PHASE_FWD);
    // E-image scale: derived from the previous step's max|E| (scale 1 right after new parameters), then checked
    // against this step's own max|E|; a minibatch whose residuals jumped out of the fp16 window triggers one
    // conditional re-run of K2 that only rewrites the image (tc_path.cu: check_escale_kernel)
    const bool img = !f64 && want_grad && !h->k2_Erow;
    if (img) {
        if (!h->tc_emax_valid) GPTB_CUDA_OK(h, cudaMemsetAsync(h->tc_state, 0, sizeof(float), h->stream));
        h->tc_emax_valid = true;
        GPTB_TRY(gptb_tc_update_escale(h));
    }
    GPTB_TRY(gptb_launch_contract(h, h->T, w_dev, h->yb, bs, h->mpad, escale, want_grad, h->fhat, h->grad,
                                  f64 ? h->E : nullptr, f64 ? nullptr : h->Eimg, f64));
    if (img) {
        GPTB_TRY(gptb_tc_check_escale(h));
        h->k2_redo = 1;
        const int st = gptb_launch_contract(h, h->T, w_dev, h->yb, bs, h->mpad, escale, want_grad, h->fhat, h->grad, nullptr,
                                            h->Eimg, false);
        h->k2_redo = 0;
        GPTB_TRY(st);
    }
    if (h->rms_active && want_grad) {
        // GPT_SGLDERM_RMSprop: w moves as soon as its gradient is complete (:1172-1188), and computeA then sees the NEW w
        // (:1194) while fhat and the residual stay those of the old w — a second K2 launch that only rewrites E
        GPTB_TRY(gptb_allreduce_range(h, h->grad, (size_t)h->cfg.Q, h->stream));
        GPTB_TRY(gptb_launch_rms_w(h, h->rms_flags, h->rms_noise_w, h->step_counter, h->stream));
        h->k2_redo = 2;
        h->k2_wA = h->w;
        const int st = gptb_launch_contract(h, h->T, h->w_old, h->yb, bs, h->mpad, escale, true, h->fhat, h->grad,
                                            f64 ? h->E : nullptr, f64 ? nullptr : h->Eimg, f64);
        h->k2_redo = 0;
        h->k2_wA = nullptr;
        GPTB_TRY(st);
    }
    if (pt) pt->done(GPTB_PHASE_CONTRACT);
    return GPTB_OK;
}

static int pipe_init(gptb_handle *h, int G);

// Node-grid step (nodes.cu): when the features can be generated at the nodes of a grid instead of at every row.
// AUTO uses it when the rank's share has many more rows than the grid has nodes (K1 + K3 over the rows cost ~6.8 ns per
// row at BASELINE config 4, the interpolation and spreading kernels ~0.7 ns per row plus ~45 us of fixed cost).
static bool nodes_usable(gptb_handle *h, int64_t lbs, bool phi_mode) {
    if (h->algorithm == GPTB_ALG_DIRECT || use_fp64(h, phi_mode) || !h->have_domain || !gptb_contract_tc_usable(h)) return false;
    if (!h->grid_valid && gptb_nodes_setup(h) != GPTB_OK) return false;
    if (h->nodes.m == 0) return false;
    if (h->algorithm == GPTB_ALG_NODES) return true;
    return lbs >= std::max<int64_t>(8192, 4 * (int64_t)h->nodes.m);
}

// Front of a node-grid step on the side stream (see common.cuh: front_stream).  dep: an event the rows wait for (the
// staged host minibatch), or null — then h->front_dep_mode says whether the front follows everything queued on the main
// stream so far (0) or nothing but the previous readers of the batch buffers (1: resident rows inside an epoch).
static int front_begin(gptb_handle *h, int64_t lbs, bool phi_mode, cudaEvent_t dep, bool *on_side) {
    *on_side = h->front_active;
    if (h->front_active) return GPTB_OK;
    if (h->profiling || !h->front_enabled || lbs <= 0 || !nodes_usable(h, lbs, phi_mode)) return GPTB_OK;
    if (!h->front_stream) {
        GPTB_CUDA_OK(h, cudaStreamCreateWithFlags(&h->front_stream, cudaStreamNonBlocking));
        GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->ev_front_dep, cudaEventDisableTiming));
        GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->ev_front_done, cudaEventDisableTiming));
        GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->ev_rows_free, cudaEventDisableTiming));
    }
    if (dep) GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->front_stream, dep, 0));
    if ((!dep && h->front_dep_mode == 0) || !h->rows_free_valid) {
        // (also the first front ever: the earlier steps' readers of the batch buffers have recorded nothing yet)
        GPTB_CUDA_OK(h, cudaEventRecord(h->ev_front_dep, h->stream));
        GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->front_stream, h->ev_front_dep, 0));
    } else {
        GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->front_stream, h->ev_rows_free, 0));
    }
    h->front_active = true;
    *on_side = true;
    return GPTB_OK;
}

// Gradient of the rank's share of the minibatch into h->grad (likelihood part only, all-reduced).
// rows_dev: device row ids of the *global* batch or null (= order[start..] / identity when staged).
int compute_grad(gptb_handle *h, double signal_var, const int32_t *rows_dev, int64_t bs, int64_t N_total,
                 bool staged_host_batch, const double *Xstage_rm, const double *ystage, PhaseTimer &pt,
                 bool skip_bwd = false, int stage = 0) {
    // stage 0: the whole gradient.  The classification driver (classify.cu) splits it around its softmax:
    // stage 1 = gather, forward pass and fhat only (T, the batch inputs and the cell-sorted records stay in place),
    // stage 2 = K2 with the residual of the targets now in yb, back-projection, reduction.
    const bool do_front = stage != 2, do_back = stage != 1;
    const gptb_config &c = h->cfg;
    const bool phi_mode = h->have_phi && !staged_host_batch;
    const int64_t lo = (int64_t)h->rank * bs / h->nranks, hi = (int64_t)(h->rank + 1) * bs / h->nranks;
    const int64_t lbs = hi - lo;
    GPTB_TRY(ensure_batch(h, std::max<int64_t>(lbs, 1)));
    GPTB_TRY(refresh_shadows(h));
    const size_t gcount = (size_t)c.Q + (size_t)c.n * c.r * c.D;
    if (h->f32_grad) GPTB_TRY(gptb_tc_ensure_gradF(h));   // FP32 gradient flow: the buffer is cleared by its reader (k_prepare)
    else if (do_back) GPTB_CUDA_OK(h, cudaMemsetAsync(h->grad, 0, gcount * sizeof(double), h->stream));
    const bool f64 = use_fp64(h, phi_mode);
    const bool nodes = lbs > 0 && nodes_usable(h, lbs, phi_mode);
    if (nodes) GPTB_TRY(gptb_nodes_ensure_batch(h));
    const int32_t *rows = rows_dev ? rows_dev + lo : nullptr;
    struct CellGuard {
        gptb_handle *h;
        ~CellGuard() { h->gather_cells = false; h->k2_Erow = nullptr; h->k2_T_rowmajor = false; }
    } cell_guard{h};
    h->gather_cells = nodes;
    bool side_front = false;
    if (do_front && stage == 0) GPTB_TRY(front_begin(h, lbs, phi_mode, nullptr, &side_front));
    struct StreamSwap {   // the front's launchers use h->stream
        gptb_handle *h;
        cudaStream_t main;
        bool on;
        ~StreamSwap() { if (on) { h->stream = main; h->front_active = false; } }
    } swap{h, h->stream, side_front};
    if (side_front) h->stream = h->front_stream;
    if (!do_front) {
        // batch inputs, yb and the sorted records are those of stage 1
    } else if (staged_host_batch) {
        const int64_t so = h->stage_share ? 0 : lo;   // (a share-only staging buffer starts at this rank's first row)
        GPTB_TRY(gptb_launch_gather_from(h, Xstage_rm + (size_t)so * c.D, ystage + so, lbs, h->mpad, f64 ? h->Xb : nullptr,
                                         f64 ? nullptr : h->XbF, h->yb));
    } else if (phi_mode) {
        int blocks = (int)std::min<int64_t>((h->mpad + 255) / 256, 1024);
        gather_y_kernel<<<blocks, 256, 0, h->stream>>>(h->y, rows, 0, lbs, h->mpad, h->yb);
        h->launches++;
    } else {
        GPTB_TRY(gptb_launch_gather(h, rows, 0, lbs, lo));
    }
    h->gather_cells = false;
    if (nodes && do_front) GPTB_TRY(gptb_nodes_sort(h, lbs, h->stream));   // cell-sorted records: both row kernels walk them
    if (side_front) {
        h->stream = swap.main;
        swap.on = false;
        h->front_active = false;
        GPTB_CUDA_OK(h, cudaEventRecord(h->ev_front_done, h->front_stream));
        GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->stream, h->ev_front_done, 0));
    }
    pt.done(GPTB_PHASE_GATHER);
    const double escale = ((double)N_total / (double)bs) / signal_var;  // (N/bs)/signal_var of :393,:407
    if (lbs > 0) {
        double *gU = h->grad + c.Q;
        if (nodes) {
            if (do_front) {
                GPTB_TRY(gptb_nodes_forward_range(h, h->Uimg, 0, c.D, h->stream));
                GPTB_TRY(gptb_nodes_interp(h, lbs));
            }
            h->k2_Erow = (skip_bwd || !do_back) ? nullptr : h->nodes.Erow;   // K2 writes plain E rows instead of the operand image
            h->k2_T_rowmajor = true;                           // ... and reads the interpolated T, row-major per sample
            {
                void *T_keep = h->T;
                h->T = h->nodes.T2;
                const int st = forward_contract(h, false, rows, h->w, h->U, h->Uimg, lbs, escale, do_back, &pt, true);
                h->T = T_keep;
                h->k2_Erow = nullptr;
                h->k2_T_rowmajor = false;
                GPTB_TRY(st);
            }
            if (!skip_bwd && do_back) {
                GPTB_TRY(gptb_nodes_spread(h, lbs));
                GPTB_TRY(gptb_nodes_backward_range(h, gU, 0, c.D));
            }
        } else {
            GPTB_TRY(forward_contract(h, phi_mode, rows, h->w, h->U, h->Uimg, lbs, escale, do_back, &pt, !do_front));
            if (skip_bwd || !do_back) {
                // GPTB_FLAG_FIX_U: the w-only sampler needs fhat and gradw only
            } else if (f64) {
                GPTB_TRY(gptb_launch_bwd_fp64(h, h->Xb, phi_mode ? h->phi : nullptr, rows, (const double *)h->E, lbs,
                                              h->mpad, gU));
            } else {
                GPTB_TRY(gptb_launch_bwd_tc(h, h->XbF, h->Eimg, lbs, h->mpad, gU));
            }
        }
    }
    if (h->front_stream) {   // the batch buffers (yb, XbF, cell records) have had their last reader
        GPTB_CUDA_OK(h, cudaEventRecord(h->ev_rows_free, h->stream));
        h->rows_free_valid = true;
    }
    pt.done(GPTB_PHASE_BWD);
    if (!do_back) return GPTB_OK;
    if (h->f32_grad)
        GPTB_TRY(gptb_allreduce_range_f32(h, h->gradFbase, (size_t)h->gradF_qpad + (size_t)c.D * h->npad * h->rpad, h->stream));
    else
        GPTB_TRY(gptb_allreduce_grad(h));
    pt.done(GPTB_PHASE_REDUCE);
    return GPTB_OK;
}


// ------------------------------------------------------------------------------------------------------------
// Dimension-group pipeline of the SGLD step (tensor-core path, Stiefel update).
//
// The update of U^(k) (K5: a latency-bound chain of small FP64 kernels that occupies a handful of SMs) depends only
// on the gradient of dimension k, and the next step's forward pass of dimension k only on the new U^(k).  The D
// dimensions are therefore split into G groups; on the main stream the step runs
//     gather, K1_0 .. K1_{G-1}, K2, K3_0 .. K3_{G-1}
// and the update of group g runs on its own side stream as soon as K3_g is done, i.e. under K3_{g+1}.. and under
// the next step's gather and K1_0 .. K1_{g-1}.  K1_g of the next step waits for the update of group g only.
// Results do not depend on G: every kernel addresses parameters, gradients and Philox streams by the global k.
// With several ranks each group's gradient slice is all-reduced on one dedicated NCCL stream.
// ------------------------------------------------------------------------------------------------------------
// G = 2 unless gptb_set_option(GPTB_OPT_PIPE_GROUPS) says otherwise (0 or 1 switches the pipeline off; setting it also
// forces the pipeline on for the small and multi-rank steps it is normally not used for — that is how the tests drive it).
static int pipe_groups_for(gptb_handle *h) {
    if (h->pipe_groups < 0) h->pipe_groups = 2;
    return h->pipe_groups;
}

// Streams and events are created on first use and only as many as needed: streams beyond the device's hardware
// queues (CUDA_DEVICE_MAX_CONNECTIONS, 8 by default) share a queue and serialise behind each other.
static int pipe_init(gptb_handle *h, int G) {
    for (int g = 0; g < G; ++g) {
        if (h->side[g]) continue;
        GPTB_CUDA_OK(h, cudaStreamCreateWithFlags(&h->side[g], cudaStreamNonBlocking));
        GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->evG[g], cudaEventDisableTiming));
        GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->evU[g], cudaEventDisableTiming));
        GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->evR[g], cudaEventDisableTiming));
    }
    if (h->nranks > 1 && !h->ar_stream) GPTB_CUDA_OK(h, cudaStreamCreateWithFlags(&h->ar_stream, cudaStreamNonBlocking));
    return GPTB_OK;
}
static int pipe_tail_init(gptb_handle *h) {
    if (h->tail_stream) return GPTB_OK;
    GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->evTail, cudaEventDisableTiming));
    GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->evTailFork, cudaEventDisableTiming));
    GPTB_CUDA_OK(h, cudaStreamCreateWithFlags(&h->tail_stream, cudaStreamNonBlocking));
    return GPTB_OK;
}

// Order the main stream after everything the pipeline still has in flight.  Every entry point that reads or
// writes parameters outside the step itself calls this first.
static int pipe_join_groups(gptb_handle *h) {
    for (int g = 0; g < gptb_handle::PIPE_MAX; ++g)
        if (h->pendU[g]) {
            GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->stream, h->evU[g], 0));
            h->pendU[g] = false;
        }
    return GPTB_OK;
}
// The read-backs of the previous step (tail stream) only have to finish before the parameters change again: the
// one-stream step waits for them right before its update, not at its start.
static int pipe_join_tail(gptb_handle *h) {
    if (h->tail_pending) {
        GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->stream, h->evTail, 0));
        h->tail_pending = false;
    }
    return GPTB_OK;
}
int pipe_join(gptb_handle *h) {
    GPTB_TRY(pipe_join_groups(h));
    return pipe_join_tail(h);
}

// Stream for work that must follow the whole step (copies of the updated parameters, the NaN flag) without
// holding the main stream back: the tail stream, ordered after every pending group update.
static int pipe_tail_begin(gptb_handle *h, cudaStream_t *out) {
    bool any = false;
    for (int g = 0; g < gptb_handle::PIPE_MAX; ++g) any = any || h->pendU[g];
    GPTB_TRY(pipe_tail_init(h));
    if (!any) {
        // one-stream step: the tail forks off the main stream here, so that two small device-to-host copies (~25 us with
        // their latencies) do not sit between this step's update and the next step's first kernel
        GPTB_CUDA_OK(h, cudaEventRecord(h->evTailFork, h->stream));
        GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->tail_stream, h->evTailFork, 0));
        *out = h->tail_stream;
        return GPTB_OK;
    }
    for (int g = 0; g < gptb_handle::PIPE_MAX; ++g)
        if (h->pendU[g]) GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->tail_stream, h->evU[g], 0));
    *out = h->tail_stream;
    return GPTB_OK;
}
static int pipe_tail_end(gptb_handle *h, cudaStream_t s) {
    if (!h->tail_stream || s != h->tail_stream) return GPTB_OK;
    GPTB_CUDA_OK(h, cudaEventRecord(h->evTail, s));
    h->tail_pending = true;
    return GPTB_OK;
}

static bool pipe_usable(gptb_handle *h, int64_t bs, bool staged_host_batch, int32_t flags) {
    const bool phi_mode = h->have_phi && !staged_host_batch;
    const int G = pipe_groups_for(h);
    if (flags & (GPTB_FLAG_FIX_U | GPTB_FLAG_RMSPROP)) return false;
    {   // the node-grid step has no row-proportional K3 to hide an update under: one stream
        const int64_t lbs = (int64_t)(h->rank + 1) * bs / h->nranks - (int64_t)h->rank * bs / h->nranks;
        if (lbs > 0 && nodes_usable(h, lbs, phi_mode)) return false;
    }
    if (!(G >= 2 && h->cfg.D >= G && !h->profiling && !use_fp64(h, phi_mode) && (flags & GPTB_FLAG_STIEFEL) &&
          bs >= h->nranks))
        return false;
    // Measured on B200 (C4, one rank): 668 -> 652 us per step with G = 2 and 5 SMs left to the update.  The update of a
    // group is slower on the few SMs it gets (its row-block kernels cannot share an SM with the register-hungry
    // persistent kernels), so it only hides under a back-projection of >= ~100 us: large single-rank batches.
    return h->pipe_forced || (h->nranks == 1 && bs >= 32768);
}

// One SGLD step, pipelined over dimension groups (see above).  Same arguments as step_common.
static int step_pipelined(gptb_handle *h, double signal_var, const int32_t *rows_dev, int64_t bs, int64_t N_total,
                          bool staged_host_batch, const double *Xstage_rm, const double *ystage, double epsw, double epsU,
                          int32_t flags, const double *nw_dev, const double *nU_dev) {
    const gptb_config &c = h->cfg;
    const int G = pipe_groups_for(h);
    GPTB_TRY(pipe_init(h, G));
    GPTB_TRY(gptb_tc_ensure_gradF(h));
    if ((flags & GPTB_FLAG_LANGEVIN) && (!nw_dev || !nU_dev)) GPTB_TRY(gptb_update_alloc_noise(h));
    const int64_t lo = (int64_t)h->rank * bs / h->nranks, hi = (int64_t)(h->rank + 1) * bs / h->nranks;
    const int64_t lbs = hi - lo;
    if (h->cap < std::max<int64_t>(lbs, 1) || !h->shadows_valid) GPTB_TRY(pipe_join(h));   // (re)allocation / repacking read U
    GPTB_TRY(ensure_batch(h, std::max<int64_t>(lbs, 1)));
    GPTB_TRY(refresh_shadows(h));
    const int32_t *rows = rows_dev ? rows_dev + lo : nullptr;
    if (staged_host_batch) {
        const int64_t so = h->stage_share ? 0 : lo;
        GPTB_TRY(gptb_launch_gather_from(h, Xstage_rm + (size_t)so * c.D, ystage + so, lbs, h->mpad, nullptr, h->XbF, h->yb));
    } else {
        GPTB_TRY(gptb_launch_gather(h, rows, 0, lbs, lo));
    }
    const double escale = ((double)N_total / (double)bs) / signal_var;  // (N/bs)/signal_var of :393,:407
    int k0s[gptb_handle::PIPE_MAX + 1];
    for (int g = 0; g <= G; ++g) k0s[g] = (int)((int64_t)c.D * g / G);
    // The expm kernel of the update takes one whole SM per dimension (200 KB of shared memory), and the forward /
    // backward kernels are persistent grids that would otherwise hold every SM until they end: they leave as many SMs
    // free as one group has dimensions.
    struct ReserveGuard {
        gptb_handle *h;
        ~ReserveGuard() { h->sm_reserve = 0; }
    } guard{h};
    {
        int rsv = h->pipe_reserve >= 0 ? h->pipe_reserve : (c.D + G - 1) / G;
        h->sm_reserve = std::max(0, std::min(rsv, h->sm_count / 4));
    }
    // forward, group by group: K1_g needs the previous step's update of group g only
    for (int g = 0; g < G; ++g) {
        if (h->pendU[g]) {
            GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->stream, h->evU[g], 0));
            h->pendU[g] = false;
        }
        GPTB_TRY(gptb_launch_fwd_tc_range(h, h->XbF, h->Uimg, lbs, h->mpad, (float *)h->T, k0s[g], k0s[g + 1] - k0s[g]));
    }
    // (group 0 carries w: from here on the previous step's w update is complete and gradw may be cleared)
    if (!h->f32_grad) GPTB_CUDA_OK(h, cudaMemsetAsync(h->grad, 0, sizeof(double) * c.Q, h->stream));
    GPTB_TRY(forward_contract(h, false, rows, h->w, h->U, h->Uimg, lbs, escale, true, nullptr, true));
    double *gU = h->grad + c.Q;
    const size_t per_k = (size_t)c.r * c.n;
    for (int g = 0; g < G; ++g) {
        const int k0 = k0s[g], kn = k0s[g + 1] - k0s[g];
        if (!h->f32_grad) GPTB_CUDA_OK(h, cudaMemsetAsync(gU + (size_t)k0 * per_k, 0, sizeof(double) * per_k * kn, h->stream));
        GPTB_TRY(gptb_launch_bwd_tc_range(h, h->XbF, h->Eimg, lbs, h->mpad, gU, k0, kn));
        GPTB_CUDA_OK(h, cudaEventRecord(h->evG[g], h->stream));
        cudaStream_t s = h->side[g];
        if (h->tail_pending) GPTB_CUDA_OK(h, cudaStreamWaitEvent(s, h->evTail, 0));   // last step's parameter copies
        if (h->nranks > 1) {
            GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->ar_stream, h->evG[g], 0));
            if (h->f32_grad) {
                const size_t per_kf = (size_t)h->npad * h->rpad;
                float *buf = g == 0 ? h->gradFbase : h->gradF + (size_t)k0 * per_kf;
                GPTB_TRY(gptb_allreduce_range_f32(h, buf, per_kf * kn + (g == 0 ? (size_t)h->gradF_qpad : 0), h->ar_stream));
            } else {
                double *buf = g == 0 ? h->grad : gU + (size_t)k0 * per_k;
                const size_t cnt = per_k * kn + (g == 0 ? (size_t)c.Q : 0);
                GPTB_TRY(gptb_allreduce_range(h, buf, cnt, h->ar_stream));
            }
            GPTB_CUDA_OK(h, cudaEventRecord(h->evR[g], h->ar_stream));
            GPTB_CUDA_OK(h, cudaStreamWaitEvent(s, h->evR[g], 0));
        } else {
            GPTB_CUDA_OK(h, cudaStreamWaitEvent(s, h->evG[g], 0));
        }
        GPTB_TRY(gptb_launch_update_range(h, epsw, epsU, flags, nw_dev, nU_dev, h->step_counter, k0, kn, g == 0, s));
        GPTB_CUDA_OK(h, cudaEventRecord(h->evU[g], s));
        h->pendU[g] = true;
    }
    h->step_counter++;
    return GPTB_OK;
}

// Resolve batch rows on the device: explicit idx (host int64) or a range of the current order.
int stage_rows(gptb_handle *h, const int64_t *idx, int64_t start, int64_t bs, const int32_t **rows_out) {
    if (bs <= 0 || (idx == nullptr && (start < 0 || start + bs > h->N))) {
        gptb_set_error(h, "batch range [%lld, %lld) outside 0..N=%lld", (long long)start, (long long)(start + bs),
                       (long long)h->N);
        return GPTB_ERR_ARG;
    }
    if (!idx) {
        *rows_out = h->order + start;
        return GPTB_OK;
    }
    GPTB_TRY(ensure_batch(h, bs));
    int64_t *tmp = nullptr;
    GPTB_CUDA_OK(h, cudaMallocAsync((void **)&tmp, sizeof(int64_t) * bs, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(tmp, idx, sizeof(int64_t) * bs, cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemsetAsync(h->nan_flag + 1, 0, sizeof(int32_t), h->stream));
    idx64_to_32_kernel<<<(int)std::min<int64_t>((bs + 255) / 256, 1024), 256, 0, h->stream>>>(tmp, h->idx_dev, bs, h->N,
                                                                                              h->nan_flag + 1);
    h->launches++;
    GPTB_CUDA_OK(h, cudaFreeAsync(tmp, h->stream));
    int32_t bad = 0;
    GPTB_CUDA_OK(h, cudaMemcpyAsync(&bad, h->nan_flag + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    if (bad) {
        gptb_set_error(h, "batch index outside 0..N-1");
        return GPTB_ERR_ARG;
    }
    *rows_out = h->idx_dev;
    return GPTB_OK;
}

int stage_noise(gptb_handle *h, const double *noise_w, const double *noise_U, const double **nw_dev,
                const double **nU_dev) {
    const gptb_config &c = h->cfg;
    *nw_dev = nullptr;
    *nU_dev = nullptr;
    if (noise_w || noise_U) GPTB_TRY(pipe_join(h));   // the previous step's update may still be reading the staged draws
    if (noise_w) {
        if (!h->noise_w) GPTB_TRY(dev_alloc(h, &h->noise_w, (size_t)c.Q));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(h->noise_w, noise_w, sizeof(double) * c.Q, cudaMemcpyHostToDevice, h->stream));
        *nw_dev = h->noise_w;
    }
    if (noise_U) {
        size_t cnt = (size_t)c.n * c.r * c.D;
        if (!h->noise_U) GPTB_TRY(dev_alloc(h, &h->noise_U, cnt));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(h->noise_U, noise_U, sizeof(double) * cnt, cudaMemcpyHostToDevice, h->stream));
        *nU_dev = h->noise_U;
    }
    return GPTB_OK;
}

int check_nan(gptb_handle *h) {
    GPTB_TRY(pipe_join(h));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->nan_flag_host, h->nan_flag, 8 * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    if (h->nan_flag_host[3]) {   // raised by apply_perm_kernel (gptb_sgld_run with host permutations)
        GPTB_CUDA_OK(h, cudaMemsetAsync(h->nan_flag + 3, 0, sizeof(int32_t), h->stream));
        gptb_set_error(h, "permutation entry outside 0..N-1");
        return GPTB_ERR_ARG;
    }
    if (h->nan_flag_host[4]) {   // raised by the gather of a node-grid step
        GPTB_CUDA_OK(h, cudaMemsetAsync(h->nan_flag + 4, 0, sizeof(int32_t), h->stream));
        gptb_set_error(h, "a minibatch row lies outside the input domain of the node grid (gptb_set_domain / gptb_set_train)");
        return GPTB_ERR_ARG;
    }
    if (*h->nan_flag_host) {
        // GPT_SGLD.jl:25
        gptb_set_error(h, "Get NaN when moving along Geodesic. Try smaller epsU");
        return GPTB_GEODESIC_NAN;
    }
    return GPTB_OK;
}

}  // namespace

// ---- hooks for the classification driver (classify.cu): the split gradient pass, noise staging, flag read-back
int gptb_i_grad_stage(gptb_handle *h, const int32_t *rows_dev, int64_t bs, int64_t N_total, int stage) {
    GPTB_TRY(pipe_join(h));
    PhaseTimer pt(h);
    h->f32_grad = false;   // the gradient is read by two updates: FP64 buffer, not cleared by its reader
    return compute_grad(h, 1.0, rows_dev, bs, N_total, false, nullptr, nullptr, pt, false, stage);
}
int gptb_i_stage_noise(gptb_handle *h, const double *noise_w, const double *noise_U, const double **nw_dev, const double **nU_dev) {
    return stage_noise(h, noise_w, noise_U, nw_dev, nU_dev);
}
int gptb_i_check_nan(gptb_handle *h) { return check_nan(h); }
int gptb_i_check_ready(gptb_handle *h) { return check_ready(h, true, true); }

extern "C" {

int gptb_version(void) { return GPTB_VERSION; }

const char *gptb_last_error(const gptb_handle *h) { return h ? h->err.c_str() : g_last_error.c_str(); }

int gptb_create(const gptb_config *cfg, gptb_handle **out) {
    if (!cfg || !out) {
        gptb_set_error(nullptr, "gptb_create: NULL argument");
        return GPTB_ERR_ARG;
    }
    *out = nullptr;
    if (cfg->D < 1 || cfg->D > 16 || cfg->n < 1 || cfg->r < 1 || cfg->r > 64 || cfg->Q < 1 || cfg->Q > 65535) {
        gptb_set_error(nullptr, "gptb_create: need 1<=D<=16, n>=1, 1<=r<=64, Q>=1 (got D=%d n=%d r=%d Q=%d)", cfg->D,
                       cfg->n, cfg->r, cfg->Q);
        return GPTB_ERR_SHAPE;
    }
    if (cfg->precision < GPTB_PREC_FP64 || cfg->precision > GPTB_PREC_TC1 || cfg->feature_kind < 0 ||
        cfg->feature_kind > 1) {
        gptb_set_error(nullptr, "gptb_create: unknown precision / feature kind");
        return GPTB_ERR_ARG;
    }
    if (cfg->feature_kind == GPTB_FEAT_SINCOS && (cfg->n % 2)) {
        gptb_set_error(nullptr, "n not even");  // GPT_SGLD.jl:90
        return GPTB_ERR_SHAPE;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        gptb_set_error(nullptr, "no CUDA device: %s (libgptb200 has no CPU fallback)", cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    gptb_handle *h = new gptb_handle();
    h->cfg = *cfg;
    int dev = cfg->device;
    if (dev < 0) cudaGetDevice(&dev);
    h->device = dev;
#define CREATE_OK(expr)                                                                   \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess) {                                                          \
            gptb_set_error(nullptr, "gptb_create: %s -> %s", #expr, cudaGetErrorString(_e)); \
            gptb_destroy(h);                                                              \
            return GPTB_ERR_CUDA;                                                         \
        }                                                                                 \
    } while (0)
    CREATE_OK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CREATE_OK(cudaGetDeviceProperties(&prop, dev));
    h->sm_count = prop.multiProcessorCount;
    if (cfg->precision != GPTB_PREC_FP64 && prop.major != 10) {
        gptb_set_error(nullptr, "tensor-core modes need an sm_100 device (found sm_%d%d)", prop.major, prop.minor);
        gptb_destroy(h);
        return GPTB_ERR_UNSUPPORTED;
    }
    CREATE_OK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    h->own_stream = true;
    const gptb_config &c = h->cfg;
    h->npad = (c.n + 127) / 128 * 128;
    h->rpad = (c.r + 15) / 16 * 16;
    int st = GPTB_OK;
    auto A = [&](int s) { if (st == GPTB_OK) st = s; };
    A(dev_alloc(h, &h->Zt, (size_t)c.D * c.n));
    A(dev_alloc(h, &h->b, (size_t)c.D * c.n));
    A(dev_alloc(h, &h->ZtF, (size_t)c.D * h->npad));
    A(dev_alloc(h, &h->bF, (size_t)c.D * h->npad));
    A(dev_alloc(h, &h->I, (size_t)c.D * c.Q));
    A(dev_alloc(h, &h->grp_ptr, (size_t)c.D * h->rpad + 1));
    A(dev_alloc(h, &h->grp_q, (size_t)c.Q * c.D));
    A(dev_alloc(h, &h->w, (size_t)c.Q));
    A(dev_alloc(h, &h->U, (size_t)c.D * c.r * c.n));
    A(dev_alloc(h, &h->G, (size_t)c.D * c.r * c.n));
    A(dev_alloc(h, &h->grad, (size_t)c.Q + (size_t)c.D * c.r * c.n));
    h->small_stride = gptb_update_small_stride(c.r);
    A(dev_alloc(h, &h->small, h->small_stride * c.D));
    A(dev_alloc(h, &h->nan_flag, 8));   // [0] geodesic NaN (sticky) [1] bad batch index [2] K2 redo [3] bad permutation [4] row outside the node-grid domain
    A(dev_alloc(h, &h->tc_state, 16));   // [10..11] max|node weight| of alternating steps (nodes.cu)
    A(dev_alloc(h, &h->pw, (size_t)c.Q));
    A(dev_alloc(h, &h->pU, (size_t)c.D * c.r * c.n));
    if (st == GPTB_OK && gptb_is_tc(h)) {
        h->Uimg_bytes = gptb_tc_uimg_bytes(h);
        A(dev_alloc(h, (unsigned char **)&h->Uimg, h->Uimg_bytes));
        A(dev_alloc(h, (unsigned char **)&h->pUimg, h->Uimg_bytes));
    }
    if (st != GPTB_OK) {
        gptb_destroy(h);
        return st;
    }
    CREATE_OK(cudaMallocHost((void **)&h->nan_flag_host, 8 * sizeof(int32_t)));
    CREATE_OK(cudaMemsetAsync(h->nan_flag, 0, 8 * sizeof(int32_t), h->stream));
    CREATE_OK(cudaMemsetAsync(h->tc_state, 0, 16 * sizeof(float), h->stream));
    CREATE_OK(cudaStreamSynchronize(h->stream));
#undef CREATE_OK
    *out = h;
    return GPTB_OK;
}

int gptb_destroy(gptb_handle *h) {
    if (!h) return GPTB_OK;
    cudaSetDevice(h->device);
    pipe_join(h);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (int g = 0; g < gptb_handle::PIPE_MAX; ++g)
        if (h->side[g]) {
            cudaStreamSynchronize(h->side[g]);
            cudaStreamDestroy(h->side[g]);
            cudaEventDestroy(h->evG[g]); cudaEventDestroy(h->evU[g]); cudaEventDestroy(h->evR[g]);
        }
    if (h->ar_stream) { cudaStreamSynchronize(h->ar_stream); cudaStreamDestroy(h->ar_stream); }
    if (h->tail_stream) {
        cudaStreamSynchronize(h->tail_stream); cudaStreamDestroy(h->tail_stream);
        cudaEventDestroy(h->evTail);
        cudaEventDestroy(h->evTailFork);
    }
    gptb_comm_destroy(h);
    gptb_nodes_free(h);
    gptb_pred_nodes_free(h);
    dev_free(&h->Zt); dev_free(&h->b); dev_free(&h->ZtF); dev_free(&h->bF); dev_free(&h->I); dev_free(&h->grp_ptr); dev_free(&h->grp_q);
    dev_free(&h->ct_off1); dev_free(&h->ct_lptr); dev_free(&h->ct_lists);
    dev_free(&h->X); dev_free(&h->y); dev_free(&h->phi); dev_free(&h->order); dev_free(&h->order_tmp);
    dev_free(&h->sort_keys); dev_free(&h->sort_keys_tmp);
    if (h->sort_tmp) cudaFree(h->sort_tmp);
    dev_free(&h->w); dev_free(&h->U); dev_free(&h->wF); dev_free(&h->pw); dev_free(&h->pU);
    if (h->Uimg) cudaFree(h->Uimg);
    if (h->pUimg) cudaFree(h->pUimg);
    if (h->ZbF) cudaFree(h->ZbF);
    dev_free(&h->stage_cm); dev_free(&h->stage_rm); dev_free(&h->stage_y);
    for (int i = 0; i < 2; ++i) {
        dev_free(&h->pipe_cm[i]); dev_free(&h->pipe_y[i]);
        if (h->pipe_staged[i]) cudaEventDestroy(h->pipe_staged[i]);
        if (h->pipe_done[i]) cudaEventDestroy(h->pipe_done[i]);
    }
    if (h->pipe_nan_host) cudaFreeHost(h->pipe_nan_host);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->front_stream) {
        cudaStreamSynchronize(h->front_stream); cudaStreamDestroy(h->front_stream);
        cudaEventDestroy(h->ev_front_dep); cudaEventDestroy(h->ev_front_done); cudaEventDestroy(h->ev_rows_free);
    }
    dev_free(&h->Xb); dev_free(&h->yb); dev_free(&h->fhat); dev_free(&h->XbF);
    if (h->T) cudaFree(h->T);
    if (h->E) cudaFree(h->E);
    if (h->Eimg) cudaFree(h->Eimg);
    dev_free(&h->idx_dev); dev_free(&h->grad); dev_free(&h->gradFbase); h->gradF = nullptr;
    if (h->tc_partials) cudaFree(h->tc_partials);
    dev_free(&h->tc_state);
    dev_free(&h->G); dev_free(&h->small); dev_free(&h->colnorm); dev_free(&h->nan_flag);
    dev_free(&h->noise_w); dev_free(&h->noise_U); dev_free(&h->pred_acc);
    dev_free(&h->gen_noise_w); dev_free(&h->gen_noise_U);
    dev_free(&h->rms_gw); dev_free(&h->rms_gU); dev_free(&h->rms_epsk); dev_free(&h->w_old);
    dev_free(&h->store_w); dev_free(&h->store_U);
    if (h->nan_flag_host) cudaFreeHost(h->nan_flag_host);
    if (h->pin) cudaFreeHost(h->pin);
    for (auto &ev : h->evpool) cudaEventDestroy(ev);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return GPTB_OK;
}

int gptb_set_stream(gptb_handle *h, void *cuda_stream) {
    if (!h) return GPTB_ERR_ARG;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    if (h->own_stream) cudaStreamDestroy(h->stream);
    if (cuda_stream) {
        h->stream = (cudaStream_t)cuda_stream;
        h->own_stream = false;
    } else {
        GPTB_CUDA_OK(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->own_stream = true;
    }
    return GPTB_OK;
}

int gptb_synchronize(gptb_handle *h) {
    if (!h) return GPTB_ERR_ARG;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    if (h->have_params) return check_nan(h);   // also reports a geodesic NaN raised by queued steps
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    return GPTB_OK;
}

int gptb_set_features(gptb_handle *h, const double *Z, const double *b, const double *length_scale, int32_t ls_len,
                      double sigma_RBF, double phi_scale) {
    if (!h || !Z || !length_scale) return GPTB_ERR_ARG;
    const gptb_config &c = h->cfg;
    if (c.feature_kind == GPTB_FEAT_COS && !b) {
        gptb_set_error(h, "b is required for cos features");
        return GPTB_ERR_ARG;
    }
    if (ls_len != 1 && ls_len != c.D) {
        gptb_set_error(h, "dimensions of X and length_scale do not match");  // GPT_SGLD.jl:326
        return GPTB_ERR_SHAPE;
    }
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const size_t cnt = (size_t)c.D * c.n;
    std::vector<double> zt(cnt), bb(cnt, 0.0);
    std::vector<float> ztf((size_t)c.D * h->npad, 0.f), bf((size_t)c.D * h->npad, 0.f);
    for (int k = 0; k < c.D; ++k) {
        const double inv = 1.0 / length_scale[ls_len == 1 ? 0 : k];  // scale(Z, 1./length_scale), :75
        h->zmax[k] = 0.0;
        for (int j = 0; j < c.n; ++j) {
            zt[(size_t)k * c.n + j] = Z[(size_t)k * c.n + j] * inv;
            h->zmax[k] = std::max(h->zmax[k], fabs(zt[(size_t)k * c.n + j]));
            if (b) bb[(size_t)k * c.n + j] = b[(size_t)k * c.n + j];
            ztf[(size_t)k * h->npad + j] = (float)zt[(size_t)k * c.n + j];
            bf[(size_t)k * h->npad + j] = (float)bb[(size_t)k * c.n + j];
        }
    }
    if (c.feature_kind == GPTB_FEAT_COS)
        h->amp = phi_scale * pow(sigma_RBF, 1.0 / c.D) * sqrt(2.0 / c.n);           // :83
    else
        h->amp = phi_scale * pow(sigma_RBF, 1.0 / c.D) / sqrt((double)(c.n / 2));   // :104
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->Zt, zt.data(), cnt * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->b, bb.data(), cnt * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->ZtF, ztf.data(), ztf.size() * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->bF, bf.data(), bf.size() * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    h->have_features = true;
    h->ZbF_valid = false;
    h->grid_valid = false;
    return GPTB_OK;
}

int gptb_set_index(gptb_handle *h, const int32_t *I) {
    if (!h || !I) return GPTB_ERR_ARG;
    const gptb_config &c = h->cfg;
    std::vector<int32_t> z((size_t)c.D * c.Q);
    for (size_t t = 0; t < z.size(); ++t) {
        if (I[t] < 1 || I[t] > c.r) {
            gptb_set_error(h, "I[%zu]=%d outside 1..r=%d", t, I[t], c.r);
            return GPTB_ERR_SHAPE;
        }
        z[t] = I[t] - 1;
    }
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->I, z.data(), z.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    // CSR over (k,l): the q with I[q,k] = l, in increasing q (the order computeA's findin visits them, :267)
    std::vector<int32_t> ptr((size_t)c.D * h->rpad + 1, 0);
    std::vector<uint16_t> gq((size_t)c.Q * c.D);
    for (int k = 0; k < c.D; ++k)
        for (int q = 0; q < c.Q; ++q) ptr[(size_t)k * h->rpad + z[(size_t)k * c.Q + q] + 1]++;
    for (size_t g = 0; g + 1 < ptr.size(); ++g) ptr[g + 1] += ptr[g];
    {
        std::vector<int32_t> fill(ptr.begin(), ptr.end() - 1);
        for (int k = 0; k < c.D; ++k)
            for (int q = 0; q < c.Q; ++q) gq[fill[(size_t)k * h->rpad + z[(size_t)k * c.Q + q]]++] = (uint16_t)q;
    }
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->grp_ptr, ptr.data(), ptr.size() * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->grp_q, gq.data(), gq.size() * sizeof(uint16_t), cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    if (gptb_is_tc(h)) GPTB_TRY(gptb_contract_tc_build_tables(h, z.data()));
    h->have_index = true;
    return GPTB_OK;
}

static int set_train_common(gptb_handle *h, const double *y, int64_t N) {
    GPTB_TRY(dev_alloc(h, &h->y, (size_t)N));
    GPTB_TRY(dev_alloc(h, &h->order, (size_t)N));
    dev_free(&h->order_tmp);
    dev_free(&h->sort_keys);
    dev_free(&h->sort_keys_tmp);
    if (h->sort_tmp) cudaFree(h->sort_tmp);
    h->sort_tmp = nullptr;
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->y, y, sizeof(double) * N, cudaMemcpyHostToDevice, h->stream));
    h->N = N;
    GPTB_TRY(gptb_reset_order(h));
    h->epoch_counter = 0;
    h->step_counter = 0;
    h->batch_cursor = 0;
    return GPTB_OK;
}

int gptb_set_train(gptb_handle *h, const double *X, const double *y, int64_t N) {
    if (!h || !X || !y || N < 1 || N > 2147483647LL) return GPTB_ERR_ARG;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const gptb_config &c = h->cfg;
    dev_free(&h->phi);
    h->have_phi = false;
    double *tmp = nullptr;
    GPTB_TRY(dev_alloc(h, &tmp, (size_t)N * c.D));
    int st = dev_alloc(h, &h->X, (size_t)N * c.D);
    if (st == GPTB_OK) {
        cudaError_t e = cudaMemcpyAsync(tmp, X, sizeof(double) * N * c.D, cudaMemcpyHostToDevice, h->stream);
        if (e != cudaSuccess) st = GPTB_ERR_CUDA;
    }
    if (st == GPTB_OK) st = gptb_launch_transpose(h, tmp, N, h->X);
    if (st == GPTB_OK) st = set_train_common(h, y, N);
    cudaStreamSynchronize(h->stream);
    cudaFree(tmp);
    if (st != GPTB_OK) return st;
    h->have_train = true;
    // input domain of the node grid: the range of every column (X[N,D] is column-major: column k is contiguous)
    for (int k = 0; k < c.D; ++k) {
        const double *col = X + (size_t)k * N;
        double lo = col[0], hi = col[0];
        for (int64_t i = 1; i < N; ++i) {
            lo = col[i] < lo ? col[i] : lo;
            hi = col[i] > hi ? col[i] : hi;
        }
        h->dom_lo[k] = lo;
        h->dom_hi[k] = hi;
    }
    h->have_domain = true;
    h->grid_valid = false;
    return GPTB_OK;
}

int gptb_set_domain(gptb_handle *h, const double *lo, const double *hi) {
    if (!h) return GPTB_ERR_ARG;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    h->grid_valid = false;
    if (!lo || !hi) {
        h->have_domain = false;
        return GPTB_OK;
    }
    for (int k = 0; k < h->cfg.D; ++k) {
        if (!(lo[k] <= hi[k]) || !std::isfinite(lo[k]) || !std::isfinite(hi[k])) {
            gptb_set_error(h, "gptb_set_domain: need finite lo[k] <= hi[k] (k=%d: %g, %g)", k, lo[k], hi[k]);
            return GPTB_ERR_ARG;
        }
        h->dom_lo[k] = lo[k];
        h->dom_hi[k] = hi[k];
    }
    h->have_domain = true;
    return GPTB_OK;
}

int gptb_set_train_phi(gptb_handle *h, const double *phi, const double *y, int64_t N) {
    if (!h || !phi || !y || N < 1 || N > 2147483647LL) return GPTB_ERR_ARG;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const gptb_config &c = h->cfg;
    dev_free(&h->X);
    h->have_train = false;
    size_t cnt = (size_t)N * c.D * c.n;
    GPTB_TRY(dev_alloc(h, &h->phi, cnt));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->phi, phi, sizeof(double) * cnt, cudaMemcpyHostToDevice, h->stream));
    GPTB_TRY(set_train_common(h, y, N));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    h->have_phi = true;
    return GPTB_OK;
}

int gptb_set_params(gptb_handle *h, const double *w, const double *U) {
    if (!h || !w || !U) return GPTB_ERR_ARG;
    const gptb_config &c = h->cfg;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->w, w, sizeof(double) * c.Q, cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->U, U, sizeof(double) * c.n * c.r * c.D, cudaMemcpyHostToDevice, h->stream));
    GPTB_CUDA_OK(h, cudaMemsetAsync(h->nan_flag, 0, sizeof(int32_t), h->stream));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    h->have_params = true;
    h->shadows_valid = false;
    h->tc_emax_valid = false;
    if (h->rms_gw) {   // a new chain starts with empty moving averages (:1143-1144)
        GPTB_CUDA_OK(h, cudaMemsetAsync(h->rms_gw, 0, sizeof(double) * c.Q, h->stream));
        GPTB_CUDA_OK(h, cudaMemsetAsync(h->rms_gU, 0, sizeof(double) * c.D * c.r * c.n, h->stream));
    }
    return GPTB_OK;
}

int gptb_get_params(gptb_handle *h, double *w, double *U) {
    if (!h) return GPTB_ERR_ARG;
    if (!h->have_params) {
        gptb_set_error(h, "parameters not set");
        return GPTB_ERR_STATE;
    }
    const gptb_config &c = h->cfg;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    if (w) GPTB_CUDA_OK(h, cudaMemcpyAsync(w, h->w, sizeof(double) * c.Q, cudaMemcpyDeviceToHost, h->stream));
    if (U) GPTB_CUDA_OK(h, cudaMemcpyAsync(U, h->U, sizeof(double) * c.n * c.r * c.D, cudaMemcpyDeviceToHost, h->stream));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    return GPTB_OK;
}

// Symmetric eigen-decomposition by cyclic Jacobi (host, r x r) for (ZZ')^(-1/2) of :366.
static void jacobi_eigh(std::vector<double> &A, int r, std::vector<double> &V) {
    V.assign((size_t)r * r, 0.0);
    for (int i = 0; i < r; ++i) V[i + (size_t)r * i] = 1.0;
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < r; ++p)
            for (int q = p + 1; q < r; ++q) off += A[p + (size_t)r * q] * A[p + (size_t)r * q];
        if (off < 1e-30) break;
        for (int p = 0; p < r; ++p)
            for (int q = p + 1; q < r; ++q) {
                double apq = A[p + (size_t)r * q];
                if (fabs(apq) < 1e-300) continue;
                double app = A[p + (size_t)r * p], aqq = A[q + (size_t)r * q];
                double tau = (aqq - app) / (2 * apq);
                double t = (tau >= 0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1 + tau * tau));
                double cs = 1 / sqrt(1 + t * t), sn = t * cs;
                for (int k = 0; k < r; ++k) {
                    double akp = A[k + (size_t)r * p], akq = A[k + (size_t)r * q];
                    A[k + (size_t)r * p] = cs * akp - sn * akq;
                    A[k + (size_t)r * q] = sn * akp + cs * akq;
                }
                for (int k = 0; k < r; ++k) {
                    double apk = A[p + (size_t)r * k], aqk = A[q + (size_t)r * k];
                    A[p + (size_t)r * k] = cs * apk - sn * aqk;
                    A[q + (size_t)r * k] = sn * apk + cs * aqk;
                }
                for (int k = 0; k < r; ++k) {
                    double vkp = V[k + (size_t)r * p], vkq = V[k + (size_t)r * q];
                    V[k + (size_t)r * p] = cs * vkp - sn * vkq;
                    V[k + (size_t)r * q] = sn * vkp + cs * vkq;
                }
            }
    }
}

int gptb_init_params(gptb_handle *h, uint64_t param_seed, int32_t stiefel) {
    if (!h) return GPTB_ERR_ARG;
    const gptb_config &c = h->cfg;
    const int n = c.n, r = c.r, D = c.D;
    if (stiefel && r > n) {
        gptb_set_error(h, "Stiefel initialisation needs r <= n");
        return GPTB_ERR_SHAPE;
    }
    std::vector<double> w(c.Q), U((size_t)n * r * D);
    for (int q = 0; q < c.Q; q += 2) {
        double a, b;
        philox_normal2(param_seed, GPTB_STREAM_INIT_W | (uint64_t)(q / 2), 0, &a, &b);
        w[q] = a;
        if (q + 1 < c.Q) w[q + 1] = b;
    }
    std::vector<double> Z((size_t)r * n), G((size_t)r * r), V, M((size_t)r * r);
    for (int k = 0; k < D; ++k) {
        // Z = randn(r,n) (:365) — or randn(n,r)/sqrt(n) (:368); same draws, column-major fill
        for (size_t e = 0; e < (size_t)r * n; e += 2) {
            double a, b;
            philox_normal2(param_seed, GPTB_STREAM_INIT_U | ((uint64_t)k << 40) | (e / 2), 0, &a, &b);
            Z[e] = a;
            if (e + 1 < (size_t)r * n) Z[e + 1] = b;
        }
        double *Uk = U.data() + (size_t)n * r * k;
        if (!stiefel) {
            for (int l = 0; l < r; ++l)
                for (int j = 0; j < n; ++j) Uk[j + (size_t)n * l] = Z[j + (size_t)n * l] / sqrt((double)n);
            continue;
        }
        for (int a = 0; a < r; ++a)
            for (int b = 0; b < r; ++b) {
                double s = 0.0;
                for (int j = 0; j < n; ++j) s += Z[a + (size_t)r * j] * Z[b + (size_t)r * j];
                G[a + (size_t)r * b] = s;
            }
        jacobi_eigh(G, r, V);  // G now ~diagonal
        for (int a = 0; a < r; ++a)
            for (int b = 0; b < r; ++b) {
                double s = 0.0;
                for (int p = 0; p < r; ++p) s += V[a + (size_t)r * p] * V[b + (size_t)r * p] / sqrt(G[p + (size_t)r * p]);
                M[a + (size_t)r * b] = s;  // (ZZ')^(-1/2)
            }
        for (int j = 0; j < n; ++j)
            for (int l = 0; l < r; ++l) {
                double s = 0.0;
                for (int a = 0; a < r; ++a) s += M[l + (size_t)r * a] * Z[a + (size_t)r * j];
                Uk[j + (size_t)n * l] = s;  // transpose of (ZZ')^(-1/2) Z
            }
    }
    return gptb_set_params(h, w.data(), U.data());
}

int gptb_features(gptb_handle *h, const double *X, int64_t N, double *phi_out) {
    if (!h || !X || !phi_out || N < 0) return GPTB_ERR_ARG;
    if (!h->have_features) {
        gptb_set_error(h, "features not set");
        return GPTB_ERR_STATE;
    }
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const gptb_config &c = h->cfg;
    const int64_t chunk = std::max<int64_t>(1, (int64_t)(256ll << 20) / ((int64_t)c.n * c.D * 8));
    double *xc = nullptr, *xr = nullptr, *ph = nullptr;
    int st = dev_alloc(h, &xc, (size_t)chunk * c.D);
    if (st == GPTB_OK) st = dev_alloc(h, &xr, (size_t)chunk * c.D);
    if (st == GPTB_OK) st = dev_alloc(h, &ph, (size_t)chunk * c.D * c.n);
    for (int64_t i0 = 0; st == GPTB_OK && i0 < N; i0 += chunk) {
        int64_t cn = std::min(chunk, N - i0);
        cudaError_t e = cudaMemcpy2DAsync(xc, cn * sizeof(double), X + i0, N * sizeof(double), cn * sizeof(double), c.D,
                                          cudaMemcpyHostToDevice, h->stream);
        if (e != cudaSuccess) { st = GPTB_ERR_CUDA; gptb_set_error(h, "features H2D: %s", cudaGetErrorString(e)); break; }
        st = gptb_launch_transpose(h, xc, cn, xr);
        // transpose kernel reads Xcm with leading dimension = its N argument (cn): matches the 2-D copy above
        if (st == GPTB_OK) st = gptb_launch_features_fp64(h, xr, cn, ph);
        if (st != GPTB_OK) break;
        e = cudaMemcpyAsync(phi_out + (size_t)i0 * c.D * c.n, ph, sizeof(double) * cn * c.D * c.n, cudaMemcpyDeviceToHost,
                            h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { st = GPTB_ERR_CUDA; gptb_set_error(h, "features D2H: %s", cudaGetErrorString(e)); }
    }
    cudaStreamSynchronize(h->stream);
    cudaFree(xc); cudaFree(xr); cudaFree(ph);
    return st;
}

// Posterior-predictive sweep on the tensor-core path (K6): the factors of Sg samples sit side by side in one operand
// image, so K1 generates each feature once per sample group, and the forward-only contraction handles the group's
// samples in one launch.  Row chunks are sized so that the group's T block stays in L2 between the two kernels.
static int pred_tc_grouped(gptb_handle *h, const double *w_store, const double *U_store, int32_t S, const double *Xtest,
                           int64_t Nt, double *mean_out, double *per_sample_out) {
    const gptb_config &c = h->cfg;
    const size_t ucnt = (size_t)c.n * c.r * c.D;
    // widest image that stays resident in shared memory (see gptb_launch_fwd_tc_cols)
    const size_t budget = 227 * 1024 - 2048;
    int nc_max = (int)(((budget / h->npad) - sizeof(float2)) / 4) / 16 * 16;
    nc_max = std::min(nc_max, 128);
    int Sg = std::max(1, nc_max / c.r);
    Sg = std::min<int>(Sg, S);
    const int NC = std::max(h->rpad, (Sg * c.r + 15) / 16 * 16);
    const int ngroups = (S + Sg - 1) / Sg;
    const size_t img_elems = (size_t)c.D * h->npad * 2 * NC;
    if ((size_t)S * ucnt * 8 > (size_t)8 << 30 || img_elems * 2 * ngroups > (size_t)8 << 30) return GPTB_ERR_UNSUPPORTED;
    int64_t chunk = (int64_t)((size_t)(48 << 20) / ((size_t)c.D * Sg * c.r * 4)) / 128 * 128;
    chunk = std::max<int64_t>(128, std::min<int64_t>(chunk, 65536));
    chunk = std::min<int64_t>(chunk, (Nt + 127) / 128 * 128);
    GPTB_TRY(ensure_batch(h, chunk));
    const int64_t mpad = h->mpad;
    double *Ud = nullptr, *wd = nullptr, *acc = nullptr, *fh = nullptr;
    float *Tg = nullptr;
    unsigned char *imgs = nullptr;
    int st = dev_alloc(h, &Ud, (size_t)S * ucnt);
    if (st == GPTB_OK) st = dev_alloc(h, &wd, (size_t)S * c.Q);
    if (st == GPTB_OK) st = dev_alloc(h, &imgs, img_elems * 2 * ngroups);
    if (st == GPTB_OK) st = dev_alloc(h, &Tg, (size_t)c.D * Sg * c.r * mpad);
    if (st == GPTB_OK) st = dev_alloc(h, &acc, (size_t)mpad);
    if (st == GPTB_OK && per_sample_out) st = dev_alloc(h, &fh, (size_t)S * mpad);
    auto cleanup = [&]() {
        cudaStreamSynchronize(h->stream);
        dev_free(&Ud); dev_free(&wd); dev_free(&imgs); dev_free(&Tg); dev_free(&acc); dev_free(&fh);
    };
    if (st != GPTB_OK) { cleanup(); return st; }
#define PRED_OK(expr)                                                                     \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess) {                                                          \
            gptb_set_error(h, "pred_samples: %s -> %s", #expr, cudaGetErrorString(_e));   \
            cleanup();                                                                    \
            return GPTB_ERR_CUDA;                                                         \
        }                                                                                 \
    } while (0)
#define PRED_TRY(expr)                 \
    do {                               \
        int _s = (expr);               \
        if (_s != GPTB_OK) { cleanup(); return _s; } \
    } while (0)
    PRED_OK(cudaMemcpyAsync(Ud, U_store, sizeof(double) * S * ucnt, cudaMemcpyHostToDevice, h->stream));
    PRED_OK(cudaMemcpyAsync(wd, w_store, sizeof(double) * S * c.Q, cudaMemcpyHostToDevice, h->stream));
    PRED_TRY(gptb_tc_pred_scale(h, Ud, (size_t)S * ucnt, true));
    PRED_TRY(gptb_tc_pred_scale_commit(h));
    for (int g = 0; g < ngroups; ++g) {
        const int sg = std::min(Sg, S - g * Sg);
        PRED_TRY(gptb_tc_pack_group(h, Ud + (size_t)g * Sg * ucnt, sg, NC, imgs + (size_t)g * img_elems * 2));
    }
    for (int64_t i0 = 0; i0 < Nt; i0 += chunk) {
        const int64_t cn = std::min(chunk, Nt - i0);
        PRED_OK(cudaMemcpy2DAsync(h->Xb, mpad * sizeof(double), Xtest + i0, Nt * sizeof(double), cn * sizeof(double), c.D,
                                  cudaMemcpyHostToDevice, h->stream));
        d2f_kernel<<<256, 256, 0, h->stream>>>(h->Xb, h->XbF, (size_t)c.D * mpad);
        h->launches++;
        PRED_OK(cudaMemsetAsync(acc, 0, sizeof(double) * mpad, h->stream));
        for (int g = 0; g < ngroups; ++g) {
            const int sg = std::min(Sg, S - g * Sg);
            PRED_TRY(gptb_launch_fwd_tc_cols(h, h->XbF, imgs + (size_t)g * img_elems * 2, h->tc_state + 7, cn, mpad, Tg,
                                             sg * c.r, NC));
            PRED_TRY(gptb_launch_contract_fwd(h, Tg, wd + (size_t)g * Sg * c.Q, sg, cn, mpad, sg * c.r,
                                              fh ? fh + (size_t)g * Sg * mpad : nullptr, mean_out ? acc : nullptr, 1.0 / S));
        }
        if (mean_out) PRED_OK(cudaMemcpyAsync(mean_out + i0, acc, sizeof(double) * cn, cudaMemcpyDeviceToHost, h->stream));
        if (per_sample_out)
            PRED_OK(cudaMemcpy2DAsync(per_sample_out + i0, Nt * sizeof(double), fh, mpad * sizeof(double), cn * sizeof(double), S,
                                      cudaMemcpyDeviceToHost, h->stream));
        PRED_OK(cudaStreamSynchronize(h->stream));
    }
#undef PRED_OK
#undef PRED_TRY
    cleanup();
    return GPTB_OK;
}

// Shared by pred / pred_phi / pred_samples: S parameter sets, rows chunked through the batch scratch.
static int pred_impl(gptb_handle *h, const double *w_store, const double *U_store, int32_t S, const double *Xtest,
                     const double *phitest, int64_t Nt, double *mean_out, double *per_sample_out) {
    const gptb_config &c = h->cfg;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const bool phi_mode = phitest != nullptr;
    if (!phi_mode && !h->have_features) {
        gptb_set_error(h, "features not set");
        return GPTB_ERR_STATE;
    }
    if (!h->have_index) {
        gptb_set_error(h, "core index table not set");
        return GPTB_ERR_STATE;
    }
    if (Nt == 0) return GPTB_OK;
    const bool f64 = use_fp64(h, phi_mode);
    if (!f64) {
        const int sn = gptb_pred_nodes(h, w_store, U_store, S, Xtest, Nt, mean_out, per_sample_out);
        if (sn != GPTB_ERR_UNSUPPORTED) return sn;   // node grid not applicable (small sweeps, GPTB_ALG_DIRECT): direct path
        const int st = pred_tc_grouped(h, w_store, U_store, S, Xtest, Nt, mean_out, per_sample_out);
        if (st != GPTB_ERR_UNSUPPORTED) return st;   // shapes it does not cover use the per-sample path below
    }
    const size_t ucnt = (size_t)c.n * c.r * c.D;
    int64_t chunk = std::min<int64_t>(Nt, 65536);
    if (phi_mode) chunk = std::min<int64_t>(chunk, std::max<int64_t>(128, (int64_t)(512ll << 20) / ((int64_t)c.n * c.D * 8)));
    GPTB_TRY(ensure_batch(h, chunk));
    double *phichunk = nullptr, *acc = nullptr, *saved_phi = h->phi;
    if (phi_mode) GPTB_TRY(dev_alloc(h, &phichunk, (size_t)chunk * c.D * c.n));
    int st = dev_alloc(h, &acc, (size_t)h->mpad);
    GPTB_TRY(ensure_pin(h, sizeof(double) * chunk));
    for (int64_t i0 = 0; st == GPTB_OK && i0 < Nt; i0 += chunk) {
        const int64_t cn = std::min(chunk, Nt - i0);
        cudaError_t e;
        if (phi_mode) {
            e = cudaMemcpyAsync(phichunk, phitest + (size_t)i0 * c.D * c.n, sizeof(double) * cn * c.D * c.n,
                                cudaMemcpyHostToDevice, h->stream);
        } else {
            e = cudaMemcpy2DAsync(h->Xb, h->mpad * sizeof(double), Xtest + i0, Nt * sizeof(double), cn * sizeof(double),
                                  c.D, cudaMemcpyHostToDevice, h->stream);
            if (e == cudaSuccess && !f64) {
                d2f_kernel<<<256, 256, 0, h->stream>>>(h->Xb, h->XbF, (size_t)c.D * h->mpad);
                h->launches++;
            }
        }
        if (e != cudaSuccess) { st = GPTB_ERR_CUDA; gptb_set_error(h, "pred H2D: %s", cudaGetErrorString(e)); break; }
        cudaMemsetAsync(acc, 0, sizeof(double) * h->mpad, h->stream);
        for (int32_t s = 0; st == GPTB_OK && s < S; ++s) {
            cudaMemcpyAsync(h->pw, w_store + (size_t)s * c.Q, sizeof(double) * c.Q, cudaMemcpyHostToDevice, h->stream);
            cudaMemcpyAsync(h->pU, U_store + (size_t)s * ucnt, sizeof(double) * ucnt, cudaMemcpyHostToDevice, h->stream);
            if (!f64) st = gptb_tc_pack_U(h, h->pU, h->pUimg);
            if (st != GPTB_OK) break;
            h->phi = phichunk;  // forward_contract reads h->phi in phi mode
            st = forward_contract(h, phi_mode, nullptr, h->pw, h->pU, h->pUimg, cn, 0.0, false, nullptr);
            h->phi = saved_phi;
            if (st != GPTB_OK) break;
            axpy_kernel<<<64, 256, 0, h->stream>>>(h->fhat, acc, 1.0 / S, (size_t)cn);
            h->launches++;
            if (per_sample_out) {
                e = cudaMemcpyAsync(per_sample_out + (size_t)s * Nt + i0, h->fhat, sizeof(double) * cn,
                                    cudaMemcpyDeviceToHost, h->stream);
                if (e != cudaSuccess) { st = GPTB_ERR_CUDA; break; }
            }
        }
        if (st != GPTB_OK) break;
        if (mean_out) {
            e = cudaMemcpyAsync(mean_out + i0, acc, sizeof(double) * cn, cudaMemcpyDeviceToHost, h->stream);
            if (e != cudaSuccess) { st = GPTB_ERR_CUDA; break; }
        }
        e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { st = GPTB_ERR_CUDA; gptb_set_error(h, "pred: %s", cudaGetErrorString(e)); }
    }
    cudaStreamSynchronize(h->stream);
    h->phi = saved_phi;
    if (phichunk) cudaFree(phichunk);
    if (acc) cudaFree(acc);
    return st;
}

int gptb_pred(gptb_handle *h, const double *w, const double *U, const double *Xtest, int64_t Nt, double *fhat_out) {
    if (!h || !w || !U || !Xtest || !fhat_out || Nt < 0) return GPTB_ERR_ARG;
    return pred_impl(h, w, U, 1, Xtest, nullptr, Nt, fhat_out, nullptr);
}

int gptb_pred_phi(gptb_handle *h, const double *w, const double *U, const double *phitest, int64_t Nt,
                  double *fhat_out) {
    if (!h || !w || !U || !phitest || !fhat_out || Nt < 0) return GPTB_ERR_ARG;
    return pred_impl(h, w, U, 1, nullptr, phitest, Nt, fhat_out, nullptr);
}

int gptb_pred_samples(gptb_handle *h, const double *w_store, const double *U_store, int32_t S, const double *Xtest,
                      int64_t Nt, double *mean_out, double *per_sample_out) {
    if (!h || !w_store || !U_store || !Xtest || S < 1 || Nt < 0 || (!mean_out && !per_sample_out)) return GPTB_ERR_ARG;
    return pred_impl(h, w_store, U_store, S, Xtest, nullptr, Nt, mean_out, per_sample_out);
}

int gptb_grad(gptb_handle *h, double signal_var, const int64_t *idx, int64_t start, int64_t bs, double *gradw_out,
              double *gradU_out, double *fhat_out, int32_t include_prior) {
    if (!h || !(signal_var > 0)) return GPTB_ERR_ARG;
    GPTB_TRY(check_ready(h, true, true));
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const gptb_config &c = h->cfg;
    const int32_t *rows = nullptr;
    GPTB_TRY(stage_rows(h, idx, start, bs, &rows));
    PhaseTimer pt(h);
    GPTB_TRY(compute_grad(h, signal_var, rows, bs, h->N, false, nullptr, nullptr, pt));
    pt.finish();
    const size_t ucnt = (size_t)c.n * c.r * c.D;
    if (gradw_out) {
        std::vector<double> w(c.Q);
        GPTB_CUDA_OK(h, cudaMemcpyAsync(gradw_out, h->grad, sizeof(double) * c.Q, cudaMemcpyDeviceToHost, h->stream));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(w.data(), h->w, sizeof(double) * c.Q, cudaMemcpyDeviceToHost, h->stream));
        GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
        if (include_prior)
            for (int q = 0; q < c.Q; ++q) gradw_out[q] -= w[q];  // - w/sigma_w^2, sigma_w = 1 (:354,:393)
    }
    if (gradU_out)
        GPTB_CUDA_OK(h, cudaMemcpyAsync(gradU_out, h->grad + c.Q, sizeof(double) * ucnt, cudaMemcpyDeviceToHost, h->stream));
    if (fhat_out) {
        // fhat of this rank's share only is resident; with one rank that is the whole batch
        const int64_t lo = (int64_t)h->rank * bs / h->nranks, hi = (int64_t)(h->rank + 1) * bs / h->nranks;
        GPTB_CUDA_OK(h, cudaMemcpyAsync(fhat_out + lo, h->fhat, sizeof(double) * (hi - lo), cudaMemcpyDeviceToHost,
                                        h->stream));
    }
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    return GPTB_OK;
}


static int proj_geod_host(int32_t device, int32_t n, int32_t r, const double *U, const double *V, double t, double *out,
                          int mode, double *out2 = nullptr) {
    if (!U || !V || !out || n < 1 || r < 1 || r > 64) {
        gptb_set_error(nullptr, "proj/geod: bad argument (need n>=1, 1<=r<=64)");
        return GPTB_ERR_ARG;
    }
    gptb_handle tmp;  // error sink + scratch owner
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        gptb_set_error(nullptr, "no CUDA device (libgptb200 has no CPU fallback)");
        return GPTB_ERR_CUDA;
    }
    if (device >= 0) cudaSetDevice(device);
    cudaStream_t s;
    if (cudaStreamCreate(&s) != cudaSuccess) return GPTB_ERR_CUDA;
    double *dU = nullptr, *dV = nullptr, *dO = nullptr, *dO2 = nullptr;
    int32_t *flag = nullptr;
    size_t bytes = sizeof(double) * (size_t)n * r;
    int st = GPTB_OK;
    if (cudaMalloc(&dU, bytes) != cudaSuccess || cudaMalloc(&dV, bytes) != cudaSuccess ||
        cudaMalloc(&dO, bytes) != cudaSuccess || cudaMalloc(&flag, sizeof(int32_t)) != cudaSuccess ||
        (out2 && cudaMalloc(&dO2, bytes) != cudaSuccess))
        st = GPTB_ERR_NOMEM;
    if (st == GPTB_OK) {
        cudaMemcpyAsync(dU, U, bytes, cudaMemcpyHostToDevice, s);
        cudaMemcpyAsync(dV, V, bytes, cudaMemcpyHostToDevice, s);
        cudaMemsetAsync(flag, 0, sizeof(int32_t), s);
        st = gptb_proj_geod_device(&tmp, s, n, r, dU, dV, t, dO, mode, flag, dO2);
    }
    int32_t bad = 0;
    if (st == GPTB_OK) {
        cudaMemcpyAsync(&bad, flag, sizeof(int32_t), cudaMemcpyDeviceToHost, s);
        cudaMemcpyAsync(out, dO, bytes, cudaMemcpyDeviceToHost, s);
        if (out2) cudaMemcpyAsync(out2, dO2, bytes, cudaMemcpyDeviceToHost, s);
        if (cudaStreamSynchronize(s) != cudaSuccess) st = GPTB_ERR_CUDA;
    }
    if (st == GPTB_OK && bad) {
        memset(out, 0, bytes);  // return zeros(n,r), GPT_SGLD.jl:26
        if (out2) memset(out2, 0, bytes);   // ... and zeros(n,r) for the momentum, :45
        gptb_set_error(nullptr, "Get NaN when moving along Geodesic. Try smaller epsU");
        st = GPTB_GEODESIC_NAN;
    }
    cudaFree(dU); cudaFree(dV); cudaFree(dO); cudaFree(dO2); cudaFree(flag);
    cudaStreamDestroy(s);
    return st;
}

int gptb_proj(int32_t device, int32_t n, int32_t r, const double *U, const double *V, double *out) {
    return proj_geod_host(device, n, r, U, V, 0.0, out, 0);
}

int gptb_geod(int32_t device, int32_t n, int32_t r, const double *U, const double *mom, double t, double *out) {
    return proj_geod_host(device, n, r, U, mom, t, out, 1);
}

int gptb_geodboth(int32_t device, int32_t n, int32_t r, const double *U, const double *mom, double t, double *U_out,
                  double *mom_out) {
    if (!mom_out) return GPTB_ERR_ARG;
    return proj_geod_host(device, n, r, U, mom, t, U_out, 1, mom_out);
}

static int step_common(gptb_handle *h, double signal_var, const int32_t *rows, int64_t bs, int64_t N_total,
                       bool staged, const double *Xstage, const double *ystage, double epsw, double epsU, int32_t flags,
                       const double *nw_dev, const double *nU_dev) {
    // FP32 gradient flow (tensor-core path): K2 sums gradw and K3 gradU into ONE FP32 buffer that the all-reduce and
    // the update read directly — no FP64 fold of the gradient (9 us at C4), no 2.6 MB memset, half the all-reduce
    // bytes.  (gptb_grad keeps the FP64 buffer: it returns the gradient to the caller.)
    struct F32Guard {
        gptb_handle *h;
        ~F32Guard() { h->f32_grad = false; }
    } f32guard{h};
    const bool rms = (flags & GPTB_FLAG_RMSPROP) != 0;
    if (rms && (!(flags & GPTB_FLAG_STIEFEL) || (flags & GPTB_FLAG_FIX_U))) {
        gptb_set_error(h, "GPTB_FLAG_RMSPROP is the Stiefel sampler of GPT_SGLD.jl:1121-1237: set GPTB_FLAG_STIEFEL, not GPTB_FLAG_FIX_U");
        return GPTB_ERR_ARG;
    }
    // (the RMSprop step reads the complete FP64 gradient twice before the update: no FP32 gradient flow)
    h->f32_grad = !rms && !use_fp64(h, h->have_phi && !staged) && gptb_contract_tc_usable(h);
    struct RmsGuard {
        gptb_handle *h;
        ~RmsGuard() { h->rms_active = false; h->rms_noise_w = nullptr; }
    } rmsguard{h};
    if (rms) {
        const gptb_config &c = h->cfg;
        if (!h->rms_gw) {
            GPTB_TRY(dev_alloc(h, &h->rms_gw, (size_t)c.Q));
            GPTB_TRY(dev_alloc(h, &h->rms_gU, (size_t)c.D * c.r * c.n));
            GPTB_TRY(dev_alloc(h, &h->rms_epsk, (size_t)c.D));
            GPTB_TRY(dev_alloc(h, &h->w_old, (size_t)c.Q));
            GPTB_CUDA_OK(h, cudaMemsetAsync(h->rms_gw, 0, sizeof(double) * c.Q, h->stream));
            GPTB_CUDA_OK(h, cudaMemsetAsync(h->rms_gU, 0, sizeof(double) * c.D * c.r * c.n, h->stream));
        }
        h->rms_active = true;
        h->rms_eps = epsw;            // the sampler's single step size epsilon (:1121)
        h->rms_N = (double)N_total;
        h->rms_noise_w = nw_dev;
        h->rms_flags = flags;
    }
    if (pipe_usable(h, bs, staged, flags))
        return step_pipelined(h, signal_var, rows, bs, N_total, staged, Xstage, ystage, epsw, epsU, flags, nw_dev, nU_dev);
    GPTB_TRY(pipe_join_groups(h));
    if (rms) GPTB_TRY(pipe_join_tail(h));   // (the RMSprop step moves w in the middle of its gradient pass)
    PhaseTimer pt(h);
    // production-mode draws of this step: generated on a side stream under the gradient kernels (they depend on the
    // step counter only; the previous step's update, earlier on the main stream, has consumed the buffer)
    const bool side_noise = (flags & GPTB_FLAG_LANGEVIN) && !nw_dev && !nU_dev && !h->profiling;
    if (side_noise) {
        GPTB_TRY(pipe_init(h, 1));
        GPTB_CUDA_OK(h, cudaEventRecord(h->evG[0], h->stream));
        GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->side[0], h->evG[0], 0));
        GPTB_TRY(gptb_launch_noise_range(h, h->step_counter, 0, (flags & GPTB_FLAG_FIX_U) ? 0 : h->cfg.D, true, h->side[0]));
        GPTB_CUDA_OK(h, cudaEventRecord(h->evR[0], h->side[0]));
    }
    const bool fix_U = (flags & GPTB_FLAG_FIX_U) != 0;
    const int kn_upd = fix_U ? 0 : h->cfg.D;          // GPT_SGLDERMw: only w moves (:1108)
    GPTB_TRY(compute_grad(h, signal_var, rows, bs, N_total, staged, Xstage, ystage, pt, fix_U));
    if (side_noise) GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->stream, h->evR[0], 0));
    if (rms) GPTB_TRY(gptb_launch_rms_U(h, h->stream));     // gU, per-dimension step sizes (:1210-1212); w has moved already
    GPTB_TRY(pipe_join_tail(h));                            // the previous step's read-backs of w / U are done
    GPTB_TRY(gptb_launch_update_range(h, epsw, epsU, flags, nw_dev, nU_dev, h->step_counter, 0, kn_upd, !rms, h->stream,
                                      side_noise));
    pt.done(GPTB_PHASE_UPDATE);
    pt.finish();
    h->step_counter++;
    return GPTB_OK;
}

int gptb_sgld_step(gptb_handle *h, double signal_var, const int64_t *idx, int64_t start, int64_t bs, double epsw,
                   double epsU, int32_t flags, const double *noise_w, const double *noise_U) {
    if (!h || !(signal_var > 0)) return GPTB_ERR_ARG;
    GPTB_TRY(check_ready(h, true, true));
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const int32_t *rows = nullptr;
    GPTB_TRY(stage_rows(h, idx, start, bs, &rows));
    const double *nw = nullptr, *nU = nullptr;
    GPTB_TRY(stage_noise(h, noise_w, noise_U, &nw, &nU));
    GPTB_TRY(step_common(h, signal_var, rows, bs, h->N, false, nullptr, nullptr, epsw, epsU, flags, nw, nU));
    if (flags & GPTB_FLAG_STIEFEL) return check_nan(h);
    GPTB_TRY(pipe_join(h));
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    return GPTB_OK;
}

int gptb_sgld_step_host(gptb_handle *h, double signal_var, const double *Xb, const double *yb, int64_t bs,
                        int64_t N_total, double epsw, double epsU, int32_t flags, const double *noise_w,
                        const double *noise_U, double *w_out) {
    if (!h || !Xb || !yb || bs < 1 || N_total < 1 || !(signal_var > 0)) return GPTB_ERR_ARG;
    if (!h->have_features || !h->have_index || !h->have_params) {
        gptb_set_error(h, "features / index / parameters must be set before gptb_sgld_step_host");
        return GPTB_ERR_STATE;
    }
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const gptb_config &c = h->cfg;
    // stage the host minibatch: Xb[bs,D] column-major -> device, then row-major rows for the sharded gather
    if (h->stage_cap < bs) {
        GPTB_TRY(dev_alloc(h, &h->stage_cm, (size_t)bs * c.D));
        GPTB_TRY(dev_alloc(h, &h->stage_rm, (size_t)bs * c.D));
        GPTB_TRY(dev_alloc(h, &h->stage_y, (size_t)bs));
        h->stage_cap = bs;
    }
    // this rank's rows [lo, hi) of the global minibatch only (see gptb_sgld_step_host_submit)
    const int64_t slo = (int64_t)h->rank * bs / h->nranks, shi = (int64_t)(h->rank + 1) * bs / h->nranks, sbs = shi - slo;
    if (sbs > 0) {
        GPTB_CUDA_OK(h, cudaMemcpy2DAsync(h->stage_cm, sizeof(double) * sbs, Xb + slo, sizeof(double) * bs, sizeof(double) * sbs,
                                          c.D, cudaMemcpyHostToDevice, h->stream));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(h->stage_y, yb + slo, sizeof(double) * sbs, cudaMemcpyHostToDevice, h->stream));
    }
    GPTB_TRY(gptb_launch_transpose(h, h->stage_cm, sbs, h->stage_rm));
    const double *nw = nullptr, *nU = nullptr;
    GPTB_TRY(stage_noise(h, noise_w, noise_U, &nw, &nU));
    h->stage_share = true;
    const int st_step = step_common(h, signal_var, nullptr, bs, N_total, true, h->stage_rm, h->stage_y, epsw, epsU, flags, nw, nU);
    h->stage_share = false;
    GPTB_TRY(st_step);
    GPTB_TRY(pipe_join(h));
    if (w_out) GPTB_CUDA_OK(h, cudaMemcpyAsync(w_out, h->w, sizeof(double) * c.Q, cudaMemcpyDeviceToHost, h->stream));
    if (flags & GPTB_FLAG_STIEFEL) return check_nan(h);
    GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
    return GPTB_OK;
}

static int pipe_wait_oldest(gptb_handle *h, int *step_status) {
    const int slot = (int)(h->pipe_waited & 1);
    GPTB_CUDA_OK(h, cudaEventSynchronize(h->pipe_done[slot]));
    ++h->pipe_waited;
    *step_status = GPTB_OK;
    if (h->pipe_nan_host[slot * 8 + 4]) {
        cudaMemsetAsync(h->nan_flag + 4, 0, sizeof(int32_t), h->stream);
        gptb_set_error(h, "a minibatch row lies outside the input domain of the node grid (gptb_set_domain / gptb_set_train)");
        *step_status = GPTB_ERR_ARG;
    } else if (h->pipe_stiefel[slot] && h->pipe_nan_host[slot * 8]) {
        gptb_set_error(h, "Get NaN when moving along Geodesic. Try smaller epsU");   // GPT_SGLD.jl:25
        *step_status = GPTB_GEODESIC_NAN;
    }
    return GPTB_OK;
}

int gptb_sgld_step_host_submit(gptb_handle *h, double signal_var, const double *Xb, const double *yb, int64_t bs,
                               int64_t N_total, double epsw, double epsU, int32_t flags, const double *noise_w,
                               const double *noise_U, double *w_out) {
    if (!h || !Xb || !yb || bs < 1 || N_total < 1 || !(signal_var > 0)) return GPTB_ERR_ARG;
    if (!h->have_features || !h->have_index || !h->have_params) {
        gptb_set_error(h, "features / index / parameters must be set before gptb_sgld_step_host_submit");
        return GPTB_ERR_STATE;
    }
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    const gptb_config &c = h->cfg;
    if (!h->copy_stream) {
        GPTB_CUDA_OK(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (int s = 0; s < 2; ++s) {
            GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->pipe_staged[s], cudaEventDisableTiming));
            GPTB_CUDA_OK(h, cudaEventCreateWithFlags(&h->pipe_done[s], cudaEventDisableTiming));
        }
        GPTB_CUDA_OK(h, cudaMallocHost((void **)&h->pipe_nan_host, 16 * sizeof(int32_t)));
        memset(h->pipe_nan_host, 0, 16 * sizeof(int32_t));
    }
    if (h->pipe_submitted - h->pipe_waited >= 2) {   // both slots busy: retire the oldest (status kept sticky on the device)
        int st_old;
        GPTB_TRY(pipe_wait_oldest(h, &st_old));
    }
    if (h->pipe_cap < bs || h->stage_cap < bs) {
        GPTB_CUDA_OK(h, cudaStreamSynchronize(h->stream));
        GPTB_CUDA_OK(h, cudaStreamSynchronize(h->copy_stream));
        for (int s = 0; s < 2; ++s) {
            GPTB_TRY(dev_alloc(h, &h->pipe_cm[s], (size_t)bs * c.D));
            GPTB_TRY(dev_alloc(h, &h->pipe_y[s], (size_t)bs));
        }
        GPTB_TRY(dev_alloc(h, &h->stage_cm, (size_t)bs * c.D));
        GPTB_TRY(dev_alloc(h, &h->stage_rm, (size_t)bs * c.D));
        GPTB_TRY(dev_alloc(h, &h->stage_y, (size_t)bs));
        h->pipe_cap = h->stage_cap = bs;
    }
    const int slot = (int)(h->pipe_submitted & 1);
    // the slot's previous occupant (two submissions ago) has been waited for or retired above, so the copy may start now
    // only this rank's rows [lo, hi) of the global minibatch cross the bus: one strided copy takes them out of the D
    // columns of the column-major Xb (with 8 ranks the full 5.8 MB batch per rank and step was the e2e bottleneck)
    const int64_t slo = (int64_t)h->rank * bs / h->nranks, shi = (int64_t)(h->rank + 1) * bs / h->nranks, sbs = shi - slo;
    if (sbs > 0) {
        GPTB_CUDA_OK(h, cudaMemcpy2DAsync(h->pipe_cm[slot], sizeof(double) * sbs, Xb + slo, sizeof(double) * bs,
                                          sizeof(double) * sbs, c.D, cudaMemcpyHostToDevice, h->copy_stream));
        GPTB_CUDA_OK(h, cudaMemcpyAsync(h->pipe_y[slot], yb + slo, sizeof(double) * sbs, cudaMemcpyHostToDevice, h->copy_stream));
    }
    GPTB_CUDA_OK(h, cudaEventRecord(h->pipe_staged[slot], h->copy_stream));
    {   // node-grid step: the batch's transpose, gather and cell sort follow the copy on the front stream, under the
        // previous step's tail; otherwise the main stream waits for the copy
        bool side = false;
        GPTB_TRY(front_begin(h, sbs, false, h->pipe_staged[slot], &side));
        if (!side) GPTB_CUDA_OK(h, cudaStreamWaitEvent(h->stream, h->pipe_staged[slot], 0));
        cudaStream_t keep = h->stream;
        if (side) h->stream = h->front_stream;
        const int st_tr = gptb_launch_transpose(h, h->pipe_cm[slot], sbs, h->stage_rm);
        h->stream = keep;
        if (st_tr != GPTB_OK) {
            h->front_active = false;
            return st_tr;
        }
    }
    const double *nw = nullptr, *nU = nullptr;
    GPTB_TRY(stage_noise(h, noise_w, noise_U, &nw, &nU));
    h->stage_share = true;
    const int st_step = step_common(h, signal_var, nullptr, bs, N_total, true, h->stage_rm, h->pipe_y[slot], epsw, epsU, flags, nw, nU);
    h->stage_share = false;
    h->front_active = false;
    GPTB_TRY(st_step);
    // the step's read-backs follow its last group update on the tail stream; the main stream is free for the next step
    cudaStream_t ts = nullptr;
    GPTB_TRY(pipe_tail_begin(h, &ts));
    if (w_out) GPTB_CUDA_OK(h, cudaMemcpyAsync(w_out, h->w, sizeof(double) * c.Q, cudaMemcpyDeviceToHost, ts));
    h->pipe_stiefel[slot] = (flags & GPTB_FLAG_STIEFEL) != 0;
    GPTB_CUDA_OK(h, cudaMemcpyAsync(h->pipe_nan_host + slot * 8, h->nan_flag, 8 * sizeof(int32_t), cudaMemcpyDeviceToHost, ts));
    GPTB_CUDA_OK(h, cudaEventRecord(h->pipe_done[slot], ts));
    GPTB_TRY(pipe_tail_end(h, ts));
    ++h->pipe_submitted;
    return GPTB_OK;
}

int gptb_sgld_step_host_wait(gptb_handle *h) {
    if (!h) return GPTB_ERR_ARG;
    if (h->pipe_submitted == h->pipe_waited) {
        gptb_set_error(h, "gptb_sgld_step_host_wait: no step in flight");
        return GPTB_ERR_STATE;
    }
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    int st;
    GPTB_TRY(pipe_wait_oldest(h, &st));
    return st;
}

int gptb_sgld_run(gptb_handle *h, double signal_var, int64_t m, double epsw, double epsU, int32_t burnin,
                  int32_t maxepoch, int32_t flags, const int64_t *perms, const double *noise_w, const double *noise_U,
                  double *w_store, double *U_store) {
    if (!h || m < 1 || burnin < 0 || maxepoch < 0 || !(signal_var > 0)) return GPTB_ERR_ARG;
    GPTB_TRY(check_ready(h, true, true));
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const gptb_config &c = h->cfg;
    const int64_t N = h->N;
    const int64_t numbatches = (N + m - 1) / m;  // :353
    const int64_t T = (int64_t)maxepoch * numbatches;
    const size_t ucnt = (size_t)c.n * c.r * c.D;
    const bool keep = (w_store || U_store) && T > 0;
    // device-side stores so the loop never waits on the host
    if (keep) {
        if (w_store) GPTB_TRY(dev_alloc(h, &h->store_w, (size_t)c.Q * T));
        if (U_store) {
            int st = dev_alloc(h, &h->store_U, ucnt * (size_t)T);
            if (st != GPTB_OK) {
                gptb_set_error(h, "U_store of %lld samples (%zu bytes) does not fit on the device", (long long)T,
                               ucnt * (size_t)T * 8);
                return st;
            }
        }
    }
    int64_t *perm_dev = nullptr;
    if (perms) GPTB_CUDA_OK(h, cudaMalloc((void **)&perm_dev, sizeof(int64_t) * N));
    // every call starts from the caller's data order, as each GPTregression call does (phi=phi[:,:,perm] rebinds a
    // local, GPT_SGLD.jl:374); gptb_sgld_advance is the entry that continues a chain's epochs
    int st = gptb_reset_order(h);
    h->batch_cursor = 0;
    if (h->rms_gw) {   // gw = zeros(Q); gU = zeros(n,r,D) at the start of every run (:1143-1144)
        cudaMemsetAsync(h->rms_gw, 0, sizeof(double) * c.Q, h->stream);
        cudaMemsetAsync(h->rms_gU, 0, sizeof(double) * ucnt, h->stream);
    }
    uint64_t step = 0;
    for (int epoch = 1; st == GPTB_OK && epoch <= burnin + maxepoch; ++epoch) {
        if (perms) {
            cudaMemcpyAsync(perm_dev, perms + (size_t)(epoch - 1) * N, sizeof(int64_t) * N, cudaMemcpyHostToDevice, h->stream);
            st = gptb_apply_perm(h, perm_dev);
        } else {
            st = gptb_shuffle_order(h, h->epoch_counter);
        }
        h->epoch_counter++;
        for (int64_t batch = 1; st == GPTB_OK && batch <= numbatches; ++batch, ++step) {
            const int64_t lo = m * (batch - 1), hi = std::min(m * batch, N);  // :379
            const double *nw = nullptr, *nU = nullptr;
            st = stage_noise(h, noise_w ? noise_w + (size_t)step * c.Q : nullptr,
                             noise_U ? noise_U + (size_t)step * ucnt : nullptr, &nw, &nU);
            if (st != GPTB_OK) break;
            h->front_dep_mode = batch > 1 ? 1 : 0;   // the epoch's row order is fixed once its permutation has run
            st = step_common(h, signal_var, h->order + lo, hi - lo, N, false, nullptr, nullptr, epsw, epsU, flags, nw, nU);
            h->front_dep_mode = 0;
            if (st != GPTB_OK) break;
            if (epoch > burnin && keep) {
                const int64_t col = (int64_t)(epoch - burnin - 1) * numbatches + batch - 1;  // :442-443
                cudaStream_t ts = nullptr;
                st = pipe_tail_begin(h, &ts);
                if (st != GPTB_OK) break;
                if (w_store) cudaMemcpyAsync(h->store_w + (size_t)col * c.Q, h->w, sizeof(double) * c.Q, cudaMemcpyDeviceToDevice, ts);
                if (U_store) cudaMemcpyAsync(h->store_U + (size_t)col * ucnt, h->U, sizeof(double) * ucnt, cudaMemcpyDeviceToDevice, ts);
                st = pipe_tail_end(h, ts);
                if (st != GPTB_OK) break;
            }
        }
        // the reference tests for the NaN abort after every geodesic (:422); the device flag is sticky and
        // freezes the parameters, so looking at it once per epoch returns the same all-zero result
        if (st == GPTB_OK && ((flags & GPTB_FLAG_STIEFEL) || perms)) st = check_nan(h);
    }
    if (pipe_join(h) != GPTB_OK && st == GPTB_OK) st = GPTB_ERR_CUDA;
    if (st == GPTB_OK) {
        cudaError_t e = cudaSuccess;
        if (keep && w_store) e = cudaMemcpyAsync(w_store, h->store_w, sizeof(double) * c.Q * T, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess && keep && U_store)
            e = cudaMemcpyAsync(U_store, h->store_U, sizeof(double) * ucnt * T, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) {
            gptb_set_error(h, "sgld_run: %s", cudaGetErrorString(e));
            st = GPTB_ERR_CUDA;
        }
    } else if (st == GPTB_GEODESIC_NAN) {
        cudaStreamSynchronize(h->stream);
        if (w_store) memset(w_store, 0, sizeof(double) * c.Q * T);       // :423
        if (U_store) memset(U_store, 0, sizeof(double) * ucnt * T);
    }
    if (perm_dev) cudaFree(perm_dev);
    dev_free(&h->store_w);
    dev_free(&h->store_U);
    return st;
}

int gptb_sgld_advance(gptb_handle *h, double signal_var, int64_t m, double epsw, double epsU, int32_t flags,
                      int64_t nsteps) {
    if (!h || m < 1 || nsteps < 0 || !(signal_var > 0)) return GPTB_ERR_ARG;
    GPTB_TRY(check_ready(h, true, true));
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(pipe_join(h));
    const int64_t N = h->N;
    const int64_t numbatches = (N + m - 1) / m;
    for (int64_t s = 0; s < nsteps; ++s) {
        if (h->batch_cursor == 0) {
            GPTB_TRY(gptb_shuffle_order(h, h->epoch_counter));   // perm=randperm(N), :373
            h->epoch_counter++;
        }
        const int64_t lo = m * h->batch_cursor, hi = std::min(m * (h->batch_cursor + 1), N);
        h->front_dep_mode = h->batch_cursor > 0 ? 1 : 0;
        const int st_step = step_common(h, signal_var, h->order + lo, hi - lo, N, false, nullptr, nullptr, epsw, epsU, flags,
                                        nullptr, nullptr);
        h->front_dep_mode = 0;
        GPTB_TRY(st_step);
        h->batch_cursor = (h->batch_cursor + 1) % numbatches;
    }
    GPTB_TRY(pipe_join(h));   // work queued on the caller's stream after this call sees the final parameters
    return GPTB_OK;
}

int gptb_set_option(gptb_handle *h, int32_t option, double value) {
    if (!h) return GPTB_ERR_ARG;
    cudaSetDevice(h->device);
    GPTB_TRY(pipe_join(h));
    switch (option) {
        case GPTB_OPT_PIPE_GROUPS:
            h->pipe_groups = std::max(0, std::min((int)value, (int)gptb_handle::PIPE_MAX));
            h->pipe_forced = true;
            return GPTB_OK;
        case GPTB_OPT_PIPE_RESERVE:
            h->pipe_reserve = (int)value;
            return GPTB_OK;
        case GPTB_OPT_PRED_LAYOUT:
            if (value != 0.0 && value != 1.0) break;
            h->pred_layout = (int)value;
            return GPTB_OK;
        case GPTB_OPT_PRED_WARPS:
            if (!(value >= 0.0 && value <= 16.0)) break;
            h->pred_warps = (int)value;
            return GPTB_OK;
        case GPTB_OPT_FRONT_PREFETCH:
            h->front_enabled = value != 0.0;
            return GPTB_OK;
        case GPTB_OPT_ALGORITHM:
            if (value != GPTB_ALG_AUTO && value != GPTB_ALG_DIRECT && value != GPTB_ALG_NODES) break;
            h->algorithm = (int)value;
            return GPTB_OK;
        case GPTB_OPT_NODE_ORDER:
            if (value < 4 || value > 16) break;
            h->node_P = (int)value;
            h->grid_valid = false;
            return GPTB_OK;
        case GPTB_OPT_PRED_CHUNK_MB:
            if (!(value >= 1.0 && value <= 65536.0)) break;
            h->pred_chunk_mb = (int)value;
            return GPTB_OK;
        case GPTB_OPT_RMS_ALPHA:
            if (!(value >= 0.0 && value < 1.0)) break;
            h->rms_alpha = value;
            return GPTB_OK;
        case GPTB_OPT_RMS_LAMBDA:
            if (!(value > 0.0)) break;
            h->rms_lambda = value;
            return GPTB_OK;
        case GPTB_OPT_NODE_THETA:
            if (!(value > 0.0) || value > 8.0) break;
            h->node_theta = value;
            h->grid_valid = false;
            return GPTB_OK;
        default:
            break;
    }
    gptb_set_error(h, "gptb_set_option: unknown option %d or value %g out of range", option, value);
    return GPTB_ERR_ARG;
}

int gptb_set_profiling(gptb_handle *h, int32_t on) {
    if (!h) return GPTB_ERR_ARG;
    cudaSetDevice(h->device);
    GPTB_TRY(pipe_join(h));   // the profiled step is the unpipelined one (phase timers live on the main stream)
    h->profiling = on != 0;
    return GPTB_OK;
}

int gptb_get_phase_ms(gptb_handle *h, float *ms_out, int64_t *launches_out, int32_t reset) {
    if (!h) return GPTB_ERR_ARG;
    cudaSetDevice(h->device);
    drain_timers(h);
    if (ms_out)
        for (int p = 0; p < GPTB_PHASE_COUNT; ++p) ms_out[p] = h->phase_ms[p];
    if (launches_out) *launches_out = h->launches;
    if (reset) {
        for (int p = 0; p < GPTB_PHASE_COUNT; ++p) h->phase_ms[p] = 0.f;
        h->launches = 0;
    }
    return GPTB_OK;
}

}  // extern "C"
