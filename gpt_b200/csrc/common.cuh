// Internal declarations shared by the translation units of libgptb200.so.
// Not part of the ABI (the ABI is include/gptb200.h).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/gptb200.h"

#define GPTB_CUDA_OK(h, expr)                                                                  \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            gptb_set_error((h), "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return GPTB_ERR_CUDA;                                                              \
        }                                                                                      \
    } while (0)

#define GPTB_TRY(expr)                 \
    do {                               \
        int _s = (expr);               \
        if (_s != GPTB_OK) return _s;  \
    } while (0)

void gptb_set_error(const gptb_handle *h, const char *fmt, ...);

// ------------------------------------------------------------------------------------------
// Device-resident model + scratch.  Layouts (device):
//   Zt, b      double [D][n]        (= Julia Z[j + n*k] order), Zt already divided by length_scale
//   ZtF, bF    float  [D][npad]     the same in FP32 for the tensor-core path, zero padded
//   I          int32  [D][Q]        0-based (= Julia I[q + Q*k] - 1)
//   X          double [N][D]        row-major rows (gather-friendly);  y double [N]
//   phi        double [N][D][n]     (= Julia phi[j + n*(k + D*i)]), compat path only
//   order      int32  [N]           current cumulative permutation (GPT_SGLD.jl:373-374)
//   w          double [Q];  U double [D][r][n]  (= Julia U[j + n*(l + r*k)])
//   batch scratch, capacity cap rows (mpad = cap rounded up to 128):
//     Xb   double [D][mpad]  / XbF float [D][mpad];  yb double [mpad]
//     T    real   [D][r][mpad]     T[k][l][i] = <phi_k(x_i), U^(k)[:,l]>            (phidotU, :193-205)
//     E    real   [D][r][mpad]     E[k][l][i] = e_i * A[l,k,i]                     (computeA :261-273 times residual)
//     fhat double [mpad]
//   grad   double [Q + n*r*D]      [gradw | gradU] in Julia order — the all-reduce buffer
// ------------------------------------------------------------------------------------------
struct gptb_pred_cache;   // pred_nodes.cu

// Node grid of the tensor-core path (nodes.cu): the features are generated at G*P Chebyshev nodes per dimension
// instead of at every minibatch row.
struct gptb_nodes {
    int G = 0, P = 0, m = 0, mnpad = 0;      // cells, nodes per cell, nodes per dimension (0 = grid not applicable), padded to 128
    double *lo = nullptr, *invh = nullptr;   // [D] lower end of the domain, 1 / cell half-width
    float *XnF = nullptr;                    // [D][mnpad] node coordinates (the "minibatch" K1 / K3 run on)
    float *Tn = nullptr;                     // [D][r][mnpad] K1 output on the nodes
    float *Cf = nullptr;                     // [D][m][rpad] Chebyshev coefficients per cell
    float *Mn = nullptr;                     // [D][mnpad][rpad] moment / node-weight table of the backward pass
    void *Enimg = nullptr;                   // fp16 hi|lo operand image of the node weights (K3's B operand)
    int32_t *cell = nullptr;                 // [D][mpad] cell of every batch row
    float *uu = nullptr;                     // [D][mpad] position inside the cell, in [-1, 1]
    int32_t *srow = nullptr, *sc = nullptr;  // [D][mpad] records (row, cell, u) of the batch in cell order (inside chunks)
    float *su = nullptr;
    float *Erow = nullptr;                   // [D][mpad][rpad] FP32 E as K2 writes it for the spreading kernel
    float *T2 = nullptr;                     // [D][mpad][rpad] interpolated T, row-major per sample (K2 transposes its tile)
    int64_t batch_mpad = 0;
    uint64_t steps = 0;
};

struct gptb_handle {
    gptb_config cfg;
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    mutable std::string err;

    // features
    bool have_features = false, have_index = false, have_train = false, have_phi = false, have_params = false;
    double amp = 0.0;
    double *Zt = nullptr, *b = nullptr;
    float *ZtF = nullptr, *bF = nullptr;
    int npad = 0;   // n rounded up to 128 (tensor-core M tile of the backward pass / K chunks of the forward)
    int rpad = 0;   // r rounded up to 16 (UMMA N granularity at M=128)
    int32_t *I = nullptr;
    int32_t *grp_ptr = nullptr;   // [D*rpad+1] CSR over (k,l): which q have I[q,k] = l (contract.cu pass 2)
    uint16_t *grp_q = nullptr;    // [Q*D]
    int32_t *ct_off1 = nullptr, *ct_lptr = nullptr, *ct_lists = nullptr;   // offset tables of contract_tc.cu
    int ct_nlists = 0;

    // training data
    int64_t N = 0;
    double *X = nullptr, *y = nullptr, *phi = nullptr;
    int32_t *order = nullptr;
    int32_t *order_tmp = nullptr;
    uint64_t *sort_keys = nullptr, *sort_keys_tmp = nullptr;
    void *sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0;

    // parameters
    double *w = nullptr, *U = nullptr;
    float *wF = nullptr;
    void *Uimg = nullptr;        // tensor-core operand image of U (fp16 hi|lo, UMMA canonical), see tc_layout.cuh
    size_t Uimg_bytes = 0;
    bool shadows_valid = false;
    double *pw = nullptr, *pU = nullptr;   // explicit parameters of pred / pred_samples calls
    void *pUimg = nullptr;
    void *ZbF = nullptr;         // float2 [D][npad] interleaved (Zt, b) for the forward generators
    bool ZbF_valid = false;
    double *store_w = nullptr, *store_U = nullptr;   // device-side chain stores of gptb_sgld_run
    double *stage_cm = nullptr, *stage_rm = nullptr, *stage_y = nullptr;  // host-minibatch staging
    int64_t stage_cap = 0;
    bool stage_share = false;          // the staged host minibatch holds this rank's rows only (gptb_sgld_step_host_submit)
    // pipelined host steps (gptb_sgld_step_host_submit / _wait): two staging slots filled on a copy stream
    cudaStream_t copy_stream = nullptr;
    double *pipe_cm[2] = {nullptr, nullptr}, *pipe_y[2] = {nullptr, nullptr};
    int64_t pipe_cap = 0;
    cudaEvent_t pipe_staged[2] = {nullptr, nullptr}, pipe_done[2] = {nullptr, nullptr};
    int32_t *pipe_nan_host = nullptr;   // pinned, one flag per slot
    uint64_t pipe_submitted = 0, pipe_waited = 0;
    bool pipe_stiefel[2] = {false, false};

    // batch scratch
    int64_t cap = 0, mpad = 0;
    double *Xb = nullptr, *yb = nullptr, *fhat = nullptr;
    float *XbF = nullptr;
    void *T = nullptr, *E = nullptr;   // real = double (FP64) or float (TC)
    void *Eimg = nullptr;              // fp16 hi|lo image of E for the backward MMA (TC)
    int32_t *idx_dev = nullptr;        // explicit batch indices
    double *grad = nullptr;            // [Q + n r D]
    float *gradF = nullptr;            // FP32 partial accumulators of the TC path: [D][npad][rpad], = gradFbase + qpad
    float *gradFbase = nullptr;        // [qpad | D npad rpad]: the first qpad floats take gradw in the FP32 gradient flow
    int gradF_qpad = 0;                // = GPTB_GW_SLOTS * gradF_qstride
    int gradF_qstride = 0;             // Q rounded up to 32
    bool f32_grad = false;             // set for the duration of a step: K2 sums gradw into gradFbase, K3's sums are not
                                       // folded into the FP64 buffer, the all-reduce and the update work on gradFbase
    void *tc_partials = nullptr;       // deterministic split-row partials of the backward MMA
    size_t tc_partials_bytes = 0;
    float *tc_state = nullptr;         // [0] max|E| of the running step, [1] escale (2^k), [2] 1/escale,
                                       // [3] uscale (2^k applied to U in Uimg), [4] 1/uscale
    bool tc_emax_valid = false;
    bool gather_cells = false;         // set around a node-grid step: the gather also writes nodes.cell / nodes.uu
    // front of the node-grid step (transpose of a staged host batch, gather, cell sort) on its own stream: it depends on
    // the minibatch rows only, so it runs under the previous step's back-projection tail, all-reduce and update
    cudaStream_t front_stream = nullptr;
    cudaEvent_t ev_front_dep = nullptr, ev_front_done = nullptr, ev_rows_free = nullptr;
    bool front_enabled = true;         // GPTB_OPT_FRONT_PREFETCH
    bool front_active = false;         // between front_begin and the end of the step's front
    bool rows_free_valid = false;      // ev_rows_free has been recorded (after the last reader of the batch buffers)
    int front_dep_mode = 0;            // 0: the front follows everything queued so far; 1: the rows need nothing new
    int k2_redo = 0;                   // set around a second K2 launch that only rewrites E: 1 = if the scale guard asks
                                       // for it (api.cu), 2 = unconditionally (RMSprop sampler: E with the updated w)
    const double *k2_wA = nullptr;     // weights of computeA when they differ from w
    // RMSprop sampler (GPT_SGLDERM_RMSprop, GPT_SGLD.jl:1121-1237)
    double rms_alpha = 0.9, rms_lambda = 1e-5;
    double *rms_gw = nullptr, *rms_gU = nullptr;   // moving averages of the squared unnormalised gradients [Q], [D][r][n]
    double *rms_epsk = nullptr;        // [D] sum over the elements of U^(k) of epsilon/(sqrt(gU)+lambda)
    double *w_old = nullptr;           // w before the step's update (fhat and the residual use it)
    bool rms_active = false;           // set for the duration of an RMSprop step
    double rms_eps = 0.0, rms_N = 1.0;
    const double *rms_noise_w = nullptr;
    int rms_flags = 0;
    bool k2_T_rowmajor = false;        // K2 reads T as [D][mpad][rpad] (nodes.cu) instead of [D][r][mpad]
    float *k2_Erow = nullptr;          // when set, K2 writes plain FP32 E [D][mpad][rpad] instead of the operand image

    // update scratch
    double *G = nullptr;               // [D][r][n]
    double *small = nullptr;           // per-k r x r reductions and expm workspace
    size_t small_stride = 0;           // doubles per k
    double *colnorm = nullptr;         // [D][r]
    int32_t *nan_flag = nullptr;       // device, sticky
    int32_t *nan_flag_host = nullptr;  // pinned
    double *noise_w = nullptr, *noise_U = nullptr;  // staged host-supplied noise
    double *gen_noise_w = nullptr, *gen_noise_U = nullptr;   // production-mode draws of the running step (k_noise)

    // dimension-group pipeline of the SGLD step (api.cu: step_pipelined): the update of group g runs on side[g] while
    // the main stream goes on with the next group's back-projection / the next step's forward pass
    static constexpr int PIPE_MAX = 4;
    int pipe_groups = -1;              // -1: not decided yet; 0/1: off
    bool pipe_forced = false;          // GPTB_OPT_PIPE_GROUPS set: use the pipeline for every eligible step
    int pipe_reserve = -1;             // GPTB_OPT_PIPE_RESERVE: SMs left to the concurrent update (-1 = one per dimension of a group)
    // node-grid algorithm (nodes.cu)
    int algorithm = GPTB_ALG_AUTO;
    int node_P = 12;                   // Chebyshev nodes per cell
    double node_theta = 2.0;           // cell half-width times the largest |frequency| of the dimension
    bool grid_valid = false;
    gptb_nodes nodes;
    int pred_layout = 0;               // GPTB_OPT_PRED_LAYOUT: 0 samples on the lanes (contract_fwd3), 1 row tiles (contract_fwd2)
    int pred_warps = 0;                // GPTB_OPT_PRED_WARPS: warps per CTA of contract_fwd3 (0 = from the row size)
    int pred_chunk_mb = 4096;          // T2 budget of the posterior sweep (GPTB_OPT_PRED_CHUNK_MB), see pred_nodes.cu
    gptb_pred_cache *pred_cache = nullptr;   // buffers of the node-grid posterior sweep (pred_nodes.cu), kept between calls
    bool have_domain = false;          // input domain known: from gptb_set_train (min / max of X) or gptb_set_domain
    double dom_lo[16] = {0}, dom_hi[16] = {0};
    double zmax[16] = {0};             // max_j |Zt[j,k]|
    cudaStream_t side[PIPE_MAX] = {nullptr, nullptr, nullptr, nullptr};
    cudaStream_t tail_stream = nullptr, ar_stream = nullptr;
    cudaEvent_t evG[PIPE_MAX] = {}, evU[PIPE_MAX] = {}, evR[PIPE_MAX] = {};
    cudaEvent_t evTail = nullptr, evTailFork = nullptr;
    bool pendU[PIPE_MAX] = {false, false, false, false};
    bool tail_pending = false;
    int sm_reserve = 0;                // SMs the persistent K1 / K3 grids leave to the concurrent group update

    // multi-sample prediction scratch
    double *pred_acc = nullptr;
    int64_t pred_cap = 0;

    // host staging
    double *pin = nullptr;
    size_t pin_bytes = 0;

    // comm
    void *nccl_comm = nullptr;
    int nranks = 1, rank = 0;

    // bookkeeping
    uint64_t step_counter = 0;
    uint64_t epoch_counter = 0;
    bool profiling = false;
    float phase_ms[GPTB_PHASE_COUNT] = {0, 0, 0, 0, 0, 0};
    int64_t launches = 0;
    // profiling: events are recorded without synchronising; gptb_get_phase_ms drains them
    std::vector<cudaEvent_t> evpool;
    std::vector<int> evtag;            // -1 = step start, else phase id
    size_t ev_used = 0;
    int64_t batch_cursor = 0;          // position inside the current epoch (gptb_sgld_advance)
    int sm_count = 148;
};

inline bool gptb_is_tc(const gptb_handle *h) { return h->cfg.precision != GPTB_PREC_FP64; }

// gradw of the FP32 gradient flow is summed into GPTB_GW_SLOTS copies (CTA b of K2 adds to copy b % SLOTS) that the
// update adds up: 1024 CTAs x 200 atomics on the 7 cache lines of one copy serialise in the L2 atomic unit.
constexpr int GPTB_GW_SLOTS = 32;

// ------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011) — counter-based, so every rank / every CTA draws the same
// number for the same (key, counter) regardless of GPU count or launch geometry.
// ------------------------------------------------------------------------------------------
struct philox4 {
    uint32_t x, y, z, w;
};

__host__ __device__ inline philox4 philox4x32_10(uint64_t key, uint64_t ctr_lo, uint64_t ctr_hi) {
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
    uint32_t k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return philox4{c0, c1, c2, c3};
}

// Two independent N(0,1) doubles from one Philox block (Box-Muller on two 53-bit uniforms... 32+21 bits each).
__host__ __device__ inline void philox_normal2(uint64_t key, uint64_t ctr_lo, uint64_t ctr_hi, double *a, double *b) {
    philox4 p = philox4x32_10(key, ctr_lo, ctr_hi);
    // u1 in (0,1], u2 in [0,1)
    double u1 = ((double)(((uint64_t)p.x << 21) | (p.y >> 11)) + 1.0) * (1.0 / 9007199254740992.0);
    double u2 = (double)(((uint64_t)p.z << 21) | (p.w >> 11)) * (1.0 / 9007199254740992.0);
    double rad = sqrt(-2.0 * log(u1));
    double s, c;
#ifdef __CUDA_ARCH__
    sincospi(2.0 * u2, &s, &c);
#else
    s = sin(6.283185307179586 * u2);
    c = cos(6.283185307179586 * u2);
#endif
    *a = rad * c;
    *b = rad * s;
}

// RNG stream ids (upper counter word): keeps noise for w, U^(k), permutations and init apart.
enum : uint64_t {
    GPTB_STREAM_NOISE_W = 1ull << 56,
    GPTB_STREAM_NOISE_U = 2ull << 56,   // + k << 40
    GPTB_STREAM_PERM = 3ull << 56,
    GPTB_STREAM_INIT_W = 4ull << 56,
    GPTB_STREAM_INIT_U = 5ull << 56
};

// ------------------------------------------------------------------------------------------
// Kernel launchers (defined in the .cu files named on the right).  All enqueue on h->stream.
// ------------------------------------------------------------------------------------------
// fp64_path.cu
int gptb_launch_features_fp64(gptb_handle *h, const double *Xrm_dev /*[N][D]*/, int64_t N, double *phi_dev);
int gptb_launch_fwd_fp64(gptb_handle *h, const double *Xb /*[D][mpad]*/, const double *phib /*[bs][D][n] or null*/,
                         const int32_t *rows /*phi row ids or null*/, const double *U, int64_t bs, int64_t mpad, double *T);
int gptb_launch_bwd_fp64(gptb_handle *h, const double *Xb, const double *phib, const int32_t *rows, const double *E,
                         int64_t bs, int64_t mpad, double *gradU /*[D][r][n], accumulated*/);
// contract.cu
int gptb_launch_contract(gptb_handle *h, const void *T, const void *w, const double *yb, int64_t bs, int64_t mpad,
                         double escale, bool want_grad, double *fhat, double *gradw /*accumulated*/, void *E, void *Eimg,
                         bool real_is_double);
// contract_tc.cu
int gptb_contract_tc_build_tables(gptb_handle *h, const int32_t *Iz /*[D][Q] 0-based, host*/);
int gptb_launch_contract_tc(gptb_handle *h, const float *T, const double *w, const double *yb, int64_t bs, int64_t mpad,
                            double escale, bool want_grad, double *fhat, double *gradw, void *Eimg);
// update.cu
int gptb_launch_update(gptb_handle *h, double epsw, double epsU, int flags, const double *noise_w,
                       const double *noise_U, uint64_t step);
constexpr int GPTB_IFLAG_PRIOR_IN_GRAD = 1 << 16;   // internal update flag: grad[0..Q) already holds gradw - w/sigma_w^2
int gptb_launch_update_range(gptb_handle *h, double epsw, double epsU, int flags, const double *noise_w,
                             const double *noise_U, uint64_t step, int k0, int kn, bool with_w, cudaStream_t s,
                             bool noise_ready = false);
int gptb_launch_noise_range(gptb_handle *h, uint64_t step, int k0, int kn, bool with_w, cudaStream_t s);
int gptb_update_alloc_noise(gptb_handle *h);
int gptb_launch_rms_w(gptb_handle *h, int flags, const double *noise_w, uint64_t step, cudaStream_t s);
int gptb_launch_rms_U(gptb_handle *h, cudaStream_t s);
int gptb_proj_geod_device(gptb_handle *scratch_owner, cudaStream_t s, int n, int r, const double *U, const double *V,
                          double t, double *out, int mode /*0 proj, 1 geod*/, int32_t *nan_flag,
                          double *out2 = nullptr /*geodboth: the transported momentum*/);
// batch.cu
int gptb_launch_gather(gptb_handle *h, const int32_t *rows, int64_t start, int64_t bs, int64_t row_off);
// (both gathers also emit the node-grid cell / offset of every row when h->gather_cells is set)
int gptb_shuffle_order(gptb_handle *h, uint64_t epoch);
// tc_path.cu
int gptb_tc_pack_params(gptb_handle *h);
int gptb_tc_pack_U(gptb_handle *h, const double *U, void *Uimg);
int gptb_tc_update_escale(gptb_handle *h);
int gptb_tc_check_escale(gptb_handle *h);
int gptb_launch_fwd_tc(gptb_handle *h, const float *XbF, const void *Uimg, int64_t bs, int64_t mpad, float *T);
int gptb_launch_fwd_tc_cols(gptb_handle *h, const float *XbF, const void *Uimg, const float *uscale, int64_t bs,
                            int64_t mpad, float *T, int ncols, int NC);
int gptb_tc_pred_scale(gptb_handle *h, const double *U, size_t cnt, bool first);
int gptb_tc_pred_scale_commit(gptb_handle *h);
int gptb_tc_pack_group(gptb_handle *h, const double *Ug, int Sg, int NC, void *img);
int gptb_tc_pack_group_scaled(gptb_handle *h, const double *Ug, int Sg, int NC, void *img, float *state, cudaStream_t s);
int gptb_launch_contract_fwd2(gptb_handle *h, const float *T2, const double *w_group, int ns, int col0, int ncols, int64_t nrows,
                              int64_t t2_stride, double *fh /*[ns][fh_stride] or null*/, int64_t fh_stride,
                              double *mean_acc /*[nrows] or null*/, double mean_scale, cudaStream_t s);
// pred_nodes.cu
int gptb_pred_nodes(gptb_handle *h, const double *w_store, const double *U_store, int32_t S, const double *Xtest, int64_t Nt,
                    double *mean_out, double *per_sample_out);
void gptb_pred_nodes_free(gptb_handle *h);
int gptb_launch_contract_fwd(gptb_handle *h, const float *T, const double *w_group, int Sg, int64_t bs, int64_t mpad,
                             int ncols_total, double *fh_group /*[Sg][mpad] or null*/, double *mean_acc /*[mpad] or null*/,
                             double mean_scale);
int gptb_launch_bwd_tc(gptb_handle *h, const float *XbF, const void *Eimg, int64_t bs, int64_t mpad, double *gradU);
// the same for the dimensions [k0, k0 + kn) only (all pointers are those of the full problem)
int gptb_launch_bwd_tc_range(gptb_handle *h, const float *XbF, const void *Eimg, int64_t bs, int64_t mpad, double *gradU,
                             int k0, int kn);
int gptb_launch_fwd_tc_range(gptb_handle *h, const float *XbF, const void *Uimg, int64_t bs, int64_t mpad, float *T, int k0,
                             int kn);
size_t gptb_tc_uimg_bytes(const gptb_handle *h);
size_t gptb_tc_eimg_bytes(const gptb_handle *h, int64_t mpad);
// nodes.cu
void gptb_nodes_free(gptb_handle *h);
int gptb_nodes_setup(gptb_handle *h);
int gptb_nodes_ensure_batch(gptb_handle *h);
int gptb_nodes_forward_range(gptb_handle *h, const void *Uimg, int k0, int kn, cudaStream_t s);
int gptb_nodes_interp(gptb_handle *h, int64_t bs);
int gptb_nodes_launch_interp(gptb_handle *h, int P, const float *Cf, const int32_t *srow, const int32_t *sc, const float *su,
                             float *T2, int ctot, int col0, int ncols, int G, int64_t row0, int64_t nrows, int64_t rec_stride,
                             int64_t t2_stride, cudaStream_t s);
int gptb_nodes_launch_coef_lanes(gptb_handle *h, const float *Tn, float *Cf, int G, int P, int ns, int tn_cols, int SL, int s0,
                                 int mnpad, cudaStream_t s);
int gptb_launch_contract_fwd3(gptb_handle *h, const float *T2, const double *w_group, int ns, int SL, int64_t nrows,
                              int64_t t2_stride, double *fh, int64_t fh_stride, double *mean_acc, double mean_scale,
                              cudaStream_t s);
int gptb_nodes_launch_coef_cols(gptb_handle *h, const float *Tn, float *Cf, int G, int P, int ncols, int tn_cols, int ctot,
                                int col0, int mnpad, cudaStream_t s);
int gptb_nodes_upload_cheb(gptb_handle *h, int P);
int gptb_nodes_launch_sort(gptb_handle *h, const int32_t *cell, const float *uu, int32_t *srow, int32_t *sc, float *su, int G,
                           int64_t bs, int64_t stride, cudaStream_t s);
int gptb_nodes_sort(gptb_handle *h, int64_t bs, cudaStream_t s);
int gptb_nodes_spread(gptb_handle *h, int64_t bs);
int gptb_nodes_backward_range(gptb_handle *h, double *gradU, int k0, int kn);
// comm.cu
int gptb_allreduce_grad(gptb_handle *h);
int gptb_allreduce_range(gptb_handle *h, double *buf, size_t count, cudaStream_t stream);
int gptb_allreduce_range_f32(gptb_handle *h, float *buf, size_t count, cudaStream_t stream);
int gptb_tc_ensure_gradF(gptb_handle *h);
bool gptb_contract_tc_usable(const gptb_handle *h);
