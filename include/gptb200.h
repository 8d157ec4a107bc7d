/*
 * gptb200.h — C ABI of libgptb200.so: the B200-native hot path of hyunjik11/GPT
 * (tensor-structured random-Fourier-feature GP regression sampled by SGLD on a Stiefel manifold).
 *
 * The reference has no FFI of its own: its boundary is the exported function set of the Julia
 * module GPT_SGLD (/root/reference/GPT_SGLD.jl:5).  Each entry point below names the reference
 * function (file:line) it replaces; julia/GPT_SGLD.jl and gpt_b200/gpt_sgld.py are the host
 * modules that keep the reference's names and argument order and `ccall`/ctypes into this ABI.
 *
 * Conventions
 *   - Plain C, no exceptions.  Every call returns an int status (GPTB_OK == 0).
 *   - All array arguments are HOST pointers owned by the caller, Julia layout: column-major
 *     Float64 (first index fastest); the core index table I is Int32, 1-based, I[q + Q*k].
 *   - The handle owns all device memory, streams and communicators.  A handle is not
 *     thread-safe; different handles are independent.
 *   - There is no CPU fallback: without a CUDA device every compute call fails with
 *     GPTB_ERR_CUDA.
 */
#ifndef GPTB200_H
#define GPTB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPTB_VERSION 100

typedef struct gptb_handle gptb_handle;

enum gptb_status {
    GPTB_OK = 0,
    GPTB_ERR_ARG = 1,         /* NULL pointer / value out of range                              */
    GPTB_ERR_SHAPE = 2,       /* inconsistent sizes (e.g. I entry outside 1..r, odd n for sin/cos) */
    GPTB_ERR_STATE = 3,       /* call order (features / index / train / params not set yet)       */
    GPTB_ERR_CUDA = 4,        /* CUDA runtime error, or no device                                */
    GPTB_ERR_NCCL = 5,        /* NCCL error / library not found                                   */
    GPTB_ERR_NOMEM = 6,
    GPTB_GEODESIC_NAN = 7,    /* expm produced NaN: GPT_SGLD.jl:24-26 prints the warning and returns
                                 zeros(n,r); :422-424 then returns all-zero stores.  Outputs are zero-filled. */
    GPTB_ERR_UNSUPPORTED = 8
};

enum gptb_precision {
    GPTB_PREC_FP64 = 0,  /* FP64 CUDA cores end to end (validation mode, <=1e-10 vs the reference)       */
    GPTB_PREC_TC3 = 1,   /* tcgen05 tensor cores, operands split hi+lo (3 MMA passes), FP32 accumulate;
                            <=1e-4 relative vs the reference.  Parameters stay FP64, update in FP64.    */
    GPTB_PREC_TC1 = 2    /* tcgen05 single pass (hi only): fastest, ~1e-3 relative (not a parity mode)   */
};

enum gptb_feature_kind {
    GPTB_FEAT_COS = 0,    /* feature  GPT_SGLD.jl:71-84 : amp*cos(x*Zt+b)                */
    GPTB_FEAT_SINCOS = 1  /* feature2 GPT_SGLD.jl:88-105: amp2*{sin,cos}(x*Zt), n even   */
};

enum gptb_flags {
    GPTB_FLAG_LANGEVIN = 1, /* keyword langevin of GPTregression (GPT_SGLD.jl:345)        */
    GPTB_FLAG_STIEFEL = 2,  /* keyword stiefel                                             */
    GPTB_FLAG_FIX_U = 4,    /* GPT_SGLDERMw (GPT_SGLD.jl:1065-1118): SGLD on w only, the factors U stay as they are
                               (no factor gradient, no U update; U_store may be NULL)      */
    GPTB_FLAG_RMSPROP = 8,  /* GPT_SGLDERM_RMSprop (GPT_SGLD.jl:1121-1237), with GPTB_FLAG_STIEFEL: `epsw` is the sampler's
                               single step size epsilon (`epsU` is ignored); every w[q] moves with epsilon/(sqrt(gw[q])+lambda),
                               every U^(k) with the mean of its elements' step sizes (:1212); gw, gU are moving averages
                               (coefficient alpha: GPTB_OPT_RMS_ALPHA, default 0.9; lambda: GPTB_OPT_RMS_LAMBDA, 1e-5 as :1146)
                               of the squared gradients scaled by 1/batch_size (:1172), zero at the start of every
                               gptb_sgld_run and after gptb_set_params; computeA sees the updated w (:1188 before :1194) */
    GPTB_FLAG_CLS_SINGLE_UPDATE = 16 /* gptb_classify_run only: skip the first of GPTclassification's two moves per
                               minibatch (GPT_SGLD.jl:624-636), i.e. update like the regression samplers do */
};

typedef struct gptb_config {
    int32_t D;            /* input dimensions                                   */
    int32_t n;            /* random Fourier features per dimension              */
    int32_t r;            /* Tucker rank (columns of each U^(k))                */
    int32_t Q;            /* non-zeros of the core w                            */
    int32_t precision;    /* enum gptb_precision                                */
    int32_t feature_kind; /* enum gptb_feature_kind                             */
    int32_t device;       /* CUDA device ordinal, -1 = current device           */
    int32_t reserved;
    uint64_t seed;        /* Philox key for production-mode noise and permutations */
} gptb_config;

/* ---- lifetime ------------------------------------------------------------------------------ */
int gptb_version(void);
int gptb_create(const gptb_config *cfg, gptb_handle **out);
int gptb_destroy(gptb_handle *h);
/* Message of the last failing call on h (h == NULL: last failing gptb_create / handle-less call). */
const char *gptb_last_error(const gptb_handle *h);
/* Run all of this handle's work on an existing cudaStream_t (pass it as void*); NULL = own stream. */
int gptb_set_stream(gptb_handle *h, void *cuda_stream);
/* Blocks until all work queued by this handle has finished; returns GPTB_GEODESIC_NAN if a queued step hit it. */
int gptb_synchronize(gptb_handle *h);

/* ---- model definition ---------------------------------------------------------------------- */
/* Random-feature draws and hyper-parameters of `feature` (GPT_SGLD.jl:71-84) / `feature2` (:88-105).
 * Z[n,D]; b[n,D] (ignored, may be NULL, for GPTB_FEAT_SINCOS); length_scale has ls_len = 1 or D
 * entries (Zt = scale(Z, 1./length_scale), :75).  The amplitude phi_scale*sigma_RBF^(1/D)*sqrt(2/n)
 * (:83; sqrt(1/(n/2)) for feature2, :104) is applied inside the kernels.
 * TGP.feature (TGP.jl:6-14) is this call with length_scale=sigmaRBF, sigma_RBF=1, phi_scale=1, b=2*pi*rand. */
int gptb_set_features(gptb_handle *h, const double *Z, const double *b, const double *length_scale,
                      int32_t ls_len, double sigma_RBF, double phi_scale);
/* Core sparsity pattern produced by `samplenz` (GPT_SGLD.jl:181-190): I[Q,D], 1-based. */
int gptb_set_index(gptb_handle *h, const int32_t *I);
/* Training set: raw inputs X[N,D] (features are generated on the fly, never stored) and y[N]. */
int gptb_set_train(gptb_handle *h, const double *X, const double *y, int64_t N);
/* Input domain of the node grid (nodes.cu): lo[D] <= hi[D] bound every input the handle will see.  gptb_set_train
 * sets it to the column ranges of X; callers that only pass host minibatches (gptb_sgld_step_host*) set it here.
 * NULL, NULL forgets the domain (direct evaluation at every row).  A row outside the domain makes the step's next
 * synchronising call return GPTB_ERR_ARG. */
int gptb_set_domain(gptb_handle *h, const double *lo, const double *hi);
/* Compatibility: a materialised phi[n,D,N] exactly as `feature` returns it (GPT_SGLD.jl:345 takes phi). */
int gptb_set_train_phi(gptb_handle *h, const double *phi, const double *y, int64_t N);

/* ---- parameters ---------------------------------------------------------------------------- */
int gptb_set_params(gptb_handle *h, const double *w /*Q*/, const double *U /*n,r,D*/);
int gptb_get_params(gptb_handle *h, double *w, double *U);
/* GPT_SGLD.jl:357-369: w = randn(Q); U^(k) = ((ZZ')^(-1/2) Z)' with Z = randn(r,n) (stiefel != 0) or
 * randn(n,r,D)/sqrt(n).  Julia's dSFMT stream is not reproducible; draws come from Philox(param_seed). */
int gptb_init_params(gptb_handle *h, uint64_t param_seed, int32_t stiefel);

/* ---- features / prediction ----------------------------------------------------------------- */
/* `feature` / `feature2` materialised: phi_out[n,D,N] (GPT_SGLD.jl:71-105). */
int gptb_features(gptb_handle *h, const double *X, int64_t N, double *phi_out);
/* `pred(w,U,I,phitest)` (GPT_SGLD.jl:233-243) with phitest = feature(Xtest,...) generated on the fly. */
int gptb_pred(gptb_handle *h, const double *w, const double *U, const double *Xtest, int64_t Nt,
              double *fhat_out);
/* Same with a materialised phitest[n,D,Nt]. */
int gptb_pred_phi(gptb_handle *h, const double *w, const double *U, const double *phitest, int64_t Nt,
                  double *fhat_out);
/* Posterior-predictive sweep (kin40kExperiment.jl:75-87, GPT_SGLD_p.jl:124-132): pred for S stored
 * samples w_store[Q,S], U_store[n,r,D,S]; mean_out[Nt] = average over samples; per_sample_out[Nt,S]
 * optional (NULL to skip). */
int gptb_pred_samples(gptb_handle *h, const double *w_store, const double *U_store, int32_t S,
                      const double *Xtest, int64_t Nt, double *mean_out, double *per_sample_out);

/* ---- gradients ----------------------------------------------------------------------------- */
/* Minibatch gradients of the log posterior at the handle's current (w,U), GPT_SGLD.jl:384-408:
 *   gradw = (N/bs) V (y_b - fhat)/signal_var [- w]      (:393, the prior term when include_prior != 0)
 *   gradU[:,:,k] = reshape((N/bs) Psi[:,:,k] (y_b - fhat)/signal_var, n, r)            (:405-408)
 * Batch rows: idx[bs] (0-based rows of the training set) or, when idx == NULL, the contiguous range
 * [start, start+bs) of the current (permuted) order.  fhat_out[bs] optional. */
int gptb_grad(gptb_handle *h, double signal_var, const int64_t *idx, int64_t start, int64_t bs,
              double *gradw_out, double *gradU_out, double *fhat_out, int32_t include_prior);

/* ---- Stiefel manifold ---------------------------------------------------------------------- */
/* proj(U,V) = V - U(U'V+V'U)/2, GPT_SGLD.jl:14-16.  Uses device `device` (-1 = current). */
int gptb_proj(int32_t device, int32_t n, int32_t r, const double *U, const double *V, double *out);
/* proj(U,V) without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernels and result as gptb_proj, but:
 *   - U_dev, V_dev, out_dev are DEVICE pointers (n x r, column-major Float64) on the calling thread's CURRENT device,
 *     and the work is only queued on cuda_stream (a cudaStream_t passed as void*; NULL = the default stream): the
 *     call does not synchronise and calls no cudaMalloc / cudaFree / allocating library routine;
 *   - the r x r Gram scratch is a static device array inside the library (98,816 bytes held by every GPU that loads
 *     the library, until exit), so r <= 64; a larger r returns GPTB_ERR_UNSUPPORTED with a message (use gptb_proj);
 *   - every call shares that array: it is cleared at the start of each call, and two calls whose work OVERLAPS on
 *     the GPU (different streams, or concurrent host threads) race — queue them on one stream or order them with
 *     events.  out_dev must not alias U_dev or V_dev. */
int gptb_proj_noalloc(int32_t n, int32_t r, const double *U_dev, const double *V_dev, double *out_dev,
                      void *cuda_stream);
/* geod(U,mom,t), GPT_SGLD.jl:19-37.  Returns GPTB_GEODESIC_NAN and zeros(n,r) when expm gives NaN. */
int gptb_geod(int32_t device, int32_t n, int32_t r, const double *U, const double *mom, double t, double *out);

/* geodboth(U,mom,t), GPT_SGLD.jl:40-59: the geodesic point U_out (columns renormalised, :51-56) and the transported
 * momentum mom_out = [U mom] E[:, r+1:2r] expm(-tA) (:50, returned as it is).  GPTB_GEODESIC_NAN and zeros for both
 * when expm gives NaN (:43-45). */
int gptb_geodboth(int32_t device, int32_t n, int32_t r, const double *U, const double *mom, double t, double *U_out,
                  double *mom_out);

/* ---- sampler ------------------------------------------------------------------------------- */
/* One loop body of GPTregression (GPT_SGLD.jl:378-439) on the resident training set.
 * noise_w[Q], noise_U[n,r,D]: the randn draws of :412/:420 supplied by the caller (test mode);
 * NULL = draw from Philox(seed, step).  Returns GPTB_GEODESIC_NAN if the geodesic failed: the factors are
 * frozen from then on (the flag is sticky; every later update is skipped), w has taken this step's update —
 * the reference likewise updates w (:412) before it tests the geodesic (:422) and then discards the chain. */
int gptb_sgld_step(gptb_handle *h, double signal_var, const int64_t *idx, int64_t start, int64_t bs,
                   double epsw, double epsU, int32_t flags, const double *noise_w, const double *noise_U);
/* Same loop body with the minibatch passed from the host: Xb[bs,D], yb[bs]; N_total is the N of :393.
 * w_out[Q] (optional) receives the updated w.  With several ranks every rank passes the same GLOBAL minibatch
 * (same composition as on one GPU) and copies only its own rows [rank*bs/nranks, (rank+1)*bs/nranks) to its device. */
int gptb_sgld_step_host(gptb_handle *h, double signal_var, const double *Xb, const double *yb, int64_t bs,
                        int64_t N_total, double epsw, double epsU, int32_t flags, const double *noise_w,
                        const double *noise_U, double *w_out);
/* Pipelined form of gptb_sgld_step_host for a caller that forms minibatch t+1 on the host while step t runs
 * (the batch of GPT_SGLD.jl:379-381 does not depend on the parameters).  _submit copies Xb, yb (pinned host
 * memory) on a copy stream into one of two device staging slots, queues the loop body behind the copy and
 * returns without waiting; Xb, yb and w_out must stay valid until the matching _wait.  _wait blocks until the
 * OLDEST submitted step has finished: its w_out is then valid and its status (GPTB_OK / GPTB_GEODESIC_NAN) is
 * returned.  At most two steps are in flight; a third _submit first waits for the oldest (its status is then
 * reported by the next _wait).  Returns GPTB_ERR_STATE from _wait when nothing is in flight.  This pair is the
 * end-to-end entry bench.py times. */
int gptb_sgld_step_host_submit(gptb_handle *h, double signal_var, const double *Xb, const double *yb, int64_t bs,
                               int64_t N_total, double epsw, double epsU, int32_t flags, const double *noise_w,
                               const double *noise_U, double *w_out);
int gptb_sgld_step_host_wait(gptb_handle *h);
/* The whole of GPTregression (GPT_SGLD.jl:345-448) on the resident training set.
 *   numbatches = ceil(N/m) (:353); each epoch permutes the data cumulatively (:373-374);
 *   perms: (burnin+maxepoch) x N 0-based permutations, row-major per epoch, or NULL (Philox shuffle);
 *   noise_w / noise_U: per-step draws, [Q] / [n,r,D] per step, steps counted over all epochs, or NULL;
 *   w_store[Q,T], U_store[n,r,D,T], T = maxepoch*numbatches, column (epoch-burnin-1)*numbatches+batch
 *   (:441-444).  Either may be NULL (nothing kept — what a burn-in epoch does).
 * On GPTB_GEODESIC_NAN the stores are zero-filled (:422-424). */
int gptb_sgld_run(gptb_handle *h, double signal_var, int64_t m, double epsw, double epsU, int32_t burnin,
                  int32_t maxepoch, int32_t flags, const int64_t *perms, const double *noise_w,
                  const double *noise_U, double *w_store, double *U_store);

/* GPT_GMC (GPT_SGLD.jl:684-805): geodesic Monte Carlo on the resident training set, starting from the handle's
 * current parameters: per epoch fresh momenta (p = randn(Q), mom_k = proj(U_k, randn(n,r)), :709-712), L leapfrog steps
 * with full-batch gradients (:715-785), Metropolis test on the Hamiltonian (:787-795).
 *   noise_p[(burnin+maxepoch)*Q], noise_mom[(burnin+maxepoch)*n*r*D] (each epoch's block in Julia order), unif[burnin+maxepoch]:
 *   the draws of :709, :711, :788 supplied by the caller, or NULL (Philox).
 *   w_store[Q,maxepoch], U_store[n,r,D,maxepoch] (either may be NULL), accept_prob[burnin+maxepoch] = exp(H - H_old) (:790).
 * restore_U_on_reject = 0 reproduces the reference: `U_old=U` (:707) aliases the array that :745 then writes into, so a
 * rejected proposal restores w only; non-zero restores U as well.  On GPTB_GEODESIC_NAN the stores are zero-filled and
 * accept_prob is NaN (:746-748).  The handle's parameters end at the chain's last state. */
int gptb_gmc_run(gptb_handle *h, double signal_var, double epsw, double epsU, int32_t burnin, int32_t maxepoch, int32_t L,
                 int32_t restore_U_on_reject, const double *noise_p, const double *noise_mom, const double *unif,
                 double *w_store, double *U_store, double *accept_prob);

/* GPT_inf (TGP.jl:37-86): Gibbs sampler of the tensor model with Gaussian factors.  The handle holds the features
 * (TGP.feature :6-14: gptb_set_features with length_scale = sigmaRBF, sigma_RBF = phi_scale = 1, b = 2 pi rand), the
 * index table, the whitened training set and the initial factors (gptb_set_params; w is ignored) — and ends at the
 * last draw.  Per iteration: w from its q x q Gaussian conditional (:62-66), then every U^(k) from its n*r x n*r
 * conditional (:73-84), sigma_u = sqrt(1/r), sigma_w = sqrt(r^D/q) (:54-55).  FP64 throughout; the row-proportional
 * work (T, V, the Gram matrices V V' and Ck Ck') and the two small Cholesky solves per conditional run on the device.
 *   noise_w[num_iterations][Q], noise_U[num_iterations][D][n*r]: the randn of :66 and :83, or NULL (Philox);
 *   W_array[Q, num_iterations-burnin], V_array[n,r,D,num_iterations-burnin] (Julia order; either may be NULL): the w of
 *   iteration m and the factors it was drawn with (:68-71).
 * As in the reference the factor draw is invSigma_U \ randn (:83), not a draw with covariance invSigma_U^-1. */
int gptb_tgp_gibbs(gptb_handle *h, double sigma, int32_t num_iterations, int32_t burnin, const double *noise_w,
                   const double *noise_U, double *W_array, double *V_array);

/* GPTclassification (GPT_SGLD.jl:452-680): C independent parameter sets (w_c, U_c) under a softmax likelihood.
 * hc[0..C): one handle per class, all on one device and one rank, with the same shape, features, core index table and
 * training inputs (gptb_set_train; the handles' y is not read) and their own parameters (gptb_init_params with
 * different seeds, or gptb_set_params) — the chain starts from them and they end at its last state.
 *   labels[N]: 0-based class of every training row (the reference's y - minimum(y), :460);
 *   the class-c gradients are the regression gradients with the residual (N/bs)(1[y_i = c] - softmax_c(fhat[i,:]))
 *   (:506-524, :546-564), signal_var plays no part;
 *   every minibatch moves (w, U) TWICE with that one gradient, as the reference does: first with noise on the Stiefel
 *   manifold whatever the flags (:624-636), then under GPTB_FLAG_LANGEVIN / GPTB_FLAG_STIEFEL (:639-671);
 *   GPTB_FLAG_CLS_SINGLE_UPDATE keeps the second move only;
 *   perms as in gptb_sgld_run (all classes walk the same order, :480-481);
 *   noise_w[step][move][C][Q], noise_U[step][move][C][n,r,D] (move = 0, 1 — or one move with
 *   GPTB_FLAG_CLS_SINGLE_UPDATE), steps counted over all epochs, or NULL (Philox, each handle's own seed);
 *   w_store[Q,C,T], U_store[n,r,D,C,T] (Julia order), T = maxepoch*numbatches (:464-465, :673-676), either may be NULL.
 * On GPTB_GEODESIC_NAN the stores are zero-filled (:631-633). */
int gptb_classify_run(gptb_handle *const *hc, int32_t C, const int32_t *labels, int64_t m, double epsw, double epsU,
                      int32_t burnin, int32_t maxepoch, int32_t flags, const int64_t *perms, const double *noise_w,
                      const double *noise_U, double *w_store, double *U_store);

/* The next nsteps loop bodies of the same sampler, continuing the epoch structure where the previous call
 * stopped (a fresh Philox permutation whenever an epoch starts), nothing stored — what burn-in epochs do.
 * Asynchronous: work is only queued; gptb_synchronize() waits and reports GPTB_GEODESIC_NAN. */
int gptb_sgld_advance(gptb_handle *h, double signal_var, int64_t m, double epsw, double epsU, int32_t flags,
                      int64_t nsteps);

/* ---- multi-GPU (one process per GPU) ---------------------------------------------------------- */
/* Row-sharded data parallelism (the CPU precedent is GPT_SGLD_p.jl:70-108,194-218): every rank holds
 * the full training set and parameters, computes the gradient of rows [g*bs/G,(g+1)*bs/G) of the global
 * minibatch, and the [gradw|gradU] buffer is all-reduced (sum) over NVLink.  All ranks then apply the
 * identical update (same Philox counters), so parameters never need a broadcast.
 * id is an opaque 128-byte NCCL unique id made by rank 0 and distributed by the host program. */
int gptb_comm_unique_id(void *id_out_128);
int gptb_comm_init(gptb_handle *h, int32_t nranks, int32_t rank, const void *id_128);

/* ---- multi-GPU from ONE process (the reference is a single Julia process: SURVEY.md 8b) ---------- */
/* hs[0..n): handles on n different devices of this process, each with the whole training set, the same parameters,
 * features, index table and seed.  gptb_comm_init_local makes them ranks 0..n-1 of one communicator
 * (ncclCommInitAll).  The gptb_multi_* entries then run the per-handle call of the same name on one host thread per
 * handle (every step contains the gradient all-reduce, which all ranks must enter); the chain is the one-GPU chain
 * up to summation order, stores come from rank 0.  Per-handle calls that contain no collective (gptb_set_params,
 * gptb_get_params, gptb_synchronize, gptb_pred*) stay usable on each handle.
 * gptb_multi_pred_samples shards the TEST ROWS over the handles (no communication; no communicator needed). */
int gptb_comm_init_local(gptb_handle *const *hs, int32_t n);
int gptb_multi_sgld_run(gptb_handle *const *hs, int32_t n, double signal_var, int64_t m, double epsw, double epsU,
                        int32_t burnin, int32_t maxepoch, int32_t flags, const int64_t *perms, const double *noise_w,
                        const double *noise_U, double *w_store, double *U_store);
int gptb_multi_sgld_advance(gptb_handle *const *hs, int32_t n, double signal_var, int64_t m, double epsw, double epsU,
                            int32_t flags, int64_t nsteps);
int gptb_multi_pred_samples(gptb_handle *const *hs, int32_t n, const double *w_store, const double *U_store, int32_t S,
                            const double *Xtest, int64_t Nt, double *mean_out, double *per_sample_out);
int gptb_comm_destroy(gptb_handle *h);

/* ---- full-theta ("no tensor") RFF model: the comparator of the tensor model ----------------------
 * featureNotensor GPT_SGLD.jl:109-120 and GPNT_SGLD GPT_SGLD.jl:809-847 (PowerPlantNoTensorExperiment.jl,
 * kin40kNoTensorExperiment.jl; BASELINE config 3 "tensor vs full-theta RFF SGLD").
 *   phi[j,i] = sqrt(2/n)*sigma_RBF*cos(sum_d X[i,d]*Z[j,d]/length_scale[d] + b[j])            (n x N, never stored)
 *   grad     = -theta/sigma_theta^2 + (N/bs)*phi_b*(y_b - phi_b'*theta)/signal_var,  eps_t = eps_theta*t^(-decay_rate)
 *   theta   += eps_t*grad/2 + sqrt(eps_t)*randn(n)
 * The features are generated on the fly in both passes of a step (forward phi_b'*theta, back-projection phi_b*e).
 * precision: GPTB_PREC_FP64 (<=1e-10 vs the reference arithmetic) or GPTB_PREC_TC3 = FP32 arithmetic with
 * MUFU.COS (<=1e-4; no tensor cores: the contraction dimension of the feature argument is D <= 16).
 * Arrays as Julia stores them: X[N,D], Z[n,D] column-major Float64; b[n]; theta_store[n,T]. */
typedef struct gptb_nt_handle gptb_nt_handle;
int gptb_nt_create(int32_t D, int32_t n, int32_t precision, int32_t device, uint64_t seed, gptb_nt_handle **out);
int gptb_nt_destroy(gptb_nt_handle *h);
const char *gptb_nt_last_error(const gptb_nt_handle *h);
/* featureNotensor's parameters (:109): length_scale scalar (ls_len = 1) or per dimension (ls_len = D). */
int gptb_nt_set_features(gptb_nt_handle *h, const double *Z, const double *b, const double *length_scale, int32_t ls_len,
                         double sigma_RBF);
/* featureNotensor materialised (:109-120): phi_out[n,N]. */
int gptb_nt_features(gptb_nt_handle *h, const double *X, int64_t N, double *phi_out);
int gptb_nt_set_train(gptb_nt_handle *h, const double *X, const double *y, int64_t N);
/* fhat[i] = phi(Xtest)[:,i]' * theta  (PowerPlantNoTensorExperiment.jl: phitest'*theta_store[:,t]). */
int gptb_nt_pred(gptb_nt_handle *h, const double *theta, const double *Xtest, int64_t Nt, double *fhat_out);
/* GPNT_SGLD (:809-847).  theta0 NULL = sigma_theta*randn(n) from Philox(param_seed) (:815-816); perms NULL = a fresh
 * permutation per epoch (:824), else int64[(maxepoch+burnin)*N] 0-based, applied cumulatively; noise NULL = Philox,
 * else double[T*n] standard normals.  theta_store_out[n,T], T = (maxepoch+burnin)*ceil(N/m): every step is stored,
 * burn-in included (:818,:838).  NaN in theta (:839-842) -> returns GPTB_GEODESIC_NAN ("Get NaN in theta. Try
 * smaller epsilon") with theta_store_out zero-filled. */
int gptb_nt_run(gptb_nt_handle *h, double signal_var, double sigma_theta, int64_t m, double eps_theta, double decay_rate,
                int32_t burnin, int32_t maxepoch, uint64_t param_seed, const double *theta0, const int64_t *perms,
                const double *noise, double *theta_store_out);
/* Device time (ms) of the steps of the last gptb_nt_run and the number of kernel launches it issued. */
int gptb_nt_last_run_stats(gptb_nt_handle *h, double *ms_out, int64_t *launches_out, int64_t *steps_out);

/* ---- introspection (bench / tests) ------------------------------------------------------------ */
/* Per-phase device time (ms) accumulated since the last reset, measured with CUDA events on the
 * handle's stream when profiling is on.  Phases: see GPTB_PHASE_*.  launches_out (optional) receives
 * the number of kernel launches issued by this library since the last reset. */
enum gptb_phase {
    GPTB_PHASE_GATHER = 0,   /* batch gather / conversions                        */
    GPTB_PHASE_FWD = 1,      /* K1 feature x factor forward                       */
    GPTB_PHASE_CONTRACT = 2, /* K2 core contraction, residual, leave-one-out      */
    GPTB_PHASE_BWD = 3,      /* K3 factor-gradient back-projection                */
    GPTB_PHASE_REDUCE = 4,   /* gradient finalisation + all-reduce                */
    GPTB_PHASE_UPDATE = 5,   /* K5 SGLD / Stiefel update                          */
    GPTB_PHASE_COUNT = 6
};
int gptb_set_profiling(gptb_handle *h, int32_t on);

/* ---- options ---------------------------------------------------------------------------------- */
enum gptb_algorithm {
    GPTB_ALG_AUTO = 0,    /* node grid when it pays (many more minibatch rows than nodes and a known input domain) */
    GPTB_ALG_DIRECT = 1,  /* features of every minibatch row generated on the fly (K1 / K3 over the rows)          */
    GPTB_ALG_NODES = 2    /* node grid wherever it is applicable                                                 */
};
enum gptb_option {
    GPTB_OPT_PIPE_GROUPS = 0,   /* dimension groups of the pipelined step (0/1 = off); also forces it on for small steps */
    GPTB_OPT_PIPE_RESERVE = 1,  /* SMs the persistent kernels leave to the concurrent group update                        */
    GPTB_OPT_ALGORITHM = 2,     /* enum gptb_algorithm                                                                    */
    GPTB_OPT_NODE_ORDER = 3,    /* Chebyshev nodes per cell of the node grid (4..16, default 12)                          */
    GPTB_OPT_NODE_THETA = 4,    /* cell half-width x max |Zt| of the dimension (default 2)                                */
    GPTB_OPT_RMS_ALPHA = 5,     /* moving-average coefficient of GPTB_FLAG_RMSPROP (`alpha` of GPT_SGLD.jl:1121), default 0.9 */
    GPTB_OPT_RMS_LAMBDA = 6,    /* its smoothing constant (`lambda` of :1146), default 1e-5                                */
    GPTB_OPT_PRED_CHUNK_MB = 7, /* MB of interpolated values the posterior sweep holds at a time (default 4096, capped at 1/8 of the free memory) */
    GPTB_OPT_PRED_LAYOUT = 9,   /* posterior sweep on the node grid: 0 (default) samples on the lanes, 1 row tiles per sample   */
    GPTB_OPT_PRED_WARPS = 10,   /* warps per tile of the lanes contraction (1..4, 0 = default 2; tuning knob)                  */
    GPTB_OPT_FRONT_PREFETCH = 8 /* 1 (default): gather and cell sort of a node-grid step run on a side stream under the
                                   previous step's tail; 0: everything on the handle's stream                              */
};
int gptb_set_option(gptb_handle *h, int32_t option, double value);
int gptb_get_phase_ms(gptb_handle *h, float *ms_out /*GPTB_PHASE_COUNT*/, int64_t *launches_out, int32_t reset);

#ifdef __cplusplus
}
#endif
#endif /* GPTB200_H */
