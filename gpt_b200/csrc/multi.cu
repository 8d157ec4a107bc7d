// Single-process multi-GPU drivers (SURVEY.md §8b "Threading"): the reference is one Julia process, so its drop-in
// GPTregression(...) call can only scale over GPUs if ONE process drives them.  The ranks are handles on different
// devices joined by gptb_comm_init_local (ncclCommInitAll); every entry here runs the per-handle call of the same name
// on one host thread per handle, because the calls contain collectives (the gradient all-reduce of every step) that
// all ranks must enter.  Data layout, sharding and the chain are those of the one-process-per-GPU mode (comm.cu):
// every handle holds the whole training set and the same parameters and seed; rank g takes rows
// [g*bs/G, (g+1)*bs/G) of each global minibatch; stores come from rank 0.
#include <string.h>

#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>

#include "common.cuh"

void gptb_comm_abort_group(gptb_handle *const *hs, int n);   // comm.cu

namespace {

int check_group(gptb_handle *const *hs, int32_t n) {
    if (!hs || n < 1 || n > 64) return GPTB_ERR_ARG;
    for (int i = 0; i < n; ++i) {
        if (!hs[i]) return GPTB_ERR_ARG;
        if (hs[i]->nranks != n || hs[i]->rank != i) {
            gptb_set_error(hs[0], "handle %d is not rank %d of a %d-rank communicator (gptb_comm_init_local)", i, i, n);
            return GPTB_ERR_STATE;
        }
        // every failure that one rank could hit before a collective while the others wait in it is checked here
        if (!hs[i]->have_params || !(hs[i]->have_train || hs[i]->have_phi) || !hs[i]->have_index ||
            !(hs[i]->have_features || hs[i]->have_phi)) {
            gptb_set_error(hs[0], "handle %d: features, index, training set and parameters must be set", i);
            return GPTB_ERR_STATE;
        }
        if (hs[i]->N != hs[0]->N || hs[i]->cfg.n != hs[0]->cfg.n ||
            hs[i]->cfg.r != hs[0]->cfg.r || hs[i]->cfg.D != hs[0]->cfg.D || hs[i]->cfg.Q != hs[0]->cfg.Q ||
            hs[i]->cfg.seed != hs[0]->cfg.seed || hs[i]->cfg.precision != hs[0]->cfg.precision) {
            gptb_set_error(hs[0], "handle %d differs from handle 0 in shape, precision, seed or training-set size", i);
            return GPTB_ERR_ARG;
        }
    }
    return GPTB_OK;
}

template <typename F>
int run_all(gptb_handle *const *hs, int32_t n, F f, bool collectives = true) {
    std::vector<int> st(n, GPTB_OK);
    std::vector<std::thread> th;
    std::atomic<bool> aborted{false};
    // a rank that fails while the others wait for it inside an all-reduce would hang them: it aborts the group instead
    auto guarded = [&](int i) {
        st[i] = f(i);
        if (st[i] != GPTB_OK && st[i] != GPTB_GEODESIC_NAN && collectives && n > 1 && !aborted.exchange(true))
            gptb_comm_abort_group(hs, n);
    };
    th.reserve(n);
    for (int i = 1; i < n; ++i) th.emplace_back([&, i]() { guarded(i); });
    guarded(0);
    for (auto &t : th) t.join();
    for (int i = 0; i < n; ++i)
        if (st[i] != GPTB_OK) {
            if (i > 0) gptb_set_error(hs[0], "rank %d: %s", i, gptb_last_error(hs[i]));
            return st[i];
        }
    return GPTB_OK;
}

}  // namespace

extern "C" int gptb_multi_sgld_run(gptb_handle *const *hs, int32_t n, double signal_var, int64_t m, double epsw, double epsU,
                                   int32_t burnin, int32_t maxepoch, int32_t flags, const int64_t *perms, const double *noise_w,
                                   const double *noise_U, double *w_store, double *U_store) {
    GPTB_TRY(check_group(hs, n));
    return run_all(hs, n, [&](int i) {
        return gptb_sgld_run(hs[i], signal_var, m, epsw, epsU, burnin, maxepoch, flags, perms, noise_w, noise_U,
                             i == 0 ? w_store : nullptr, i == 0 ? U_store : nullptr);
    });
}

extern "C" int gptb_multi_sgld_advance(gptb_handle *const *hs, int32_t n, double signal_var, int64_t m, double epsw, double epsU,
                                       int32_t flags, int64_t nsteps) {
    GPTB_TRY(check_group(hs, n));
    return run_all(hs, n, [&](int i) { return gptb_sgld_advance(hs[i], signal_var, m, epsw, epsU, flags, nsteps); });
}

// Posterior sweep with the test rows sharded over the handles (no communication): rank g predicts rows
// [g*Nt/G, (g+1)*Nt/G) of the column-major Xtest[Nt, D]; mean_out[Nt] and per_sample_out[Nt, S] are assembled on the host.
extern "C" int gptb_multi_pred_samples(gptb_handle *const *hs, int32_t n, const double *w_store, const double *U_store, int32_t S,
                                       const double *Xtest, int64_t Nt, double *mean_out, double *per_sample_out) {
    if (!hs || n < 1 || n > 64 || !Xtest || Nt < 1 || S < 1) return GPTB_ERR_ARG;
    for (int i = 0; i < n; ++i)
        if (!hs[i]) return GPTB_ERR_ARG;
    const int D = hs[0]->cfg.D;
    return run_all(hs, n, [&](int i) -> int {
        const int64_t lo = (int64_t)i * Nt / n, hi = (int64_t)(i + 1) * Nt / n, cnt = hi - lo;
        if (cnt == 0) return GPTB_OK;
        std::vector<double> Xs((size_t)cnt * D), per;
        for (int k = 0; k < D; ++k) memcpy(Xs.data() + (size_t)k * cnt, Xtest + (size_t)k * Nt + lo, sizeof(double) * cnt);
        if (per_sample_out) per.resize((size_t)cnt * S);
        const int st = gptb_pred_samples(hs[i], w_store, U_store, S, Xs.data(), cnt, mean_out ? mean_out + lo : nullptr,
                                         per_sample_out ? per.data() : nullptr);
        if (st == GPTB_OK && per_sample_out)
            for (int s = 0; s < S; ++s) memcpy(per_sample_out + (size_t)s * Nt + lo, per.data() + (size_t)s * cnt, sizeof(double) * cnt);
        return st;
    }, false);
}
