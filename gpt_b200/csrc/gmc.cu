// GPT_GMC — geodesic Monte Carlo, GPT_SGLD.jl:684-805: full-batch HMC on the core w and on the Stiefel factors U^(k),
// L leapfrog steps per epoch, Metropolis accept / reject.
//
// The loop runs on the host (as the reference's does) and drives the library's own device entries: two full-batch
// gradients per leapfrog step (gptb_grad over all N resident rows: K1 / K2 / K3 or their node-grid forms), the tangent
// projection (gptb_proj, :14-16) and the geodesic with its transported momentum (gptb_geodboth, :40-59) per dimension.
// Only O(Q + n r D) numbers cross the bus per call; the N-proportional work stays on the device.
#include <math.h>
#include <string.h>

#include <vector>

#include "common.cuh"

int gptb_reset_order(gptb_handle *h);   // batch.cu

namespace {
constexpr uint64_t STREAM_GMC_UNIF = 6ull << 56;
}

extern "C" int gptb_gmc_run(gptb_handle *h, double signal_var, double epsw, double epsU, int32_t burnin, int32_t maxepoch,
                            int32_t L, int32_t restore_U_on_reject, const double *noise_p, const double *noise_mom,
                            const double *unif, double *w_store, double *U_store, double *accept_prob) {
    if (!h || !(signal_var > 0) || burnin < 0 || maxepoch < 0 || L < 1 || !accept_prob) return GPTB_ERR_ARG;
    if (!h->have_params || !(h->have_train || h->have_phi) || !h->have_index) {
        gptb_set_error(h, "gptb_gmc_run: features, index, training set and parameters must be set");
        return GPTB_ERR_STATE;
    }
    const gptb_config &c = h->cfg;
    const int n = c.n, r = c.r, D = c.D, Q = c.Q;
    const int64_t N = h->N;
    const size_t per_k = (size_t)n * r, ucnt = per_k * D;
    GPTB_CUDA_OK(h, cudaSetDevice(h->device));
    GPTB_TRY(gptb_reset_order(h));                       // fhat[i] must belong to y[i]
    GPTB_TRY(gptb_synchronize(h));
    std::vector<double> y(N), w(Q), U(ucnt), w_old(Q), U_old(ucnt), p(Q), mom(ucnt), gw(Q), gU(ucnt), fhat(N), tmp(per_k),
        tmp2(per_k);
    GPTB_CUDA_OK(h, cudaMemcpy(y.data(), h->y, sizeof(double) * N, cudaMemcpyDeviceToHost));
    GPTB_TRY(gptb_get_params(h, w.data(), U.data()));
    const int T = burnin + maxepoch;
    const double sw = sqrt(epsw), sU = sqrt(epsU);
    auto sumsq = [](const std::vector<double> &v) {
        double s = 0.0;
        for (double x : v) s += x * x;
        return s;
    };
    auto resid2 = [&]() {
        double s = 0.0;
        for (int64_t i = 0; i < N; ++i) s += (y[i] - fhat[i]) * (y[i] - fhat[i]);
        return s;
    };
    // gradients of the log posterior at (w, U): gradw = V (y - fhat)/signal_var - w  (:723),  gradU (:733-736); bs = N
    auto grads = [&]() -> int {
        GPTB_TRY(gptb_set_params(h, w.data(), U.data()));
        return gptb_grad(h, signal_var, nullptr, 0, N, gw.data(), gU.data(), fhat.data(), 1);
    };
    // mom_k <- proj(U_k, mom_k + sqrt(epsU) gradU_k / 2)   (:738-740, :777-779)
    auto kick = [&]() -> int {
        for (int k = 0; k < D; ++k) {
            double *mk = mom.data() + k * per_k;
            const double *gk = gU.data() + k * per_k;
            for (size_t e = 0; e < per_k; ++e) tmp[e] = mk[e] + sU * gk[e] / 2;
            GPTB_TRY(gptb_proj(h->device, n, r, U.data() + k * per_k, tmp.data(), mk));
        }
        for (int q = 0; q < Q; ++q) p[q] += sw * gw[q] / 2;     // :725 / :764
        return GPTB_OK;
    };
    auto fail_nan = [&]() {
        if (w_store) memset(w_store, 0, sizeof(double) * (size_t)Q * maxepoch);          // :746-748
        if (U_store) memset(U_store, 0, sizeof(double) * ucnt * (size_t)maxepoch);
        for (int e = 0; e < T; ++e) accept_prob[e] = NAN;
        return GPTB_GEODESIC_NAN;
    };
    for (int epoch = 1; epoch <= T; ++epoch) {
        w_old = w;
        U_old = U;
        // p = randn(Q); mom_k = proj(U_k, randn(n,r))   (:709-712)
        for (int q = 0; q < Q; q += 2) {
            double a, b;
            if (noise_p) { a = noise_p[(size_t)(epoch - 1) * Q + q]; b = q + 1 < Q ? noise_p[(size_t)(epoch - 1) * Q + q + 1] : 0.0; }
            else philox_normal2(c.seed, GPTB_STREAM_NOISE_W | (uint64_t)(q / 2), (uint64_t)epoch | (1ull << 40), &a, &b);
            p[q] = a;
            if (q + 1 < Q) p[q + 1] = b;
        }
        for (int k = 0; k < D; ++k) {
            for (size_t e = 0; e < per_k; e += 2) {
                double a, b;
                if (noise_mom) {
                    a = noise_mom[(size_t)(epoch - 1) * ucnt + k * per_k + e];
                    b = e + 1 < per_k ? noise_mom[(size_t)(epoch - 1) * ucnt + k * per_k + e + 1] : 0.0;
                } else {
                    philox_normal2(c.seed, GPTB_STREAM_NOISE_U | ((uint64_t)k << 40) | (e / 2), (uint64_t)epoch | (1ull << 40), &a, &b);
                }
                tmp[e] = a;
                if (e + 1 < per_k) tmp[e + 1] = b;
            }
            GPTB_TRY(gptb_proj(h->device, n, r, U.data() + k * per_k, tmp.data(), mom.data() + k * per_k));
        }
        GPTB_TRY(grads());                                   // fhat = pred(w, U, I, phi); its gradient starts the leapfrog
        const double H_old = -sumsq(w) / 2 - resid2() / (2 * signal_var) - sumsq(mom) / 2 - sumsq(p) / 2;   // :713
        for (int l = 0; l < L; ++l) {
            // (:716-740 re-evaluate the gradients at the (w, U) the previous half step ended on — gw, gU, fhat still hold
            // exactly those, from the evaluation above for l = 0 and from the end of step l-1 otherwise)
            GPTB_TRY(kick());
            for (int q = 0; q < Q; ++q) w[q] += sw * p[q];   // :742
            for (int k = 0; k < D; ++k) {                    // :744-749
                const int st = gptb_geodboth(h->device, n, r, U.data() + k * per_k, mom.data() + k * per_k, sU, tmp.data(), tmp2.data());
                if (st == GPTB_GEODESIC_NAN) return fail_nan();
                GPTB_TRY(st);
                memcpy(U.data() + k * per_k, tmp.data(), sizeof(double) * per_k);
                memcpy(mom.data() + k * per_k, tmp2.data(), sizeof(double) * per_k);
            }
            GPTB_TRY(grads());                               // :752-775 at the new (w, U); the last one is pred_new (:782)
            GPTB_TRY(kick());
        }
        const double H = -sumsq(w) / 2 - resid2() / (2 * signal_var) - sumsq(mom) / 2 - sumsq(p) / 2;       // :787
        accept_prob[epoch - 1] = exp(H - H_old);                                                           // :790
        double u;
        if (unif) u = unif[epoch - 1];
        else {
            philox4 x = philox4x32_10(c.seed, STREAM_GMC_UNIF | (uint64_t)epoch, 0);
            u = (double)(((uint64_t)x.x << 21) | (x.y >> 11)) * (1.0 / 9007199254740992.0);
        }
        if (u > accept_prob[epoch - 1]) {                    // reject (:793-795)
            w = w_old;
            // `U_old=U` (:707) is an alias in Julia and :745 writes into that array: the reference restores w only
            if (restore_U_on_reject) U = U_old;
        }
        if (epoch > burnin) {
            if (w_store) memcpy(w_store + (size_t)(epoch - burnin - 1) * Q, w.data(), sizeof(double) * Q);
            if (U_store) memcpy(U_store + (size_t)(epoch - burnin - 1) * ucnt, U.data(), sizeof(double) * ucnt);
        }
    }
    return gptb_set_params(h, w.data(), U.data());
}
