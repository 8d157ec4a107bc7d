/*
 * CPU restatement (plain C, Float64, loop-for-loop) of the reference's tensor-RFF GP SGLD path.
 *
 * TEST INFRASTRUCTURE ONLY.  Built by oracle/Makefile into oracle/_build/libgptoracle.so and
 * loaded by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  The product (gpt_b200/) never links, loads or calls it.
 *
 * Every function follows one function of /root/reference/GPT_SGLD.jl (Julia 0.3) and cites it.
 * All arrays are Julia column-major Float64; I is Int32, 1-based, I[q + Q*k].
 *
 *   X[N,D] Z[n,D] b[n,D] phi[n,D,N] U[n,r,D] w[Q] temp[D,r,N] V[Q,N] A[r,D,N]
 *
 * Two deliberate departures, both value-preserving up to rounding and stated in DESIGN.md:
 *   - computeA (GPT_SGLD.jl:261-273) re-scans I with unique()/findin() inside its i,k loops;
 *     here the same sums are formed by one scatter pass over q (same terms, same order in q).
 *   - Psi (GPT_SGLD.jl:276-286, :405-408) is never materialised: Psi[:,:,k]*e is accumulated
 *     directly as sum_i e_i A[l,k,i] phi[j,k,i].
 * expm: Julia 0.3 Base expm (Higham 2005 Pade-13 scaling & squaring) is restated without the
 * gebal balancing step; parity for expm/sqrtm/RNG is UNPINNED (no reference known-answer test).
 *
 * OpenMP (-fopenmp) parallelises over minibatch rows, mirroring GPT_SGLD_p.jl:70-108,194-218.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define PHI(j, k, i) phi[(size_t)(j) + (size_t)n * ((size_t)(k) + (size_t)D * (size_t)(i))]
#define UU(j, l, k) U[(size_t)(j) + (size_t)n * ((size_t)(l) + (size_t)r * (size_t)(k))]
#define TEMP(k, l, i) temp[(size_t)(k) + (size_t)D * ((size_t)(l) + (size_t)r * (size_t)(i))]
#define VV(q, i) V[(size_t)(q) + (size_t)Q * (size_t)(i)]
#define AA(l, k, i) A[(size_t)(l) + (size_t)r * ((size_t)(k) + (size_t)D * (size_t)(i))]
#define II(q, k) (I[(size_t)(q) + (size_t)Q * (size_t)(k)] - 1)

int gpo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* GPT_SGLD.jl:71-84.  ls_len==1: scalar length scale; else ls[D]. */
void gpo_feature(const double *X, int N, int D, const double *ls, int ls_len, double sigma_RBF,
                 double phi_scale, const double *Z, const double *b, int n, double *phi, int nthreads) {
    double amp = phi_scale * pow(sigma_RBF, 1.0 / D) * sqrt(2.0 / n);
    double *Zt = (double *)malloc(sizeof(double) * n * D);
    for (int k = 0; k < D; ++k)
        for (int j = 0; j < n; ++j) Zt[j + n * k] = Z[j + n * k] * (1.0 / ls[ls_len == 1 ? 0 : k]);
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int i = 0; i < N; ++i)
        for (int k = 0; k < D; ++k)
            for (int j = 0; j < n; ++j)
                PHI(j, k, i) = amp * cos(X[i + (size_t)N * k] * Zt[j + n * k] + b[j + n * k]);
    free(Zt);
}

/* GPT_SGLD.jl:88-105 (n = 2*half_n rows of output, first half_n rows of Z used). */
int gpo_feature2(const double *X, int N, int D, const double *ls, int ls_len, double sigma_RBF,
                 double phi_scale, const double *Z, int n, double *phi, int nthreads) {
    if (n % 2 == 1) return 1; /* error("n not even") */
    int half_n = n / 2;
    double amp = phi_scale * pow(sigma_RBF, 1.0 / D) / sqrt((double)half_n);
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int i = 0; i < N; ++i)
        for (int k = 0; k < D; ++k)
            for (int j = 0; j < half_n; ++j) {
                double a = X[i + (size_t)N * k] * (Z[j + n * k] * (1.0 / ls[ls_len == 1 ? 0 : k]));
                PHI(2 * j, k, i) = amp * sin(a);
                PHI(2 * j + 1, k, i) = amp * cos(a);
            }
    return 0;
}

/* GPT_SGLD.jl:193-205 */
void gpo_phidotU(const double *U, const double *phi, int n, int r, int D, int N, double *temp, int nthreads) {
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int i = 0; i < N; ++i)
        for (int l = 0; l < r; ++l)
            for (int k = 0; k < D; ++k) {
                double s = 0.0;
                for (int j = 0; j < n; ++j) s += PHI(j, k, i) * UU(j, l, k);
                TEMP(k, l, i) = s;
            }
}

/* GPT_SGLD.jl:208-220 */
void gpo_computeV(const double *temp, const int *I, int Q, int D, int r, int N, double *V, int nthreads) {
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int i = 0; i < N; ++i)
        for (int q = 0; q < Q; ++q) {
            double p = 1.0;
            for (int k = 0; k < D; ++k) p *= TEMP(k, II(q, k), i);
            VV(q, i) = p;
        }
}

/* GPT_SGLD.jl:223-230 */
void gpo_computefhat(const double *V, const double *w, int Q, int N, double *fhat) {
    for (int i = 0; i < N; ++i) {
        double s = 0.0;
        for (int q = 0; q < Q; ++q) s += VV(q, i) * w[q];
        fhat[i] = s;
    }
}

/* GPT_SGLD.jl:233-243 */
void gpo_pred(const double *w, const double *U, const int *I, const double *phi, int n, int r, int D, int Q,
              int N, double *fhat, int nthreads) {
    double *temp = (double *)malloc(sizeof(double) * D * r * (size_t)N);
    double *V = (double *)malloc(sizeof(double) * Q * (size_t)N);
    gpo_phidotU(U, phi, n, r, D, N, temp, nthreads);
    gpo_computeV(temp, I, Q, D, r, N, V, nthreads);
    gpo_computefhat(V, w, Q, N, fhat);
    free(temp);
    free(V);
}

/* GPT_SGLD.jl:384-408: minibatch gradients of the log posterior.
 * gradw[Q], gradU[n,r,D], fhat[bs].  N_total is the training-set size N of :393. */
void gpo_gradients(const double *phi, const double *y, int bs, const double *w, const double *U, const int *I,
                   int n, int r, int D, int Q, double N_total, double signal_var, double *gradw,
                   double *gradU, double *fhat, int nthreads) {
    int N = bs;
    double *temp = (double *)malloc(sizeof(double) * D * r * (size_t)N);
    double *V = (double *)malloc(sizeof(double) * Q * (size_t)N);
    double *A = (double *)calloc((size_t)r * D * N, sizeof(double));
    double *e = (double *)malloc(sizeof(double) * N);
    gpo_phidotU(U, phi, n, r, D, N, temp, nthreads); /* :384 */
    gpo_computeV(temp, I, Q, D, r, N, V, nthreads);  /* :387 */
    gpo_computefhat(V, w, Q, N, fhat);               /* :390 */
    for (int i = 0; i < N; ++i) e[i] = (N_total / bs) * (y[i] - fhat[i]) / signal_var;
    /* :393  gradw = (N/bs) V (y-fhat)/signal_var - w/sigma_w^2, sigma_w = 1 (:354) */
    for (int q = 0; q < Q; ++q) {
        double s = 0.0;
        for (int i = 0; i < N; ++i) s += VV(q, i) * e[i];
        gradw[q] = s - w[q];
    }
    /* :396-399  U_phi[q,i,k] = V[q,i]/temp[k,I[q,k],i];  A[l,k,i] = sum_{q:I[q,k]=l} U_phi*w[q] */
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int i = 0; i < N; ++i)
        for (int k = 0; k < D; ++k)
            for (int q = 0; q < Q; ++q) AA(II(q, k), k, i) += VV(q, i) / TEMP(k, II(q, k), i) * w[q];
    /* :402-408  gradU[:,:,k] = reshape(Psi[:,:,k] e, n, r),  Psi[:,i,k] = kron(A[:,k,i], phi[:,k,i]) */
    memset(gradU, 0, sizeof(double) * (size_t)n * r * D);
#pragma omp parallel for num_threads(nthreads) schedule(static) collapse(2)
    for (int k = 0; k < D; ++k)
        for (int l = 0; l < r; ++l) {
            double *g = gradU + (size_t)n * (l + (size_t)r * k);
            for (int i = 0; i < N; ++i) {
                double c = AA(l, k, i) * e[i];
                if (c == 0.0) continue;
                const double *p = &PHI(0, k, i);
                for (int j = 0; j < n; ++j) g[j] += c * p[j];
            }
        }
    free(temp);
    free(V);
    free(A);
    free(e);
}

/* ---------------------------------------------------------------- small dense helpers (col-major) */
static void mm(const double *A, const double *B, double *C, int m, int k, int n2) { /* C[m,n2]=A[m,k]B[k,n2] */
    for (int j = 0; j < n2; ++j)
        for (int i = 0; i < m; ++i) {
            double s = 0.0;
            for (int p = 0; p < k; ++p) s += A[i + (size_t)m * p] * B[p + (size_t)k * j];
            C[i + (size_t)m * j] = s;
        }
}

static int solve(double *A, double *B, int n2, int nrhs) { /* A X = B, partial pivoting, in place */
    for (int c = 0; c < n2; ++c) {
        int p = c;
        for (int i = c + 1; i < n2; ++i)
            if (fabs(A[i + n2 * c]) > fabs(A[p + n2 * c])) p = i;
        if (!(fabs(A[p + n2 * c]) > 0.0)) return 1;
        if (p != c) {
            for (int j = 0; j < n2; ++j) { double t = A[c + n2 * j]; A[c + n2 * j] = A[p + n2 * j]; A[p + n2 * j] = t; }
            for (int j = 0; j < nrhs; ++j) { double t = B[c + n2 * j]; B[c + n2 * j] = B[p + n2 * j]; B[p + n2 * j] = t; }
        }
        for (int i = c + 1; i < n2; ++i) {
            double f = A[i + n2 * c] / A[c + n2 * c];
            if (f == 0.0) continue;
            for (int j = c; j < n2; ++j) A[i + n2 * j] -= f * A[c + n2 * j];
            for (int j = 0; j < nrhs; ++j) B[i + n2 * j] -= f * B[c + n2 * j];
        }
    }
    for (int j = 0; j < nrhs; ++j)
        for (int i = n2 - 1; i >= 0; --i) {
            double s = B[i + n2 * j];
            for (int p = i + 1; p < n2; ++p) s -= A[i + n2 * p] * B[p + n2 * j];
            B[i + n2 * j] = s / A[i + n2 * i];
        }
    return 0;
}

/* Matrix exponential, Higham (2005) scaling & squaring with the degree-13 Pade approximant — the
 * algorithm of Julia 0.3 Base.expm used at GPT_SGLD.jl:23,28 (without the gebal balancing). */
void gpo_expm(const double *Ain, int n2, double *E) {
    static const double c[14] = {64764752532480000., 32382376266240000., 7771770303897600., 1187353796428800.,
                                 129060195264000.,   10559470521600.,    670442572800.,     33522128640.,
                                 1323241920.,        40840800.,          960960.,           16380.,
                                 182.,               1.};
    size_t sz = (size_t)n2 * n2;
    double *A = (double *)malloc(sizeof(double) * sz * 8);
    double *A2 = A + sz, *A4 = A2 + sz, *A6 = A4 + sz, *Um = A6 + sz, *Vm = Um + sz, *T1 = Vm + sz, *T2 = T1 + sz;
    memcpy(A, Ain, sizeof(double) * sz);
    double nrm = 0.0;
    for (int j = 0; j < n2; ++j) {
        double s = 0.0;
        for (int i = 0; i < n2; ++i) s += fabs(A[i + (size_t)n2 * j]);
        if (s > nrm || s != s) nrm = s;
    }
    if (nrm != nrm || isinf(nrm)) { /* NaN/Inf input propagates, as LAPACK-backed expm does */
        for (size_t i = 0; i < sz; ++i) E[i] = NAN;
        free(A);
        return;
    }
    int s = 0;
    const double theta13 = 5.371920351148152;
    if (nrm > theta13) {
        s = (int)ceil(log2(nrm / theta13));
        if (s < 0) s = 0;
        double f = ldexp(1.0, -s);
        for (size_t i = 0; i < sz; ++i) A[i] *= f;
    }
    mm(A, A, A2, n2, n2, n2);
    mm(A2, A2, A4, n2, n2, n2);
    mm(A4, A2, A6, n2, n2, n2);
    /* U = A*(A6*(c13 A6 + c11 A4 + c9 A2) + c7 A6 + c5 A4 + c3 A2 + c1 I) */
    for (size_t i = 0; i < sz; ++i) T1[i] = c[13] * A6[i] + c[11] * A4[i] + c[9] * A2[i];
    mm(A6, T1, T2, n2, n2, n2);
    for (size_t i = 0; i < sz; ++i) T2[i] += c[7] * A6[i] + c[5] * A4[i] + c[3] * A2[i];
    for (int i = 0; i < n2; ++i) T2[i + (size_t)n2 * i] += c[1];
    mm(A, T2, Um, n2, n2, n2);
    /* V = A6*(c12 A6 + c10 A4 + c8 A2) + c6 A6 + c4 A4 + c2 A2 + c0 I */
    for (size_t i = 0; i < sz; ++i) T1[i] = c[12] * A6[i] + c[10] * A4[i] + c[8] * A2[i];
    mm(A6, T1, Vm, n2, n2, n2);
    for (size_t i = 0; i < sz; ++i) Vm[i] += c[6] * A6[i] + c[4] * A4[i] + c[2] * A2[i];
    for (int i = 0; i < n2; ++i) Vm[i + (size_t)n2 * i] += c[0];
    /* (V - U) X = (V + U) */
    for (size_t i = 0; i < sz; ++i) { T1[i] = Vm[i] - Um[i]; T2[i] = Vm[i] + Um[i]; }
    if (solve(T1, T2, n2, n2)) {
        for (size_t i = 0; i < sz; ++i) E[i] = NAN;
        free(A);
        return;
    }
    for (int q = 0; q < s; ++q) {
        mm(T2, T2, T1, n2, n2, n2);
        memcpy(T2, T1, sizeof(double) * sz);
    }
    memcpy(E, T2, sizeof(double) * sz);
    free(A);
}

/* GPT_SGLD.jl:14-16:  out = V - U (U'V + V'U)/2 */
void gpo_proj(const double *U, const double *V, int n, int r, double *out) {
    double *S = (double *)malloc(sizeof(double) * r * r);
    for (int a = 0; a < r; ++a)
        for (int b2 = 0; b2 < r; ++b2) {
            double s1 = 0.0, s2 = 0.0;
            for (int j = 0; j < n; ++j) {
                s1 += U[j + (size_t)n * a] * V[j + (size_t)n * b2];
                s2 += V[j + (size_t)n * a] * U[j + (size_t)n * b2];
            }
            S[a + r * b2] = (s1 + s2) / 2;
        }
    for (int l = 0; l < r; ++l)
        for (int j = 0; j < n; ++j) {
            double s = 0.0;
            for (int a = 0; a < r; ++a) s += U[j + (size_t)n * a] * S[a + r * l];
            out[j + (size_t)n * l] = V[j + (size_t)n * l] - s;
        }
    free(S);
}

/* GPT_SGLD.jl:19-37.  Returns 1 (and zeros in out) when expm produced NaN (:24-26). */
int gpo_geod(const double *U, const double *mom, int n, int r, double t, double *out) {
    int r2 = 2 * r;
    double *A = (double *)calloc((size_t)r * r, sizeof(double));
    double *T = (double *)calloc((size_t)r2 * r2, sizeof(double));
    double *E = (double *)malloc(sizeof(double) * r2 * r2);
    double *mexp = (double *)malloc(sizeof(double) * r * r);
    double *W = (double *)malloc(sizeof(double) * r2 * r);
    for (int a = 0; a < r; ++a)
        for (int b2 = 0; b2 < r; ++b2) {
            double s1 = 0.0, s2 = 0.0;
            for (int j = 0; j < n; ++j) {
                s1 += U[j + (size_t)n * a] * mom[j + (size_t)n * b2];
                s2 += mom[j + (size_t)n * a] * mom[j + (size_t)n * b2];
            }
            A[a + r * b2] = s1;
            T[a + r2 * b2] = t * s1;              /* [A  -mom'mom] */
            T[a + r2 * (b2 + r)] = -t * s2;
            T[(a + r) + r2 * (b2 + r)] = t * s1;  /* [I   A      ] */
        }
    for (int a = 0; a < r; ++a) T[(a + r) + r2 * a] = t;
    gpo_expm(T, r2, E);
    int bad = 0;
    for (int i = 0; i < r2 * r2; ++i)
        if (E[i] != E[i]) bad = 1;
    if (bad) {
        memset(out, 0, sizeof(double) * (size_t)n * r);
    } else {
        for (int i = 0; i < r * r; ++i) A[i] *= -t;
        gpo_expm(A, r, mexp);
        mm(E, mexp, W, r2, r, r); /* E[:,1:r] is the first r columns (col-major: first r2*r entries) */
        for (int l = 0; l < r; ++l) {
            double nn = 0.0;
            for (int j = 0; j < n; ++j) {
                double s = 0.0;
                for (int a = 0; a < r; ++a)
                    s += U[j + (size_t)n * a] * W[a + r2 * l] + mom[j + (size_t)n * a] * W[(a + r) + r2 * l];
                out[j + (size_t)n * l] = s;
                nn += s * s;
            }
            nn = sqrt(nn);
            for (int j = 0; j < n; ++j) out[j + (size_t)n * l] /= nn;
        }
    }
    free(A); free(T); free(E); free(mexp); free(W);
    return bad;
}

/* One iteration of GPT_SGLD.jl:378-439 with the randn draws supplied (noise_w[Q], noise_U[n,r,D];
 * NULL = no noise term, i.e. langevin=false).  w and U are updated in place.  Returns 1 on NaN. */
int gpo_sgld_step(const double *phi, const double *y, int bs, double *w, double *U, const int *I, int n, int r,
                  int D, int Q, double N_total, double signal_var, double epsw, double epsU,
                  const double *noise_w, const double *noise_U, int stiefel, double *fhat, int nthreads) {
    double *gradw = (double *)malloc(sizeof(double) * Q);
    double *gradU = (double *)malloc(sizeof(double) * (size_t)n * r * D);
    double *G = (double *)malloc(sizeof(double) * (size_t)n * r);
    double *mom = (double *)malloc(sizeof(double) * (size_t)n * r);
    double *Un = (double *)malloc(sizeof(double) * (size_t)n * r);
    int bad = 0;
    gpo_gradients(phi, y, bs, w, U, I, n, r, D, Q, N_total, signal_var, gradw, gradU, fhat, nthreads);
    for (int q = 0; q < Q; ++q) w[q] += epsw * gradw[q] / 2 + (noise_w ? sqrt(epsw) * noise_w[q] : 0.0);
    size_t nr = (size_t)n * r;
    for (int k = 0; k < D && !bad; ++k) {
        double *Uk = U + nr * k;
        const double *gk = gradU + nr * k;
        if (stiefel) {
            for (size_t i = 0; i < nr; ++i) G[i] = sqrt(epsU) * gk[i] / 2 + (noise_U ? noise_U[nr * k + i] : 0.0);
            gpo_proj(Uk, G, n, r, mom);
            bad = gpo_geod(Uk, mom, n, r, sqrt(epsU), Un);
            if (!bad) memcpy(Uk, Un, sizeof(double) * nr);
        } else {
            for (size_t i = 0; i < nr; ++i)
                Uk[i] += epsU * (gk[i] - n * Uk[i]) / 2 + (noise_U ? sqrt(epsU) * noise_U[nr * k + i] : 0.0);
        }
    }
    free(gradw); free(gradU); free(G); free(mom); free(Un);
    return bad;
}
