// K5 — SGLD step on w and SGLD-ERM step on every U^(k) (FP64 throughout).
//
// Replaces (all /root/reference/GPT_SGLD.jl):
//   :411-414  w += epsw*gradw/2 + sqrt(epsw)*randn(Q)
//   :417-439  mom = proj(U_k, sqrt(epsU)*gradU_k/2 + randn(n,r));  U_k = geod(U_k, mom, sqrt(epsU))
//             (and the Euclidean / no-noise variants)
//   proj :14-16,  geod :19-37 (expm of the 2r x 2r block matrix, expm(-tA), column renormalisation,
//             NaN -> zeros + abort).
//
// The n x r work is spread over (row-block, k) CTAs in four phases; only r x r quantities cross
// phases, so the step never serialises on one CTA per factor matrix:
//   P1 prepare : G = sqrt(epsU)/2*gradU + noise;  partial S1 = U'G, GG = G'G, UU = U'U  (+ the w update)
//   P2 small   : sym = (S1+S1')/2;  A = U'M = S1 - UU*sym;  M'M = GG - S1'sym - sym S1 + sym UU sym
//                E = expm(t[[A,-M'M],[I,A]]);  X = expm(-tA);  W = E[:,1:r] X;  C1 = W1 - sym W2, C2 = W2
//   P3 apply   : Unew = [U M] W = U C1 + G C2   (M = G - U sym is never stored);  column norms
//   P4 finish  : U = Unew / ||column||, refresh the tensor-core operand image
// M = proj(U,G) is formed implicitly; the algebra is the reference's, re-associated (agreement ~1e-15).
#include <cuda_fp16.h>

#include <atomic>

#include "common.cuh"
#include "tc_layout.cuh"

namespace {

constexpr int UP_ROWS = 64;      // rows of U per CTA in P1/P3/P4
constexpr int UP_THREADS = 256;

// layout of the per-k small workspace (doubles)
struct SmallLayout {
    int r, r2;
    size_t S1, GG, UU, CN, C1, C2, sym, A, MM, Tm, E, X, W, ws;  // offsets
    size_t total;
    __host__ __device__ explicit SmallLayout(int r_) : r(r_), r2(2 * r_) {
        size_t o = 0, rr = (size_t)r * r, r22 = (size_t)r2 * r2;
        S1 = o; o += rr;
        GG = o; o += rr;
        UU = o; o += rr;
        CN = o; o += r;          // column norms^2  (S1..CN are zeroed every step)
        C1 = o; o += rr;
        C2 = o; o += rr;
        sym = o; o += rr;
        A = o; o += rr;
        MM = o; o += rr;
        Tm = o; o += r22;
        E = o; o += r22;
        X = o; o += rr;
        W = o; o += (size_t)r2 * r;
        ws = o; o += 8 * r22;    // expm workspace
        total = o;
    }
    __host__ __device__ size_t zero_count() const { return 3 * (size_t)r * r + r; }
};

__device__ __forceinline__ double noise_at(uint64_t key, uint64_t stream, uint64_t step, uint64_t e) {
    double a, b;
    philox_normal2(key, stream | (e >> 1), step, &a, &b);
    return (e & 1) ? b : a;
}

// ---- FP64 tensor-core products for the r x r / 2r x 2r algebra of the Stiefel step ----------------
// DMMA m8n8k4 issues 256 FMAs per instruction (64 FMA/clk/SM, the same peak as DFMA) with a fraction of the
// instruction and shared-memory traffic of a scalar product loop.
__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}
// leading dimension = 4 (mod 16): the 8x4 / 4x8 fragment loads then touch every bank pair once per half warp
__host__ __device__ inline int dmm_ld(int np) { return np + 4; }

// C[M x N] = A[M x K] * B[K x N]: col-major in shared memory, leading dimensions from dmm_ld; M a multiple of
// 8*TM, N of 8, K of 4 (zero padded).  Tiles of (8*TM) x 8, one per warp and pass over the warps [w0, w0+nw) of
// the CTA, two accumulator chains per tile row block.  No barrier inside: independent products share a stage.
template <int TM>
__device__ __forceinline__ void dmm_t(const double *__restrict__ A, int lda, const double *__restrict__ B, int ldb,
                                      double *__restrict__ C, int ldc, int M, int N, int K, int w0, int nw,
                                      const double *__restrict__ Dm = nullptr, int ldd = 0 /* C = Dm - A*B */) {
    const int warp = (int)(threadIdx.x >> 5) - w0, lane = threadIdx.x & 31, g = lane >> 2, tg = lane & 3;
    if (warp < 0 || warp >= nw) return;
    const int mb = M / (8 * TM), nt = mb * (N >> 3);
    for (int t = warp; t < nt; t += nw) {
        const int tj = t / mb, i0 = (t - tj * mb) * 8 * TM, j0 = tj << 3;
        const double *Ap = A + i0 + g + (size_t)lda * tg;
        const double *Bp = B + tg + (size_t)ldb * (j0 + g);
        double ca[2][TM], cb[2][TM];
#pragma unroll
        for (int i = 0; i < TM; ++i) ca[0][i] = cb[0][i] = ca[1][i] = cb[1][i] = 0.0;
        int p0 = 0;
#pragma unroll 2
        for (; p0 + 8 <= K; p0 += 8) {
            const double b0 = Bp[p0], b1 = Bp[p0 + 4];
#pragma unroll
            for (int i = 0; i < TM; ++i) {
                dmma884(ca[0][i], cb[0][i], Ap[(size_t)lda * p0 + 8 * i], b0);
                dmma884(ca[1][i], cb[1][i], Ap[(size_t)lda * (p0 + 4) + 8 * i], b1);
            }
        }
        if (p0 < K) {
            const double b0 = Bp[p0];
#pragma unroll
            for (int i = 0; i < TM; ++i) dmma884(ca[0][i], cb[0][i], Ap[(size_t)lda * p0 + 8 * i], b0);
        }
        double *Cp = C + i0 + g + (size_t)ldc * (j0 + 2 * tg);
#pragma unroll
        for (int i = 0; i < TM; ++i) {
            double va = ca[0][i] + ca[1][i], vb = cb[0][i] + cb[1][i];
            if (Dm) {
                const double *Dp = Dm + i0 + g + 8 * i + (size_t)ldd * (j0 + 2 * tg);
                va = Dp[0] - va;
                vb = Dp[ldd] - vb;
            }
            Cp[8 * i] = va;
            Cp[8 * i + ldc] = vb;
        }
    }
}

// ---------------------------------------------------------------- P1
// grid (ceil(n/UP_ROWS), D + 1); blockIdx.y == D handles w (block x == 0 only).
__global__ void __launch_bounds__(UP_THREADS)
k_prepare(const double *__restrict__ grad /*[Q | D r n]*/, double *__restrict__ w, double *__restrict__ U,
          double *__restrict__ G, double *__restrict__ small, size_t small_stride, int n, int r, int D, int Q,
          double epsw, double epsU, int flags, const double *__restrict__ noise_w,
          const double *__restrict__ noise_U, uint64_t key, uint64_t step, int have_grad, int k0, int kn,
          float *__restrict__ gradFbase, int qpad, int qstride, int npad, int rpad,
          const double *__restrict__ epsk_sum, double inv_cnt) {
    // epsk_sum != null (RMSprop sampler): dimension k moves with its own step size epsk_sum[k] * inv_cnt (:1212)
    // gradFbase != null: the FP32 gradient flow — gradw in gradFbase[0..Q), gradU as K3 left it in
    // gradFbase + qpad ([D][npad][rpad]); every entry is read once here and cleared for the next step
    // grid.y = kn (+ 1): rows 0..kn-1 are the dimensions k0.., row kn (if launched) is w
    extern __shared__ double sm[];
    const int tid = threadIdx.x;
    const bool langevin = flags & GPTB_FLAG_LANGEVIN, stiefel = flags & GPTB_FLAG_STIEFEL;
    (void)D;
    if ((int)blockIdx.y == kn) {
        if (blockIdx.x != 0 || !have_grad) return;
        const double sq = sqrt(epsw);
        for (int q = tid; q < Q; q += UP_THREADS) {
            double gl = grad[q];
            if (gradFbase) {   // the GPTB_GW_SLOTS partial copies K2 summed into
                gl = 0.0;
                for (int s_ = 0; s_ < GPTB_GW_SLOTS; ++s_) {
                    gl += (double)gradFbase[(size_t)s_ * qstride + q];
                    gradFbase[(size_t)s_ * qstride + q] = 0.f;
                }
            }
            // prior term of :393 (sigma_w = 1, :354), added once after the all-reduce — unless the caller has folded it
            // into grad already (GPTclassification's second move uses the prior of the w before its first move, :524)
            double g = (flags & GPTB_IFLAG_PRIOR_IN_GRAD) ? gl : gl - w[q];
            double z = 0.0;
            if (langevin) z = noise_w ? noise_w[q] : noise_at(key, GPTB_STREAM_NOISE_W, step, (uint64_t)q);
            w[q] += epsw * g / 2 + sq * z;
        }
        return;
    }
    const int k = k0 + blockIdx.y;
    const int j0 = blockIdx.x * UP_ROWS;
    const int rows = min(UP_ROWS, n - j0);
    const double *gU = grad + Q + (size_t)k * r * n;
    double *Uk = U + (size_t)k * r * n, *Gk = G + (size_t)k * r * n;
    if (epsk_sum) epsU = epsk_sum[k] * inv_cnt;
    const double sqU = sqrt(epsU);
    const int npr = (r + 15) & ~15, ldt = dmm_ld(npr);   // tile layout [jj][l], zero padded to npr columns
    double *Us = sm, *Gs = sm + (size_t)UP_ROWS * ldt;
    // FP32 gradient flow: this CTA's rows of gradF ([j][l], contiguous) are staged through shared memory with
    // coalesced 16-byte loads and cleared the same way (per-element strided loads + 4-byte clears took 60 us).  The
    // loads are issued first and land while pass A below fetches U and the noise.
    float *gst = reinterpret_cast<float *>(sm + 2 * (size_t)UP_ROWS * ldt);   // [UP_ROWS][rpad + 1]
    float4 pre[4];
    float4 *src = nullptr;
    int cnt4 = 0;
    if (gradFbase) {
        src = reinterpret_cast<float4 *>(gradFbase + qpad + ((size_t)k * npad + j0) * rpad);
        cnt4 = min(UP_ROWS, npad - j0) * rpad / 4;      // <= 4 * UP_THREADS (rpad <= 64)
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (tid + i * UP_THREADS < cnt4) pre[i] = src[tid + i * UP_THREADS];
    }
    // pass A: U and the step's draw into the tiles
    for (int t = tid; t < UP_ROWS * npr; t += UP_THREADS) {
        const int jj = t % UP_ROWS, l = t / UP_ROWS, j = j0 + jj;
        double u = 0.0, z = 0.0;
        if (jj < rows && l < r) {
            const size_t e = (size_t)l * n + j;
            u = Uk[e];
            if (langevin)
                z = noise_U ? noise_U[(size_t)k * r * n + e]
                            : noise_at(key, GPTB_STREAM_NOISE_U | ((uint64_t)k << 40), step, (uint64_t)e);
        }
        Us[(size_t)jj * ldt + l] = u;
        Gs[(size_t)jj * ldt + l] = z;
    }
    if (gradFbase) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int t = tid + i * UP_THREADS;
            if (t < cnt4) {
                src[t] = make_float4(0.f, 0.f, 0.f, 0.f);
                const int row = (4 * t) / rpad, col = (4 * t) - row * rpad;
                float *d = gst + row * (rpad + 1) + col;
                d[0] = pre[i].x; d[1] = pre[i].y; d[2] = pre[i].z; d[3] = pre[i].w;
            }
        }
        __syncthreads();
    }
    // pass B: G = sqrt(epsU)/2 * gradient + draw (:420), or the Euclidean update (:426 / :437); a thread revisits the
    // tile entries it wrote in pass A
    for (int t = tid; t < UP_ROWS * npr; t += UP_THREADS) {
        const int jj = t % UP_ROWS, l = t / UP_ROWS, j = j0 + jj;
        double g = 0.0;
        if (jj < rows && l < r) {
            const size_t e = (size_t)l * n + j;
            const double z = Gs[(size_t)jj * ldt + l];
            const double gr = gradFbase ? (double)gst[jj * (rpad + 1) + l] : (have_grad ? gU[e] : 0.0);
            if (stiefel) {
                g = sqU * gr / 2 + z;                       // :420 argument of proj
                Gk[e] = g;
            } else {
                const double u = Us[(size_t)jj * ldt + l];
                Uk[e] = u + epsU * (gr - n * u) / 2 + sqU * z;  // :426 / :437
            }
        }
        Gs[(size_t)jj * ldt + l] = g;
    }
    if (!stiefel) return;
    __syncthreads();
    double *sk = small + (size_t)k * small_stride;
    SmallLayout L(r);
    // Gram sums of this row block on the FP64 tensor cores (8x8 tiles, K = UP_ROWS): S1 = U'G in full, GG = G'G and
    // UU = U'U by their upper triangles, mirrored on the way out.  (The scalar loop this replaces was bound by its
    // four shared-memory loads per three FMAs.)
    const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, tg = lane & 3;
    const int nbt = npr >> 3, ntri = nbt * (nbt + 1) / 2, nitems = nbt * nbt + 2 * ntri;
    for (int it = warp; it < nitems; it += UP_THREADS / 32) {
        int prod = 0, ti, tj;
        if (it < nbt * nbt) {
            tj = it / nbt;
            ti = it - tj * nbt;
        } else {
            int u = it - nbt * nbt;
            if (u >= ntri) { prod = 2; u -= ntri; } else prod = 1;
            tj = 0;
            while (u > tj) { u -= tj + 1; ++tj; }
            ti = u;                                   // ti <= tj
        }
        const double *At = prod == 1 ? Gs : Us;       // S1 = U'G, GG = G'G, UU = U'U
        const double *Bt = prod == 2 ? Us : Gs;
        const double *Ap = At + ti * 8 + g + (size_t)ldt * tg;   // A[m][k] = At[m + ldt k],  B[k][n] = Bt[n + ldt k]
        const double *Bp = Bt + tj * 8 + g + (size_t)ldt * tg;
        double c0a = 0.0, c0b = 0.0, c1a = 0.0, c1b = 0.0;
#pragma unroll 4
        for (int p0 = 0; p0 < UP_ROWS; p0 += 8) {
            dmma884(c0a, c0b, Ap[(size_t)ldt * p0], Bp[(size_t)ldt * p0]);
            dmma884(c1a, c1b, Ap[(size_t)ldt * (p0 + 4)], Bp[(size_t)ldt * (p0 + 4)]);
        }
        const double va = c0a + c1a, vb = c0b + c1b;
        const int a = ti * 8 + g, b = tj * 8 + 2 * tg;
        double *o = sk + (prod == 0 ? L.S1 : prod == 1 ? L.GG : L.UU);
        if (a < r && b < r) atomicAdd(&o[a + (size_t)r * b], va);
        if (a < r && b + 1 < r) atomicAdd(&o[a + (size_t)r * (b + 1)], vb);
        if (prod != 0 && ti != tj) {
            if (a < r && b < r) atomicAdd(&o[b + (size_t)r * a], va);
            if (a < r && b + 1 < r) atomicAdd(&o[b + 1 + (size_t)r * a], vb);
        }
    }
}

// ---------------------------------------------------------------- P2 helpers (one CTA, col-major)
__device__ void bmm(const double *A, const double *B, double *C, int m, int kk, int n2) {
    for (int t = threadIdx.x; t < m * n2; t += blockDim.x) {
        int i = t % m, j = t / m;
        double s = 0.0;
        for (int p = 0; p < kk; ++p) s = fma(A[i + (size_t)m * p], B[p + (size_t)kk * j], s);
        C[t] = s;
    }
    __syncthreads();
}

// In-place Gaussian elimination with partial pivoting: A X = B (n2 x n2 each).  Returns via *bad.
__device__ void bsolve(double *A, double *B, int n2, int *bad_s) {
    __shared__ int piv;
    for (int c = 0; c < n2; ++c) {
        if (threadIdx.x == 0) {
            int p = c;
            double best = fabs(A[c + (size_t)n2 * c]);
            for (int i = c + 1; i < n2; ++i) {
                double v = fabs(A[i + (size_t)n2 * c]);
                if (v > best) { best = v; p = i; }
            }
            piv = p;
            if (!(best > 0.0)) *bad_s = 1;
        }
        __syncthreads();
        int p = piv;
        if (p != c) {
            for (int j = threadIdx.x; j < 2 * n2; j += blockDim.x) {
                double *M = j < n2 ? A : B;
                int jc = j < n2 ? j : j - n2;
                double tmp = M[c + (size_t)n2 * jc];
                M[c + (size_t)n2 * jc] = M[p + (size_t)n2 * jc];
                M[p + (size_t)n2 * jc] = tmp;
            }
        }
        __syncthreads();
        const double d = A[c + (size_t)n2 * c];
        // eliminate below the pivot: columns c+1.. of A and all of B (column c of A is left as is)
        int nrow = n2 - c - 1, ncolA = n2 - c - 1;
        for (int t = threadIdx.x; t < nrow * (ncolA + n2); t += blockDim.x) {
            int i = c + 1 + t % nrow, jj = t / nrow;
            double f = A[i + (size_t)n2 * c] / d;
            if (jj < ncolA) {
                int j = c + 1 + jj;
                A[i + (size_t)n2 * j] -= f * A[c + (size_t)n2 * j];
            } else {
                int j = jj - ncolA;
                B[i + (size_t)n2 * j] -= f * B[c + (size_t)n2 * j];
            }
        }
        __syncthreads();
    }
    for (int j = threadIdx.x; j < n2; j += blockDim.x) {
        for (int i = n2 - 1; i >= 0; --i) {
            double s = B[i + (size_t)n2 * j];
            for (int p = i + 1; p < n2; ++p) s -= A[i + (size_t)n2 * p] * B[p + (size_t)n2 * j];
            B[i + (size_t)n2 * j] = s / A[i + (size_t)n2 * i];
        }
    }
    __syncthreads();
}

// expm by scaling and squaring with a diagonal Pade approximant (Higham 2005) — the algorithm behind
// Julia 0.3 Base.expm at GPT_SGLD.jl:23,28: degree 3/5/7/9 for small 1-norms, degree 13 with scaling
// above 2.1.  M (n2 x n2, col-major) is overwritten by exp(M).  ws: 8*n2*n2 doubles.
__device__ void bexpm(double *M, int n2, double *ws, int *bad_s) {
    __shared__ double nrm_s;
    __shared__ int s_s;
    const size_t sz = (size_t)n2 * n2;
    double *A2 = ws, *A4 = A2 + sz, *A6 = A4 + sz, *Um = A6 + sz, *Vm = Um + sz, *T1 = Vm + sz, *T2 = T1 + sz,
           *A8 = T2 + sz;
    if (threadIdx.x == 0) {
        double nrm = 0.0;
        bool nan = false;
        for (int j = 0; j < n2; ++j) {
            double s = 0.0;
            for (int i = 0; i < n2; ++i) s += fabs(M[i + (size_t)n2 * j]);
            if (s != s || isinf(s)) nan = true;
            nrm = fmax(nrm, s);
        }
        if (nan) *bad_s = 1;
        nrm_s = nrm;
        int s = 0;
        if (nrm > 5.371920351148152) s = max(0, (int)ceil(log2(nrm / 5.371920351148152)));
        s_s = s;
    }
    __syncthreads();
    if (*bad_s) return;
    const double nrm = nrm_s;
    const int s = s_s;
    int deg = 13;
    if (nrm <= 1.495585217958292e-2) deg = 3;
    else if (nrm <= 2.539398330063230e-1) deg = 5;
    else if (nrm <= 9.504178996162932e-1) deg = 7;
    else if (nrm <= 2.097847961257068e0) deg = 9;
    if (s > 0) {
        double f = ldexp(1.0, -s);
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) M[t] *= f;
        __syncthreads();
    }
    bmm(M, M, A2, n2, n2, n2);
    if (deg >= 5) bmm(A2, A2, A4, n2, n2, n2);
    if (deg >= 7) bmm(A4, A2, A6, n2, n2, n2);
    if (deg == 9) bmm(A6, A2, A8, n2, n2, n2);
    if (deg == 13) {
        const double c[14] = {64764752532480000., 32382376266240000., 7771770303897600., 1187353796428800.,
                              129060195264000.,   10559470521600.,    670442572800.,     33522128640.,
                              1323241920.,        40840800.,          960960.,           16380.,
                              182.,               1.};
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) T1[t] = c[13] * A6[t] + c[11] * A4[t] + c[9] * A2[t];
        __syncthreads();
        bmm(A6, T1, T2, n2, n2, n2);
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) {
            int i = t % n2, j = t / n2;
            T2[t] += c[7] * A6[t] + c[5] * A4[t] + c[3] * A2[t] + (i == j ? c[1] : 0.0);
        }
        __syncthreads();
        bmm(M, T2, Um, n2, n2, n2);
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) T1[t] = c[12] * A6[t] + c[10] * A4[t] + c[8] * A2[t];
        __syncthreads();
        bmm(A6, T1, Vm, n2, n2, n2);
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) {
            int i = t % n2, j = t / n2;
            Vm[t] += c[6] * A6[t] + c[4] * A4[t] + c[2] * A2[t] + (i == j ? c[0] : 0.0);
        }
        __syncthreads();
    } else {
        // U = M * (sum_{odd} c_i M^(i-1)),  V = sum_{even} c_i M^i
        double c[10];
        if (deg == 3) { const double cc[4] = {120., 60., 12., 1.}; for (int i = 0; i < 4; ++i) c[i] = cc[i]; }
        else if (deg == 5) { const double cc[6] = {30240., 15120., 3360., 420., 30., 1.}; for (int i = 0; i < 6; ++i) c[i] = cc[i]; }
        else if (deg == 7) { const double cc[8] = {17297280., 8648640., 1995840., 277200., 25200., 1512., 56., 1.}; for (int i = 0; i < 8; ++i) c[i] = cc[i]; }
        else { const double cc[10] = {17643225600., 8821612800., 2075673600., 302702400., 30270240., 2162160., 110880., 3960., 90., 1.}; for (int i = 0; i < 10; ++i) c[i] = cc[i]; }
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) {
            int i = t % n2, j = t / n2;
            double u = (i == j ? c[1] : 0.0) + c[3] * A2[t];
            double v = (i == j ? c[0] : 0.0) + c[2] * A2[t];
            if (deg >= 5) { u += c[5] * A4[t]; v += c[4] * A4[t]; }
            if (deg >= 7) { u += c[7] * A6[t]; v += c[6] * A6[t]; }
            if (deg >= 9) { u += c[9] * A8[t]; v += c[8] * A8[t]; }
            T2[t] = u;
            Vm[t] = v;
        }
        __syncthreads();
        bmm(M, T2, Um, n2, n2, n2);
    }
    for (size_t t = threadIdx.x; t < sz; t += blockDim.x) {
        T1[t] = Vm[t] - Um[t];
        T2[t] = Vm[t] + Um[t];
    }
    __syncthreads();
    bsolve(T1, T2, n2, bad_s);
    for (int q = 0; q < s; ++q) {
        bmm(T2, T2, T1, n2, n2, n2);
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) T2[t] = T1[t];
        __syncthreads();
    }
    for (size_t t = threadIdx.x; t < sz; t += blockDim.x) M[t] = T2[t];
    __syncthreads();
}

// ---------------------------------------------------------------- shared-memory expm (n2 <= 96)
// C = A*B, all ld x ld col-major in shared memory (ld multiple of 4, zero padded); 4x4 register tiles.
__device__ void smm(const double *__restrict__ A, const double *__restrict__ B, double *__restrict__ C, int ld, int kk) {
    const int nt = ld >> 2;
    for (int t = threadIdx.x; t < nt * nt; t += blockDim.x) {
        const int i0 = (t % nt) * 4, j0 = (t / nt) * 4;
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        for (int p = 0; p < kk; ++p) {
            const double2 a01 = *reinterpret_cast<const double2 *>(A + i0 + (size_t)ld * p);
            const double2 a23 = *reinterpret_cast<const double2 *>(A + i0 + 2 + (size_t)ld * p);
            const double av[4] = {a01.x, a01.y, a23.x, a23.y};
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const double bv = B[p + (size_t)ld * (j0 + b)];
#pragma unroll
                for (int a = 0; a < 4; ++a) acc[a][b] = fma(av[a], bv, acc[a][b]);
            }
        }
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int a = 0; a < 4; ++a) C[i0 + a + (size_t)ld * (j0 + b)] = acc[a][b];
    }
    __syncthreads();
}

// exp(M) for an n2 x n2 matrix in global memory, evaluated entirely in shared memory (6 buffers of ld*ld doubles):
// scaling and squaring around a Taylor polynomial of degree m in {4,8,12,16,20} evaluated by Paterson-Stockmeyer
// in powers of X^4 (3 + m/4 - 1 matrix products, no linear solve — the Pade solve is a 64-step serial chain on a
// single CTA).  m and the scaling s are chosen so the truncation error is below 1.1e-16 * ||exp||; agreement with
// Julia/SciPy's Pade-13 expm (GPT_SGLD.jl:23,28) is ~1e-15 relative.  NaN/Inf input sets *bad_s (:24).
__device__ void sexpm(double *M, int n2, double *buf, int *bad_s) {
    __shared__ double nrm_s;
    __shared__ int s_s, m_s;
    const int ld = (n2 + 3) & ~3;
    const size_t sz = (size_t)ld * ld;
    double *X = buf, *X2 = X + sz, *X3 = X2 + sz, *X4 = X3 + sz, *P = X4 + sz, *Tm = P + sz;
    for (size_t t = threadIdx.x; t < sz; t += blockDim.x) {
        int i = t % ld, j = t / ld;
        X[t] = (i < n2 && j < n2) ? M[i + (size_t)n2 * j] : 0.0;
    }
    if (threadIdx.x == 0) nrm_s = 0.0;
    __syncthreads();
    if (threadIdx.x < n2) {
        double s = 0.0;
        for (int i = 0; i < n2; ++i) s += fabs(X[i + (size_t)ld * threadIdx.x]);
        if (s != s || isinf(s)) *bad_s = 1;
        else atomicMax(reinterpret_cast<unsigned long long *>(&nrm_s), (unsigned long long)__double_as_longlong(s));
    }
    __syncthreads();
    if (*bad_s) return;
    if (threadIdx.x == 0) {
        const double nrm = nrm_s;
        int m = 20, s = 0;
        if (nrm <= 1.7e-3) m = 4;
        else if (nrm <= 7.0e-2) m = 8;
        else if (nrm <= 3.3e-1) m = 12;
        else if (nrm <= 8.2e-1) m = 16;
        else if (nrm > 1.5) s = max(0, (int)ceil(log2(nrm / 1.5)));
        m_s = m;
        s_s = s;
    }
    __syncthreads();
    const int m = m_s, s = s_s;
    if (s > 0) {
        const double f = ldexp(1.0, -s);
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) X[t] *= f;
        __syncthreads();
    }
    smm(X, X, X2, ld, n2);
    smm(X2, X, X3, ld, n2);
    smm(X2, X2, X4, ld, n2);
    double c[21];
    c[0] = 1.0;
    for (int j = 1; j <= 20; ++j) c[j] = c[j - 1] / j;
    const int nb = m / 4;
    // block i: B_i = c[4i] I + c[4i+1] X + c[4i+2] X2 + c[4i+3] X3
    for (size_t t = threadIdx.x; t < sz; t += blockDim.x) {
        const int i = t % ld, j = t / ld, b0 = 4 * (nb - 1);
        P[t] = ((i == j && i < n2) ? c[b0] : 0.0) + c[b0 + 1] * X[t] + c[b0 + 2] * X2[t] + c[b0 + 3] * X3[t] + c[m] * X4[t];
    }
    __syncthreads();
    for (int bi = nb - 2; bi >= 0; --bi) {
        smm(X4, P, Tm, ld, n2);
        for (size_t t = threadIdx.x; t < sz; t += blockDim.x) {
            const int i = t % ld, j = t / ld, b0 = 4 * bi;
            P[t] = Tm[t] + ((i == j && i < n2) ? c[b0] : 0.0) + c[b0 + 1] * X[t] + c[b0 + 2] * X2[t] + c[b0 + 3] * X3[t];
        }
        __syncthreads();
    }
    double *cur = P, *oth = Tm;
    for (int q = 0; q < s; ++q) {
        smm(cur, cur, oth, ld, n2);
        double *sw = cur; cur = oth; oth = sw;
    }
    for (size_t t = threadIdx.x; t < (size_t)n2 * n2; t += blockDim.x) {
        int i = t % n2, j = t / n2;
        M[t] = cur[i + (size_t)ld * j];
    }
    __syncthreads();
}

// ---------------------------------------------------------------- P2
// grid (D).  use_sym = 1: G is the un-projected direction (SGLD-ERM step);  0: G already is the
// tangent vector `mom` of geod(U,mom,t) (GPT_SGLD.jl:19-37 called directly).
__global__ void __launch_bounds__(1024)
k_small(double *__restrict__ small, size_t small_stride, int r, double t, int use_sym, int32_t *nan_flag,
        int use_smem, const double *__restrict__ epsk_sum = nullptr, double inv_cnt = 0.0, int both = 0) {
    // both != 0 (geodboth, GPT_SGLD.jl:40-59): also leaves C1' = W'[0:r], C2' = W'[r:2r] of W' = E[:, r+1:2r] expm(-tA) in
    // the A / MM slots, so that the transported momentum is U C1' + mom C2' (:50)
    extern __shared__ double sm[];
    __shared__ int bad;
    if (epsk_sum) t = sqrt(epsk_sum[blockIdx.x] * inv_cnt);
    double *sk = small + (size_t)blockIdx.x * small_stride;
    SmallLayout L(r);
    const int r2 = 2 * r, rr = r * r;
    if (threadIdx.x == 0) bad = 0;
    __syncthreads();
    double *S1 = sk + L.S1, *GG = sk + L.GG, *UU = sk + L.UU, *sym = sk + L.sym, *A = sk + L.A, *MM = sk + L.MM;
    double *Tm = sk + L.Tm, *X = sk + L.X, *W = sk + L.W, *C1 = sk + L.C1, *C2 = sk + L.C2, *ws = sk + L.ws;
    double *tmp = sk + L.E, *tmp2 = sk + L.E + rr;  // E region doubles as scratch before expm
    for (int i = threadIdx.x; i < rr; i += blockDim.x) {
        int a = i % r, b = i / r;
        sym[i] = use_sym ? (S1[a + r * b] + S1[b + r * a]) / 2 : 0.0;
    }
    __syncthreads();
    bmm(UU, sym, tmp, r, r, r);        // UU*sym
    bmm(sym, tmp, tmp2, r, r, r);      // sym*UU*sym
    for (int i = threadIdx.x; i < rr; i += blockDim.x) A[i] = S1[i] - tmp[i];
    __syncthreads();
    // S1'*sym  -> tmp ; sym*S1 is its transpose (sym symmetric)
    for (int i = threadIdx.x; i < rr; i += blockDim.x) {
        int a = i % r, b = i / r;
        double s = 0.0;
        for (int p = 0; p < r; ++p) s = fma(S1[p + r * a], sym[p + r * b], s);
        tmp[i] = s;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < rr; i += blockDim.x) {
        int a = i % r, b = i / r;
        MM[i] = GG[i] - tmp[a + r * b] - tmp[b + r * a] + tmp2[i];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < r2 * r2; i += blockDim.x) {
        int a = i % r2, b = i / r2;
        double v;
        if (a < r && b < r) v = A[a + r * b];
        else if (a < r) v = -MM[a + r * (b - r)];
        else if (b < r) v = (a - r == b) ? 1.0 : 0.0;
        else v = A[(a - r) + r * (b - r)];
        Tm[i] = t * v;
    }
    for (int i = threadIdx.x; i < rr; i += blockDim.x) X[i] = -t * A[i];
    __syncthreads();
    if (use_smem) sexpm(Tm, r2, sm, &bad);
    else bexpm(Tm, r2, ws, &bad);
    if (!bad) {
        // "if sum(isnan(E))>0" (:24)
        int local = 0;
        for (int i = threadIdx.x; i < r2 * r2; i += blockDim.x)
            if (Tm[i] != Tm[i]) local = 1;
        if (local) bad = 1;
    }
    __syncthreads();
    if (bad) {
        if (threadIdx.x == 0) atomicExch(nan_flag, 1);
        return;
    }
    if (use_smem) sexpm(X, r, sm, &bad);
    else bexpm(X, r, ws, &bad);
    __syncthreads();
    if (bad) {
        if (threadIdx.x == 0) atomicExch(nan_flag, 1);
        return;
    }
    bmm(Tm, X, W, r2, r, r);  // first r columns of E are the first r2*r entries (col-major)
    // C1 = W1 - sym*W2, C2 = W2
    for (int i = threadIdx.x; i < rr; i += blockDim.x) {
        int a = i % r, b = i / r;
        double s = 0.0;
        for (int p = 0; p < r; ++p) s = fma(sym[a + r * p], W[(p + r) + r2 * b], s);
        C1[i] = W[a + r2 * b] - s;
        C2[i] = W[(a + r) + r2 * b];
    }
    if (both) {
        __syncthreads();
        double *Wp = sk + L.E;                       // scratch again now that E sits in Tm
        bmm(Tm + (size_t)r2 * r, X, Wp, r2, r, r);   // columns r..2r-1 of E (col-major: entries from r2*r on)
        for (int i = threadIdx.x; i < rr; i += blockDim.x) {
            int a = i % r, b = i / r;
            A[i] = Wp[a + r2 * b];
            MM[i] = Wp[(a + r) + r2 * b];
        }
    }
}

// ---------------------------------------------------------------- P3: Unew = U C1 + G C2 -> G ; column norms
__global__ void __launch_bounds__(UP_THREADS)
k_apply(const double *__restrict__ U, double *__restrict__ G, double *__restrict__ small, size_t small_stride, int n,
        int r, const int32_t *__restrict__ nan_flag) {
    if (*nan_flag) return;
    extern __shared__ double sm[];
    const int k = blockIdx.y, j0 = blockIdx.x * UP_ROWS, tid = threadIdx.x;
    const int rows = min(UP_ROWS, n - j0);
    double *sk = small + (size_t)k * small_stride;
    SmallLayout L(r);
    double *Us = sm, *Gs = Us + (size_t)UP_ROWS * r, *C1 = Gs + (size_t)UP_ROWS * r, *C2 = C1 + r * r,
           *cn = C2 + r * r;
    const double *Uk = U + (size_t)k * r * n;
    double *Gk = G + (size_t)k * r * n;
    for (int t = tid; t < UP_ROWS * r; t += UP_THREADS) {
        int jj = t % UP_ROWS, l = t / UP_ROWS;
        bool ok = jj < rows;
        Us[t] = ok ? Uk[(size_t)l * n + j0 + jj] : 0.0;
        Gs[t] = ok ? Gk[(size_t)l * n + j0 + jj] : 0.0;
    }
    for (int t = tid; t < r * r; t += UP_THREADS) {
        C1[t] = sk[L.C1 + t];
        C2[t] = sk[L.C2 + t];
    }
    for (int t = tid; t < r; t += UP_THREADS) cn[t] = 0.0;
    __syncthreads();
    for (int t = tid; t < UP_ROWS * r; t += UP_THREADS) {
        int jj = t % UP_ROWS, l = t / UP_ROWS;
        double s = 0.0;
        for (int a = 0; a < r; ++a)
            s += Us[(size_t)a * UP_ROWS + jj] * C1[a + r * l] + Gs[(size_t)a * UP_ROWS + jj] * C2[a + r * l];
        if (jj < rows) {
            Gk[(size_t)l * n + j0 + jj] = s;
            atomicAdd(&cn[l], s * s);
        }
    }
    __syncthreads();
    for (int t = tid; t < r; t += UP_THREADS) atomicAdd(&sk[L.CN + t], cn[t]);
}

// ---------------------------------------------------------------- P4: U = Unew / ||col||
__global__ void k_finish(double *__restrict__ U, const double *__restrict__ G, const double *__restrict__ small,
                         size_t small_stride, int n, int r, const int32_t *__restrict__ nan_flag) {
    if (*nan_flag) return;
    const int k = blockIdx.y;
    SmallLayout L(r);
    const double *cn = small + (size_t)k * small_stride + L.CN;
    size_t total = (size_t)r * n;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int l = (int)(e / n);
        U[(size_t)k * total + e] = G[(size_t)k * total + e] / sqrt(cn[l]);
    }
}

// out = U C1 + V C2 for r x r matrices C1, C2 in global memory (the transported momentum of geodboth, :50)
__global__ void k_apply_plain(const double *__restrict__ U, const double *__restrict__ V, const double *__restrict__ C1,
                              const double *__restrict__ C2, int n, int r, double *__restrict__ out) {
    const size_t total = (size_t)r * n;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int j = (int)(e % n), l = (int)(e / n);
        double s = 0.0;
        for (int a = 0; a < r; ++a) s += U[(size_t)a * n + j] * C1[a + r * l] + V[(size_t)a * n + j] * C2[a + r * l];
        out[e] = s;
    }
}

// out = V - U * (S1 + S1')/2   (proj, :14-16), S1 = U'V accumulated by k_prepare-style reduction
__global__ void k_proj_apply(const double *__restrict__ U, const double *__restrict__ V, const double *__restrict__ S1,
                             int n, int r, double *__restrict__ out) {
    size_t total = (size_t)r * n;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        int j = (int)(e % n), l = (int)(e / n);
        double s = 0.0;
        for (int a = 0; a < r; ++a) s += U[(size_t)a * n + j] * (S1[a + r * l] + S1[l + r * a]) / 2;
        out[e] = V[e] - s;
    }
}

// reductions only (S1 = U'G, GG = G'G, UU = U'U) for the standalone proj/geod entries
__global__ void __launch_bounds__(UP_THREADS)
k_reduce_only(const double *__restrict__ U, const double *__restrict__ G, double *__restrict__ small, int n, int r) {
    extern __shared__ double sm[];
    const int tid = threadIdx.x, j0 = blockIdx.x * UP_ROWS;
    const int rows = min(UP_ROWS, n - j0);
    double *Us = sm, *Gs = sm + (size_t)UP_ROWS * r;
    for (int t = tid; t < UP_ROWS * r; t += UP_THREADS) {
        int jj = t % UP_ROWS, l = t / UP_ROWS;
        bool ok = jj < rows;
        Us[t] = ok ? U[(size_t)l * n + j0 + jj] : 0.0;
        Gs[t] = ok ? G[(size_t)l * n + j0 + jj] : 0.0;
    }
    __syncthreads();
    SmallLayout L(r);
    for (int t = tid; t < r * r; t += UP_THREADS) {
        int a = t % r, b = t / r;
        double s1 = 0.0, gg = 0.0, uu = 0.0;
        for (int jj = 0; jj < UP_ROWS; ++jj) {
            double ua = Us[(size_t)a * UP_ROWS + jj], ub = Us[(size_t)b * UP_ROWS + jj];
            double ga = Gs[(size_t)a * UP_ROWS + jj], gb = Gs[(size_t)b * UP_ROWS + jj];
            s1 = fma(ua, gb, s1);
            gg = fma(ga, gb, gg);
            uu = fma(ua, ub, uu);
        }
        atomicAdd(&small[L.S1 + t], s1);
        atomicAdd(&small[L.GG + t], gg);
        atomicAdd(&small[L.UU + t], uu);
    }
}


// ================================================================================================
// RMSprop sampler (GPT_SGLDERM_RMSprop, GPT_SGLD.jl:1121-1237): step sizes from moving averages of the squared
// UNNORMALISED gradients (the reference's (1/batch_size)-scaled ones, :1172/:1202: our likelihood gradients / N).
// ================================================================================================
// w: gw = alpha gw + (1-alpha) raw^2;  epsw = eps/(sqrt(gw)+lambda);  w += epsw (N raw - w)/2 + sqrt(epsw) z   (:1175-1188).
// The pre-update w is kept in w_old: fhat and the residual of this step belong to it, computeA sees the new w.
__global__ void k_rms_w(const double *__restrict__ grad, double *__restrict__ w, double *__restrict__ w_old,
                        double *__restrict__ gw, int Q, double N, double eps, double alpha, double lambda,
                        const double *__restrict__ noise_w, uint64_t key, uint64_t step, int langevin) {
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < Q; q += gridDim.x * blockDim.x) {
        const double lik = grad[q], raw = lik / N;
        const double g2 = alpha * gw[q] + (1.0 - alpha) * raw * raw;
        gw[q] = g2;
        const double e = eps / (sqrt(g2) + lambda);
        const double wq = w[q];
        w_old[q] = wq;
        double z = 0.0;
        if (langevin) z = noise_w ? noise_w[q] : noise_at(key, GPTB_STREAM_NOISE_W, step, (uint64_t)q);
        w[q] = wq + e * (lik - wq) / 2 + sqrt(e) * z;
    }
}

// U: gU = alpha gU + (1-alpha) raw^2 element-wise, and the sum over each U^(k) of eps/(sqrt(gU)+lambda) (:1210-1212).
// grid (blocks, D); gradU and gU are [D][r][n].
__global__ void __launch_bounds__(256) k_rms_U(const double *__restrict__ gradU, double *__restrict__ gU,
                                               double *__restrict__ epsk_sum, size_t per_k, double N, double eps, double alpha,
                                               double lambda) {
    const int k = blockIdx.y;
    double acc = 0.0;
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < per_k; e += (size_t)gridDim.x * blockDim.x) {
        const double raw = gradU[(size_t)k * per_k + e] / N;
        const double g2 = alpha * gU[(size_t)k * per_k + e] + (1.0 - alpha) * raw * raw;
        gU[(size_t)k * per_k + e] = g2;
        acc += eps / (sqrt(g2) + lambda);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ double part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) acc += part[w];
        atomicAdd(&epsk_sum[k], acc);
    }
}

// ================================================================================================
// Fast path of the sampler's update (r <= 32): noise -> prepare -> small2 -> apply2.
// ================================================================================================

// Production-mode noise of one step, two draws per Philox block (same counters as noise_at()).
// grid.y = D + 1; row D is the w noise.
__global__ void k_noise(double *__restrict__ noise_w, double *__restrict__ noise_U, int Q, size_t per_k, int k0, int kn,
                        uint64_t key, uint64_t step) {
    // grid.y = kn (+ 1): rows 0..kn-1 are the dimensions k0.., row kn (if launched) is the w noise
    const bool is_w = (int)blockIdx.y == kn;
    const int k = k0 + blockIdx.y;
    const size_t cnt = is_w ? (size_t)Q : per_k;
    const uint64_t stream = is_w ? GPTB_STREAM_NOISE_W : (GPTB_STREAM_NOISE_U | ((uint64_t)k << 40));
    double *out = is_w ? noise_w : noise_U + (size_t)k * per_k;
    for (size_t pe = blockIdx.x * (size_t)blockDim.x + threadIdx.x; 2 * pe < cnt; pe += (size_t)gridDim.x * blockDim.x) {
        double a, b;
        philox_normal2(key, stream | pe, step, &a, &b);
        out[2 * pe] = a;
        if (2 * pe + 1 < cnt) out[2 * pe + 1] = b;
    }
}

__constant__ double c_inv_fact[21] = {1.0, 1.0, 0.5, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040, 1.0 / 40320,
                                      1.0 / 362880, 1.0 / 3628800, 1.0 / 39916800, 1.0 / 479001600, 1.0 / 6227020800.0,
                                      1.0 / 87178291200.0, 1.0 / 1307674368000.0, 1.0 / 20922789888000.0,
                                      1.0 / 355687428096000.0, 1.0 / 6402373705728000.0, 1.0 / 121645100408832000.0,
                                      1.0 / 2432902008176640000.0};

// exp(X) for the n2 x n2 matrix held zero padded to np x np (np a multiple of 8*TM, leading dimension ld) in buf[0];
// the result is left in the buffer whose pointer is returned.  Taylor / Paterson-Stockmeyer + squaring as in sexpm;
// buf holds six np*ld blocks; slot selects the function's shared scalars.  NaN/Inf input sets *bad_s.
template <int TM>
__device__ double *sexpm3(double *buf, int n2, int np, int ld, int *bad_s) {
    __shared__ double nrm3_s;
    __shared__ int s3_s, m3_s;
    const size_t sz = (size_t)ld * np;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    double *X = buf, *X2 = X + sz, *X3 = X2 + sz, *X4 = X3 + sz, *P = X4 + sz, *Tm = P + sz;
#ifdef GPTB_K5_DBG
    long long te[8];
    int ne = 0;
#define K5E() do { te[ne++] = clock64(); } while (0)
#else
#define K5E() do {} while (0)
#endif
    K5E();
    if (threadIdx.x == 0) nrm3_s = 0.0;
    __syncthreads();
    for (int c = warp; c < n2; c += nw) {   // 1-norm: one warp per column
        double sc = 0.0;
        for (int i = lane; i < n2; i += 32) sc += fabs(X[i + (size_t)ld * c]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sc += __shfl_xor_sync(0xffffffffu, sc, o);
        if (lane == 0) {
            if (sc != sc || isinf(sc)) *bad_s = 1;
            else atomicMax(reinterpret_cast<unsigned long long *>(&nrm3_s), (unsigned long long)__double_as_longlong(sc));
        }
    }
    __syncthreads();
    if (*bad_s) return X;
    if (threadIdx.x == 0) {
        const double nrm = nrm3_s;
        int m = 20, sq = 0;
        if (nrm <= 1.7e-3) m = 4;
        else if (nrm <= 7.0e-2) m = 8;
        else if (nrm <= 3.3e-1) m = 12;
        else if (nrm <= 8.2e-1) m = 16;
        else if (nrm > 1.5) sq = max(0, (int)ceil(log2(nrm / 1.5)));
        m3_s = m;
        s3_s = sq;
    }
    __syncthreads();
    const int m = m3_s, sq = s3_s;
    K5E();
    if (sq > 0) {
        const double f = ldexp(1.0, -sq);
        for (int c = warp; c < n2; c += nw)
            for (int i = lane; i < n2; i += 32) X[i + (size_t)ld * c] *= f;
        __syncthreads();
    }
    dmm_t<TM>(X, ld, X, ld, X2, ld, np, np, np, 0, nw);
    __syncthreads();
    K5E();
    if (m == 4) {
        // the common case of a small step (||X||_1 <= 1.7e-3): exp(X) ~ I + X + X2/2 + X2 (X/6 + X2/24) — one more
        // product on all warps instead of X3 and X4 side by side on half of them each
        const double c3 = c_inv_fact[3], c4 = c_inv_fact[4];
        for (int c = warp; c < np; c += nw)
            for (int i = lane; i < np; i += 32) {
                const size_t t = i + (size_t)ld * c;
                X3[t] = c3 * X[t] + c4 * X2[t];
            }
        __syncthreads();
        dmm_t<TM>(X2, ld, X3, ld, Tm, ld, np, np, np, 0, nw);
        __syncthreads();
        for (int c = warp; c < np; c += nw)
            for (int i = lane; i < np; i += 32) {
                const size_t t = i + (size_t)ld * c;
                P[t] = ((i == c && i < n2) ? 1.0 : 0.0) + X[t] + 0.5 * X2[t] + Tm[t];
            }
        __syncthreads();
        double *cur4 = P, *oth4 = Tm;
        for (int q = 0; q < sq; ++q) {   // (sq == 0 whenever m == 4; kept for symmetry with the general path)
            dmm_t<TM>(cur4, ld, cur4, ld, oth4, ld, np, np, np, 0, nw);
            __syncthreads();
            double *sw = cur4; cur4 = oth4; oth4 = sw;
        }
        return cur4;
    }
    {   // X3 = X2 X and X4 = X2 X2 in one stage
        const int h = nw / 2;
        dmm_t<TM>(X2, ld, X, ld, X3, ld, np, np, np, 0, h);
        dmm_t<TM>(X2, ld, X2, ld, X4, ld, np, np, np, h, nw - h);
    }
    __syncthreads();
    K5E();
    const double *inv_fact = c_inv_fact;   // (a per-thread copy of the table in local memory cost ~6k clk per call)
    const int nbk = m / 4;
    {   // top block, with the X^m term folded in
        const int b0 = 4 * (nbk - 1);
        const double c0 = inv_fact[b0], c1 = inv_fact[b0 + 1], c2 = inv_fact[b0 + 2], c3 = inv_fact[b0 + 3], c4 = inv_fact[m];
        for (int c = warp; c < np; c += nw)
            for (int i = lane; i < np; i += 32) {
                const size_t t = i + (size_t)ld * c;
                P[t] = ((i == c && i < n2) ? c0 : 0.0) + c1 * X[t] + c2 * X2[t] + c3 * X3[t] + c4 * X4[t];
            }
        __syncthreads();
    }
    K5E();
    for (int bi = nbk - 2; bi >= 0; --bi) {
        dmm_t<TM>(X4, ld, P, ld, Tm, ld, np, np, np, 0, nw);
        __syncthreads();
        const int b0 = 4 * bi;
        const double c0 = inv_fact[b0], c1 = inv_fact[b0 + 1], c2 = inv_fact[b0 + 2], c3 = inv_fact[b0 + 3];
        for (int c = warp; c < np; c += nw)
            for (int i = lane; i < np; i += 32) {
                const size_t t = i + (size_t)ld * c;
                P[t] = Tm[t] + ((i == c && i < n2) ? c0 : 0.0) + c1 * X[t] + c2 * X2[t] + c3 * X3[t];
            }
        __syncthreads();
    }
    K5E();
#ifdef GPTB_K5_DBG
    if (blockIdx.x == 0 && threadIdx.x == 0)
        printf("[k5 dbg] expm n=%d m=%d: norm %lld, X2 %lld, X3|X4 %lld, top %lld, horner %lld\n", n2, m, te[1] - te[0],
               te[2] - te[1], te[3] - te[2], te[4] - te[3], te[5] - te[4]);
#endif
#undef K5E
    double *cur = P, *oth = Tm;
    for (int q = 0; q < sq; ++q) {
        dmm_t<TM>(cur, ld, cur, ld, oth, ld, np, np, np, 0, nw);
        __syncthreads();
        double *sw = cur; cur = oth; oth = sw;
    }
    return cur;
}

// P2' : the r x r / 2r x 2r algebra of geod() (GPT_SGLD.jl:19-37) for one k per CTA, entirely in shared memory on
// the FP64 tensor cores.  The column norms of Unew = U C1 + G C2 come from the r x r Gram matrices, so C1, C2 leave
// already divided by the norms (:30-35) and the apply kernel writes the final U in one pass.
//   in : small[S1 = U'G, UU = U'U, GG = G'G]          out : small[C1, C2]
struct Small2Geo {
    int r2, np2, ld2, npr, ldr;
    size_t sz2, szr;
    __host__ __device__ explicit Small2Geo(int r) {
        r2 = 2 * r;
        npr = (r + 15) & ~15; ldr = dmm_ld(npr); szr = (size_t)ldr * npr;
        np2 = (r2 + 15) & ~15; ld2 = dmm_ld(np2); sz2 = (size_t)ld2 * np2;
        // the r x r scratch of phases 1, 3, 4 lives in the 2r x 2r blocks: when 2r rounds up much less than r does
        // (r = 20: 48 vs 32) the blocks are too small, and a wider zero padding of the 2r x 2r problem is far cheaper
        // than the scalar fallback (k_small + k_apply + k_finish: 82 us against 35 us at r = 20)
        while (!ok(r) && np2 < 2 * npr) {
            np2 += 16; ld2 = dmm_ld(np2); sz2 = (size_t)ld2 * np2;
        }
    }
    // phase 1 keeps eight r x r blocks beside X, the small exponential six beside E, phase 4 nine plus W (np2 x npr)
    __host__ __device__ bool ok(int) const {
        return 8 * szr <= 5 * sz2 && 6 * szr <= 4 * sz2 && 9 * szr + (size_t)ld2 * npr <= 4 * sz2 &&
               npr * npr <= 2 * 1024 /* -tA waits in two registers per thread */;
    }
    __host__ __device__ size_t bytes() const { return 6 * sz2 * sizeof(double); }
};

__global__ void __launch_bounds__(1024)
k_small2(double *__restrict__ small, size_t small_stride, int r, double t, int32_t *nan_flag,
         const double *__restrict__ epsk_sum, double inv_cnt) {
    extern __shared__ double sm[];
    __shared__ int bad;
    __shared__ double cn[64];
    if (epsk_sum) t = sqrt(epsk_sum[blockIdx.x] * inv_cnt);
    double *sk = small + (size_t)blockIdx.x * small_stride;
    SmallLayout L(r);
    const Small2Geo g(r);
    const int r2 = g.r2, npr = g.npr, ldr = g.ldr, np2 = g.np2, ld2 = g.ld2;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
    if (tid == 0) bad = 0;
#ifdef GPTB_K5_DBG   // per-phase clocks of CTA 0 (build with EXTRA_NVCC_FLAGS=-DGPTB_K5_DBG)
    long long tq[8];
    int nq = 0;
#define K5T() do { __syncthreads(); tq[nq++] = clock64(); } while (0)
#else
#define K5T() do {} while (0)
#endif
    K5T();
    // ---- phase 1: r x r algebra in blocks 1..5 (padded, leading dimension ldr), X straight into block 0
    double *Xb = sm;
    double *S1 = sm + g.sz2, *S1t = S1 + g.szr, *UU = S1t + g.szr, *GG = UU + g.szr, *sym = GG + g.szr, *t1 = sym + g.szr,
           *t2 = t1 + g.szr, *t3 = t2 + g.szr;
    for (int c = warp; c < npr; c += nw)
        for (int a = lane; a < npr; a += 32) {
            const bool in = a < r && c < r;
            const double s1 = in ? sk[L.S1 + a + r * c] : 0.0, s1t = in ? sk[L.S1 + c + r * a] : 0.0;
            const size_t o = a + (size_t)ldr * c;
            S1[o] = s1;
            S1t[o] = s1t;
            sym[o] = (s1 + s1t) / 2;
            UU[o] = in ? sk[L.UU + a + r * c] : 0.0;
            GG[o] = in ? sk[L.GG + a + r * c] : 0.0;
        }
    __syncthreads();
    K5T();
    {
        const int h = nw / 2;
        dmm_t<1>(UU, ldr, sym, ldr, t1, ldr, npr, npr, npr, 0, h);         // UU*sym
        dmm_t<1>(S1t, ldr, sym, ldr, t3, ldr, npr, npr, npr, h, nw - h);   // S1'*sym  (sym*S1 is its transpose)
    }
    __syncthreads();
    dmm_t<1>(sym, ldr, t1, ldr, t2, ldr, npr, npr, npr, 0, nw);            // sym*UU*sym
    __syncthreads();
    for (int b = warp; b < np2; b += nw)
        for (int a = lane; a < np2; a += 32) {
            double v = 0.0;
            if (a < r2 && b < r2) {
                if (a < r && b < r) v = S1[a + ldr * b] - t1[a + ldr * b];                              // A = U'M
                else if (a < r) {                                                                       // -M'M
                    const int bb = b - r;
                    v = -(GG[a + ldr * bb] - t3[a + ldr * bb] - t3[bb + ldr * a] + t2[a + ldr * bb]);
                } else if (b < r) v = (a - r == b) ? 1.0 : 0.0;
                else v = S1[(a - r) + ldr * (b - r)] - t1[(a - r) + ldr * (b - r)];
            }
            Xb[a + (size_t)ld2 * b] = t * v;
        }
    // -tA for phase 3 waits in the registers of the threads that will store it (blocks 0..3 are busy until then)
    double xs_keep[2] = {0.0, 0.0};
    {
        int q = 0;
        for (int c = warp; c < npr; c += nw)
            for (int a = lane; a < npr; a += 32, ++q)
                if (q < 2) xs_keep[q] = (a < r && c < r) ? -t * (S1[a + ldr * c] - t1[a + ldr * c]) : 0.0;
    }
    __syncthreads();
    K5T();
    // ---- phase 2: E = expm(t [[A, -M'M], [I, A]])
    double *E = sexpm3<2>(sm, r2, np2, ld2, &bad);
    if (!bad) {
        int local = 0;                 // "if sum(isnan(E))>0" (:24)
        for (int c = warp; c < r2; c += nw)
            for (int a = lane; a < r2; a += 32) {
                const double v = E[a + (size_t)ld2 * c];
                if (v != v) local = 1;
            }
        if (local) bad = 1;
    }
    __syncthreads();
    if (bad) {
        if (tid == 0) atomicExch(nan_flag, 1);
        return;
    }
    K5T();
    // ---- phase 3: expm(-tA) in blocks 0..3 (E sits in block 4 or 5), parked in the other of blocks 4/5
    {
        int q = 0;
        for (int c = warp; c < npr; c += nw)
            for (int a = lane; a < npr; a += 32, ++q) sm[a + (size_t)ldr * c] = q < 2 ? xs_keep[q] : 0.0;
    }
    __syncthreads();
    double *Xe0 = sexpm3<1>(sm, r, npr, ldr, &bad);
    if (bad) {
        if (tid == 0) atomicExch(nan_flag, 1);
        return;
    }
    double *Xe = (E == sm + 4 * g.sz2) ? sm + 5 * g.sz2 : sm + 4 * g.sz2;
    for (int c = warp; c < npr; c += nw)
        for (int a = lane; a < npr; a += 32) Xe[a + (size_t)ldr * c] = Xe0[a + (size_t)ldr * c];
    __syncthreads();
    K5T();
    // ---- phase 4: W = E[:,1:r] * expm(-tA); C1 = W1 - sym*W2, C2 = W2; column norms; blocks 0..3 are scratch
    double *W = sm, *W2 = W + (size_t)ld2 * npr, *C1 = W2 + g.szr, *symb = C1 + g.szr, *M1 = symb + g.szr, *M2 = M1 + g.szr,
           *M3 = M2 + g.szr, *P1 = M3 + g.szr, *P2 = P1 + g.szr, *P3 = P2 + g.szr;
    for (int c = warp; c < npr; c += nw)
        for (int a = lane; a < npr; a += 32) {
            const bool in = a < r && c < r;
            const size_t o = a + (size_t)ldr * c;
            const double s1 = in ? sk[L.S1 + a + r * c] : 0.0, s1t = in ? sk[L.S1 + c + r * a] : 0.0;
            symb[o] = (s1 + s1t) / 2;
            M1[o] = in ? sk[L.UU + a + r * c] : 0.0;
            M2[o] = s1;
            M3[o] = in ? sk[L.GG + a + r * c] : 0.0;
        }
    dmm_t<1>(E, ld2, Xe, ldr, W, ld2, np2, npr, npr, 0, nw);   // (Xe is zero outside r x r, so the extra columns of E drop out)
    __syncthreads();
    int ldw2 = ldr;
    if (r == npr) {                                            // rows r..2r-1 of W are a padded block as they stand
        W2 = W + r;
        ldw2 = ld2;
    } else {
        for (int c = warp; c < npr; c += nw)
            for (int a = lane; a < npr; a += 32)
                W2[a + (size_t)ldr * c] = (a < r && c < r) ? W[r + a + (size_t)ld2 * c] : 0.0;
        __syncthreads();
    }
    // C1 = W1 - sym*W2 (rows/columns >= r of W1 are zero: E and Xe are zero padded)
    dmm_t<1>(symb, ldr, W2, ldw2, C1, ldr, npr, npr, npr, 0, nw, W, ld2);
    __syncthreads();
    {   // squared column norms of U C1 + G C2:  diag(C1'UU C1 + 2 C1'S1 C2 + C2'GG C2)
        const int h = nw / 3;
        dmm_t<1>(M1, ldr, C1, ldr, P1, ldr, npr, npr, npr, 0, h);
        dmm_t<1>(M2, ldr, W2, ldw2, P2, ldr, npr, npr, npr, h, h);
        dmm_t<1>(M3, ldr, W2, ldw2, P3, ldr, npr, npr, npr, 2 * h, nw - 2 * h);
    }
    __syncthreads();
    for (int l = warp; l < r; l += nw) {
        double sc = 0.0;
        for (int a = lane; a < r; a += 32) {
            const size_t o = a + (size_t)ldr * l;
            sc += C1[o] * (P1[o] + 2.0 * P2[o]) + W2[a + (size_t)ldw2 * l] * P3[o];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sc += __shfl_xor_sync(0xffffffffu, sc, o);
        if (lane == 0) cn[l] = 1.0 / sqrt(sc);
    }
    __syncthreads();
    for (int l = warp; l < r; l += nw)
        for (int a = lane; a < r; a += 32) {
            sk[L.C1 + a + r * l] = C1[a + (size_t)ldr * l] * cn[l];
            sk[L.C2 + a + r * l] = W2[a + (size_t)ldw2 * l] * cn[l];
        }
    K5T();
#ifdef GPTB_K5_DBG
    if (blockIdx.x == 0 && tid == 0)
        printf("[k5 dbg] clk: load %lld, phase1 %lld, expm(2r) %lld, expm(r) %lld, phase4 %lld\n", tq[1] - tq[0],
               tq[2] - tq[1], tq[3] - tq[2], tq[4] - tq[3], tq[5] - tq[4]);
#endif
#undef K5T
}

// P3' : U <- U C1 + G C2 (C1, C2 pre-divided by the column norms), plus the fp16 hi|lo operand image of the new U
// with the fixed scale 2^8 (|U| <= 1 once the columns are unit vectors).
constexpr float UIMG_FIXED_SCALE = 256.f;
__global__ void __launch_bounds__(UP_THREADS)
k_apply2(double *__restrict__ U, const double *__restrict__ G, const double *__restrict__ small, size_t small_stride, int n,
         int r, const int32_t *__restrict__ nan_flag, __half *__restrict__ Uimg, int npad, int rpad,
         float *__restrict__ uscale_out) {
    if (*nan_flag) return;
    extern __shared__ double sm[];
    const int tid = threadIdx.x, k = blockIdx.y, j0 = blockIdx.x * UP_ROWS;
    const int rows = min(UP_ROWS, n - j0);
    const double *sk = small + (size_t)k * small_stride;
    SmallLayout L(r);
    // one product [U | G] (UP_ROWS x 2 npr) * [C1 ; C2] (2 npr x npr) on the FP64 tensor cores
    const int npr = (r + 15) & ~15, lda = dmm_ld(UP_ROWS), ldb = dmm_ld(2 * npr), ldc = lda;
    double *As = sm, *Bs = As + (size_t)lda * 2 * npr, *Cs = Bs + (size_t)ldb * npr;
    double *Uk = U + (size_t)k * r * n;
    const double *Gk = G + (size_t)k * r * n;
    if (uscale_out && blockIdx.x == 0 && k == 0 && tid == 0) {
        uscale_out[0] = UIMG_FIXED_SCALE;
        uscale_out[1] = 1.f / UIMG_FIXED_SCALE;
    }
    for (int t = tid; t < UP_ROWS * npr; t += UP_THREADS) {
        const int jj = t % UP_ROWS, l = t / UP_ROWS;
        const bool ok = jj < rows && l < r;
        As[jj + (size_t)lda * l] = ok ? Uk[(size_t)l * n + j0 + jj] : 0.0;
        As[jj + (size_t)lda * (npr + l)] = ok ? Gk[(size_t)l * n + j0 + jj] : 0.0;
    }
    for (int t = tid; t < npr * npr; t += UP_THREADS) {
        const int a = t % npr, l = t / npr;
        const bool ok = a < r && l < r;
        Bs[a + (size_t)ldb * l] = ok ? sk[L.C1 + a + r * l] : 0.0;
        Bs[npr + a + (size_t)ldb * l] = ok ? sk[L.C2 + a + r * l] : 0.0;
    }
    __syncthreads();
    dmm_t<1>(As, lda, Bs, ldb, Cs, ldc, UP_ROWS, npr, 2 * npr, 0, UP_THREADS / 32);
    __syncthreads();
    for (int t = tid; t < UP_ROWS * r; t += UP_THREADS) {
        const int jj = t % UP_ROWS, l = t / UP_ROWS;
        if (jj < rows) Uk[(size_t)l * n + j0 + jj] = Cs[jj + (size_t)ldc * l];
    }
    if (Uimg) {
        // image rows [j/8][nn/8][nn%8][j%8]: one 16-byte store per (8-row block, nn); rows past n and columns past r
        // are zero in Cs (zero padded operands), which is what the image's padding must hold
        for (int t = tid; t < (UP_ROWS / 8) * rpad; t += UP_THREADS) {
            const int l = t % rpad, jb = t / rpad;
            if (j0 + jb * 8 >= npad) continue;
            __align__(16) __half hi[8], lo[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float v = l < npr ? (float)Cs[jb * 8 + e + (size_t)ldc * l] * UIMG_FIXED_SCALE : 0.f;
                hi[e] = __float2half_rn(v);
                lo[e] = __float2half_rn(v - __half2float(hi[e]));
            }
            *reinterpret_cast<uint4 *>(Uimg + gptb_uimg_offset(k, j0 + jb * 8, l, npad, rpad)) = *reinterpret_cast<uint4 *>(hi);
            *reinterpret_cast<uint4 *>(Uimg + gptb_uimg_offset(k, j0 + jb * 8, rpad + l, npad, rpad)) = *reinterpret_cast<uint4 *>(lo);
        }
    }
}

}  // namespace

size_t gptb_update_small_stride(int r) { return SmallLayout(r).total; }

// k_prepare / k_apply / k_reduce_only use up to ~99 KB of dynamic shared memory at r = 64: opt in once.
static void optin_smem() {
    // (function attributes are per device: a process that drives several GPUs opts in on each)
    static std::atomic<uint64_t> done{0};
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_acquire) & bit) return;
    cudaFuncSetAttribute(k_prepare, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(k_apply, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(k_reduce_only, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    cudaFuncSetAttribute(k_small, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_small2, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024);
    cudaFuncSetAttribute(k_apply2, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    done.fetch_or(bit, std::memory_order_release);
}

// shared-memory expm workspace for the 2r x 2r exponential (0 = does not fit, use the global-memory Pade path)
static size_t small_smem_bytes(int r) {
    size_t ld = (size_t)((2 * r + 3) & ~3);
    size_t b = 6 * ld * ld * sizeof(double);
    return b <= 200 * 1024 ? b : 0;
}

int gptb_update_alloc_noise(gptb_handle *h) {
    const gptb_config &c = h->cfg;
    const size_t per_k = (size_t)c.r * c.n;
    if (!h->gen_noise_w) GPTB_CUDA_OK(h, cudaMalloc((void **)&h->gen_noise_w, sizeof(double) * (c.Q + 2)));
    if (!h->gen_noise_U) GPTB_CUDA_OK(h, cudaMalloc((void **)&h->gen_noise_U, sizeof(double) * (per_k * c.D + 2)));
    return GPTB_OK;
}

// The step's production-mode draws for the dimensions [k0, k0 + kn) (and w): they depend on the step counter only, so
// the one-stream step generates them on a side stream under the forward / backward kernels.
int gptb_launch_noise_range(gptb_handle *h, uint64_t step, int k0, int kn, bool with_w, cudaStream_t s) {
    const gptb_config &c = h->cfg;
    const size_t per_k = (size_t)c.r * c.n;
    GPTB_TRY(gptb_update_alloc_noise(h));
    const int gx = (int)std::min<size_t>((per_k / 2 + 255) / 256, 64);
    k_noise<<<dim3(gx, kn + (with_w ? 1 : 0)), 256, 0, s>>>(h->gen_noise_w, h->gen_noise_U, c.Q, per_k, k0, kn, c.seed, step);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// The update of the dimensions [k0, k0 + kn) (and of w when with_w) on stream s.  The kernels address everything by
// the global k, so that the Philox streams and hence the chain do not depend on how the dimensions are grouped.
int gptb_launch_update_range(gptb_handle *h, double epsw, double epsU, int flags, const double *noise_w,
                             const double *noise_U, uint64_t step, int k0, int kn, bool with_w, cudaStream_t s,
                             bool noise_ready) {
    const gptb_config &c = h->cfg;
    const int nb = (c.n + UP_ROWS - 1) / UP_ROWS;
    const bool stiefel = flags & GPTB_FLAG_STIEFEL, langevin = flags & GPTB_FLAG_LANGEVIN;
    const int gy = kn + (with_w ? 1 : 0);
    optin_smem();
    double *small_g = h->small + (size_t)k0 * h->small_stride;
    if (stiefel && kn > 0) {
        SmallLayout L(c.r);
        GPTB_CUDA_OK(h, cudaMemset2DAsync(small_g, h->small_stride * sizeof(double), 0, L.zero_count() * sizeof(double), kn, s));
    }
    if (langevin && (!noise_w || !noise_U)) {
        // production mode: the step's draws are generated up front by a grid-wide kernel (two per Philox block)
        if (!noise_ready) GPTB_TRY(gptb_launch_noise_range(h, step, k0, kn, with_w, s));
        if (!noise_w) noise_w = h->gen_noise_w;
        if (!noise_U) noise_U = h->gen_noise_U;
    }
    // RMSprop step: every dimension has its own step size (the kernels index the array by their global / local k)
    const double *epsk = h->rms_active ? h->rms_epsk : nullptr;
    const double inv_cnt = 1.0 / ((double)c.n * c.r);
    size_t sm1 = 2 * (size_t)UP_ROWS * dmm_ld((c.r + 15) & ~15) * sizeof(double) +
                 (h->f32_grad ? (size_t)UP_ROWS * (h->rpad + 1) * sizeof(float) : 0);
    k_prepare<<<dim3(nb, gy), UP_THREADS, sm1, s>>>(h->grad, h->w, h->U, h->G, h->small, h->small_stride, c.n, c.r, c.D, c.Q,
                                                   epsw, epsU, flags, noise_w, noise_U, c.seed, step, 1, k0, kn,
                                                   h->f32_grad ? h->gradFbase : nullptr, h->gradF_qpad, h->gradF_qstride,
                                                   h->npad, h->rpad, epsk, inv_cnt);
    h->launches++;
    if (kn > 0) h->shadows_valid = false;
    if (stiefel && kn > 0) {
        const size_t sb = small_smem_bytes(c.r);
        const Small2Geo sg(c.r);
        const size_t koff = (size_t)k0 * c.r * c.n;
        if (sg.ok(c.r) && sg.bytes() <= 225 * 1024) {
            k_small2<<<kn, 1024, sg.bytes(), s>>>(small_g, h->small_stride, c.r, sqrt(epsU), h->nan_flag, epsk ? epsk + k0 : nullptr,
                                                  inv_cnt);
            const bool tc = gptb_is_tc(h);
            const int npr = (c.r + 15) & ~15;
            size_t sm3 = ((size_t)dmm_ld(UP_ROWS) * 3 * npr + (size_t)dmm_ld(2 * npr) * npr) * sizeof(double);
            __half *img = tc ? (__half *)h->Uimg + (size_t)k0 * h->npad * 2 * h->rpad : nullptr;
            k_apply2<<<dim3(nb, kn), UP_THREADS, sm3, s>>>(h->U + koff, h->G + koff, small_g, h->small_stride, c.n, c.r,
                                                          h->nan_flag, img, h->npad, h->rpad,
                                                          tc ? h->tc_state + 4 : nullptr);
            h->launches += 2;
            if (tc) h->shadows_valid = true;   // k_apply2 refreshed the operand image
        } else {
            k_small<<<kn, 1024, sb, s>>>(small_g, h->small_stride, c.r, sqrt(epsU), 1, h->nan_flag, sb ? 1 : 0,
                                         epsk ? epsk + k0 : nullptr, inv_cnt);
            size_t sm3 = (2 * (size_t)UP_ROWS * c.r + 2 * (size_t)c.r * c.r + c.r) * sizeof(double);
            k_apply<<<dim3(nb, kn), UP_THREADS, sm3, s>>>(h->U + koff, h->G + koff, small_g, h->small_stride, c.n, c.r,
                                                         h->nan_flag);
            int fb = (int)std::min<size_t>(((size_t)c.r * c.n + 255) / 256, 64);
            k_finish<<<dim3(fb, kn), 256, 0, s>>>(h->U + koff, h->G + koff, small_g, h->small_stride, c.n, c.r, h->nan_flag);
            h->launches += 3;
        }
    }
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// RMSprop: the w half of the step (after the gradient of w is complete in h->grad, before computeA).
int gptb_launch_rms_w(gptb_handle *h, int flags, const double *noise_w, uint64_t step, cudaStream_t s) {
    const gptb_config &c = h->cfg;
    k_rms_w<<<(c.Q + 255) / 256, 256, 0, s>>>(h->grad, h->w, h->w_old, h->rms_gw, c.Q, h->rms_N, h->rms_eps, h->rms_alpha,
                                              h->rms_lambda, noise_w, c.seed, step, (flags & GPTB_FLAG_LANGEVIN) ? 1 : 0);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

// RMSprop: moving average of gradU^2 and the per-dimension step-size sums (before the update kernels).
int gptb_launch_rms_U(gptb_handle *h, cudaStream_t s) {
    const gptb_config &c = h->cfg;
    const size_t per_k = (size_t)c.r * c.n;
    GPTB_CUDA_OK(h, cudaMemsetAsync(h->rms_epsk, 0, sizeof(double) * c.D, s));
    k_rms_U<<<dim3((unsigned)std::min<size_t>((per_k + 255) / 256, 64), c.D), 256, 0, s>>>(h->grad + c.Q, h->rms_gU, h->rms_epsk, per_k,
                                                                                          h->rms_N, h->rms_eps, h->rms_alpha,
                                                                                          h->rms_lambda);
    h->launches++;
    GPTB_CUDA_OK(h, cudaGetLastError());
    return GPTB_OK;
}

int gptb_launch_update(gptb_handle *h, double epsw, double epsU, int flags, const double *noise_w,
                       const double *noise_U, uint64_t step) {
    return gptb_launch_update_range(h, epsw, epsU, flags, noise_w, noise_U, step, 0, h->cfg.D, true, h->stream, false);
}

// Standalone proj (mode 0) / geod (mode 1) on device buffers U, V (n x r col-major); scratch from `h`.
int gptb_proj_geod_device(gptb_handle *h, cudaStream_t s, int n, int r, const double *U, const double *V, double t,
                          double *out, int mode, int32_t *nan_flag, double *out2) {
    SmallLayout L(r);
    double *small = nullptr;
    optin_smem();
    GPTB_CUDA_OK(h, cudaMalloc(&small, L.total * sizeof(double)));
    cudaMemsetAsync(small, 0, L.zero_count() * sizeof(double), s);
    const int nb = (n + UP_ROWS - 1) / UP_ROWS;
    size_t sm1 = 2 * (size_t)UP_ROWS * r * sizeof(double);
    k_reduce_only<<<nb, UP_THREADS, sm1, s>>>(U, V, small, n, r);
    if (mode == 0) {
        int fb = (int)std::min<size_t>(((size_t)r * n + 255) / 256, 1024);
        k_proj_apply<<<fb, 256, 0, s>>>(U, V, small + L.S1, n, r, out);
    } else {
        {
            size_t sb = small_smem_bytes(r);
            k_small<<<1, 1024, sb, s>>>(small, L.total, r, t, 0, nan_flag, sb ? 1 : 0, nullptr, 0.0, out2 ? 1 : 0);
        }
        if (out2) {   // geodboth: tmpV = [U mom] E[:, r+1:2r] expm(-tA), not normalised (:50)
            int fb2 = (int)std::min<size_t>(((size_t)r * n + 255) / 256, 1024);
            k_apply_plain<<<fb2, 256, 0, s>>>(U, V, small + L.A, small + L.MM, n, r, out2);
        }
        // Unew = U C1 + V C2 written over a copy of V (k_apply overwrites its G argument)
        cudaMemcpyAsync(out, V, sizeof(double) * (size_t)n * r, cudaMemcpyDeviceToDevice, s);
        size_t sm3 = (2 * (size_t)UP_ROWS * r + 2 * (size_t)r * r + r) * sizeof(double);
        k_apply<<<dim3(nb, 1), UP_THREADS, sm3, s>>>(U, out, small, L.total, n, r, nan_flag);
        // normalise in place: k_finish reads G (= out) and writes U (= out)
        int fb = (int)std::min<size_t>(((size_t)r * n + 255) / 256, 64);
        k_finish<<<dim3(fb, 1), 256, 0, s>>>(out, out, small, L.total, n, r, nan_flag);
    }
    cudaError_t e = cudaGetLastError();
    cudaError_t e2 = cudaStreamSynchronize(s);
    cudaFree(small);
    if (e != cudaSuccess || e2 != cudaSuccess) {
        gptb_set_error(h, "proj/geod kernels: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
        return GPTB_ERR_CUDA;
    }
    return GPTB_OK;
}

// ================================================================================================
// gptb_proj_noalloc — proj(U,V) for callers that must not allocate (CUDA graph capture, serving loops).
// No counterpart in the reference (see include/gptb200.h).
//
// Same two kernels as gptb_proj (k_reduce_only -> k_proj_apply), but the r x r Gram scratch lives in the file-scope
// device array below instead of a cudaMalloc'd block.  Every GPU that loads the library holds the array until exit.
// In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, taken once per device with that device current.
// ================================================================================================
constexpr int NOALLOC_MAX_R = 64;
constexpr size_t NOALLOC_SCRATCH_DOUBLES = 3 * (size_t)NOALLOC_MAX_R * NOALLOC_MAX_R + NOALLOC_MAX_R;   // 98,816 bytes
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
double *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};   // device address of g_noalloc_scratch, per device ordinal
}  // namespace

// Set-up, once per device (the calling thread's current device): resolves the scratch address and opts the reduction
// kernel into the dynamic shared memory it needs at r = 64.  Not thread-safe against a concurrent first call on the
// same device (like the handles of this library: one caller at a time).
static int noalloc_setup(int dev, double **out) {
    if (g_noalloc_ptr[dev]) {
        *out = g_noalloc_ptr[dev];
        return GPTB_OK;
    }
    double *ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
        gptb_set_error(nullptr, "gptb_proj_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    e = cudaFuncSetAttribute(k_reduce_only, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    if (e != cudaSuccess) {
        gptb_set_error(nullptr, "gptb_proj_noalloc: cudaFuncSetAttribute on device %d: %s", dev, cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    g_noalloc_ptr[dev] = ptr;
    *out = ptr;
    return GPTB_OK;
}

extern "C" int gptb_proj_noalloc(int32_t n, int32_t r, const double *U_dev, const double *V_dev, double *out_dev,
                                 void *cuda_stream) {
    if (!U_dev || !V_dev || !out_dev || n < 1 || r < 1) {
        gptb_set_error(nullptr, "gptb_proj_noalloc: NULL pointer or n < 1 or r < 1 (n=%d, r=%d)", n, r);
        return GPTB_ERR_ARG;
    }
    if (r > NOALLOC_MAX_R) {
        gptb_set_error(nullptr, "gptb_proj_noalloc: r=%d exceeds the static scratch (r <= %d: %zu bytes per GPU); use gptb_proj",
                       r, NOALLOC_MAX_R, NOALLOC_SCRATCH_DOUBLES * sizeof(double));
        return GPTB_ERR_UNSUPPORTED;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
        gptb_set_error(nullptr, "gptb_proj_noalloc: no usable current CUDA device (%s)", cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    double *scratch = nullptr;
    GPTB_TRY(noalloc_setup(dev, &scratch));
    cudaStream_t s = (cudaStream_t)cuda_stream;
    SmallLayout L(r);
    // the array keeps the previous call's sums: the accumulators this call adds into must start at zero
    e = cudaMemsetAsync(scratch, 0, L.zero_count() * sizeof(double), s);
    if (e != cudaSuccess) {
        gptb_set_error(nullptr, "gptb_proj_noalloc: clearing the scratch: %s", cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    const int nb = (n + UP_ROWS - 1) / UP_ROWS;
    const size_t sm1 = 2 * (size_t)UP_ROWS * r * sizeof(double);
    k_reduce_only<<<nb, UP_THREADS, sm1, s>>>(U_dev, V_dev, scratch, n, r);
    const int fb = (int)std::min<size_t>(((size_t)r * n + 255) / 256, 1024);
    k_proj_apply<<<fb, 256, 0, s>>>(U_dev, V_dev, scratch + L.S1, n, r, out_dev);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        gptb_set_error(nullptr, "gptb_proj_noalloc: kernel launch: %s", cudaGetErrorString(e));
        return GPTB_ERR_CUDA;
    }
    return GPTB_OK;
}
