"""CPU restatement (NumPy/SciPy, Float64) of the reference's tensor-RFF GP SGLD path.

TEST INFRASTRUCTURE ONLY.  Nothing under gpt_b200/ may import this module; only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do, and there
only as the checker (or as the timed CPU baseline), never as the product.

Every function follows one function of /root/reference/GPT_SGLD.jl (Julia 0.3, column-major
Float64, 1-based Int32 index table I) and cites it.  Arrays keep the *Julia* shapes:

    X[N,D]  Z[n,D]  b[n,D]  phi[n,D,N]  U[n,r,D]  w[Q]  I[Q,D] (1-based)  temp[D,r,N]
    V[Q,N]  U_phi[Q,N,D]  A[r,D,N]  Psi[n*r,N,D]

Pinning status (see tests/test_oracle_golden.py):
  * `pred`        — pinned by TensorSynthData{5,10}D100N.h5 (y_j - pred(w,U,I,phi) is the
                    N(0, 0.1/0.01/0.001) noise MakeSynthData.jl:11-13 added).
  * `feature`     — functional form and amplitude pinned by the same files (phi[j,k,:] is an
                    exact cosine in X[:,k] of amplitude phi_scale*sqrt(2/n)); Z,b were not saved.
  * gradients     — pinned against finite differences of the reference's own log-likelihood
                    restatement (Diagnostic_gradients.jl:5-37, :133-150).
  * `geod`/`expm` — the reference holds no known-answer test for these call sites (GPT_SGLD.jl:23,28) and
                    Julia 0.3 Base `expm` (Pade-13 scaling and squaring) is replaced by scipy.linalg.expm
                    (same algorithm family).  Pinned independently instead: against the closed-form great
                    circle for r = 1 (`geod_great_circle`) and against a 40-digit mpmath evaluation of the
                    same lines (`expm_mp`, `geod_mp`) at r in {2, 10, 32}, t in {1e-7, 1e-3, 0.3}.
  * `sqrtm`/RNG   — **parity unpinned**: `sqrtm` (:366) only enters the initial draw, which the tests inject;
                    Julia's dSFMT streams cannot be reproduced (draws are injected on both sides).
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import expm as _expm


# --------------------------------------------------------------------------- features
def feature(X, length_scale, sigma_RBF, phi_scale, Z, b):
    """GPT_SGLD.jl:71-84.  phi[j,k,i] = phi_scale*sigma_RBF^(1/D)*sqrt(2/n)*cos(X[i,k]*Zt[j,k]+b[j,k]),
    Zt = scale(Z, 1./length_scale) (:75) — scalar length_scale divides every column."""
    X = np.asarray(X, dtype=np.float64)
    Z = np.asarray(Z, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    N, D = X.shape
    n = Z.shape[0]
    Zt = Z * (1.0 / np.broadcast_to(np.asarray(length_scale, dtype=np.float64), (D,)))[None, :]
    # phi[j,k,i] : (n,D,N)
    phi = np.cos(X.T[None, :, :] * Zt[:, :, None] + b[:, :, None])
    return phi_scale * sigma_RBF ** (1.0 / D) * np.sqrt(2.0 / n) * phi


def feature2(X, length_scale, sigma_RBF, phi_scale, Z):
    """GPT_SGLD.jl:88-105.  sin/cos pairs from the first n/2 rows of Z; errors when n is odd (:89-91)."""
    Z = np.asarray(Z, dtype=np.float64)
    if Z.shape[0] % 2 == 1:
        raise ValueError("n not even")
    half_n = Z.shape[0] // 2
    X = np.asarray(X, dtype=np.float64)
    N, D = X.shape
    Zt = Z * (1.0 / np.broadcast_to(np.asarray(length_scale, dtype=np.float64), (D,)))[None, :]
    arg = X.T[None, :, :] * Zt[:half_n, :, None]          # (half_n, D, N)
    phi = np.empty((2 * half_n, D, N))
    phi[0::2] = np.sin(arg)                               # Julia rows 2j-1
    phi[1::2] = np.cos(arg)                               # Julia rows 2j
    return phi_scale * sigma_RBF ** (1.0 / D) * phi / np.sqrt(half_n)


def tgp_feature(x, Z_over_sigma, b01):
    """TGP.jl:6-14 for one row x[D]: sqrt(2/n)*cos(x .* Z + 2*pi*b) with Z=randn(n,D)/sigmaRBF and
    b=rand(n,D) supplied by the caller (the reference reseeds Julia's RNG; its stream cannot be
    reproduced here)."""
    Z = np.asarray(Z_over_sigma, dtype=np.float64)
    n = Z.shape[0]
    return np.sqrt(2.0 / n) * np.cos(np.asarray(x, dtype=np.float64)[None, :] * Z + 2 * np.pi * np.asarray(b01))


def datawhitening(X):
    """GPT_SGLD.jl:62-67: per-column (x-mean)/std with Julia's std (n-1 denominator); in place."""
    X = np.asarray(X, dtype=np.float64)
    for i in range(X.shape[1]):
        X[:, i] = (X[:, i] - X[:, i].mean()) / X[:, i].std(ddof=1)
    return X


def samplenz(r, D, Q, rng):
    """GPT_SGLD.jl:181-190: Q distinct cells L of 0:(r^D-1); I[q,:] = digits(L[q], r, D)+1
    (little-endian base-r digits).  `rng` is a numpy Generator (Julia's sample() stream is not
    reproducible); the digit convention and distinctness are what is restated."""
    total = r ** D
    if total <= 2 ** 62 and total <= 10_000_000:
        L = rng.choice(total, size=Q, replace=False)
    else:
        seen = set()
        while len(seen) < Q:
            seen.add(int(rng.integers(0, total)))
        L = np.array(sorted(seen), dtype=object)
        rng.shuffle(L)
    I = np.empty((Q, D), dtype=np.int32)
    for q in range(Q):
        v = int(L[q])
        for k in range(D):
            I[q, k] = v % r + 1
            v //= r
    return I


def stiefel_from_gaussian(Zrn):
    """GPT_SGLD.jl:365-366: U = transpose(sqrtm(Z*Z') \\ Z) for Z = randn(r,n): uniform on V_{n,r}."""
    Zrn = np.asarray(Zrn, dtype=np.float64)
    G = Zrn @ Zrn.T
    lam, Vv = np.linalg.eigh(G)
    inv_sqrt = (Vv / np.sqrt(lam)) @ Vv.T
    return (inv_sqrt @ Zrn).T


# --------------------------------------------------------------------------- forward model
def phidotU(U, phi):
    """GPT_SGLD.jl:193-205: temp[k,l,i] = dot(phi[:,k,i], U[:,l,k])."""
    return np.einsum("jki,jlk->kli", phi, U, optimize=True)


def computeV(temp, I):
    """GPT_SGLD.jl:208-220: V[q,i] = prod_k temp[k, I[q,k], i]."""
    Q, D = I.shape
    V = np.ones((Q, temp.shape[2]))
    for k in range(D):
        V *= temp[k, I[:, k] - 1, :]
    return V


def computefhat(V, w):
    """GPT_SGLD.jl:223-230: fhat[i] = dot(V[:,i], w)."""
    return V.T @ w


def pred(w, U, I, phitest):
    """GPT_SGLD.jl:233-243."""
    return computefhat(computeV(phidotU(U, phitest), I), w)


# --------------------------------------------------------------------------- gradient pieces
def computeU_phi(V, temp, I):
    """GPT_SGLD.jl:246-258: U_phi[q,i,k] = V[q,i] / temp[k, I[q,k], i]  (division, as the reference)."""
    Q, D = I.shape
    out = np.empty((Q, V.shape[1], D))
    for k in range(D):
        out[:, :, k] = V / temp[k, I[:, k] - 1, :]
    return out


def computeA(U_phi, w, I, r):
    """GPT_SGLD.jl:261-273: A[l,k,i] = sum_{q: I[q,k]=l} U_phi[q,i,k]*w[q]; unused l stay 0 (:263)."""
    Q, N, D = U_phi.shape
    A = np.zeros((r, D, N))
    for k in range(D):
        np.add.at(A[:, k, :], I[:, k] - 1, U_phi[:, :, k] * w[:, None])
    return A


def computePsi(A, phi):
    """GPT_SGLD.jl:276-286: Psi[:,i,k] = kron(A[:,k,i], phi[:,k,i])  (block l = A[l]*phi)."""
    r, D, N = A.shape
    n = phi.shape[0]
    return np.einsum("lki,jki->ljik", A, phi).reshape(r * n, N, D)   # row index l*n + j


def gradients(phi_batch, y_batch, w, U, I, N_total, signal_var, sigma_w=1.0):
    """GPT_SGLD.jl:384-408 for one minibatch.  Returns (gradw, gradU, fhat).

    gradw = (N/bs) V (y-fhat)/signal_var - w/sigma_w^2                      (:393)
    gradU[:,:,k] = reshape((N/bs) Psi[:,:,k] (y-fhat)/signal_var, n, r)     (:405-408)
    Psi is never formed: Psi[:,:,k] e == vec(phi_k diag(e) A_k') (block l = sum_i A[l,k,i] e_i phi[:,k,i])."""
    n, D, bs = phi_batch.shape
    r = U.shape[1]
    temp = phidotU(U, phi_batch)
    V = computeV(temp, I)
    fhat = computefhat(V, w)
    e = (N_total / bs) * (y_batch - fhat) / signal_var
    gradw = V @ e - w / sigma_w ** 2
    A = computeA(computeU_phi(V, temp, I), w, I, r)
    gradU = np.einsum("jki,lki,i->jlk", phi_batch, A, e, optimize=True)
    return gradw, gradU, fhat


def gradients_literal(phi_batch, y_batch, w, U, I, N_total, signal_var, sigma_w=1.0):
    """Same as `gradients` but through the literal Psi array (GPT_SGLD.jl:402,405-408); small cases only."""
    n, D, bs = phi_batch.shape
    r = U.shape[1]
    temp = phidotU(U, phi_batch)
    V = computeV(temp, I)
    fhat = computefhat(V, w)
    gradw = (N_total / bs) * V @ (y_batch - fhat) / signal_var - w / sigma_w ** 2
    A = computeA(computeU_phi(V, temp, I), w, I, r)
    Psi = np.empty((n * r, bs, D))
    for i in range(bs):
        for k in range(D):
            Psi[:, i, k] = np.kron(A[:, k, i], phi_batch[:, k, i])
    gradU = np.empty((n, r, D))
    for k in range(D):
        gradU[:, :, k] = ((N_total / bs) * Psi[:, :, k] @ (y_batch - fhat) / signal_var).reshape(n, r, order="F")
    return gradw, gradU, fhat


def loglik(phi, y, w, U, I, sigma):
    """Diagnostic_gradients.jl:5-37: log p(y|x,w,U) = -||y - fhat||^2 / (2 sigma^2)."""
    return -np.sum((y - pred(w, U, I, phi)) ** 2) / (2 * sigma ** 2)


# --------------------------------------------------------------------------- Stiefel manifold
def proj(U, V):
    """GPT_SGLD.jl:14-16: V - U (U'V + V'U)/2."""
    return V - U @ (U.T @ V + V.T @ U) / 2


def geod(U, mom, t):
    """GPT_SGLD.jl:19-37.  Returns (Unew, nan_flag); on NaN in expm the reference prints a warning and
    returns zeros(n,r) (:24-26)."""
    n, r = U.shape
    A = U.T @ mom
    temp = np.block([[A, -mom.T @ mom], [np.eye(r), A]])
    with np.errstate(all="ignore"):
        E = _expm(t * temp)
    if np.isnan(E).any():
        return np.zeros((n, r)), True
    mexp = _expm(-t * A)
    tmpU = np.hstack([U, mom]) @ E[:, :r] @ mexp
    return tmpU / np.linalg.norm(tmpU, axis=0, keepdims=True), False


def geod_great_circle(U, mom, t):
    """Closed form of geod for r = 1 (V_{n,1} is the unit sphere): with u'm = 0 the geodesic of GPT_SGLD.jl:19-37 is the
    great circle  u cos(t|m|) + (m/|m|) sin(t|m|).  An independent pin of `geod` (no expm involved)."""
    u = np.asarray(U, dtype=np.float64).reshape(-1)
    m = np.asarray(mom, dtype=np.float64).reshape(-1)
    nm = np.linalg.norm(m)
    return (u * np.cos(t * nm) + (m / nm) * np.sin(t * nm)).reshape(-1, 1)


def expm_mp(A, dps=40):
    """Matrix exponential in arbitrary precision (mpmath, `dps` digits) — the pin for the expm calls of GPT_SGLD.jl:23,28
    that is independent of SciPy / LAPACK and of every double-precision algorithm."""
    import mpmath as mp
    with mp.workdps(dps):
        E = mp.expm(mp.matrix(np.asarray(A, dtype=np.float64).tolist()))
        return np.array([[float(E[i, j]) for j in range(E.cols)] for i in range(E.rows)])


def geod_mp(U, mom, t, dps=40):
    """GPT_SGLD.jl:19-37 with both exponentials and all products carried in `dps`-digit arithmetic (mpmath)."""
    import mpmath as mp
    n, r = U.shape
    with mp.workdps(dps):
        Um, Mm = mp.matrix(U.tolist()), mp.matrix(mom.tolist())
        A = Um.T * Mm                                                   # :20
        T = mp.zeros(2 * r)
        MM = Mm.T * Mm
        for a in range(r):
            for b in range(r):
                T[a, b] = A[a, b]; T[a, r + b] = -MM[a, b]; T[r + a, r + b] = A[a, b]
            T[r + a, a] = 1                                             # :21-22
        E = mp.expm(mp.mpf(t) * T)                                      # :23
        mexp = mp.expm(-mp.mpf(t) * A)                                  # :28
        UM = mp.zeros(n, 2 * r)
        for j in range(n):
            for a in range(r):
                UM[j, a] = Um[j, a]; UM[j, r + a] = Mm[j, a]
        tmp = UM * E[:, 0:r] * mexp                                     # :29
        out = np.array([[float(tmp[j, a]) for a in range(r)] for j in range(n)])
    return out / np.linalg.norm(out, axis=0, keepdims=True)            # :30-35


# --------------------------------------------------------------------------- sampler
def sgld_update(w, U, gradw, gradU, epsw, epsU, noise_w=None, noise_U=None, langevin=True, stiefel=True):
    """The update half of the loop body, GPT_SGLD.jl:411-439, from given gradients.  Returns (w, U, nan_flag)."""
    n, r, D = U.shape
    if langevin:
        w_new = w + epsw * gradw / 2 + np.sqrt(epsw) * noise_w          # :412
    else:
        w_new = w + epsw * gradw / 2                                    # :414
    U_new = U.copy()
    if stiefel:
        for k in range(D):
            G = np.sqrt(epsU) * gradU[:, :, k] / 2
            if langevin:
                G = G + noise_U[:, :, k]
            mom = proj(U[:, :, k], G)                                   # :420 / :431
            Uk, bad = geod(U[:, :, k], mom, np.sqrt(epsU))              # :421 / :432
            if bad:
                return w_new, U_new, True                               # :422-424
            U_new[:, :, k] = Uk
    else:
        U_new = U + epsU * (gradU - n * U) / 2                          # :437
        if langevin:
            U_new = U_new + np.sqrt(epsU) * noise_U                     # :426
    return w_new, U_new, False


def sgld_step(phi_batch, y_batch, w, U, I, N_total, signal_var, epsw, epsU,
              noise_w=None, noise_U=None, langevin=True, stiefel=True):
    """One iteration of the loop body GPT_SGLD.jl:378-439.  noise_w[Q], noise_U[n,r,D] are the
    randn draws of :412 and :420/:426 (supplied by the caller).  Returns (w, U, fhat, nan_flag).
    Gradients use the pre-update w (:399 precedes :412)."""
    gradw, gradU, fhat = gradients(phi_batch, y_batch, w, U, I, N_total, signal_var)
    w_new, U_new, bad = sgld_update(w, U, gradw, gradU, epsw, epsU, noise_w, noise_U, langevin, stiefel)
    return w_new, U_new, fhat, bad


def gradients_chunked(phi_fn, X_batch, y_batch, w, U, I, N_total, signal_var, chunk=4096):
    """`gradients` for a minibatch too large to hold phi (n x D x bs): the sums over rows of :393 and :405-408 are
    accumulated over row chunks in FP64; phi_fn(X_chunk) returns the chunk's feature tensor.  Same result as
    gradients(phi_fn(X_batch), ...) up to summation order."""
    bs = len(y_batch)
    gw = np.zeros_like(np.asarray(w, dtype=np.float64))
    gU = np.zeros_like(np.asarray(U, dtype=np.float64))
    fhat = np.empty(bs)
    for i0 in range(0, bs, chunk):
        sl = slice(i0, min(i0 + chunk, bs))
        cb = sl.stop - sl.start
        # with N_total = chunk length the (N/bs) factor is 1: raw sums of this chunk (the prior term is taken out again)
        g1, g2, fh = gradients(phi_fn(X_batch[sl]), y_batch[sl], w, U, I, cb, signal_var)
        gw += g1 + w
        gU += g2
        fhat[sl] = fh
    scale = N_total / bs
    return scale * gw - w, scale * gU, fhat


def GPTregression(phi, y, signal_var, I, r, Q, m, epsw, epsU, burnin, maxepoch,
                  w0, U0, perms, noise_w=None, noise_U=None, langevin=True, stiefel=True):
    """GPT_SGLD.jl:345-448 with the random draws made explicit.

    w0[Q], U0[n,r,D] replace the srand/randn initialisation (:357-369);
    perms[epoch] (0-based permutation of N) replaces randperm (:373) and is applied cumulatively (:374);
    noise_w[step][Q], noise_U[step][n,r,D] replace the per-step randn calls, step counted from 0
    over all burnin+maxepoch epochs.  Returns (w_store[Q,T], U_store[n,r,D,T]) with
    T = maxepoch*numbatches, column (epoch-burnin-1)*numbatches+batch (:441-444); all zeros if the
    geodesic produced NaN (:422-424)."""
    phi = np.array(phi, dtype=np.float64, order="F")
    y = np.array(y, dtype=np.float64)
    n, D, N = phi.shape
    numbatches = int(np.ceil(N / m))
    T = maxepoch * numbatches
    w_store = np.zeros((Q, T))
    U_store = np.zeros((n, r, D, T))
    w = np.array(w0, dtype=np.float64)
    U = np.array(U0, dtype=np.float64)
    step = 0
    for epoch in range(1, burnin + maxepoch + 1):
        perm = perms[epoch - 1]
        phi = phi[:, :, perm]
        y = y[perm]
        for batch in range(1, numbatches + 1):
            lo, hi = m * (batch - 1), min(m * batch, N)
            nw = None if noise_w is None else noise_w[step]
            nU = None if noise_U is None else noise_U[step]
            w, U, _, bad = sgld_step(phi[:, :, lo:hi], y[lo:hi], w, U, I, N, signal_var,
                                     epsw, epsU, nw, nU, langevin, stiefel)
            if bad:
                return np.zeros((Q, T)), np.zeros((n, r, D, T))
            if epoch > burnin:
                col = (epoch - burnin - 1) * numbatches + batch - 1
                w_store[:, col] = w
                U_store[:, :, :, col] = U
            step += 1
    return w_store, U_store


def posterior_mean_pred(w_store, U_store, I, phitest, cols):
    """kin40kExperiment.jl:75-87: mean over the listed store columns of pred(w_s,U_s,I,phitest)."""
    acc = np.zeros(phitest.shape[2])
    for c in cols:
        acc += pred(w_store[:, c], U_store[:, :, :, c], I, phitest)
    return acc / len(cols)


# ------------------------------------------------------------------------------------------------
# Full-theta ("no tensor") RFF model — the comparator of the tensor model.
# Shapes: Z[n,D], b[n], phi[n,N], theta[n].  Pinning: **parity unpinned** beyond the formulas — the reference
# holds no fixture or known-answer test for these two functions (only plots); the restatement is checked against
# finite differences of its own log posterior (tests/test_oracle_golden.py).
# ------------------------------------------------------------------------------------------------
def featureNotensor(X, length_scale, sigma_RBF, Z, b):
    """GPT_SGLD.jl:109-120: phi[j,i] = sqrt(2/n)*sigma_RBF*cos(sum_d X[i,d]*Z[j,d]/length_scale[d] + b[j])."""
    X = np.asarray(X, dtype=np.float64)
    Z = np.asarray(Z, dtype=np.float64)
    n = Z.shape[0]
    Zt = Z / np.asarray(length_scale, dtype=np.float64)          # scale(Z, 1./length_scale), :113
    return np.sqrt(2.0 / n) * sigma_RBF * np.cos(Zt @ X.T + np.asarray(b, dtype=np.float64)[:, None])


def nt_grad(phi_batch, y_batch, theta, N_total, signal_var, sigma_theta):
    """GPT_SGLD.jl:833: -theta/sigma_theta^2 + (N/batch_size)*phi_batch*(y_batch - phi_batch'*theta)/signal_var."""
    bs = y_batch.size
    return -theta / sigma_theta ** 2 + (N_total / bs) * (phi_batch @ (y_batch - phi_batch.T @ theta)) / signal_var


def nt_logpost(phi, y, theta, signal_var, sigma_theta):
    """Log posterior whose gradient nt_grad is at full batch (up to constants)."""
    res = y - phi.T @ theta
    return -0.5 * res @ res / signal_var - 0.5 * theta @ theta / sigma_theta ** 2


def GPNT_SGLD(phi, y, signal_var, sigma_theta, m, eps_theta, decay_rate, burnin, maxepoch, theta0, perms, noise):
    """GPT_SGLD.jl:809-847 with the random draws injected: theta0[n] (= sigma_theta*randn(n), :816), one permutation
    per epoch applied cumulatively (:824-825), noise[T, n] standard normals (:835).  Every step is stored, burn-in
    included (:818, :838); returns theta_store[n, (maxepoch+burnin)*numbatches], or zeros(n) after a NaN (:839-842)."""
    phi = np.array(phi, dtype=np.float64)
    y = np.array(y, dtype=np.float64).ravel()
    n, N = phi.shape
    numbatches = int(np.ceil(N / m))
    theta = np.array(theta0, dtype=np.float64).copy()
    store = np.zeros((n, (maxepoch + burnin) * numbatches))
    t = 0
    for epoch in range(maxepoch + burnin):
        perm = np.asarray(perms[epoch])
        phi = phi[:, perm]
        y = y[perm]
        for batch in range(numbatches):
            t += 1
            lo, hi = m * batch, min(m * (batch + 1), N)
            epsilon = eps_theta * float(t) ** (-decay_rate)                       # :832
            g = nt_grad(phi[:, lo:hi], y[lo:hi], theta, N, signal_var, sigma_theta)
            theta = theta + epsilon * g / 2 + np.sqrt(epsilon) * np.asarray(noise[t - 1])
            store[:, t - 1] = theta
            if np.isnan(theta).any():
                return np.zeros(n)
    return store


def GPT_SGLDERMw(phi, y, signal_var, I, r, Q, m, epsw, burnin, maxepoch, w0, U0, perms, noise_w):
    """GPT_SGLD.jl:1065-1118 with the draws injected: SGLD on w only, U fixed at U0 (= the Stiefel draw of :1080-1084).
    Returns (w_store[Q, maxepoch*numbatches], U)."""
    phi = np.array(phi, dtype=np.float64, order="F")
    y = np.array(y, dtype=np.float64)
    n, D, N = phi.shape
    numbatches = int(np.ceil(N / m))
    w = np.array(w0, dtype=np.float64)
    U = np.array(U0, dtype=np.float64)
    w_store = np.zeros((Q, maxepoch * numbatches))
    step = 0
    for epoch in range(1, burnin + maxepoch + 1):
        perm = perms[epoch - 1]
        phi = phi[:, :, perm]
        y = y[perm]
        for batch in range(1, numbatches + 1):
            lo, hi = m * (batch - 1), min(m * batch, N)
            V = computeV(phidotU(U, phi[:, :, lo:hi]), I)                                   # :1097-1100
            fhat = computefhat(V, w)                                                        # :1103
            gradw = (N / (hi - lo)) * (V @ (y[lo:hi] - fhat)) / signal_var - w              # :1106 (sigma_w = 1)
            w = w + epsw * gradw / 2 + np.sqrt(epsw) * np.asarray(noise_w[step])            # :1109
            if epoch > burnin:
                w_store[:, (epoch - burnin - 1) * numbatches + batch - 1] = w               # :1111-1113
            step += 1
    return w_store, U


def GPT_SGLDERM_RMSprop(phi, y, signal_var, I, r, Q, m, epsilon, alpha, burnin, maxepoch, w0, U0, perms, noise_w, noise_U,
                        lam=1e-5):
    """GPT_SGLD.jl:1121-1237 with the draws injected (w0/U0 for :1135-1141, perms for :1151, noise_w[step][Q] for :1188,
    noise_U[step][n,r,D] for :1221).  RMSprop-preconditioned SGLD-ERM:
      * gw, gU are moving averages of the squared UNNORMALISED gradients (1/batch_size) ... (:1172-1176, :1204-1211);
      * w takes a per-element step size epsw = epsilon/(sqrt(gw)+lambda) (:1177, :1188);
      * every U^(k) takes the mean of its elements' step sizes (:1212-1213), since a geodesic needs a scalar time;
      * computeA is called AFTER the w update and therefore sees the new w (:1188 precedes :1194), while the residual
        y_batch - fhat still comes from the old w (:1169).
    Returns (w_store, U_store) like GPTregression; all zeros after a NaN in the geodesic (:1223-1225)."""
    phi = np.array(phi, dtype=np.float64, order="F")
    y = np.array(y, dtype=np.float64)
    n, D, N = phi.shape
    numbatches = int(np.ceil(N / m))
    T = maxepoch * numbatches
    w_store = np.zeros((Q, T)); U_store = np.zeros((n, r, D, T))
    w = np.array(w0, dtype=np.float64); U = np.array(U0, dtype=np.float64)
    gw = np.zeros(Q); gU = np.zeros((n, r, D))
    step = 0
    for epoch in range(1, burnin + maxepoch + 1):
        perm = perms[epoch - 1]
        phi = phi[:, :, perm]; y = y[perm]
        for batch in range(1, numbatches + 1):
            lo, hi = m * (batch - 1), min(m * batch, N)
            pb, yb = phi[:, :, lo:hi], y[lo:hi]
            bs = hi - lo
            temp = phidotU(U, pb)
            V = computeV(temp, I)
            fhat = computefhat(V, w)
            gradw = (1.0 / bs) * (V @ (yb - fhat)) / signal_var                      # :1172
            gw = alpha * gw + (1 - alpha) * gradw ** 2                                # :1175
            epsw = epsilon / (np.sqrt(gw) + lam)                                      # :1176
            gradw = N * gradw - w                                                     # :1186 (sigma_w = 1)
            w = w + epsw * gradw / 2 + np.sqrt(epsw) * np.asarray(noise_w[step])      # :1188
            A = computeA(computeU_phi(V, temp, I), w, I, r)                           # :1191-1194: the NEW w
            gradU = (1.0 / bs) * np.einsum("jki,lki,i->jlk", pb, A, yb - fhat, optimize=True) / signal_var   # :1200-1203
            gU = alpha * gU + (1 - alpha) * gradU ** 2                                # :1210
            epsU = epsilon / (np.sqrt(gU) + lam)                                      # :1211
            meanepsU = epsU.mean(axis=(0, 1))                                         # :1212
            gradU = N * gradU                                                         # :1221
            for k in range(D):
                t = np.sqrt(meanepsU[k])
                mom = proj(U[:, :, k], t * gradU[:, :, k] / 2 + np.asarray(noise_U[step])[:, :, k])   # :1225
                Uk, bad = geod(U[:, :, k], mom, t)                                    # :1226
                if bad:
                    return np.zeros((Q, T)), np.zeros((n, r, D, T))                   # :1227-1229
                U[:, :, k] = Uk
            if epoch > burnin:
                col = (epoch - burnin - 1) * numbatches + batch - 1                   # :1231-1234
                w_store[:, col] = w
                U_store[:, :, :, col] = U
            step += 1
    return w_store, U_store


def geodboth(U, mom, t):
    """GPT_SGLD.jl:40-59: the geodesic and the parallel-transported momentum.  Returns (Unew, momnew, nan_flag);
    Unew has its columns renormalised (:51-56), momnew = [U mom] E[:, r+1:2r] expm(-tA) is returned as it is (:50)."""
    n, r = U.shape
    A = U.T @ mom
    temp = np.block([[A, -mom.T @ mom], [np.eye(r), A]])
    with np.errstate(all="ignore"):
        E = _expm(t * temp)
    if np.isnan(E).any():
        return np.zeros((n, r)), np.zeros((n, r)), True
    mexp = _expm(-t * A)
    UM = np.hstack([U, mom])
    tmpU = UM @ E[:, :r] @ mexp
    tmpV = UM @ E[:, r:] @ mexp
    return tmpU / np.linalg.norm(tmpU, axis=0, keepdims=True), tmpV, False


def GPT_GMC(phi, y, signal_var, I, r, Q, epsw, epsU, burnin, maxepoch, L, w0, U0, noise_p, noise_mom, unif,
            restore_U_on_reject=False):
    """GPT_SGLD.jl:684-805, geodesic Monte Carlo (full-batch HMC on w and on the Stiefel factors), draws injected:
    w0/U0 for :697-704, noise_p[epoch][Q] for :709, noise_mom[epoch][n,r,D] for :711, unif[epoch] for :787.
    Returns (w_store[Q,maxepoch], U_store[n,r,D,maxepoch], accept_prob[burnin+maxepoch]).

    Reference quirk kept by default: `w_old=w; U_old=U` (:707) are aliases in Julia; `w+=...` (:742) rebinds w, but
    `U[:,:,k],mom[:,:,k]=geodboth(...)` (:745) writes into the array U_old also names — so a rejected proposal (:793-795)
    restores w and leaves U at the proposed value.  restore_U_on_reject=True gives the textbook behaviour instead."""
    phi = np.asarray(phi, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    n, D, N = phi.shape
    w = np.array(w0, dtype=np.float64); U = np.array(U0, dtype=np.float64)
    w_store = np.zeros((Q, maxepoch)); U_store = np.zeros((n, r, D, maxepoch))
    accept = np.zeros(burnin + maxepoch)

    def grads(w_, U_):
        temp = phidotU(U_, phi); V = computeV(temp, I); fhat = computefhat(V, w_)
        gradw = V @ (y - fhat) / signal_var - w_                                              # :723 / :762
        A = computeA(computeU_phi(V, temp, I), w_, I, r)
        gradU = np.einsum("jki,lki,i->jlk", phi, A, (y - fhat) / signal_var, optimize=True)   # :733-736 / :772-775
        return gradw, gradU, fhat

    for epoch in range(1, burnin + maxepoch + 1):
        w_old = w.copy(); U_old = U.copy()
        p = np.array(noise_p[epoch - 1], dtype=np.float64)                                    # :709
        mom = np.stack([proj(U[:, :, k], np.asarray(noise_mom[epoch - 1])[:, :, k]) for k in range(D)], 2)   # :710-712
        H_old = (-np.sum(w * w) / 2 - np.linalg.norm(y - pred(w, U, I, phi)) ** 2 / (2 * signal_var)
                 - np.sum(mom * mom) / 2 - np.sum(p * p) / 2)                                 # :713
        pred_new = None
        for l in range(L):
            gradw, gradU, _ = grads(w, U)
            p = p + np.sqrt(epsw) * gradw / 2                                                 # :725
            for k in range(D):
                mom[:, :, k] = proj(U[:, :, k], mom[:, :, k] + np.sqrt(epsU) * gradU[:, :, k] / 2)   # :738-740
            w = w + np.sqrt(epsw) * p                                                         # :742
            for k in range(D):
                Uk, mk, bad = geodboth(U[:, :, k], mom[:, :, k], np.sqrt(epsU))               # :745
                if bad:
                    return np.zeros((Q, maxepoch)), np.zeros((n, r, D, maxepoch)), np.full(burnin + maxepoch, np.nan)   # :746-748
                U[:, :, k] = Uk; mom[:, :, k] = mk
            gradw, gradU, fhat = grads(w, U)
            p = p + np.sqrt(epsw) * gradw / 2                                                 # :764
            for k in range(D):
                mom[:, :, k] = proj(U[:, :, k], mom[:, :, k] + np.sqrt(epsU) * gradU[:, :, k] / 2)   # :777-779
            if l == L - 1:
                pred_new = fhat                                                               # :782-784
        H = (-np.sum(w * w) / 2 - np.linalg.norm(y - pred_new) ** 2 / (2 * signal_var)
             - np.sum(mom * mom) / 2 - np.sum(p * p) / 2)                                     # :787
        with np.errstate(over="ignore"):
            accept[epoch - 1] = np.exp(H - H_old)                                             # :790
        if unif[epoch - 1] > accept[epoch - 1]:                                               # :793: reject
            w = w_old
            if restore_U_on_reject:
                U = U_old
        if epoch > burnin:
            w_store[:, epoch - burnin - 1] = w
            U_store[:, :, :, epoch - burnin - 1] = U
    return w_store, U_store, accept


# ------------------------------------------------------------------------------------------------
# Classification: C independent (w, U) sets under a softmax likelihood.
# Pinning: **parity unpinned** — the reference holds no caller, fixture or known-answer test for
# GPTclassification; the restatement is checked against finite differences of its own log posterior (the check the
# reference keeps commented out at GPT_SGLD.jl:566-622).
# ------------------------------------------------------------------------------------------------
def logsumexp(x):
    """BloodTransfusionExperiment.jl:53-56 / GPT_SGLD.jl export: log(sum(exp(x))) with the maximum taken out."""
    x = np.asarray(x, dtype=np.float64)
    a = np.max(x)
    return a + np.log(np.sum(np.exp(x - a)))


def class_gradients(phi_batch, y_batch, w, U, I, N_total, sigma_w=1.0):
    """GPT_SGLD.jl:487-563 for one minibatch.  w[Q,C], U[n,r,D,C], y_batch 0-based class labels.
    Returns (gradw[Q,C], gradU[n,r,D,C], fhat[bs,C]).

    With p[i,c] = exp(fhat[i,c] - logsumexp(fhat[i,:])) (:506-516) the class-c gradients are the regression ones
    with the residual e[i,c] = (N/bs) (1[y_i = c] - p[i,c]):
        gradw[:,c] = sum_i e[i,c] V[:,i,c] - w[:,c]/sigma_w^2          (:519-524)
        gradU[:,:,k,c] = sum_i e[i,c] Psi[:,i,k,c]                      (:546-564)"""
    n, D, bs = phi_batch.shape
    Q, C = w.shape
    r = U.shape[1]
    temp = [phidotU(U[:, :, :, c], phi_batch) for c in range(C)]                          # :490-493
    V = [computeV(temp[c], I) for c in range(C)]                                           # :496-499
    fhat = np.stack([computefhat(V[c], w[:, c]) for c in range(C)], axis=1)                # :502-505
    lse = np.array([logsumexp(fhat[i, :]) for i in range(bs)])                             # :508-511
    p = np.exp(fhat - lse[:, None])
    onehot = np.zeros((bs, C))
    onehot[np.arange(bs), np.asarray(y_batch, dtype=np.int64)] = 1.0
    e = (N_total / bs) * (onehot - p)
    gradw = np.empty((Q, C))
    gradU = np.empty((n, r, D, C))
    for c in range(C):
        gradw[:, c] = V[c] @ e[:, c] - w[:, c] / sigma_w ** 2
        A = computeA(computeU_phi(V[c], temp[c], I), w[:, c], I, r)                        # :527-536
        gradU[:, :, :, c] = np.einsum("jki,lki,i->jlk", phi_batch, A, e[:, c], optimize=True)
    return gradw, gradU, fhat


def class_loglik(phi, y, w, U, I):
    """sum_i fhat[i,y_i] - logsumexp(fhat[i,:])  (the quantity the commented-out check differentiates, :611-616)."""
    C = w.shape[1]
    fhat = np.stack([pred(w[:, c], U[:, :, :, c], I, phi) for c in range(C)], axis=1)
    return float(sum(fhat[i, int(y[i])] - logsumexp(fhat[i, :]) for i in range(fhat.shape[0])))


def GPTclassification(phi, y, I, r, Q, m, epsw, epsU, burnin, maxepoch, w0, U0, perms, noise_w, noise_U,
                      langevin=True, stiefel=True, double_update=True):
    """GPT_SGLD.jl:452-680 with the random draws made explicit.  y: 0-based labels in 0..C-1 (the reference's 1..C).

    w0[Q,C], U0[n,r,D,C] replace :463-476; perms[epoch] replaces randperm (:480), applied cumulatively (:481);
    noise_w[step][2][Q,C], noise_U[step][2][n,r,D,C]: the draws of the reference's TWO updates per minibatch — it moves
    w and U once unconditionally with noise on the Stiefel manifold (:624-636) and then again under the langevin /
    stiefel switches (:639-671), both times with the gradient evaluated before the first move.
    double_update=False skips the first (what the regression samplers do).  Stores: w_store[Q,C,T], U_store[n,r,D,C,T]."""
    phi = np.array(phi, dtype=np.float64, order="F")
    y = np.array(y, dtype=np.int64)
    n, D, N = phi.shape
    C = w0.shape[1]
    numbatches = int(np.ceil(N / m))
    T = maxepoch * numbatches
    w_store = np.zeros((Q, C, T))
    U_store = np.zeros((n, r, D, C, T))
    w = np.array(w0, dtype=np.float64)
    U = np.array(U0, dtype=np.float64)
    step = 0

    def move(w, U, gradw, gradU, nw, nU, lang, stief):
        w2 = np.empty_like(w)
        U2 = np.empty_like(U)
        for c in range(C):
            wc, Uc, bad = sgld_update(w[:, c], U[:, :, :, c], gradw[:, c], gradU[:, :, :, c], epsw, epsU,
                                      None if nw is None else nw[:, c], None if nU is None else nU[:, :, :, c], lang, stief)
            if bad:
                return None, None, True
            w2[:, c] = wc
            U2[:, :, :, c] = Uc
        return w2, U2, False

    for epoch in range(1, burnin + maxepoch + 1):
        perm = perms[epoch - 1]
        phi = phi[:, :, perm]
        y = y[perm]
        for batch in range(1, numbatches + 1):
            lo, hi = m * (batch - 1), min(m * batch, N)
            gradw, gradU, _ = class_gradients(phi[:, :, lo:hi], y[lo:hi], w, U, I, N)
            # (:524: gradw -= w/sigma_w^2 uses the w before either move; class_gradients holds it)
            if double_update:
                w, U, bad = move(w, U, gradw, gradU, noise_w[step][0], noise_U[step][0], True, True)     # :624-636
                if bad:
                    return np.zeros((Q, C, T)), np.zeros((n, r, D, C, T))
            w, U, bad = move(w, U, gradw, gradU, None if noise_w is None else noise_w[step][1],
                             None if noise_U is None else noise_U[step][1], langevin, stiefel)           # :639-671
            if bad:
                return np.zeros((Q, C, T)), np.zeros((n, r, D, C, T))
            if epoch > burnin:
                col = (epoch - burnin - 1) * numbatches + batch - 1
                w_store[:, :, col] = w
                U_store[:, :, :, :, col] = U
            step += 1
    return w_store, U_store


# ------------------------------------------------------------------------------------------------
# Gibbs sampler of TGP.jl (last row).  Pinning: **parity unpinned** — the reference holds no fixture
# or known-answer test for GPT_inf; the restatement follows TGP.jl:37-86 line by line, and its conditional means are
# checked against the normal equations they solve (tests/test_oracle_golden.py).
# ------------------------------------------------------------------------------------------------
def tgp_gibbs(b, y, sigma, r, I, U0, num_iterations, burnin, noise_w, noise_U):
    """TGP.jl:37-86 (GPT_inf) with the random draws made explicit.

    b[n,D,N]: features (TGP.feature, :6-14) of the whitened inputs; y[N] whitened targets; I[q,D] 0-based (:57);
    U0[n,r,D] replaces sigma_u*randn(n,r) (:56); noise_w[m][q] and noise_U[m][D][n*r] replace the randn of :66 and :83.
    Returns (W_array[q, T], V_array[n,r,D,T]), T = num_iterations-burnin: W of iteration m and the factors it was drawn
    with (:68-71 store before the factors move).
    Quirks kept: the factor draw is invSigma_U \\ randn (covariance invSigma_U^-2, :83), where the w draw uses the
    Cholesky factor (:66); rows of C for values of l that no I[:,k] takes are uninitialised in the reference (:76-79) and
    zero here."""
    b = np.asarray(b, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    n, D, N = b.shape
    q = I.shape[0]
    sigma_u = np.sqrt(1.0 / r)                                             # :54
    sigma_w = np.sqrt(float(r) ** D / q)                                   # :55
    U = np.array(U0, dtype=np.float64)
    T = num_iterations - burnin
    W_array = np.zeros((q, T))
    V_array = np.zeros((n, r, D, T))
    for m in range(1, num_iterations + 1):
        t = np.einsum("jlk,jki->kli", U, b)                                # t[k,l,i] = U_k[:,l]' b[:,k,i]
        V = np.ones((q, N))
        for k in range(D):
            V *= t[k, I[:, k], :]                                          # :62
        invSigma_w = V @ V.T / sigma ** 2 + np.eye(q) / sigma_w ** 2       # :64
        Mu_w = np.linalg.solve(invSigma_w, V @ y / sigma ** 2)             # :65
        R = np.linalg.cholesky(invSigma_w).T                               # chol(.,:U): invSigma_w = R'R
        W = np.linalg.solve(R, noise_w[m - 1]) + Mu_w                      # :66
        if m > burnin:
            W_array[:, m - burnin - 1] = W                                 # :69
            V_array[:, :, :, m - burnin - 1] = U                           # :70
        for k in range(D):
            tk = np.einsum("jl,ji->li", U[:, :, k], b[:, k, :])
            V_k = V / tk[I[:, k], :]                                       # :74
            V_kk = W[:, None] * V_k                                        # :75
            C = np.zeros((r, N))
            for l in np.unique(I[:, k]):
                C[l, :] = V_kk[I[:, k] == l, :].sum(axis=0)                # :77-79
            Ck = np.repeat(C, n, axis=0) * np.tile(b[:, k, :], (r, 1))     # :80: row l*n + j = C[l,:] .* b[j,k,:]
            invSigma_U = Ck @ Ck.T / sigma ** 2 + np.eye(n * r) / sigma_u ** 2          # :81
            Mu_U = np.linalg.solve(invSigma_U, Ck @ y / sigma ** 2)        # :82
            U[:, :, k] = (np.linalg.solve(invSigma_U, noise_U[m - 1][k]) + Mu_U).reshape(n, r, order="F")   # :83
            tk = np.einsum("jl,ji->li", U[:, :, k], b[:, k, :])
            V = V_k * tk[I[:, k], :]                                       # :84
    return W_array, V_array
