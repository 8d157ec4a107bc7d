"""Pins the CPU oracle (oracle/gpt_oracle.py + oracle/gpt_oracle.c) against the reference's own fixtures
 — runs without a GPU."""
import numpy as np
import pytest
from scipy.linalg import expm

import c_oracle as co
import gpt_oracle as o
from util import load_golden, make_problem, relerr


@pytest.mark.parametrize("name,D,Q", [("tensorsynth5d.npz", 5, 32), ("tensorsynth10d.npz", 10, 100)])
def test_pred_known_answer(name, D, Q):
    """y_j = pred(w,U,I,phi) + sqrt(v_j) randn (MakeSynthData.jl:10-13): residuals must be that noise."""
    d = load_golden(name)
    for impl in (o.pred, co.pred):
        p = impl(d["w"], d["U"], d["I"], d["phi"])
        for key, v in (("y1", 0.1), ("y2", 0.01), ("y3", 0.001)):
            res = d[key] - p
            assert abs(np.std(res) / np.sqrt(v) - 1) < 0.25, (key, np.std(res))
            assert abs(np.mean(res)) < 4 * np.sqrt(v / res.size)
        # the three noisy copies share one noiseless y: differences are pure noise of known variance
        assert abs(np.std(d["y1"] - d["y2"]) / np.sqrt(0.11) - 1) < 0.25


@pytest.mark.parametrize("name", ["tensorsynth5d.npz", "tensorsynth10d.npz"])
def test_feature_known_answer(name):
    """feature(X, length_scale, 1, phi_scale, Z_rec, b_rec) reproduces the phi the reference stored
    (Z,b recovered by tests/golden/make_golden.py; GPT_SGLD.jl:71-84)."""
    d = load_golden(name)
    for impl in (o.feature, co.feature):
        phi = impl(d["X"], d["length_scale"], 1.0, float(d["phi_scale"]), d["Z_rec"], d["b_rec"])
        assert np.abs(phi - d["phi"]).max() < 1e-12
    n = d["phi"].shape[0]
    assert np.abs(d["phi"]).max() <= float(d["phi_scale"]) * np.sqrt(2.0 / n) * (1 + 1e-12)


@pytest.mark.parametrize("name,r", [("tensorsynth5d.npz", 2), ("tensorsynth10d.npz", 2)])
def test_fixture_invariants(name, r):
    d = load_golden(name)
    Q, D = d["I"].shape
    assert d["I"].min() >= 1 and d["I"].max() <= r                       # samplenz range (:185)
    assert len({tuple(row) for row in d["I"]}) == Q                      # sampled without replacement (:182)
    for k in range(D):
        assert np.abs(d["U"][:, :, k].T @ d["U"][:, :, k] - np.eye(r)).max() < 1e-13   # Stiefel init (:366)


def test_samplenz_digits():
    rng = np.random.default_rng(3)
    I = o.samplenz(3, 4, 50, rng)
    assert I.dtype == np.int32 and I.shape == (50, 4)
    L = ((I - 1) * 3 ** np.arange(4)).sum(1)                            # little-endian base-r digits
    assert len(set(L.tolist())) == 50 and L.max() < 3 ** 4
    I = o.samplenz(32, 10, 200, rng)                                     # r^D = 2^50 lattice
    assert I.min() >= 1 and I.max() <= 32 and len({tuple(r_) for r_ in I}) == 200


def test_c_oracle_matches_numpy_oracle():
    P = make_problem(4, 37, 12, 3, Q=20, seed=1)
    phi = P["phi_fn"](P["X"])
    assert np.abs(co.feature(P["X"], P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"], 2) - phi).max() < 1e-14
    assert np.abs(co.feature2(P["X"], P["ls"], 1.3, P["phi_scale"], P["Z"]) -
                  o.feature2(P["X"], P["ls"], 1.3, P["phi_scale"], P["Z"])).max() < 1e-14
    assert relerr(co.pred(P["w0"], P["U0"], P["I"], phi, 3), o.pred(P["w0"], P["U0"], P["I"], phi)) < 1e-13
    g_np = o.gradients(phi, P["y"], P["w0"], P["U0"], P["I"], 1000, 0.01)
    g_lit = o.gradients_literal(phi, P["y"], P["w0"], P["U0"], P["I"], 1000, 0.01)
    g_c = co.gradients(phi, P["y"], P["w0"], P["U0"], P["I"], 1000, 0.01, 4)
    for a, b, c in zip(g_np, g_lit, g_c):
        assert relerr(a, b) < 1e-12 and relerr(c, b) < 1e-12
    with pytest.raises(ValueError):
        o.feature2(P["X"], P["ls"], 1.0, 1.0, P["Z"][:11])               # "n not even" (:90)
    with pytest.raises(ValueError):
        co.feature2(P["X"], P["ls"], 1.0, 1.0, P["Z"][:11])


def test_gradients_vs_finite_differences():
    """Diagnostic_gradients.jl:107-158 on its own data (first 200 rows of 10000SynthData.h5), n=10,r=5,Q=10,
    sigma=0.2, forward differences h=1e-6."""
    d = load_golden("synth_fd.npz")
    rng = np.random.default_rng(17)
    X, y = d["Xtrain"], d["ytrain"]
    N, D = X.shape
    n, r, Q, sigma = 10, 5, 10, 0.2
    Z = rng.standard_normal((n, D)); b = 2 * np.pi * rng.random((n, D))
    phi = o.feature(X, 1.4, 1.0, 1.0, Z, b)
    I = o.samplenz(r, D, Q, rng)
    U = np.stack([o.stiefel_from_gaussian(rng.standard_normal((r, n))) for _ in range(D)], 2)
    w = np.sqrt(n ** D / Q) * rng.standard_normal(Q)                    # Diagnostic_gradients.jl:116,130
    ll = o.loglik(phi, y, w, U, I, sigma)
    gw, gU, _ = o.gradients(phi, y, w, U, I, N, sigma ** 2)             # N/bs = 1: full-batch likelihood gradient
    gw = gw + w                                                          # remove the prior term of :393
    h = 1e-6
    fdw = np.array([(o.loglik(phi, y, w + h * np.eye(Q)[q], U, I, sigma) - ll) / h for q in range(Q)])
    assert relerr(gw, fdw) < 1e-4
    fdU = np.zeros_like(U)
    for k in range(D):
        for l in range(r):
            for j in range(n):
                U2 = U.copy(); U2[j, l, k] += h
                fdU[j, l, k] = (o.loglik(phi, y, w, U2, I, sigma) - ll) / h
    assert relerr(gU, fdU) < 1e-4
    gwc, gUc, _ = co.gradients(phi, y, w, U, I, N, sigma ** 2)
    assert relerr(gwc + w, fdw) < 1e-4 and relerr(gUc, fdU) < 1e-4


def test_expm_c_vs_scipy():
    rng = np.random.default_rng(5)
    for scale in (1e-3, 0.3, 2.0, 30.0):
        A = scale * rng.standard_normal((12, 12))
        assert relerr(co.expm(A), expm(A)) < 1e-11
    assert np.isnan(co.expm(np.full((4, 4), np.nan))).all()


def test_proj_geod_invariants_and_c():
    rng = np.random.default_rng(7)
    n, r = 40, 6
    U = o.stiefel_from_gaussian(rng.standard_normal((r, n)))
    V = rng.standard_normal((n, r))
    M = o.proj(U, V)
    assert np.abs(U.T @ M + M.T @ U).max() < 1e-12                       # tangent: U'M skew
    assert np.abs(o.proj(U, M) - M).max() < 1e-12                        # idempotent
    assert np.abs(co.proj(U, V) - M).max() < 1e-13
    for t in (1e-4, 1e-2, 0.3):
        Un, bad = o.geod(U, M, t)
        assert not bad and np.abs(Un.T @ Un - np.eye(r)).max() < 1e-10   # stays on the manifold
        Uc, badc = co.geod(U, M, t)
        assert not badc and np.abs(Uc - Un).max() < 1e-11
    # small-t series: U(t) = U + t M + O(t^2)
    Un, _ = o.geod(U, M, 1e-6)
    assert np.abs(Un - U - 1e-6 * M).max() < 1e-9
    Z, bad = o.geod(U, M * 1e200, 1e200)
    assert bad and not Z.any()                                           # NaN -> zeros (:24-26)
    Zc, badc = co.geod(U, M * 1e200, 1e200)
    assert badc and not Zc.any()


def test_sgld_step_c_vs_numpy_and_store_layout():
    P = make_problem(3, 50, 8, 3, Q=12, seed=2)
    phi = P["phi_fn"](P["X"])
    rng = np.random.default_rng(0)
    nw = rng.standard_normal(P["Q"]); nU = rng.standard_normal((P["n"], P["r"], P["D"]))
    for stiefel in (True, False):
        a = o.sgld_step(phi, P["y"], P["w0"], P["U0"], P["I"], 500, 0.01, 1e-5, 1e-8, nw, nU, True, stiefel)
        c = co.sgld_step(phi, P["y"], P["w0"], P["U0"], P["I"], 500, 0.01, 1e-5, 1e-8, nw, nU, stiefel)
        assert relerr(c[0], a[0]) < 1e-13 and relerr(c[1], a[1]) < 1e-13
    # GPTregression: T = maxepoch*ceil(N/m), short last batch, burn-in epochs not stored (:353,:379-381,:441-444)
    N, m, burnin, maxepoch = 50, 16, 1, 2
    nb = 4
    perms = [rng.permutation(N) for _ in range(burnin + maxepoch)]
    steps = (burnin + maxepoch) * nb
    NW = rng.standard_normal((steps, P["Q"])); NU = rng.standard_normal((steps, P["n"], P["r"], P["D"]))
    ws, Us = o.GPTregression(phi, P["y"], 0.01, P["I"], P["r"], P["Q"], m, 1e-5, 1e-8, burnin, maxepoch,
                             P["w0"], P["U0"], perms, NW, NU)
    assert ws.shape == (P["Q"], maxepoch * nb) and Us.shape == (P["n"], P["r"], P["D"], maxepoch * nb)
    # replay by hand
    w, U = P["w0"].copy(), P["U0"].copy()
    ph, yy = phi.copy(), P["y"].copy()
    s = 0
    for ep in range(burnin + maxepoch):
        ph, yy = ph[:, :, perms[ep]], yy[perms[ep]]
        for bidx in range(nb):
            lo, hi = bidx * m, min((bidx + 1) * m, N)
            w, U, _, _ = o.sgld_step(ph[:, :, lo:hi], yy[lo:hi], w, U, P["I"], N, 0.01, 1e-5, 1e-8, NW[s], NU[s])
            if ep >= burnin:
                col = (ep - burnin) * nb + bidx
                assert relerr(ws[:, col], w) < 1e-12 and relerr(Us[:, :, :, col], U) < 1e-12
            s += 1
    # NaN abort returns all-zero stores (:422-424)
    ws, Us = o.GPTregression(phi, P["y"], 0.01, P["I"], P["r"], P["Q"], m, 1e-5, 1e300, 0, 1, P["w0"], P["U0"],
                             perms, NW, NU)
    assert not ws.any() and not Us.any()


def test_fulltheta_oracle_gradient_and_chain():
    """Full-theta comparator (GPT_SGLD.jl:109-120, :809-847): no reference fixture exists for it (parity unpinned), so
    the restatement is checked against finite differences of its own log posterior and for the store layout."""
    rng = np.random.default_rng(5)
    N, D, n = 60, 3, 24
    X = rng.standard_normal((N, D)); Z = rng.standard_normal((n, D)); b = 2 * np.pi * rng.random(n)
    phi = o.featureNotensor(X, np.array([1.0, 0.7, 1.3]), 1.2, Z, b)
    assert phi.shape == (n, N) and np.abs(phi).max() <= np.sqrt(2.0 / n) * 1.2 + 1e-15
    # literal loop of :114-118 on a few entries
    for (j, i) in [(0, 0), (5, 17), (23, 59)]:
        want = np.sqrt(2.0 / n) * 1.2 * np.cos(np.sum(X[i] * Z[j] / np.array([1.0, 0.7, 1.3])) + b[j])
        assert abs(phi[j, i] - want) < 1e-14
    y = rng.standard_normal(N); th = rng.standard_normal(n)
    g = o.nt_grad(phi, y, th, N, 0.1, 1.5)
    eye = np.eye(n)
    fd = np.array([(o.nt_logpost(phi, y, th + 1e-6 * eye[j], 0.1, 1.5) - o.nt_logpost(phi, y, th - 1e-6 * eye[j], 0.1, 1.5)) / 2e-6
                   for j in range(n)])
    assert np.abs(g - fd).max() / np.abs(g).max() < 1e-7
    perms = [rng.permutation(N) for _ in range(3)]
    noise = rng.standard_normal((3 * 4, n))
    st = o.GPNT_SGLD(phi, y, 0.1, 1.5, 16, 1e-4, 0.5, 1, 2, th, perms, noise)
    assert st.shape == (n, 12)
    # first step by hand: eps = eps_theta * 1^(-decay) (:832), batch = first 16 columns of phi[:, perm]
    p0 = perms[0][:16]
    g0 = o.nt_grad(phi[:, p0], y[p0], th, N, 0.1, 1.5)
    assert np.allclose(st[:, 0], th + 1e-4 * g0 / 2 + np.sqrt(1e-4) * noise[0], rtol=0, atol=1e-14)
    assert not o.GPNT_SGLD(phi, y, 0.1, 1.5, 16, 1e300, 0.0, 0, 1, th, perms, noise).any()      # NaN -> zeros(n), :839-842


def test_chunked_gradients_and_update_split_match_the_one_piece_oracle():
    """gradients_chunked (used by the GPU tests at the benchmarked minibatch sizes, where phi does not fit in memory)
    and sgld_update (the update half of sgld_step) restate the same lines as gradients / sgld_step."""
    from util import make_problem
    P = make_problem(3, 300, 12, 3, Q=15, seed=33)
    idx = np.arange(37, 260)
    phi = P["phi_fn"](P["X"][idx])
    a = o.gradients(phi, P["y"][idx], P["w0"], P["U0"], P["I"], P["N"], P["signal_var"])
    b = o.gradients_chunked(P["phi_fn"], P["X"][idx], P["y"][idx], P["w0"], P["U0"], P["I"], P["N"], P["signal_var"], chunk=50)
    for x, y_ in zip(a, b):
        assert np.allclose(x, y_, rtol=1e-11, atol=1e-9 * np.abs(x).max())
    rng = np.random.default_rng(0)
    nw, nU = rng.standard_normal(P["Q"]), rng.standard_normal(P["U0"].shape)
    s1 = o.sgld_step(phi, P["y"][idx], P["w0"], P["U0"], P["I"], P["N"], P["signal_var"], 1e-5, 1e-8, nw, nU)
    s2 = o.sgld_update(P["w0"], P["U0"], a[0], a[1], 1e-5, 1e-8, nw, nU)
    assert np.array_equal(s1[0], s2[0]) and np.array_equal(s1[1], s2[1])


def test_geod_pinned_by_closed_form_and_arbitrary_precision():
    """`geod` (GPT_SGLD.jl:19-37) has no known-answer test in the reference.  Two independent pins: the great circle for
    r = 1, and 40-digit mpmath evaluations of the same lines (tests/golden/geod_mp.npz, made by make_geod_golden.py) —
    for the NumPy oracle (SciPy expm), the C oracle (its own Pade-13) and the 2r x 2r exponential by itself."""
    import c_oracle as co
    rng = np.random.default_rng(4)
    for n in (2, 7, 300):
        u = rng.standard_normal((n, 1)); u /= np.linalg.norm(u)
        m = o.proj(u, rng.standard_normal((n, 1)))
        for t in (1e-7, 1e-3, 0.3, 2.5):
            want = o.geod_great_circle(u, m, t)
            assert np.abs(o.geod(u, m, t)[0] - want).max() < 1e-14
            assert np.abs(co.geod(u, m, t)[0] - want).max() < 1e-14
    g = load_golden("geod_mp.npz")
    for c in range(int(g["ncases"])):
        U, M, t = g[f"U{c}"], g[f"M{c}"], float(g[f"t{c}"])
        r = U.shape[1]
        A = U.T @ M
        blk = t * np.block([[A, -M.T @ M], [np.eye(r), A]])
        from scipy.linalg import expm
        assert relerr(expm(blk), g[f"expm{c}"]) < 1e-14, c
        assert relerr(co.expm(blk), g[f"expm{c}"]) < 1e-13, c
        assert relerr(o.geod(U, M, t)[0], g[f"geod{c}"]) < 1e-14, c
        assert relerr(co.geod(U, M, t)[0], g[f"geod{c}"]) < 1e-13, c


def test_class_gradients_finite_differences():
    """GPTclassification's gradients (GPT_SGLD.jl:487-563) against finite differences of the softmax log likelihood —
    the check the reference keeps commented out at :566-622 — and the softmax residual's sum-to-zero over classes."""
    rng = np.random.default_rng(0)
    n, D, N, r, Q, C = 6, 3, 20, 2, 5, 3
    phi = rng.standard_normal((n, D, N)); y = rng.integers(0, C, N)
    I = rng.integers(0, r, (Q, D))
    w = rng.standard_normal((Q, C))
    U = np.stack([np.stack([np.linalg.qr(rng.standard_normal((n, r)))[0] for _ in range(D)], axis=2) for _ in range(C)], axis=3)
    gw, gU, fhat = o.class_gradients(phi, y, w, U, I, N)
    l0 = o.class_loglik(phi, y, w, U, I)
    h = 1e-6
    for (q, c) in [(2, 1), (0, 0), (4, 2)]:
        wp = w.copy(); wp[q, c] += h
        assert abs((o.class_loglik(phi, y, wp, U, I) - l0) / h - (gw[q, c] + w[q, c])) < 1e-4
    for (j, l, k, c) in [(3, 1, 2, 0), (0, 0, 0, 2), (5, 1, 1, 1)]:
        Up = U.copy(); Up[j, l, k, c] += h
        assert abs((o.class_loglik(phi, y, w, Up, I) - l0) / h - gU[j, l, k, c]) < 1e-4
    assert abs(o.logsumexp(np.array([1000.0, 1000.0])) - (1000.0 + np.log(2.0))) < 1e-12
    # double update with zero noise and zero step sizes is the identity; stores have the reference's shapes
    perms = [np.arange(N)]
    NW = np.zeros((2, 2, Q, C)); NU = np.zeros((2, 2, n, r, D, C))
    ws, Us = o.GPTclassification(phi, y, I, r, Q, 10, 0.0, 0.0, 0, 1, w, U, perms, NW, NU)
    assert ws.shape == (Q, C, 2) and Us.shape == (n, r, D, C, 2)
    assert np.allclose(ws[:, :, -1], w) and np.allclose(Us[..., -1], U, atol=1e-12)


def test_tgp_gibbs_restatement():
    """TGP.jl:37-86 restated: with zero noise the w draw is the ridge solution of V'w ~ y (its normal equations hold),
    stores are taken before the factors move, and a second call continues from stored factors identically."""
    rng = np.random.default_rng(0)
    n, D, N, r, q, M = 5, 3, 40, 2, 6, 3
    X = rng.standard_normal((N, D)); y = rng.standard_normal(N)
    Z = rng.standard_normal((n, D)) / 1.3; b01 = rng.random((n, D))
    b = np.stack([o.tgp_feature(X[i], Z, b01) for i in range(N)], axis=2)
    I = rng.integers(0, r, (q, D)); I[:r, 0] = np.arange(r)
    U0 = np.sqrt(1 / r) * rng.standard_normal((n, r, D))
    sigma = 0.5
    W, V = o.tgp_gibbs(b, y, sigma, r, I, U0, M, 0, np.zeros((M, q)), np.zeros((M, D, n * r)))
    assert W.shape == (q, M) and V.shape == (n, r, D, M)
    assert np.array_equal(V[..., 0], U0)                                   # :68-71 store before the factors move
    t = np.einsum("jlk,jki->kli", U0, b)
    Vm = np.prod([t[k, I[:, k], :] for k in range(D)], axis=0)
    sw2 = float(r) ** D / q
    lhs = (Vm @ Vm.T / sigma ** 2 + np.eye(q) / sw2) @ W[:, 0]
    assert np.allclose(lhs, Vm @ y / sigma ** 2, rtol=1e-10, atol=1e-12)
    # chain property: running 3 iterations == running 1 and then 2 more from the stored state (noise-free)
    W1, V1 = o.tgp_gibbs(b, y, sigma, r, I, U0, 2, 0, np.zeros((2, q)), np.zeros((2, D, n * r)))
    W2, V2 = o.tgp_gibbs(b, y, sigma, r, I, V1[..., 1], 2, 0, np.zeros((2, q)), np.zeros((2, D, n * r)))
    assert np.allclose(W2[:, 1], W[:, 2], rtol=1e-9) and np.allclose(V2[..., 1], V[..., 2], rtol=1e-9)
