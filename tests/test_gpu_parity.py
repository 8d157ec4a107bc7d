"""Parity of the CUDA path (through the C ABI) with the CPU oracle — needs a B200.

Tolerances (BASELINE.json north_star): <=1e-10 relative in GPTB_PREC_FP64, <=1e-4 relative in the
tensor-core parity mode GPTB_PREC_TC3; norm-wise per array."""
import numpy as np
import pytest

import gpt_oracle as o
from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G
from util import load_golden, make_problem, relerr

pytestmark = pytest.mark.gpu

TOL = {L.PREC_FP64: 1e-10, L.PREC_TC3: 1e-4}
PRECS = [L.PREC_FP64, L.PREC_TC3]
IDS = ["fp64", "tc3"]


def _feat(P, X=None):
    X = P["X"] if X is None else X
    if P["feature2"]:
        return G.feature2(X, P["ls"], P["sigma_RBF"], P["phi_scale"], P["Z"])
    return G.feature(X, P["ls"], P["sigma_RBF"], P["phi_scale"], P["Z"], P["b"])


def _model(P, prec, train=True, phi_mode=False):
    kind = L.FEAT_SINCOS if P["feature2"] else L.FEAT_COS
    m = G.Model(P["D"], P["n"], P["r"], P["Q"], precision=prec, feature_kind=kind, seed=123)
    m.set_features(_feat(P))
    m.set_index(P["I"])
    if train:
        if phi_mode:
            m.set_train_phi(P["phi_fn"](P["X"]), P["y"])
        else:
            m.set_train(P["X"], P["y"])
    m.set_params(P["w0"], P["U0"])
    return m


# ------------------------------------------------------------------ features
@pytest.mark.parametrize("name", ["tensorsynth5d.npz", "tensorsynth10d.npz"])
def test_feature_known_answer(name):
    d = load_golden(name)
    phi = G.feature(d["X"], d["length_scale"], 1.0, float(d["phi_scale"]), d["Z_rec"], d["b_rec"]).materialize()
    assert phi.shape == d["phi"].shape
    assert np.abs(phi - d["phi"]).max() < 1e-12


def test_feature_variants_vs_oracle():
    P = make_problem(3, 301, 14, 3, Q=10, seed=5)
    assert np.abs(_feat(P).materialize() - P["phi_fn"](P["X"])).max() < 1e-13
    f2 = G.feature2(P["X"], P["ls"], 1.7, P["phi_scale"], P["Z"]).materialize()
    assert np.abs(f2 - o.feature2(P["X"], P["ls"], 1.7, P["phi_scale"], P["Z"])).max() < 1e-13
    fs = G.feature(P["X"], 0.8, 1.0, 1.0, P["Z"], P["b"]).materialize()   # scalar length scale
    assert np.abs(fs - o.feature(P["X"], 0.8, 1.0, 1.0, P["Z"], P["b"])).max() < 1e-13
    # TGP.feature (TGP.jl:6-14): sqrt(2/n)*cos(x.*Z/sigma + 2 pi b)
    b01 = P["b"] / (2 * np.pi)
    t = G.tgp_feature(P["X"][:1], P["n"], 1.4, P["Z"], b01).materialize()[:, :, 0]
    assert np.abs(t - o.tgp_feature(P["X"][0], P["Z"] / 1.4, b01)).max() < 1e-13


# ------------------------------------------------------------------ prediction
@pytest.mark.parametrize("prec", PRECS, ids=IDS)
@pytest.mark.parametrize("name", ["tensorsynth5d.npz", "tensorsynth10d.npz"])
def test_pred_known_answer(name, prec):
    """pred on the reference's fixture, lazy features (recovered Z,b) and dense phi; residuals are the noise."""
    d = load_golden(name)
    want = o.pred(d["w"], d["U"], d["I"], d["phi"])
    f = G.feature(d["X"], d["length_scale"], 1.0, float(d["phi_scale"]), d["Z_rec"], d["b_rec"])
    got = G.pred(d["w"], d["U"], d["I"], f, precision=prec)
    assert relerr(got, want) < TOL[prec]
    dense = G.pred(d["w"], d["U"], d["I"], d["phi"], precision=prec)
    assert relerr(dense, want) < 1e-12                                   # dense-phi path is FP64
    assert abs(np.std(d["y3"] - got) / np.sqrt(1e-3) - 1) < 0.25


SHAPES = {  # reduced-N versions of BASELINE configs C1..C4 + odd sizes
    "c1": dict(D=4, N=700, n=50, r=10),
    "c2": dict(D=8, N=1100, n=100, r=20),
    "c3": dict(D=10, N=300, n=500, r=20),
    "c4": dict(D=10, N=260, n=1000, r=32),
    "ragged": dict(D=3, N=131, n=37, r=5, Q=40),
    "one_row": dict(D=2, N=1, n=8, r=2, Q=4),
}


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
@pytest.mark.parametrize("shape", list(SHAPES))
def test_pred_vs_oracle(shape, prec):
    P = make_problem(seed=1, **SHAPES[shape])
    want = o.pred(P["w0"], P["U0"], P["I"], P["phi_fn"](P["X"]))
    got = G.pred(P["w0"], P["U0"], P["I"], _feat(P), precision=prec)
    assert relerr(got, want) < TOL[prec]


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
def test_pred_feature2_and_samples(prec):
    P = make_problem(4, 500, 24, 4, Q=30, seed=3, feature2=True)
    want = o.pred(P["w0"], P["U0"], P["I"], P["phi_fn"](P["X"]))
    assert relerr(G.pred(P["w0"], P["U0"], P["I"], _feat(P), precision=prec), want) < TOL[prec]
    rng = np.random.default_rng(0)
    S = 5
    ws = np.asfortranarray(rng.standard_normal((P["Q"], S)))
    Us = np.asfortranarray(np.stack([np.stack([o.stiefel_from_gaussian(rng.standard_normal((P["r"], P["n"])))
                                               for _ in range(P["D"])], 2) for _ in range(S)], 3))
    phi = P["phi_fn"](P["X"])
    per_want = np.stack([o.pred(ws[:, s], Us[:, :, :, s], P["I"], phi) for s in range(S)], 1)
    mean, per = G.pred_samples(ws, Us, P["I"], _feat(P), per_sample=True, precision=prec)
    assert relerr(per, per_want) < TOL[prec] and relerr(mean, per_want.mean(1)) < TOL[prec]
    sub = G.pred_samples(ws, Us, P["I"], _feat(P), cols=[1, 3], precision=prec)
    assert relerr(sub, per_want[:, [1, 3]].mean(1)) < TOL[prec]


# ------------------------------------------------------------------ gradients
@pytest.mark.parametrize("prec", PRECS, ids=IDS)
@pytest.mark.parametrize("shape", list(SHAPES))
def test_grad_vs_oracle(shape, prec):
    P = make_problem(seed=2, **SHAPES[shape])
    N = P["N"]
    with _model(P, prec) as m:
        # explicit row list (ragged minibatch) and contiguous range
        rng = np.random.default_rng(0)
        idx = rng.permutation(N)[: max(1, (2 * N) // 3)]
        phi = P["phi_fn"](P["X"][idx])
        want = o.gradients(phi, P["y"][idx], P["w0"], P["U0"], P["I"], N, P["signal_var"])
        gw, gU, fh = m.grad(P["signal_var"], idx=idx)
        assert relerr(fh, want[2]) < TOL[prec]
        assert relerr(gw, want[0]) < TOL[prec]
        assert relerr(gU, want[1]) < TOL[prec]
        gw2, _, _ = m.grad(P["signal_var"], idx=idx, include_prior=False)
        assert relerr(gw2, want[0] + P["w0"]) < TOL[prec]
        lo, bs = 0, N
        want = o.gradients(P["phi_fn"](P["X"][lo:lo + bs]), P["y"][lo:lo + bs], P["w0"], P["U0"], P["I"], N,
                           P["signal_var"])
        gw, gU, fh = m.grad(P["signal_var"], start=lo, bs=bs)
        assert relerr(gw, want[0]) < TOL[prec] and relerr(gU, want[1]) < TOL[prec]


def test_grad_dense_phi_and_errors():
    P = make_problem(3, 90, 10, 3, Q=12, seed=6)
    with _model(P, L.PREC_FP64, phi_mode=True) as m:
        want = o.gradients(P["phi_fn"](P["X"][10:70]), P["y"][10:70], P["w0"], P["U0"], P["I"], P["N"], 1e-3)
        gw, gU, fh = m.grad(1e-3, start=10, bs=60)
        assert relerr(gw, want[0]) < 1e-10 and relerr(gU, want[1]) < 1e-10 and relerr(fh, want[2]) < 1e-10
        with pytest.raises(G.GptbError):
            m.grad(1e-3, start=80, bs=20)                                 # range outside N
        with pytest.raises(G.GptbError):
            m.grad(1e-3, idx=np.array([0, 95]))                           # index outside N
    with G.Model(3, 10, 3, 12, precision=L.PREC_FP64) as m:
        with pytest.raises(G.GptbError):
            m.grad(1e-3, start=0, bs=4)                                   # nothing set yet
        with pytest.raises(G.GptbError):
            m.set_index(np.full((12, 3), 4, dtype=np.int32))              # I outside 1..r


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
def test_grad_vs_finite_differences(prec):
    """Diagnostic_gradients.jl on its own data: analytic (GPU) vs forward differences of the log-likelihood."""
    d = load_golden("synth_fd.npz")
    rng = np.random.default_rng(17)
    X, y = d["Xtrain"], d["ytrain"]
    N, D = X.shape
    n, r, Q, sigma = 10, 5, 10, 0.2
    Z = rng.standard_normal((n, D)); b = 2 * np.pi * rng.random((n, D))
    I = o.samplenz(r, D, Q, rng)
    U = np.stack([o.stiefel_from_gaussian(rng.standard_normal((r, n))) for _ in range(D)], 2)
    w = rng.standard_normal(Q)
    phi = o.feature(X, 1.4, 1.0, 1.0, Z, b)
    with G.Model(D, n, r, Q, precision=prec) as m:
        m.set_features(G.feature(X, 1.4, 1.0, 1.0, Z, b)); m.set_index(I); m.set_train(X, y); m.set_params(w, U)
        gw, gU, _ = m.grad(sigma ** 2, start=0, bs=N, include_prior=False)
    ll = o.loglik(phi, y, w, U, I, sigma)
    h = 1e-6
    fdw = np.array([(o.loglik(phi, y, w + h * np.eye(Q)[q], U, I, sigma) - ll) / h for q in range(Q)])
    assert relerr(gw, fdw) < 2e-4
    fdU = np.zeros_like(U)
    for k in range(D):
        for l in range(r):
            for j in range(n):
                U2 = U.copy(); U2[j, l, k] += h
                fdU[j, l, k] = (o.loglik(phi, y, w, U2, I, sigma) - ll) / h
    assert relerr(gU, fdU) < 2e-4


# ------------------------------------------------------------------ Stiefel ops
def test_proj_geod_vs_oracle():
    rng = np.random.default_rng(9)
    for n, r in ((50, 10), (1000, 32), (37, 5), (8, 8)):
        U = o.stiefel_from_gaussian(rng.standard_normal((r, n)))
        V = rng.standard_normal((n, r))
        M = G.proj(U, V)
        assert relerr(M, o.proj(U, V)) < 1e-12
        assert np.abs(U.T @ M + M.T @ U).max() < 1e-11
        for t in (1e-4, 3e-3, 0.2):
            Un = G.geod(U, M, t)
            want, _ = o.geod(U, M, t)
            assert relerr(Un, want) < 1e-10
            assert np.abs(Un.T @ Un - np.eye(r)).max() < 1e-10
    Z = G.geod(U, M * 1e200, 1e200)                                       # NaN -> warning + zeros (:24-26)
    assert Z.shape == U.shape and not Z.any()


# ------------------------------------------------------------------ sampler
@pytest.mark.parametrize("prec", PRECS, ids=IDS)
@pytest.mark.parametrize("langevin,stiefel", [(True, True), (False, True), (True, False), (False, False)])
def test_sgld_step_vs_oracle(prec, langevin, stiefel):
    P = make_problem(4, 400, 50, 10, Q=60, seed=7)
    rng = np.random.default_rng(1)
    nw = rng.standard_normal(P["Q"]); nU = np.asfortranarray(rng.standard_normal((P["n"], P["r"], P["D"])))
    epsw, epsU = 1e-5, 1e-8
    idx = rng.permutation(P["N"])[:150]
    want = o.sgld_step(P["phi_fn"](P["X"][idx]), P["y"][idx], P["w0"], P["U0"], P["I"], P["N"], P["signal_var"],
                       epsw, epsU, nw, nU, langevin, stiefel)
    with _model(P, prec) as m:
        m.sgld_step(P["signal_var"], epsw, epsU, idx=idx, langevin=langevin, stiefel=stiefel, noise_w=nw, noise_U=nU)
        w, U = m.get_params()
    # the update is FP64 in every mode; in TC3 only the gradient carries 1e-4, scaled by eps/2
    dw = relerr(w - P["w0"], want[0] - P["w0"]); dU = relerr(U - P["U0"], want[1] - P["U0"])
    assert dw < max(TOL[prec], 1e-9) and dU < max(TOL[prec], 1e-7)
    assert relerr(w, want[0]) < 1e-9 if prec == L.PREC_FP64 else relerr(w, want[0]) < 1e-5
    if stiefel:
        for k in range(P["D"]):
            assert np.abs(U[:, :, k].T @ U[:, :, k] - np.eye(P["r"])).max() < 1e-10


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
def test_sgld_step_host_and_pipelined(prec):
    """Host-minibatch loop body (gptb_sgld_step_host) against an oracle chain, and the pipelined submit / wait
    form against the blocking one (same operations in the same order)."""
    P = make_problem(4, 400, 50, 10, Q=60, seed=9)
    rng = np.random.default_rng(3)
    T, bs, epsw, epsU = 5, 150, 1e-5, 1e-8
    idx = [rng.permutation(P["N"])[:bs] for _ in range(T)]
    nw = [rng.standard_normal(P["Q"]) for _ in range(T)]
    nU = [np.asfortranarray(rng.standard_normal((P["n"], P["r"], P["D"]))) for _ in range(T)]
    Xb = [np.asfortranarray(P["X"][i]) for i in idx]
    yb = [np.ascontiguousarray(P["y"][i]) for i in idx]
    w, U = P["w0"], P["U0"]
    for t in range(T):
        w, U = o.sgld_step(P["phi_fn"](P["X"][idx[t]]), P["y"][idx[t]], w, U, P["I"], P["N"], P["signal_var"], epsw, epsU,
                           nw[t], nU[t], True, True)[:2]
    wo = np.zeros(P["Q"])
    with _model(P, prec, train=False) as m:
        for t in range(T):
            m.sgld_step_host(P["signal_var"], Xb[t], yb[t], P["N"], epsw, epsU, noise_w=nw[t], noise_U=nU[t], w_out=wo)
        w1, U1 = m.get_params()
    assert np.array_equal(wo, w1)
    assert relerr(w1 - P["w0"], w - P["w0"]) < max(TOL[prec], 1e-8) * 10
    assert relerr(U1 - P["U0"], U - P["U0"]) < max(TOL[prec], 1e-7) * 10
    wouts = [np.zeros(P["Q"]) for _ in range(T)]
    with _model(P, prec, train=False) as m:
        with pytest.raises(G.GptbError):
            m.sgld_step_host_wait()                       # nothing in flight
        m.sgld_step_host_submit(P["signal_var"], Xb[0], yb[0], P["N"], epsw, epsU, noise_w=nw[0], noise_U=nU[0], w_out=wouts[0])
        for t in range(1, T):
            m.sgld_step_host_submit(P["signal_var"], Xb[t], yb[t], P["N"], epsw, epsU, noise_w=nw[t], noise_U=nU[t],
                                    w_out=wouts[t])
            if t != 2:                                    # once, let a third submit retire the oldest by itself
                m.sgld_step_host_wait()
        m.sgld_step_host_wait()
        with pytest.raises(G.GptbError):
            m.sgld_step_host_wait()
        w2, U2 = m.get_params()
    # (sums over rows use floating-point atomics, so two runs agree to rounding, not bit for bit)
    tol = 1e-9 if prec == L.PREC_FP64 else 1e-5     # of the accumulated update; TC3 sums the gradient with fp32 atomics
    assert relerr(w2 - P["w0"], w1 - P["w0"]) < tol and relerr(U2 - P["U0"], U1 - P["U0"]) < tol
    assert np.array_equal(wouts[-1], w2)


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
@pytest.mark.parametrize("phi_mode", [False, True], ids=["lazy", "densephi"])
def test_GPTregression_vs_oracle(prec, phi_mode):
    """Whole sampler with injected draws: epochs, cumulative permutations, short last batch, burn-in, store layout."""
    P = make_problem(3, 150, 16, 4, Q=20, seed=8)
    N, m, burnin, maxepoch = 150, 64, 1, 2
    nb = 3
    rng = np.random.default_rng(2)
    perms = [rng.permutation(N) for _ in range(burnin + maxepoch)]
    steps = (burnin + maxepoch) * nb
    NW = rng.standard_normal((steps, P["Q"])); NU = rng.standard_normal((steps, P["n"], P["r"], P["D"]))
    phi = P["phi_fn"](P["X"])
    ws, Us = o.GPTregression(phi, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, 1e-8, burnin, maxepoch,
                             P["w0"], P["U0"], perms, NW, NU)
    src = phi if phi_mode else _feat(P)
    gw, gU = G.GPTregression(src, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, 1e-8, burnin, maxepoch, 1,
                             precision=prec, w0=P["w0"], U0=P["U0"], perms=np.stack(perms), noise_w=NW, noise_U=NU)
    assert gw.shape == ws.shape and gU.shape == Us.shape
    tol = 1e-8 if (prec == L.PREC_FP64 or phi_mode) else 1e-4
    assert relerr(gw, ws) < tol and relerr(gU, Us) < tol


def test_GPTregression_production_mode_and_nan_abort():
    P = make_problem(3, 300, 16, 4, Q=20, seed=10)
    f = _feat(P)
    a = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], 100, 1e-5, 1e-8, 1, 2, 7, precision=L.PREC_FP64)
    b = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], 100, 1e-5, 1e-8, 1, 2, 7, precision=L.PREC_FP64)
    assert a[0].shape == (P["Q"], 6) and a[1].shape == (P["n"], P["r"], P["D"], 6)
    assert relerr(a[0], b[0]) < 1e-12 and relerr(a[1], b[1]) < 1e-12      # counter-based RNG: reproducible
    assert np.isfinite(a[0]).all() and a[0].any()
    c = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], 100, 1e-5, 1e-8, 1, 2, 8, precision=L.PREC_FP64)
    assert relerr(c[0], a[0]) > 1e-3                                      # different param_seed, different chain
    for k in range(P["D"]):
        Uk = a[1][:, :, k, -1]
        assert np.abs(Uk.T @ Uk - np.eye(P["r"])).max() < 1e-10
    z = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], 100, 1e-5, 1e300, 0, 1, 7, precision=L.PREC_FP64)
    assert not z[0].any() and not z[1].any()                               # :422-424


@pytest.mark.parametrize("groups", [2, 3])
def test_dimension_group_pipeline(groups):
    """The dimension-group pipeline of the step (api.cu: step_pipelined — the update of a group of dimensions runs on
    a side stream under the next group's back-projection and the next step's forward pass) must not change the
    chain: same Philox draws, same parameters as the one-stream step, in every driver of the step."""
    P = make_problem(6, 700, 64, 8, Q=50, seed=21)
    f = _feat(P)
    epsw, epsU, m_ = 1e-6, 1e-9, 256
    rng = np.random.default_rng(4)
    idx = [rng.permutation(P["N"])[:200] for _ in range(4)]
    Xb = [np.asfortranarray(P["X"][i]) for i in idx]
    yb = [np.ascontiguousarray(P["y"][i]) for i in idx]

    def drive(g):
        opts = {L.OPT_PIPE_GROUPS: g, L.OPT_ALGORITHM: L.ALG_DIRECT}
        out = {}
        with _model(P, L.PREC_TC3) as m:                       # resident data, library-side loop
            for k_, v_ in opts.items():
                m.set_option(k_, v_)
            m.sgld_advance(P["signal_var"], m_, epsw, epsU, 7)
            out["advance"] = m.get_params()
        wouts = [np.zeros(P["Q"]) for _ in idx]
        with _model(P, L.PREC_TC3, train=False) as m:          # host minibatches, submit / wait
            for k_, v_ in opts.items():
                m.set_option(k_, v_)
            m.sgld_step_host_submit(P["signal_var"], Xb[0], yb[0], P["N"], epsw, epsU, w_out=wouts[0])
            for t in range(1, len(idx)):
                m.sgld_step_host_submit(P["signal_var"], Xb[t], yb[t], P["N"], epsw, epsU, w_out=wouts[t])
                m.sgld_step_host_wait()
            m.sgld_step_host_wait()
            out["host"] = m.get_params()
            assert np.array_equal(wouts[-1], out["host"][0])
        out["host_w"] = np.stack(wouts)
        out["run"] = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m_, epsw, epsU, 1, 2, 7,
                                     precision=L.PREC_TC3, options=opts)
        z = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m_, epsw, 1e300, 0, 1, 7,
                            precision=L.PREC_TC3, options=opts)
        assert not z[0].any() and not z[1].any()               # NaN abort (:422-424) through the pipeline
        return out

    ref, got = drive(0), drive(groups)
    tol = 2e-5   # of the accumulated update: the gradient sums use fp32 atomics, two runs agree to rounding only
    for key in ("advance", "host"):
        assert relerr(got[key][0] - P["w0"], ref[key][0] - P["w0"]) < tol, key
        assert relerr(got[key][1] - P["U0"], ref[key][1] - P["U0"]) < tol, key
    assert relerr(got["host_w"] - P["w0"], ref["host_w"] - P["w0"]) < tol
    assert got["run"][0].shape == ref["run"][0].shape
    assert relerr(got["run"][0], ref["run"][0]) < 1e-5 and relerr(got["run"][1], ref["run"][1]) < 1e-5
    for k in range(P["D"]):
        Uk = got["advance"][1][:, :, k]
        assert np.abs(Uk.T @ Uk - np.eye(P["r"])).max() < 1e-10


def test_philox_noise_statistics():
    """Production-mode noise: w_{t+1}-w_t with zero step size on the gradient is sqrt(epsw)*N(0,1)."""
    P = make_problem(2, 64, 8, 2, Q=4, seed=11)
    with G.Model(2, 8, 2, 4, precision=L.PREC_FP64, seed=99) as m:
        m.set_features(_feat(P)); m.set_index(P["I"]); m.set_train(P["X"], P["y"] * 0)
        draws = []
        for _ in range(400):
            m.set_params(np.zeros(4), P["U0"])
            m.sgld_step(1e30, 1.0, 1e-30, start=0, bs=64)                 # signal_var huge: gradient ~ 0
            draws.append(m.get_params()[0])
        z = np.concatenate(draws)
    assert abs(z.mean()) < 0.1 and abs(z.std() - 1) < 0.06
    assert abs(np.mean(z ** 4) - 3) < 0.6


# ------------------------------------------------------------------ size-independent properties at full scale
@pytest.mark.parametrize("prec", [L.PREC_TC3], ids=["tc3"])
def test_full_size_properties_c2(prec):
    """C2 at full size (D=8, N=10000, n=100, r=20, m=1000): f-hat is linear in w; the minibatch gradient over a
    partition of the data adds up to the full-batch gradient; parameters stay on the manifold."""
    P = make_problem(8, 10000, 100, 20, seed=12)
    with _model(P, prec) as m:
        f1 = m.pred(P["w0"], P["U0"], P["X"])
        f2 = m.pred(2.5 * P["w0"], P["U0"], P["X"])
        assert relerr(f2, 2.5 * f1) < 1e-6
        full = m.grad(P["signal_var"], start=0, bs=10000, include_prior=False, want_fhat=False)
        acc_w = np.zeros(P["Q"]); acc_U = np.zeros_like(P["U0"])
        for s in range(0, 10000, 1000):
            gw, gU, _ = m.grad(P["signal_var"], start=s, bs=1000, include_prior=False, want_fhat=False)
            acc_w += gw / 10; acc_U += gU / 10                             # each batch is scaled by N/bs = 10
        assert relerr(acc_w, full[0]) < 1e-4 and relerr(acc_U, full[1]) < 1e-4
        for _ in range(3):
            m.sgld_step(P["signal_var"], 1e-5, 1e-8, start=0, bs=1000)
        _, U = m.get_params()
        for k in range(8):
            assert np.abs(U[:, :, k].T @ U[:, :, k] - np.eye(20)).max() < 1e-9


# ------------------------------------------------------------------ full-theta comparator (GPT_SGLD.jl:109-120, :809-847)
def _nt_problem(seed=31, N=700, D=5, n=300):
    rng = np.random.default_rng(seed)
    X = np.asfortranarray(rng.standard_normal((N, D)))
    Z = np.asfortranarray(rng.standard_normal((n, D)))
    b = 2 * np.pi * rng.random(n)
    ls = np.maximum(1.0 + 0.2 * rng.standard_normal(D), 0.3)
    th_true = rng.standard_normal(n)
    phi = o.featureNotensor(X, ls, 1.3, Z, b)
    y = phi.T @ th_true + 0.05 * rng.standard_normal(N)
    return dict(X=X, Z=Z, b=b, ls=ls, sig=1.3, phi=phi, y=y, rng=rng, N=N, D=D, n=n)


@pytest.mark.parametrize("prec", PRECS, ids=["fp64", "fp32"])
def test_fulltheta_features_pred_and_chain(prec):
    P = _nt_problem()
    tol = 1e-10 if prec == L.PREC_FP64 else 1e-4
    f = G.featureNotensor(P["X"], P["ls"], P["sig"], P["Z"], P["b"])
    assert f.shape == (P["n"], P["N"])
    assert relerr(np.asarray(f), P["phi"]) < 1e-12                      # materialisation is FP64 in both modes
    th = P["rng"].standard_normal(P["n"])
    assert relerr(G.pred_notensor(th, f, precision=prec), P["phi"].T @ th) < tol
    burnin, maxepoch, m = 1, 2, 256                                     # 3 batches per epoch, the last one short
    nb = -(-P["N"] // m)
    perms = np.stack([P["rng"].permutation(P["N"]) for _ in range(burnin + maxepoch)])
    noise = P["rng"].standard_normal(((burnin + maxepoch) * nb, P["n"]))
    th0 = 0.5 * P["rng"].standard_normal(P["n"])
    want = o.GPNT_SGLD(P["phi"], P["y"], 0.01, 1.0, m, 2e-6, 0.3, burnin, maxepoch, th0, perms, noise)
    got = G.GPNT_SGLD(f, P["y"], 0.01, 1.0, m, 2e-6, 0.3, burnin, maxepoch, 7, precision=prec, theta0=th0, perms=perms,
                      noise=noise)
    assert got.shape == want.shape
    assert relerr(got - th0[:, None], want - th0[:, None]) < tol
    # scalar length scale and a single-row batch
    f1 = G.featureNotensor(P["X"][:129], 0.8, 1.0, P["Z"], P["b"])
    want1 = o.GPNT_SGLD(o.featureNotensor(P["X"][:129], 0.8, 1.0, P["Z"], P["b"]), P["y"][:129], 0.01, 2.0, 128, 1e-6, 0.0, 0, 1,
                        th0, [np.arange(129)], noise[:2])
    got1 = G.GPNT_SGLD(f1, P["y"][:129], 0.01, 2.0, 128, 1e-6, 0.0, 0, 1, 7, precision=prec, theta0=th0,
                       perms=np.arange(129)[None, :], noise=noise[:2])
    assert relerr(got1 - th0[:, None], want1 - th0[:, None]) < tol


def test_fulltheta_production_mode_and_nan_abort():
    P = _nt_problem(seed=32, N=500, D=4, n=130)
    f = G.featureNotensor(P["X"], P["ls"], P["sig"], P["Z"], P["b"])
    a = G.GPNT_SGLD(f, P["y"], 0.01, 1.0, 200, 1e-6, 0.2, 1, 2, 11, precision=L.PREC_FP64)
    b = G.GPNT_SGLD(f, P["y"], 0.01, 1.0, 200, 1e-6, 0.2, 1, 2, 11, precision=L.PREC_FP64)
    assert a.shape == (P["n"], 9) and np.isfinite(a).all()
    assert relerr(a, b) < 1e-12                                            # counter-based RNG: reproducible
    c = G.GPNT_SGLD(f, P["y"], 0.01, 1.0, 200, 1e-6, 0.2, 1, 2, 12, precision=L.PREC_FP64)
    assert relerr(c, a) > 1e-3
    # the chain moves towards the data: residual of the last sample well below that of the first
    r0 = np.linalg.norm(P["y"] - P["phi"].T @ a[:, 0]); r1 = np.linalg.norm(P["y"] - P["phi"].T @ a[:, -1])
    assert r1 < r0
    z = G.GPNT_SGLD(f, P["y"], 0.01, 1.0, 200, 1e300, 0.0, 0, 1, 11, precision=L.PREC_FP64)
    assert z.shape == (P["n"],) and not z.any()                            # :839-842
    with pytest.raises(ValueError):
        G.featureNotensor(P["X"], np.ones(3), 1.0, P["Z"], P["b"])


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
def test_GPT_SGLDERMw_vs_oracle(prec):
    """w-only sampler (GPT_SGLD.jl:1065-1118, GPTB_FLAG_FIX_U): chain of w against the oracle with injected draws; U
    must come back untouched."""
    P = make_problem(3, 230, 16, 4, Q=20, seed=12)
    N, m, burnin, maxepoch = 230, 100, 1, 2
    nb = 3
    rng = np.random.default_rng(6)
    perms = [rng.permutation(N) for _ in range(burnin + maxepoch)]
    NW = rng.standard_normal(((burnin + maxepoch) * nb, P["Q"]))
    phi = P["phi_fn"](P["X"])
    ws, U = o.GPT_SGLDERMw(phi, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, burnin, maxepoch, P["w0"], P["U0"],
                           perms, NW)
    gw, gU = G.GPT_SGLDERMw(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, burnin, maxepoch, 3,
                            precision=prec, w0=P["w0"], U0=P["U0"], perms=np.stack(perms), noise_w=NW)
    assert gw.shape == ws.shape == (P["Q"], maxepoch * nb)
    assert relerr(gw, ws) < (1e-9 if prec == L.PREC_FP64 else 1e-4)
    assert np.array_equal(gU, P["U0"])
    # production mode: reproducible, U stays on the manifold it was drawn from
    a = G.GPT_SGLDERMw(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, 0, 2, 5, precision=prec)
    b = G.GPT_SGLDERMw(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, 0, 2, 5, precision=prec)
    assert relerr(a[0], b[0]) < 1e-5 and np.array_equal(a[1], b[1])
    for k in range(P["D"]):
        assert np.abs(a[1][:, :, k].T @ a[1][:, :, k] - np.eye(P["r"])).max() < 1e-10


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
def test_rmse_trajectory_powerplant(prec):
    """North-star criterion "test RMSE trajectories of the SGLD chains within 1 % over the run": the reference's
    PowerPlant experiment (whitened CCPP data of PPdata_full.h5, length_scale and signal_var of
    PowerPlantDataExperiment.jl:20-24, the BASELINE config-1 shape n=50, r=10, Q=200, minibatch 500) run by the oracle
    and by the library with the same draws; per epoch, pred of the epoch's last sample on held-out rows and
    RMSE = ||ytest - fhat|| / sqrt(Ntest) as in kin40kExperiment.jl:75-87."""
    pp = load_golden("powerplant.npz")
    X = np.asfortranarray(pp["Xtrain"].astype(np.float64)); y = pp["ytrain"].astype(np.float64)
    Ntr, Nte = 2500, 1500
    Xtr, ytr, Xte, yte = X[:Ntr], y[:Ntr], X[Ntr:Ntr + Nte], y[Ntr:Ntr + Nte]
    D, n, r, Q, m = 4, 50, 10, 200, 500
    rng = np.random.default_rng(18)
    Z = np.asfortranarray(rng.standard_normal((n, D))); b = np.asfortranarray(2 * np.pi * rng.random((n, D)))
    ls, sig, signal_var = 1.4332, 1.0, 0.2299 ** 2
    phi_scale = np.sqrt(n / Q ** (1.0 / D))
    I = o.samplenz(r, D, Q, rng)
    w0 = rng.standard_normal(Q)
    U0 = np.asfortranarray(np.stack([o.stiefel_from_gaussian(rng.standard_normal((r, n))) for _ in range(D)], 2))
    epochs, nb = 8, Ntr // m
    perms = [rng.permutation(Ntr) for _ in range(epochs)]
    NW = rng.standard_normal((epochs * nb, Q)); NU = rng.standard_normal((epochs * nb, n, r, D))
    epsw, epsU = 2e-4, 2e-7
    phitr = o.feature(Xtr, ls, sig, phi_scale, Z, b); phite = o.feature(Xte, ls, sig, phi_scale, Z, b)
    ws, Us = o.GPTregression(phitr, ytr, signal_var, I, r, Q, m, epsw, epsU, 0, epochs, w0, U0, perms, NW, NU)
    ftr = G.feature(Xtr, ls, sig, phi_scale, Z, b); fte = G.feature(Xte, ls, sig, phi_scale, Z, b)
    gw, gU = G.GPTregression(ftr, ytr, signal_var, I, r, Q, m, epsw, epsU, 0, epochs, 1, precision=prec, w0=w0, U0=U0,
                             perms=np.stack(perms), noise_w=NW, noise_U=NU)
    ref_curve, got_curve = [], []
    for ep in range(epochs):
        col = (ep + 1) * nb - 1
        ref_curve.append(np.linalg.norm(yte - o.pred(ws[:, col], Us[..., col], I, phite)) / np.sqrt(Nte))
        got_curve.append(np.linalg.norm(yte - G.pred(gw[:, col], gU[..., col], I, fte, precision=prec)) / np.sqrt(Nte))
    ref_curve, got_curve = np.array(ref_curve), np.array(got_curve)
    assert np.isfinite(ref_curve).all() and ref_curve[-1] < 0.7 * ref_curve[0]        # the chain learns the data (0.60 -> 0.38)
    assert np.abs(got_curve / ref_curve - 1).max() < (1e-7 if prec == L.PREC_FP64 else 1e-2)


def test_sgld_run_starts_from_data_order_and_rejects_bad_perms():
    """Each gptb_sgld_run starts from the caller's data order, as each GPTregression call does (GPT_SGLD.jl:374 rebinds
    a local): the same host permutations give the same minibatches on a second run of the same handle.  A permutation
    entry outside 0..N-1 is an argument error, not an out-of-bounds read."""
    P = make_problem(3, 120, 12, 3, Q=10, seed=14)
    rng = np.random.default_rng(5)
    perms = np.stack([rng.permutation(120) for _ in range(2)])
    with _model(P, L.PREC_FP64) as m:
        a = m.sgld_run(P["signal_var"], 50, 1e-5, 1e-8, 0, 2, langevin=False, perms=perms, N=120)
        m.set_params(P["w0"], P["U0"])
        b = m.sgld_run(P["signal_var"], 50, 1e-5, 1e-8, 0, 2, langevin=False, perms=perms, N=120)
        assert relerr(b[0], a[0]) < 1e-12 and relerr(b[1], a[1]) < 1e-12
        bad = perms.copy(); bad[1, 7] = 120
        m.set_params(P["w0"], P["U0"])
        with pytest.raises(G.GptbError):
            m.sgld_run(P["signal_var"], 50, 1e-5, 1e-8, 0, 2, langevin=False, perms=bad, N=120)


# ------------------------------------------------------------------ node-grid algorithm (csrc/nodes.cu)
@pytest.mark.parametrize("shape", list(SHAPES))
def test_nodes_grad_and_step_vs_oracle(shape):
    """GPTB_ALG_NODES: features generated at the Chebyshev nodes of a grid over the input domain, T interpolated to the
    rows, E spread back to the nodes (nodes.cu).  Same tolerances as the direct tensor-core path."""
    P = make_problem(seed=2, **SHAPES[shape])
    N = P["N"]
    rng = np.random.default_rng(0)
    idx = rng.permutation(N)[: max(1, (2 * N) // 3)]
    phi = P["phi_fn"](P["X"][idx])
    want = o.gradients(phi, P["y"][idx], P["w0"], P["U0"], P["I"], N, P["signal_var"])
    nw = rng.standard_normal(P["Q"]); nU = np.asfortranarray(rng.standard_normal((P["n"], P["r"], P["D"])))
    step = o.sgld_step(phi, P["y"][idx], P["w0"], P["U0"], P["I"], N, P["signal_var"], 1e-5, 1e-8, nw, nU)
    with _model(P, L.PREC_TC3) as m:
        m.set_option(L.OPT_ALGORITHM, L.ALG_NODES)
        gw, gU, fh = m.grad(P["signal_var"], idx=idx)
        assert relerr(fh, want[2]) < 1e-4 and relerr(gw, want[0]) < 1e-4 and relerr(gU, want[1]) < 1e-4
        full = o.gradients(P["phi_fn"](P["X"]), P["y"], P["w0"], P["U0"], P["I"], N, P["signal_var"])
        gw, gU, fh = m.grad(P["signal_var"], start=0, bs=N)
        assert relerr(fh, full[2]) < 1e-4 and relerr(gw, full[0]) < 1e-4 and relerr(gU, full[1]) < 1e-4
        m.sgld_step(P["signal_var"], 1e-5, 1e-8, idx=idx, noise_w=nw, noise_U=nU)
        w, U = m.get_params()
    assert relerr(w - P["w0"], step[0] - P["w0"]) < 1e-4 and relerr(U - P["U0"], step[1] - P["U0"]) < 1e-4


@pytest.mark.parametrize("order,theta,tol", [(12, 2.0, 1e-4), (16, 2.0, 1e-4), (8, 1.0, 1e-4), (4, 4.0, None)])
def test_nodes_orders_and_host_domain(order, theta, tol):
    """Grid options (nodes per cell, cell width), feature2, and the host-minibatch entry with an explicit domain; a
    grid that is far too coarse (4 nodes on cells of half-width 4/zmax) must visibly miss the tolerance — i.e. the
    options really select the grid."""
    P = make_problem(5, 900, 60, 6, Q=40, seed=15, feature2=True)
    idx = np.arange(100, 800)
    want = o.gradients(P["phi_fn"](P["X"][idx]), P["y"][idx], P["w0"], P["U0"], P["I"], P["N"], P["signal_var"])
    with _model(P, L.PREC_TC3, train=False) as m:
        m.set_option(L.OPT_ALGORITHM, L.ALG_NODES)
        m.set_option(L.OPT_NODE_ORDER, order)
        m.set_option(L.OPT_NODE_THETA, theta)
        m.set_domain(P["X"].min(0) - 0.1, P["X"].max(0) + 0.1)
        m.set_train(P["X"], P["y"])                     # (set_train replaces the domain by the exact column ranges)
        gw, gU, fh = m.grad(P["signal_var"], idx=idx)
        err = max(relerr(gw, want[0]), relerr(gU, want[1]), relerr(fh, want[2]))
        if tol is None:
            assert err > 1e-4
            return
        assert err < tol
    # host minibatches: the domain comes from gptb_set_domain; a row outside it is an argument error
    rng = np.random.default_rng(2)
    nw = rng.standard_normal(P["Q"]); nU = np.asfortranarray(rng.standard_normal((P["n"], P["r"], P["D"])))
    Xb = np.asfortranarray(P["X"][idx]); yb = np.ascontiguousarray(P["y"][idx])
    step = o.sgld_step(P["phi_fn"](P["X"][idx]), yb, P["w0"], P["U0"], P["I"], P["N"], P["signal_var"], 1e-5, 1e-8, nw, nU)
    with _model(P, L.PREC_TC3, train=False) as m:
        m.set_option(L.OPT_ALGORITHM, L.ALG_NODES)
        m.set_option(L.OPT_NODE_ORDER, order)
        m.set_option(L.OPT_NODE_THETA, theta)
        m.set_domain(P["X"].min(0), P["X"].max(0))
        m.sgld_step_host(P["signal_var"], Xb, yb, P["N"], 1e-5, 1e-8, noise_w=nw, noise_U=nU)
        w, U = m.get_params()
        assert relerr(w - P["w0"], step[0] - P["w0"]) < 1e-4 and relerr(U - P["U0"], step[1] - P["U0"]) < 1e-4
        Xbad = Xb.copy(); Xbad[3, 1] = P["X"][:, 1].max() + 1.0
        with pytest.raises(G.GptbError):
            m.sgld_step_host(P["signal_var"], Xbad, yb, P["N"], 1e-5, 1e-8, noise_w=nw, noise_U=nU)


def test_nodes_GPTregression_vs_oracle():
    P = make_problem(3, 150, 16, 4, Q=20, seed=8)
    N, m, burnin, maxepoch = 150, 64, 1, 2
    rng = np.random.default_rng(2)
    perms = [rng.permutation(N) for _ in range(burnin + maxepoch)]
    steps = (burnin + maxepoch) * 3
    NW = rng.standard_normal((steps, P["Q"])); NU = rng.standard_normal((steps, P["n"], P["r"], P["D"]))
    ws, Us = o.GPTregression(P["phi_fn"](P["X"]), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, 1e-8, burnin,
                             maxepoch, P["w0"], P["U0"], perms, NW, NU)
    gw, gU = G.GPTregression(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-5, 1e-8, burnin, maxepoch, 1,
                             precision=L.PREC_TC3, w0=P["w0"], U0=P["U0"], perms=np.stack(perms), noise_w=NW, noise_U=NU,
                             options={L.OPT_ALGORITHM: L.ALG_NODES})
    assert relerr(gw, ws) < 1e-4 and relerr(gU, Us) < 1e-4


@pytest.mark.parametrize("layout", [0, 1], ids=["lanes", "tiles"])
@pytest.mark.parametrize("r,S,order,feature2", [(4, 5, 12, True), (5, 37, 12, False), (20, 11, 8, False), (33, 3, 16, False)])
def test_nodes_pred_samples_vs_oracle(r, S, order, feature2, layout):
    """Posterior sweep on the node grid (pred_nodes.cu) forced on at small sizes: r not a multiple of 4 (scalar tile
    loads), more samples than one column group (ragged last group), orders other than 12 (64-column groups), r > 32;
    mean, per-sample output and the mean-only call against oracle.pred."""
    P = make_problem(4, 3000, 24 if feature2 else 40, r, Q=30, seed=31, feature2=feature2)
    rng = np.random.default_rng(1)
    ws = np.asfortranarray(rng.standard_normal((P["Q"], S)))
    Us = np.asfortranarray(np.stack([np.stack([o.stiefel_from_gaussian(rng.standard_normal((P["r"], P["n"])))
                                               for _ in range(P["D"])], 2) for _ in range(S)], 3))
    phi = P["phi_fn"](P["X"])
    want = np.stack([o.pred(ws[:, s], Us[:, :, :, s], P["I"], phi) for s in range(S)], 1)
    # both column layouts of a sample group: samples on the lanes (8 / 16 / 32 side by side: S = 5, 11, 37 reach all
    # three and a ragged last group) and one row tile per sample
    opts = {L.OPT_ALGORITHM: L.ALG_NODES, L.OPT_NODE_ORDER: order, L.OPT_PRED_LAYOUT: layout}
    mean, per = G.pred_samples(ws, Us, P["I"], _feat(P), per_sample=True, precision=L.PREC_TC3, options=opts)
    assert relerr(per, want) < 1e-4 and relerr(mean, want.mean(1)) < 1e-4
    for s in (0, S - 1):
        assert relerr(per[:, s], want[:, s]) < 1e-4, s
    only_mean = G.pred_samples(ws, Us, P["I"], _feat(P), precision=L.PREC_TC3, options=opts)
    assert relerr(only_mean, want.mean(1)) < 1e-4
    # a single parameter set through pred(), and a second call on the same handle with fewer rows (cached buffers)
    with _model(P, L.PREC_TC3, train=False) as m:
        m.set_option(L.OPT_ALGORITHM, L.ALG_NODES)
        m.set_option(L.OPT_NODE_ORDER, order)
        m.set_option(L.OPT_PRED_LAYOUT, layout)
        a = m.pred(ws[:, 0], Us[:, :, :, 0], P["X"])
        b = m.pred(ws[:, 1], Us[:, :, :, 1], P["X"][:777])
    assert relerr(a, want[:, 0]) < 1e-4 and relerr(b, want[:777, 1]) < 1e-4


def test_geod_vs_independent_pins():
    """gptb_geod against the two pins of `geod` that do not involve the oracle's expm: the great circle for r = 1 and
    the 40-digit mpmath evaluations of GPT_SGLD.jl:19-37 in tests/golden/geod_mp.npz (r in {2,10,32}, t in {1e-7,1e-3,0.3})."""
    rng = np.random.default_rng(4)
    for n in (2, 7, 300):
        u = rng.standard_normal((n, 1)); u /= np.linalg.norm(u)
        m = o.proj(u, rng.standard_normal((n, 1)))
        for t in (1e-7, 1e-3, 0.3, 2.5):
            assert np.abs(G.geod(u, m, t) - o.geod_great_circle(u, m, t)).max() < 1e-12
    g = load_golden("geod_mp.npz")
    for c in range(int(g["ncases"])):
        got = G.geod(g[f"U{c}"], g[f"M{c}"], float(g[f"t{c}"]))
        assert relerr(got, g[f"geod{c}"]) < 1e-10, c


@pytest.mark.parametrize("prec,alg,phi_mode", [(L.PREC_FP64, L.ALG_DIRECT, False), (L.PREC_FP64, L.ALG_DIRECT, True),
                                               (L.PREC_TC3, L.ALG_DIRECT, False), (L.PREC_TC3, L.ALG_NODES, False)],
                         ids=["fp64", "fp64-densephi", "tc3-direct", "tc3-nodes"])
def test_GPT_SGLDERM_RMSprop_vs_oracle(prec, alg, phi_mode):
    """RMSprop-preconditioned sampler (GPT_SGLD.jl:1121-1237, GPTB_FLAG_RMSPROP) with injected draws: per-element step
    sizes for w, one mean step size per U^(k), computeA on the already-updated w, moving averages carried over the
    steps (short last batch, burn-in, store layout as in GPTregression)."""
    P = make_problem(3, 150, 16, 4, Q=20, seed=8)
    N, m, burnin, maxepoch, eps, alpha = 150, 64, 1, 2, 2e-5, 0.9
    rng = np.random.default_rng(2)
    perms = [rng.permutation(N) for _ in range(burnin + maxepoch)]
    steps = (burnin + maxepoch) * 3
    NW = rng.standard_normal((steps, P["Q"])); NU = rng.standard_normal((steps, P["n"], P["r"], P["D"]))
    phi = P["phi_fn"](P["X"])
    ws, Us = o.GPT_SGLDERM_RMSprop(phi, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, eps, alpha, burnin, maxepoch,
                                   P["w0"], P["U0"], perms, NW, NU)
    assert np.isfinite(ws).all() and ws.any()
    gw, gU = G.GPT_SGLDERM_RMSprop(phi if phi_mode else _feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, eps, alpha,
                                   burnin, maxepoch, 1, precision=prec, w0=P["w0"], U0=P["U0"], perms=np.stack(perms),
                                   noise_w=NW, noise_U=NU, options={L.OPT_ALGORITHM: alg})
    assert gw.shape == ws.shape and gU.shape == Us.shape
    # FP64 pins the logic.  In TC3 the gradient is within 1e-4 NORM-wise, and RMSprop divides every element by the running
    # magnitude of its own gradient: elements with small gradients carry a larger relative error into the update, so
    # the accumulated update is compared at 1e-3 and the chain values themselves at 1e-4
    tol = 1e-8 if prec == L.PREC_FP64 else 1e-3
    assert relerr(gw - P["w0"][:, None], ws - P["w0"][:, None]) < tol
    assert relerr(gU - P["U0"][..., None], Us - P["U0"][..., None]) < tol
    assert relerr(gw, ws) < min(tol, 1e-4) and relerr(gU, Us) < min(tol, 1e-4)
    # production mode: reproducible, on the manifold, and a second run on the same seed starts from fresh averages
    a = G.GPT_SGLDERM_RMSprop(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, eps, alpha, 0, 2, 5, precision=prec)
    b = G.GPT_SGLDERM_RMSprop(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, eps, alpha, 0, 2, 5, precision=prec)
    assert relerr(a[0], b[0]) < 1e-5 and relerr(a[1], b[1]) < 1e-5
    for k in range(P["D"]):
        assert np.abs(a[1][:, :, k, -1].T @ a[1][:, :, k, -1] - np.eye(P["r"])).max() < 1e-10


def test_geodboth_vs_oracle():
    """gptb_geodboth (GPT_SGLD.jl:40-59): geodesic point and transported momentum; the momentum stays tangent and keeps
    its norm; the point equals geod's."""
    rng = np.random.default_rng(12)
    for n, r in ((50, 10), (300, 32), (9, 2), (40, 40)):
        U = o.stiefel_from_gaussian(rng.standard_normal((r, n)))
        M = o.proj(U, rng.standard_normal((n, r)))
        for t in (1e-4, 0.05):
            Un, Mn = G.geodboth(U, M, t)
            wU, wM, _ = o.geodboth(U, M, t)
            assert relerr(Un, wU) < 1e-10 and relerr(Mn, wM) < 1e-10
            assert relerr(Un, G.geod(U, M, t)) < 1e-12
            assert np.abs(Un.T @ Mn + Mn.T @ Un).max() < 1e-9 * max(1.0, np.abs(Mn).max() ** 2)
    Z, Zm = G.geodboth(U, M * 1e200, 1e200)                               # NaN -> zeros for both (:43-45)
    assert not Z.any() and not Zm.any()


@pytest.mark.parametrize("prec", PRECS, ids=IDS)
def test_GPT_GMC_vs_oracle(prec):
    """Geodesic Monte Carlo (GPT_SGLD.jl:684-805) with injected momenta and uniforms: acceptance probabilities and the
    stored chain against the oracle, including a REJECTED epoch (where the reference, through its aliasing of U_old,
    restores w only) in both settings of restore_U_on_reject."""
    P = make_problem(3, 120, 12, 3, Q=15, seed=8, signal_var=0.05)
    rng = np.random.default_rng(2)
    burnin, maxepoch, Lf = 1, 4, 3
    T = burnin + maxepoch
    NP = rng.standard_normal((T, P["Q"])); NM = rng.standard_normal((T, P["n"], P["r"], P["D"]))
    un = rng.random(T)
    un[2] = 1e9                                   # epoch 3 is rejected whatever its acceptance probability
    phi = P["phi_fn"](P["X"])
    tol = 1e-8 if prec == L.PREC_FP64 else 1e-4
    for restore in (False, True):
        ws, Us, acc = o.GPT_GMC(phi, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], 1e-5, 1e-7, burnin, maxepoch, Lf, P["w0"],
                                P["U0"], NP, NM, un, restore_U_on_reject=restore)
        gw, gU, gacc = G.GPT_GMC(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], 1e-5, 1e-7, burnin, maxepoch, Lf, 3,
                                 precision=prec, w0=P["w0"], U0=P["U0"], noise_p=NP, noise_mom=NM, unif=un,
                                 restore_U_on_reject=restore)
        assert gw.shape == ws.shape and gU.shape == Us.shape and gacc.shape == acc.shape
        # H is a difference of sums of ~N/signal_var: the acceptance probability inherits the gradient tolerance
        assert np.abs(gacc / acc - 1).max() < (1e-7 if prec == L.PREC_FP64 else 2e-2)
        assert relerr(gw, ws) < tol and relerr(gU, Us) < tol
        # the rejected epoch: w equals the previous column; U does so only when restore_U_on_reject
        assert np.array_equal(gw[:, 1], gw[:, 0])
        assert np.array_equal(gU[..., 1], gU[..., 0]) == restore
    z = G.GPT_GMC(_feat(P), P["y"], P["signal_var"], P["I"], P["r"], P["Q"], 1e-5, 1e300, 0, 2, 1, 3, precision=L.PREC_FP64)
    assert not z[0].any() and not z[1].any() and np.isnan(z[2]).all()      # :746-748


# ------------------------------------------------------------------------------------------------ classification (§8 f3)
def _class_problem(C=3, seed=12):
    P = make_problem(3, 150, 16, 4, Q=20, seed=seed)
    rng = np.random.default_rng(seed + 100)
    P["labels"] = rng.integers(0, C, P["N"])
    P["labels"][:C] = np.arange(C)                        # every class present
    P["W0"] = rng.standard_normal((P["Q"], C))
    P["UU0"] = np.stack([np.stack([np.linalg.qr(rng.standard_normal((P["n"], P["r"])))[0] for _ in range(P["D"])], axis=2)
                         for _ in range(C)], axis=3)
    return P, C


@pytest.mark.parametrize("prec,alg", [(L.PREC_FP64, L.ALG_DIRECT), (L.PREC_TC3, L.ALG_DIRECT), (L.PREC_TC3, L.ALG_NODES)],
                         ids=["fp64", "tc3-direct", "tc3-nodes"])
@pytest.mark.parametrize("langevin,stiefel,double", [(True, True, True), (False, True, True), (True, False, True), (True, True, False)],
                         ids=["ref", "nolangevin", "euclid", "single"])
def test_GPTclassification_vs_oracle(prec, alg, langevin, stiefel, double):
    """GPT_SGLD.jl:452-680 with injected permutations and draws: softmax residual, the reference's two moves per minibatch
    (the second under the langevin / stiefel switches), short last batch, burn-in, store layout [Q,C,T] / [n,r,D,C,T]."""
    P, C = _class_problem()
    N, m, burnin, maxepoch = P["N"], 64, 1, 2
    rng = np.random.default_rng(4)
    perms = [rng.permutation(N) for _ in range(burnin + maxepoch)]
    steps = (burnin + maxepoch) * 3
    NW = rng.standard_normal((steps, 2, P["Q"], C)); NU = rng.standard_normal((steps, 2, P["n"], P["r"], P["D"], C))
    ws, Us = o.GPTclassification(P["phi_fn"](P["X"]), P["labels"], P["I"], P["r"], P["Q"], m, 1e-4, 1e-6, burnin, maxepoch,
                                 P["W0"], P["UU0"], perms, NW, NU, langevin, stiefel, double)
    gw, gU = G.GPTclassification(_feat(P), P["labels"] + 1, P["I"], P["r"], P["Q"], m, 1e-4, 1e-6, burnin, maxepoch, 5,
                                 langevin, stiefel, precision=prec, w0=P["W0"], U0=P["UU0"], perms=np.stack(perms),
                                 noise_w=NW, noise_U=NU, double_update=double, options={L.OPT_ALGORITHM: alg})
    assert gw.shape == ws.shape == (P["Q"], C, 6) and gU.shape == Us.shape
    tol = 1e-9 if prec == L.PREC_FP64 else 1e-4
    assert relerr(gw, ws) < tol and relerr(gU, Us) < tol
    # the update relative to the start, too (w0 dominates w after six small steps)
    d_ref = ws[:, :, -1] - P["W0"]
    assert relerr(gw[:, :, -1] - P["W0"], d_ref) < (1e-7 if prec == L.PREC_FP64 else 2e-3)


def test_GPTclassification_philox_and_nan():
    """Production draws (Philox, one seed per class): finite chain, factors on the manifold, classes differ; a huge epsU
    trips the geodesic's NaN abort and zero-fills the stores (:631-633)."""
    P, C = _class_problem(C=2, seed=21)
    gw, gU = G.GPTclassification(_feat(P), P["labels"], P["I"], P["r"], P["Q"], 50, 1e-4, 1e-6, 0, 2, 9, precision=L.PREC_TC3)
    assert gw.shape == (P["Q"], C, 6) and np.isfinite(gw).all() and np.isfinite(gU).all()
    for c in range(C):
        Uc = gU[:, :, 1, c, -1]
        assert np.abs(Uc.T @ Uc - np.eye(P["r"])).max() < 1e-9
    assert np.abs(gw[:, 0, -1] - gw[:, 1, -1]).max() > 1e-3
    zw, zU = G.GPTclassification(_feat(P), P["labels"], P["I"], P["r"], P["Q"], 50, 1e-4, 1e300, 0, 1, 9, precision=L.PREC_FP64)
    assert not zw.any() and not zU.any()


def test_nodes_front_prefetch_same_chain():
    """The node-grid step's gather and cell sort run on a side stream under the previous step's tail
    (GPTB_OPT_FRONT_PREFETCH): resident epochs (gptb_sgld_run, gptb_sgld_advance) and pipelined host minibatches
    (gptb_sgld_step_host_submit) give the chain of the one-stream step (FP32 atomics reorder sums: 1e-6, not bits)."""
    P = make_problem(4, 6000, 48, 8, Q=40, seed=17)
    m = 1500
    out = {}
    for pre in (1, 0):
        with _model(P, L.PREC_TC3) as mdl:
            mdl.set_option(L.OPT_ALGORITHM, L.ALG_NODES)
            mdl.set_option(L.OPT_FRONT_PREFETCH, pre)
            mdl.set_params(P["w0"], P["U0"])
            ws, Us = mdl.sgld_run(P["signal_var"], m, 1e-6, 1e-9, 0, 2, True, True, None, None, None, store=True, N=P["N"])
            mdl.sgld_advance(P["signal_var"], m, 1e-6, 1e-9, 7)
            mdl.synchronize()
            wa, Ua = mdl.get_params()
            rng = np.random.default_rng(5)
            keep = []
            for s_ in range(5):
                idx = rng.permutation(P["N"])[:m]
                keep.append((np.asfortranarray(P["X"][idx]), np.ascontiguousarray(P["y"][idx])))
                mdl.sgld_step_host_submit(P["signal_var"], keep[-1][0], keep[-1][1], P["N"], 1e-6, 1e-9)
                if s_ >= 1:
                    mdl.sgld_step_host_wait()
            mdl.sgld_step_host_wait()
            wb, Ub = mdl.get_params()
            out[pre] = (ws, Us, wa, Ua, wb, Ub)
    for a, b in zip(out[1], out[0]):
        assert np.isfinite(a).all() and relerr(a, b) < 1e-6


def test_tgp_gibbs_vs_oracle():
    """GPT_inf (TGP.jl:37-86) through gptb_tgp_gibbs with injected draws: W_array, V_array against the oracle, burn-in
    and store layout; Philox mode stays finite."""
    rng = np.random.default_rng(3)
    n, D, N, r, q, M, burnin = 12, 3, 300, 4, 20, 5, 2
    X = rng.standard_normal((N, D)); y = rng.standard_normal(N)
    Z = rng.standard_normal((n, D)); b01 = rng.random((n, D))
    I = rng.integers(1, r + 1, (q, D)).astype(np.int32)
    U0 = np.sqrt(1 / r) * rng.standard_normal((n, r, D))
    NW = rng.standard_normal((M, q)); NU = rng.standard_normal((M, D, n * r))
    sigma, sigmaRBF = 0.3, 1.4
    Xw = G.datawhitening(np.array(X, order="F")); yw = G.datawhitening(y.reshape(-1, 1).copy()).ravel()
    b = np.stack([o.tgp_feature(Xw[i], Z / sigmaRBF, b01) for i in range(N)], axis=2)
    Wr, Vr = o.tgp_gibbs(b, yw, sigma, r, I - 1, U0, M, burnin, NW, NU)
    W, V, Iout = G.GPT_inf(X, y, sigma, n, r, sigmaRBF, q, 7, M, burnin, Z=Z, b01=b01, I=I, U0=U0, noise_w=NW, noise_U=NU)
    assert W.shape == (q, M - burnin) and V.shape == (n, r, D, M - burnin) and np.array_equal(Iout, I)
    assert relerr(W, Wr) < 1e-8 and relerr(V, Vr) < 1e-8
    W2, V2, _ = G.GPT_inf(X, y, sigma, n, r, sigmaRBF, q, 7, 3, 1)
    assert np.isfinite(W2).all() and np.isfinite(V2).all() and W2.shape == (q, 2)
