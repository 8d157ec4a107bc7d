"""Parity at the sizes bench.py measures (BASELINE.json configs 4 and 5) — needs a B200.

The reduced-N cases of test_gpu_parity.py never make a persistent K1/K3 CTA process a second unit, reload a U image on
a change of dimension, or run K3's feature-block group loop over more than one row chunk; and the posterior sweep there
has a single sample group.  Here the CUDA path runs the real minibatch (m = 65536 at D=10, n=1000, r=32) and the real
sample count (S = 200 at D=8, n=500, r=20) and is checked against the oracle evaluated in row chunks
(oracle.gradients_chunked; GPT_SGLD.jl:384-408, :233-243)."""
import numpy as np
import pytest

import gpt_oracle as o
from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G
from util import make_problem, relerr

pytestmark = pytest.mark.gpu

TOL = {L.PREC_FP64: 1e-10, L.PREC_TC3: 1e-4}


@pytest.fixture(scope="module")
def c4():
    """C4 at its real minibatch: N = m = 65536 rows (the resident training set of the bench has 10M rows; the step
    only ever sees the minibatch), D=10, n=1000, r=32, Q=200, and the oracle's chunked gradients at (w0, U0)."""
    P = make_problem(10, 65536, 1000, 32, Q=200, seed=44)
    P["ref"] = o.gradients_chunked(P["phi_fn"], P["X"], P["y"], P["w0"], P["U0"], P["I"], P["N"], P["signal_var"])
    return P


def _model(P, prec, alg=None):
    m = G.Model(P["D"], P["n"], P["r"], P["Q"], precision=prec, seed=5)
    m.set_features(G.feature(P["X"], P["ls"], P["sigma_RBF"], P["phi_scale"], P["Z"], P["b"]))
    m.set_index(P["I"])
    m.set_train(P["X"], P["y"])
    m.set_params(P["w0"], P["U0"])
    if alg is not None:
        m.set_option(L.OPT_ALGORITHM, alg)
    return m


@pytest.mark.parametrize("prec,alg", [(L.PREC_FP64, L.ALG_DIRECT), (L.PREC_TC3, L.ALG_DIRECT), (L.PREC_TC3, L.ALG_NODES)],
                         ids=["fp64", "tc3-direct", "tc3-nodes"])
def test_c4_minibatch_65536_grad_and_step(c4, prec, alg):
    P = c4
    gw_ref, gU_ref, fh_ref = P["ref"]
    rng = np.random.default_rng(3)
    nw = rng.standard_normal(P["Q"]); nU = np.asfortranarray(rng.standard_normal((P["n"], P["r"], P["D"])))
    epsw, epsU = 1e-5 * 50 / P["N"], 1e-8 * 50 / P["N"]
    w_ref, U_ref, bad = o.sgld_update(P["w0"], P["U0"], gw_ref, gU_ref, epsw, epsU, nw, nU)
    assert not bad
    with _model(P, prec, alg) as m:
        gw, gU, fh = m.grad(P["signal_var"], start=0, bs=P["N"])
        assert relerr(fh, fh_ref) < TOL[prec]
        assert relerr(gw, gw_ref) < TOL[prec]
        assert relerr(gU, gU_ref) < TOL[prec]
        for k in range(P["D"]):       # per dimension too: a wrong U-image reload would hide in the global norm
            assert relerr(gU[:, :, k], gU_ref[:, :, k]) < TOL[prec], k
        m.sgld_step(P["signal_var"], epsw, epsU, start=0, bs=P["N"], noise_w=nw, noise_U=nU)
        w, U = m.get_params()
    assert relerr(w - P["w0"], w_ref - P["w0"]) < max(TOL[prec], 1e-9)
    assert relerr(U - P["U0"], U_ref - P["U0"]) < max(TOL[prec], 1e-7)
    for k in range(P["D"]):
        assert np.abs(U[:, :, k].T @ U[:, :, k] - np.eye(P["r"])).max() < 1e-10


@pytest.mark.parametrize("alg", [L.ALG_DIRECT, L.ALG_NODES], ids=["direct", "nodes"])
def test_c4_escale_guard_residual_jumps(c4, alg):
    """The fp16 scale of the E image comes from the previous step's max|E|.  Consecutive minibatches whose residuals
    differ by x1000 and /1000 (y of the batch rescaled) must neither overflow the image (Inf -> NaN gradient -> a
    spurious geodesic-NaN abort) nor lose its low parts: every step's gradient stays within 1e-4 of the oracle's."""
    P = c4
    bs = 8192
    with _model(P, L.PREC_TC3, alg) as m:
        for step, fac in enumerate((1.0, 1000.0, 1e-3, 1e3)):
            y2 = P["y"].copy()
            lo = step * bs
            y2[lo:lo + bs] = fac * P["y"][lo:lo + bs] + (fac - 1.0) * 3.0
            m.set_train(P["X"], y2)
            if step == 0:
                m.set_params(P["w0"], P["U0"])
            want = o.gradients_chunked(P["phi_fn"], P["X"][lo:lo + bs], y2[lo:lo + bs], P["w0"], P["U0"], P["I"], P["N"],
                                       P["signal_var"])
            # the gradient call inherits the image scale of the PREVIOUS minibatch (x1000 or /1000 off) ...
            gw, gU, fh = m.grad(P["signal_var"], start=lo, bs=bs)
            assert np.isfinite(gU).all()
            assert relerr(gw, want[0]) < 1e-4 and relerr(gU, want[1]) < 1e-4, (step, fac)
            # ... and so does a sampler step (zero step sizes: a non-finite gradient would still raise the NaN abort)
            m.sgld_step(P["signal_var"], 0.0, 0.0, start=lo, bs=bs, langevin=False)
            w, U = m.get_params()
            assert np.isfinite(w).all() and np.isfinite(U).all()


@pytest.mark.parametrize("prec,alg", [(L.PREC_TC3, L.ALG_DIRECT), (L.PREC_TC3, L.ALG_NODES), (L.PREC_FP64, L.ALG_DIRECT)],
                         ids=["tc3-direct", "tc3-nodes", "fp64"])
def test_c5_pred_samples_200(prec, alg):
    """C5 geometry: D=8, n=500, r=20, Q=200, S=200 stored samples (50 sample groups on the direct tensor-core path, 25
    column groups of 8 samples on the node grid), Nt = 40000 test rows (several row chunks on both paths); mean and
    per-sample predictions against oracle.pred on a row subset."""
    D, n, r, Q, S, Nt = 8, 500, 20, 200, 200, 40000
    if prec == L.PREC_FP64:
        S, Nt = 24, 6000                       # the FP64 path is the per-sample validation loop: keep it short
    P = make_problem(D, 256, n, r, Q=Q, seed=55)
    rng = np.random.default_rng(9)
    Xt = np.asfortranarray(rng.standard_normal((Nt, D)))
    ws = np.asfortranarray(P["w_true"][:, None] + 0.05 * rng.standard_normal((Q, S)))
    Us = np.empty((n, r, D, S), order="F")
    for s in range(S):
        for k in range(D):
            Us[:, :, k, s] = np.linalg.qr(P["U_true"][:, :, k] + 0.01 * rng.standard_normal((n, r)))[0]
    f = G.feature(Xt, P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"])
    mean, per = G.pred_samples(ws, Us, P["I"], f, per_sample=True, precision=prec, options={L.OPT_ALGORITHM: alg})
    assert mean.shape == (Nt,) and per.shape == (Nt, S)
    sub = np.arange(0, Nt, 20)                 # 2000 rows spread over all row chunks
    phi = P["phi_fn"](Xt[sub])
    want = np.stack([o.pred(ws[:, s], Us[:, :, :, s], P["I"], phi) for s in range(S)], 1)
    assert relerr(per[sub], want) < TOL[prec]
    for s in (0, 3, 4, S - 1):                 # first/last member of a sample group, last group
        assert relerr(per[sub, s], want[:, s]) < TOL[prec], s
    assert relerr(mean[sub], want.mean(1)) < TOL[prec]
    assert relerr(per.mean(1), mean) < 1e-12 if prec == L.PREC_FP64 else relerr(per.mean(1), mean) < 1e-6
