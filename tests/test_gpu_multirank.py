"""Two ranks on two real GPUs through libgptb200's own communicator entry (gptb_comm_init): the row-sharded chain must
equal the one-rank chain (same global minibatch) and every rank must hold bit-identical parameters.
Skipped on a box with fewer than two GPUs (run it with `gpurun --gpus 2`; the log is kept under profiles/)."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from util import relerr

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _run(world, tmp):
    idfile = os.path.join(tmp, f"ncclid_w{world}")
    outs = [os.path.join(tmp, f"out_w{world}_r{r}.npz") for r in range(world)]
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "_mr_worker.py"), str(r), str(world), idfile, outs[r]])
             for r in range(world)]
    for p in procs:
        assert p.wait(timeout=600) == 0
    return [dict(np.load(o)) for o in outs]


@pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs")
def test_two_rank_chain_equals_one_rank_chain():
    from _mr_worker import problem
    P = problem()
    with tempfile.TemporaryDirectory() as tmp:
        ref = _run(1, tmp)[0]
        got = _run(2, tmp)
    w0, U0 = P["w0"], P["U0"]
    tol = 2e-5   # of the accumulated update: FP32 atomics / a different summation split, not a different chain
    for key in ("host", "adv", "s1", "pipe", "nodes", "nodes_host"):
        assert relerr(got[0][f"{key}_w"] - w0, ref[f"{key}_w"] - w0) < tol, key
        assert relerr(got[0][f"{key}_U"] - U0, ref[f"{key}_U"] - U0) < tol, key
    assert relerr(got[0]["host_wout"] - w0, ref["host_wout"] - w0) < tol
    for key in ("g1", "g333"):
        assert relerr(got[0][f"{key}_w"], ref[f"{key}_w"]) < 1e-5 and relerr(got[0][f"{key}_U"], ref[f"{key}_U"]) < 1e-5, key
    assert relerr(got[0]["nodes_gw"], ref["nodes_gw"]) < 1e-5 and relerr(got[0]["nodes_gU"], ref["nodes_gU"]) < 1e-5
    # no broadcast ever happens: the ranks stay together only if every reduced buffer and update is bit-identical
    for key in got[0]:
        assert np.array_equal(got[0][key], got[1][key]), f"{key} differs between ranks"


@pytest.mark.gpu
@pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs")
def test_single_process_two_gpus_through_the_c_abi():
    """gptb_comm_init_local + gptb_multi_*: ONE process drives two GPUs (what the drop-in GPTregression call of a
    single Julia process needs).  The row-sharded chain equals the one-GPU chain, both handles end with
    bit-identical parameters, and the posterior sweep with its test rows sharded equals the one-GPU sweep."""
    import numpy as np
    from gpt_b200 import _lib as L
    from gpt_b200 import gpt_sgld as G
    from _mr_worker import problem
    P = problem()
    f = G.feature(P["X"], P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"])
    N, m, burnin, maxepoch = P["N"], 1001, 1, 2
    rng = np.random.default_rng(5)
    perms = np.stack([rng.permutation(N) for _ in range(burnin + maxepoch)])
    steps = (burnin + maxepoch) * 3
    NW = rng.standard_normal((steps, P["Q"])); NU = rng.standard_normal((steps, P["n"], P["r"], P["D"]))
    kw = dict(precision=L.PREC_TC3, w0=P["w0"], U0=P["U0"], perms=perms, noise_w=NW, noise_U=NU)
    for alg in (L.ALG_DIRECT, L.ALG_NODES):
        one = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-6, 1e-9, burnin, maxepoch, 1,
                              options={L.OPT_ALGORITHM: alg}, **kw)
        two = G.GPTregression(f, P["y"], P["signal_var"], P["I"], P["r"], P["Q"], m, 1e-6, 1e-9, burnin, maxepoch, 1,
                              options={L.OPT_ALGORITHM: alg}, devices=[0, 1], **kw)
        assert relerr(two[0] - P["w0"][:, None], one[0] - P["w0"][:, None]) < 2e-5, alg
        assert relerr(two[1] - P["U0"][..., None], one[1] - P["U0"][..., None]) < 2e-5, alg
    with G.ModelGroup([0, 1], P["D"], P["n"], P["r"], P["Q"], precision=L.PREC_TC3, seed=99) as grp:
        grp.set_features(f); grp.set_index(P["I"]); grp.set_train(P["X"], P["y"]); grp.set_params(P["w0"], P["U0"])
        grp.sgld_advance(P["signal_var"], 700, 1e-6, 1e-9, 6)
        grp.synchronize()
        a, b = grp.models[0].get_params(), grp.models[1].get_params()
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert relerr(a[0], P["w0"]) > 1e-7
    S = 6
    ws = np.asfortranarray(one[0][:, -S:]); Us = np.asfortranarray(one[1][..., -S:])
    m1, p1 = G.pred_samples(ws, Us, P["I"], f, per_sample=True, precision=L.PREC_TC3)
    m2, p2 = G.pred_samples(ws, Us, P["I"], f, per_sample=True, precision=L.PREC_TC3, devices=[0, 1])
    assert relerr(m2, m1) < 1e-5 and relerr(p2, p1) < 1e-5
