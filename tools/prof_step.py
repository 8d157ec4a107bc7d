#!/usr/bin/env python
"""Short profiling target: a few SGLD steps (and optionally a small prediction sweep) at a BASELINE shape, with
random y and a reduced N so set-up is quick.  Used under ncu (see profiles/README.md); never a bench number.

    python tools/prof_step.py [--workload C4] [--steps 2] [--warmup 3] [--N 1000000] [--pred]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C4")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--N", type=int, default=1_000_000)
    ap.add_argument("--precision", default="tc3")
    ap.add_argument("--alg", default="auto", choices=["auto", "direct", "nodes"])
    ap.add_argument("--pred", action="store_true")
    a = ap.parse_args()
    from gpt_b200 import _lib as L
    from gpt_b200 import gpt_sgld as G
    prec = {"tc3": L.PREC_TC3, "tc1": L.PREC_TC1, "fp64": L.PREC_FP64}[a.precision]
    if a.pred:
        C = B.C5
        pm = B.synth_model(C["D"], C["n"], C["r"], seed=5)
        rng = np.random.default_rng(13)
        S, Nt = 8, 65536
        Xt = np.asfortranarray(rng.standard_normal((Nt, C["D"])))
        ws = np.asfortranarray(pm["w_true"][:, None] + 0.05 * rng.standard_normal((B.Q, S)))
        Us = np.asfortranarray(np.repeat(pm["U_true"][:, :, :, None], S, axis=3))
        pf = G.feature(Xt, pm["ls"], 1.0, pm["phi_scale"], pm["Z"], pm["b"])
        with G.Model(C["D"], C["n"], C["r"], B.Q, precision=prec, device=0) as pmodel:
            pmodel.set_features(pf); pmodel.set_index(pm["I"])
            for _ in range(a.warmup + a.steps):
                pmodel.pred_samples(ws, Us, Xt)
        print("prof_step pred done")
        return
    W = B.WORKLOADS[a.workload]
    D, n, r, m = W["D"], W["n"], W["r"], W["m"]
    N = min(a.N, W["N"])
    mdl = B.synth_model(D, n, r, seed=3)
    rng = np.random.default_rng(11)
    X = np.asfortranarray(rng.standard_normal((N, D)))
    y = rng.standard_normal(N)
    feat = G.feature(X, mdl["ls"], 1.0, mdl["phi_scale"], mdl["Z"], mdl["b"])
    EPSW, EPSU = B.eps_for(W["N"])
    with G.Model(D, n, r, B.Q, precision=prec, device=0, seed=1234) as model:
        model.set_features(feat); model.set_index(mdl["I"]); model.set_train(X, y); model.set_params(mdl["w0"], mdl["U0"])
        model.set_option(L.OPT_ALGORITHM, {"auto": L.ALG_AUTO, "direct": L.ALG_DIRECT, "nodes": L.ALG_NODES}[a.alg])
        model.sgld_advance(B.SIGNAL_VAR, m, EPSW, EPSU, a.warmup)
        model.synchronize()
        model.sgld_advance(B.SIGNAL_VAR, m, EPSW, EPSU, a.steps)
        model.synchronize()
    print("prof_step done")


if __name__ == "__main__":
    main()
