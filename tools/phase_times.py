#!/usr/bin/env python
"""Per-phase device time of the SGLD step at a BASELINE shape (CUDA events inside the library).

    python tools/phase_times.py [--workload C4] [--steps 30] [--N 2000000] [--precision tc3,tc1]
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
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--N", type=int, default=2_000_000)
    ap.add_argument("--precision", default="tc3")
    ap.add_argument("--alg", default="auto", choices=["auto", "direct", "nodes"])
    a = ap.parse_args()
    from gpt_b200 import _lib as L
    from gpt_b200 import gpt_sgld as G
    W = B.WORKLOADS[a.workload]
    D, n, r, m = W["D"], W["n"], W["r"], W["m"]
    N = min(a.N, W["N"])
    mdl = B.synth_model(D, n, r, seed=3)
    rng = np.random.default_rng(11)
    X = np.asfortranarray(rng.standard_normal((N, D)))
    y = rng.standard_normal(N)
    feat = G.feature(X, mdl["ls"], 1.0, mdl["phi_scale"], mdl["Z"], mdl["b"])
    EPSW, EPSU = B.eps_for(W["N"])
    for pname in a.precision.split(","):
        prec = {"tc3": L.PREC_TC3, "tc1": L.PREC_TC1, "fp64": L.PREC_FP64}[pname]
        with G.Model(D, n, r, B.Q, precision=prec, device=0, seed=1234) as model:
            model.set_features(feat); model.set_index(mdl["I"]); model.set_train(X, y)
            model.set_params(mdl["w0"], mdl["U0"])
            model.set_option(L.OPT_ALGORITHM, {"auto": L.ALG_AUTO, "direct": L.ALG_DIRECT, "nodes": L.ALG_NODES}[a.alg])
            model.sgld_advance(B.SIGNAL_VAR, m, EPSW, EPSU, 5)
            model.synchronize()
            import time
            t0 = time.perf_counter()
            model.sgld_advance(B.SIGNAL_VAR, m, EPSW, EPSU, a.steps)
            model.synchronize()
            print(a.alg, pname, "unprofiled ms/step %.4f (pipelined when GPTB_PIPE_GROUPS >= 2, default 2)" %
                  ((time.perf_counter() - t0) * 1e3 / a.steps), flush=True)
            model.set_profiling(True)
            model.phase_ms(reset=True)
            model.sgld_advance(B.SIGNAL_VAR, m, EPSW, EPSU, a.steps)
            model.synchronize()
            ms, launches = model.phase_ms(reset=True)
            tot = sum(ms.values()) / a.steps
            print(a.alg, pname, "ms/step %.4f" % tot, {k: round(v / a.steps, 4) for k, v in ms.items()},
                  "launches/step", launches / a.steps, flush=True)


if __name__ == "__main__":
    main()
