#!/usr/bin/env python
"""Full-theta RFF SGLD (GPNT_SGLD, GPT_SGLD.jl:809-847) at the BASELINE config-3 shape (tensorsynth 10-D, N=32000,
minibatch 1000): steps/s on the GPU (device time of the step loop, gptb_nt_last_run_stats) beside the NumPy
restatement on the host cores, plus the parity of the two chains with shared draws.

    python tools/nt_time.py [--n 500,4000,16000] [--epochs 2]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import gpt_oracle as o  # noqa: E402  (checker / CPU baseline only)
from gpt_b200 import _lib as L  # noqa: E402
from gpt_b200 import gpt_sgld as G  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", default="500,4000,16000")
    ap.add_argument("--epochs", type=int, default=2)
    a = ap.parse_args()
    N, D, m = 32000, 10, 1000
    rng = np.random.default_rng(3)
    X = np.asfortranarray(rng.standard_normal((N, D)))
    ls = np.maximum(1.0 + 0.2 * rng.standard_normal(D), 0.3)
    for n in [int(v) for v in a.n.split(",")]:
        Z = np.asfortranarray(rng.standard_normal((n, D))); b = 2 * np.pi * rng.random(n)
        f = G.featureNotensor(X, ls, 1.0, Z, b)
        th_true = rng.standard_normal(n)
        with G.NTModel(f, precision=L.PREC_TC3) as mdl:
            y = mdl.pred(th_true, X) + np.sqrt(1e-3) * rng.standard_normal(N)
        eps = 1e-5 * 50.0 / N
        out = {"workload": f"C3 full-theta: D={D} N={N} n={n} minibatch={m}", "epochs": a.epochs}
        for name, prec in (("fp32_mufu", L.PREC_TC3), ("fp64", L.PREC_FP64)):
            with G.NTModel(f, precision=prec, seed=1) as mdl:
                mdl.set_train(X, y)
                mdl.run(1e-3, 1.0, m, eps, 0.0, 0, 1, 5)                     # warm-up epoch
                mdl.run(1e-3, 1.0, m, eps, 0.0, 0, a.epochs, 5)
                ms, launches, steps = mdl.last_run_stats()
            out[name] = {"steps_per_s": steps / (ms * 1e-3), "us_per_step": ms * 1e3 / steps, "launches_per_step": launches / steps,
                         "cos_per_s": 2.0 * m * n * steps / (ms * 1e-3)}
        # CPU restatement, features precomputed as the reference does; bounded sample of steps
        phi = o.featureNotensor(X[:4000], ls, 1.0, Z, b)
        th = np.zeros(n)
        ks = 8
        t0 = time.perf_counter()
        for s in range(ks):
            lo = (s % 4) * m
            th = th + eps * o.nt_grad(phi[:, lo:lo + m], y[lo:lo + m], th, N, 1e-3, 1.0) / 2
        out["cpu_numpy"] = {"steps_per_s": ks / (time.perf_counter() - t0), "sample": f"{ks} steps, features precomputed"}
        # parity of one epoch with shared draws
        sub = 3000
        fs = f.with_X(X[:sub])
        perms = np.stack([rng.permutation(sub)]); noise = rng.standard_normal((3, n)); th0 = rng.standard_normal(n)
        want = o.GPNT_SGLD(o.featureNotensor(X[:sub], ls, 1.0, Z, b), y[:sub], 1e-3, 1.0, m, eps, 0.1, 0, 1, th0, perms, noise)
        for name, prec in (("fp32_mufu", L.PREC_TC3), ("fp64", L.PREC_FP64)):
            got = G.GPNT_SGLD(fs, y[:sub], 1e-3, 1.0, m, eps, 0.1, 0, 1, 5, precision=prec, theta0=th0, perms=perms, noise=noise)
            out[name]["chain_rel_err"] = float(np.linalg.norm((got - want)) / np.linalg.norm(want - th0[:, None]))
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
