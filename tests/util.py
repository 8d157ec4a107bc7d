"""Seeded synthetic problems in the reference's conventions (MakeSynthData.jl:6-13)."""
import os

import numpy as np

import gpt_oracle as o

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# BASELINE.json configs: (D, N, n, r, m[, Nt, S])
CONFIGS = {
    "C1": dict(D=4, N=9568, n=50, r=10, m=500),
    "C2": dict(D=8, N=10000, n=100, r=20, m=1000, Nt=30000),
    "C3": dict(D=10, N=32000, n=500, r=20, m=1000),
    "C4": dict(D=10, N=10_000_000, n=1000, r=32, m=65536),
    "C5": dict(D=8, Nt=1_000_000, S=200, n=500, r=20),
}


def make_problem(D, N, n, r, Q=200, seed=0, signal_var=1e-3, feature2=False, scalar_ls=False):
    rng = np.random.default_rng(seed)
    Q = min(Q, r ** D)
    X = rng.standard_normal((N, D))
    ls = np.maximum(1.0 + 0.2 * rng.standard_normal(D), 0.3)      # MakeSynthData.jl:7
    if scalar_ls:
        ls = ls[:1]
    Z = rng.standard_normal((n, D))
    b = 2 * np.pi * rng.random((n, D))
    phi_scale = np.sqrt(n / Q ** (1.0 / D))                       # GPT_SGLD.jl:306
    I = o.samplenz(r, D, Q, rng)
    w_true = rng.standard_normal(Q)
    U_true = np.stack([o.stiefel_from_gaussian(rng.standard_normal((r, n))) for _ in range(D)], 2)
    w0 = rng.standard_normal(Q)
    U0 = np.stack([o.stiefel_from_gaussian(rng.standard_normal((r, n))) for _ in range(D)], 2)
    if feature2:
        phi_fn = lambda XX: o.feature2(XX, ls, 1.0, phi_scale, Z)
    else:
        phi_fn = lambda XX: o.feature(XX, ls, 1.0, phi_scale, Z, b)
    # y = fhat*(X) + sqrt(signal_var) noise, row-chunked so phi is never large
    y = np.empty(N)
    for i0 in range(0, N, 4096):
        y[i0:i0 + 4096] = o.pred(w_true, U_true, I, phi_fn(X[i0:i0 + 4096]))
    y += np.sqrt(signal_var) * rng.standard_normal(N)
    return dict(D=D, N=N, n=n, r=r, Q=Q, X=np.asfortranarray(X), y=y, ls=ls, Z=np.asfortranarray(Z),
                b=np.asfortranarray(b), phi_scale=phi_scale, sigma_RBF=1.0, I=np.asfortranarray(I), w_true=w_true,
                U_true=np.asfortranarray(U_true), w0=w0, U0=np.asfortranarray(U0), signal_var=signal_var,
                phi_fn=phi_fn, feature2=feature2, rng=rng)


def relerr(a, b):
    """norm-wise relative error |a-b|/|b| (SURVEY.md §8c tolerance definition)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = np.linalg.norm(b.ravel())
    return np.linalg.norm((a - b).ravel()) / (den if den > 0 else 1.0)


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name)))
