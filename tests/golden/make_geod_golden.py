#!/usr/bin/env python
"""Generates tests/golden/geod_mp.npz: inputs (U, mom, t) and 40-digit mpmath results of geod (GPT_SGLD.jl:19-37) and
of the 2r x 2r exponential inside it (:23), at r in {2, 10, 32}, t in {1e-7, 1e-3, 0.3}.  The arithmetic is
oracle/gpt_oracle.py: geod_mp / expm_mp (mpmath only: independent of SciPy, LAPACK and any double-precision expm).

    python tests/golden/make_geod_golden.py        # ~1 minute
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import gpt_oracle as o  # noqa: E402

rng = np.random.default_rng(20260101)
out = {}
case = 0
for r, n in ((2, 9), (10, 50), (32, 200)):
    U = o.stiefel_from_gaussian(rng.standard_normal((r, n)))
    M = o.proj(U, rng.standard_normal((n, r)))
    A = U.T @ M
    blk = np.block([[A, -M.T @ M], [np.eye(r), A]])
    for t in (1e-7, 1e-3, 0.3):
        out[f"U{case}"], out[f"M{case}"], out[f"t{case}"] = U, M, np.float64(t)
        out[f"geod{case}"] = o.geod_mp(U, M, t)
        out[f"expm{case}"] = o.expm_mp(t * blk)
        case += 1
        print("case", case, r, n, t, flush=True)
out["ncases"] = np.int64(case)
np.savez_compressed(os.path.join(HERE, "geod_mp.npz"), **out)
