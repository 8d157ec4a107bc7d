#!/usr/bin/env python
"""A reduced posterior sweep at the C5 geometry for ncu: 64 samples (two lane groups) x 120k rows."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B
from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G
C = B.C5
S, Nt = 64, 120000
pm = B.synth_model(C["D"], C["n"], C["r"], seed=5)
rng = np.random.default_rng(13)
Xt = np.asfortranarray(rng.standard_normal((Nt, C["D"])))
ws = np.asfortranarray(pm["w_true"][:, None] + 0.05 * rng.standard_normal((B.Q, S)))
Us = np.asfortranarray(np.repeat(pm["U_true"][:, :, :, None], S, axis=3))
pf = G.feature(Xt, pm["ls"], 1.0, pm["phi_scale"], pm["Z"], pm["b"])
with G.Model(C["D"], C["n"], C["r"], B.Q, precision=L.PREC_TC3, device=0) as m:
    m.set_features(pf); m.set_index(pm["I"]); m.set_option(L.OPT_ALGORITHM, L.ALG_NODES)
    mean = m.pred_samples(ws, Us, Xt)
    print("checksum", float(np.abs(mean).sum()))
