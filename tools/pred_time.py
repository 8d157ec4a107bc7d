#!/usr/bin/env python
"""Timing of the C5 posterior-predictive sweep (gptb_pred_samples), repeated calls.
    python tools/pred_time.py [S] [Nt] [auto|direct|nodes] [chunk MB] [layout 0|1] [warps]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B
from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G
C = B.C5
S = int(sys.argv[1]) if len(sys.argv) > 1 else 200
Nt = int(sys.argv[2]) if len(sys.argv) > 2 else C["Nt"]
pm = B.synth_model(C["D"], C["n"], C["r"], seed=5)
rng = np.random.default_rng(13)
Xt = np.asfortranarray(rng.standard_normal((Nt, C["D"])))
ws = np.asfortranarray(pm["w_true"][:, None] + 0.05 * rng.standard_normal((B.Q, S)))
Us = np.asfortranarray(np.repeat(pm["U_true"][:, :, :, None], S, axis=3))
pf = G.feature(Xt, pm["ls"], 1.0, pm["phi_scale"], pm["Z"], pm["b"])
ALG = {"auto": L.ALG_AUTO, "direct": L.ALG_DIRECT, "nodes": L.ALG_NODES}[sys.argv[3] if len(sys.argv) > 3 else "auto"]
with G.Model(C["D"], C["n"], C["r"], B.Q, precision=L.PREC_TC3, device=0) as m:
    m.set_features(pf); m.set_index(pm["I"])
    m.set_option(L.OPT_ALGORITHM, ALG)
    if len(sys.argv) > 4: m.set_option(L.OPT_PRED_CHUNK_MB, float(sys.argv[4]))
    if len(sys.argv) > 5: m.set_option(L.OPT_PRED_LAYOUT, float(sys.argv[5]))      # 0 lanes, 1 tiles
    if len(sys.argv) > 6: m.set_option(L.OPT_PRED_WARPS, float(sys.argv[6]))
    for it in range(3):
        t0 = time.perf_counter()
        mean = m.pred_samples(ws, Us, Xt)
        dt = time.perf_counter() - t0
        print(f"call {it}: {dt*1e3:.1f} ms  {Nt*S/dt/1e6:.1f} M pred/s  checksum {np.abs(mean).sum():.4f}", flush=True)
