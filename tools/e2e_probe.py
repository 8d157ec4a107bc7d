#!/usr/bin/env python
"""Host-side cost of the pipelined host-minibatch entries at C4: time spent inside submit() and wait() per step."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench as B
from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G
W = B.WORKLOADS["C4"]; D, n, r, m = W["D"], W["n"], W["r"], W["m"]; N = 2_000_000
mdl = B.synth_model(D, n, r, seed=3); rng = np.random.default_rng(11)
X = np.asfortranarray(rng.standard_normal((N, D))); y = rng.standard_normal(N)
feat = G.feature(X, mdl["ls"], 1.0, mdl["phi_scale"], mdl["Z"], mdl["b"])
EPSW, EPSU = B.eps_for(W["N"])
inflight = int(sys.argv[1]) if len(sys.argv) > 1 else 1
with G.Model(D, n, r, B.Q, precision=L.PREC_TC3, device=0, seed=1) as model:
    model.set_features(feat); model.set_index(mdl["I"]); model.set_train(X, y); model.set_params(mdl["w0"], mdl["U0"])
    for kv in sys.argv[2:]:
        k_, v_ = kv.split("="); model.set_option(int(k_), float(v_))
    Xb = [torch.empty((D, m), dtype=torch.float64).pin_memory() for _ in range(3)]
    yb = [torch.empty(m, dtype=torch.float64).pin_memory() for _ in range(3)]
    for s in range(3):
        rows = rng.choice(N, size=m, replace=False)
        Xb[s].numpy()[...] = X[rows].T; yb[s].numpy()[...] = y[rows]
    xb = [np.asfortranarray(Xb[s].numpy().T) for s in range(3)]; ybn = [t.numpy() for t in yb]
    def loop(k, rec):
        model.sgld_step_host_submit(B.SIGNAL_VAR, xb[0], ybn[0], N, EPSW, EPSU)
        for s in range(1, k):
            t0 = time.perf_counter()
            model.sgld_step_host_submit(B.SIGNAL_VAR, xb[s % 2], ybn[s % 2], N, EPSW, EPSU)
            t1 = time.perf_counter()
            model.sgld_step_host_wait()
            t2 = time.perf_counter()
            rec.append((t1 - t0, t2 - t1))
        model.sgld_step_host_wait()
    loop(5, [])
    torch.cuda.synchronize()
    rec = []
    t0 = time.perf_counter(); loop(200, rec); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    a = np.array(rec) * 1e6
    print(f"e2e {dt/200*1e6:.1f} us/step; submit() mean {a[:,0].mean():.1f} us (p90 {np.percentile(a[:,0],90):.1f}), wait() mean {a[:,1].mean():.1f} us")
    t0 = time.perf_counter(); model.sgld_advance(B.SIGNAL_VAR, m, EPSW, EPSU, 200); t1 = time.perf_counter(); model.synchronize(); t2 = time.perf_counter()
    print(f"resident: enqueue {(t1-t0)/200*1e6:.1f} us/step, total {(t2-t0)/200*1e6:.1f} us/step")
