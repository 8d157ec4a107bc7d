import sys, os
sys.path[:0]=['/root/repo','/root/repo/oracle','/root/repo/tests']
import numpy as np
import gpt_oracle as o
from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G
from util import make_problem, relerr
P = make_problem(4, 600, 50, 10, Q=60, seed=0)
f = G.feature(P["X"], P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"])
idx = np.arange(512)
ref = o.gradients(P["phi_fn"](P["X"][idx]), P["y"][idx], P["w0"], P["U0"], P["I"], P["N"], P["signal_var"])
for prec in (L.PREC_TC1, L.PREC_TC3):
    with G.Model(P["D"], P["n"], P["r"], P["Q"], precision=prec, device=0) as m:
        m.set_features(f); m.set_index(P["I"]); m.set_train(P["X"], P["y"]); m.set_params(P["w0"], P["U0"])
        gw, gU, fh = m.grad(P["signal_var"], idx=idx)
        print(prec, relerr(gw, ref[0]), relerr(gU, ref[1]), relerr(fh, ref[2]))
        print(np.abs(gU).max(), np.abs(ref[1]).max())
        r = gU[:, :, 0] / ref[1][:, :, 0]
        print(np.round(r[:6, :6], 3))
