"""Generates the committed golden fixtures from the reference's own data files.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

Outputs (tests/golden/*.npz, Julia-shaped arrays):
  tensorsynth5d.npz / tensorsynth10d.npz
      Everything MakeSynthData.jl:17-27 wrote (X w U I phi length_scale y1 y2 y3) plus the
      random-feature draws Z, b that MakeSynthData.jl did NOT save: every phi[j,k,:] is, to
      1e-15, phi_scale*sqrt(2/n)*cos(X[:,k]*omega + beta) (GPT_SGLD.jl:79,83), so
      omega = Z[j,k]/length_scale[k] and beta = b[j,k] mod 2pi are recovered by a nonlinear fit
      and stored as Z_rec = omega*length_scale, b_rec.  feature(X, length_scale, 1, phi_scale,
      Z_rec, b_rec) then reproduces the reference's own phi: a known-answer test for `feature`.
  synth_fd.npz
      First 200 whitened training rows of 10000SynthData.h5 — the input of
      Diagnostic_gradients.jl (finite-difference gradient check, n=10,r=5,Q=10,sigma=0.2).
  rmse_curves.npz
      testRMSE_kin40k.h5 / testRMSE_PP.h5: 200-epoch test-RMSE trajectories (statistical targets).
  powerplant.npz
      PPdata_full.h5 (whitened CCPP inputs, raw targets) for trajectory runs on real data.
"""
import os
import sys

import numpy as np
from scipy.optimize import least_squares

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import h5lite  # noqa: E402

REF = "/root/reference"


def recover_Zb(X, phi, length_scale, amp):
    n, D, N = phi.shape
    Z = np.empty((n, D))
    b = np.empty((n, D))
    worst = 0.0
    grid = np.linspace(-8, 8, 641)
    for k in range(D):
        x = X[:, k]
        for j in range(n):
            t = phi[j, k, :]
            best = (np.inf, 0.0, 0.0)
            for om in grid:
                M = np.stack([np.cos(om * x), np.sin(om * x)], 1)
                co = np.linalg.lstsq(M, t, rcond=None)[0]
                rr = np.linalg.norm(M @ co - t)
                if rr < best[0]:
                    best = (rr, om, np.arctan2(-co[1], co[0]))
            s = least_squares(lambda p: amp * np.cos(x * p[0] + p[1]) - t, [best[1], best[2]],
                              xtol=1e-15, ftol=1e-15, gtol=1e-15)
            om, be = s.x
            worst = max(worst, np.abs(s.fun).max())
            if om < 0:  # cos is even: (om,be) ~ (-om,-be); Z's sign is not identifiable, keep om>=0
                om, be = -om, -be
            Z[j, k] = om * length_scale[k]
            b[j, k] = np.mod(be, 2 * np.pi)
    return Z, b, worst


def main():
    for name, out in [("TensorSynthData5D100N.h5", "tensorsynth5d.npz"),
                      ("TensorSynthData10D100N.h5", "tensorsynth10d.npz")]:
        d = h5lite.read_all(os.path.join(REF, name))
        n, D, N = d["phi"].shape
        Q = d["w"].size
        phi_scale = np.sqrt(n / Q ** (1.0 / D))       # GPT_SGLD.jl:328
        amp = phi_scale * np.sqrt(2.0 / n)
        Z, b, worst = recover_Zb(d["X"], d["phi"], d["length_scale"], amp)
        print(name, "worst cosine-fit residual", worst)
        assert worst < 1e-12
        np.savez_compressed(os.path.join(HERE, out), Z_rec=Z, b_rec=b, phi_scale=phi_scale, **d)
    d = h5lite.read_all(os.path.join(REF, "10000SynthData.h5"))
    np.savez_compressed(os.path.join(HERE, "synth_fd.npz"), Xtrain=d["Xtrain"][:200], ytrain=d["ytrain"][:200])
    k = h5lite.read_all(os.path.join(REF, "testRMSE_kin40k.h5"))
    p = h5lite.read_all(os.path.join(REF, "testRMSE_PP.h5"))
    np.savez_compressed(os.path.join(HERE, "rmse_curves.npz"), kin40k=k["testRMSE"], pp_sgld=p["testRMSE"],
                        pp_rmsprop=p["testRMSE2"])
    pp = h5lite.read_all(os.path.join(REF, "PPdata_full.h5"))
    np.savez_compressed(os.path.join(HERE, "powerplant.npz"), Xtrain=pp["Xtrain"].astype(np.float32),
                        ytrain=pp["ytrain"].astype(np.float32))


if __name__ == "__main__":
    main()
