"""ctypes binding of oracle/_build/libgptoracle.so (the plain-C restatement, gpt_oracle.c).

TEST INFRASTRUCTURE ONLY — same rule as gpt_oracle.py.  Arrays are Julia-shaped, Fortran order.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libgptoracle.so")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "gpt_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "_build/libgptoracle.so"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.gpo_max_threads.restype = C.c_int
    return _lib


def _d(a):
    a = np.asfortranarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


def _i(a):
    a = np.asfortranarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(C.POINTER(C.c_int))


def max_threads() -> int:
    return int(lib().gpo_max_threads())


def feature(X, length_scale, sigma_RBF, phi_scale, Z, b, nthreads=1):
    X, pX = _d(X); Z, pZ = _d(Z); b, pb = _d(b)
    ls, pls = _d(np.atleast_1d(length_scale))
    N, D = X.shape; n = Z.shape[0]
    phi = np.empty((n, D, N), order="F"); _, pphi = _d(phi)
    lib().gpo_feature(pX, N, D, pls, ls.size, C.c_double(sigma_RBF), C.c_double(phi_scale), pZ, pb, n, pphi, nthreads)
    return phi


def feature2(X, length_scale, sigma_RBF, phi_scale, Z, nthreads=1):
    X, pX = _d(X); Z, pZ = _d(Z)
    ls, pls = _d(np.atleast_1d(length_scale))
    N, D = X.shape; n = Z.shape[0]
    phi = np.empty((n, D, N), order="F"); _, pphi = _d(phi)
    rc = lib().gpo_feature2(pX, N, D, pls, ls.size, C.c_double(sigma_RBF), C.c_double(phi_scale), pZ, n, pphi, nthreads)
    if rc:
        raise ValueError("n not even")
    return phi


def pred(w, U, I, phi, nthreads=1):
    w, pw = _d(w); U, pU = _d(U); I, pI = _i(I); phi, pphi = _d(phi)
    n, r, D = U.shape; Q = w.size; N = phi.shape[2]
    out = np.empty(N); _, po = _d(out)
    lib().gpo_pred(pw, pU, pI, pphi, n, r, D, Q, N, po, nthreads)
    return out


def gradients(phi_batch, y_batch, w, U, I, N_total, signal_var, nthreads=1):
    w, pw = _d(w); U, pU = _d(U); I, pI = _i(I); phi, pphi = _d(phi_batch); y, py = _d(y_batch)
    n, r, D = U.shape; Q = w.size; bs = phi.shape[2]
    gw = np.empty(Q); gU = np.empty((n, r, D), order="F"); fh = np.empty(bs)
    lib().gpo_gradients(pphi, py, bs, pw, pU, pI, n, r, D, Q, C.c_double(N_total), C.c_double(signal_var),
                        _d(gw)[1], _d(gU)[1], _d(fh)[1], nthreads)
    return gw, gU, fh


def expm(A):
    A, pA = _d(A); n = A.shape[0]
    E = np.empty((n, n), order="F")
    lib().gpo_expm(pA, n, _d(E)[1])
    return E


def proj(U, V):
    U, pU = _d(U); V, pV = _d(V); n, r = U.shape
    out = np.empty((n, r), order="F")
    lib().gpo_proj(pU, pV, n, r, _d(out)[1])
    return out


def geod(U, mom, t):
    U, pU = _d(U); mom, pm = _d(mom); n, r = U.shape
    out = np.empty((n, r), order="F")
    bad = lib().gpo_geod(pU, pm, n, r, C.c_double(t), _d(out)[1])
    return out, bool(bad)


def sgld_step(phi_batch, y_batch, w, U, I, N_total, signal_var, epsw, epsU, noise_w=None, noise_U=None,
              stiefel=True, nthreads=1):
    """In-place on copies; returns (w, U, fhat, nan_flag)."""
    w = np.array(w, dtype=np.float64); U = np.array(U, dtype=np.float64, order="F")
    I, pI = _i(I); phi, pphi = _d(phi_batch); y, py = _d(y_batch)
    n, r, D = U.shape; Q = w.size; bs = phi.shape[2]
    fh = np.empty(bs)
    pnw = _d(noise_w)[1] if noise_w is not None else None
    nU = np.asfortranarray(noise_U, dtype=np.float64) if noise_U is not None else None
    pnU = nU.ctypes.data_as(C.POINTER(C.c_double)) if nU is not None else None
    nwk = np.asfortranarray(noise_w, dtype=np.float64) if noise_w is not None else None
    pnw = nwk.ctypes.data_as(C.POINTER(C.c_double)) if nwk is not None else None
    bad = lib().gpo_sgld_step(pphi, py, bs, w.ctypes.data_as(C.POINTER(C.c_double)),
                              U.ctypes.data_as(C.POINTER(C.c_double)), pI, n, r, D, Q, C.c_double(N_total),
                              C.c_double(signal_var), C.c_double(epsw), C.c_double(epsU), pnw, pnU,
                              int(bool(stiefel)), _d(fh)[1], nthreads)
    return w, U, fh, bool(bad)
