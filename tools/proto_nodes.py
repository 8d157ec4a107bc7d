"""Prototype (NumPy, CPU): node-grid evaluation of the 1-D feature sums of one dimension.

T(x)[l] = amp * sum_j cos(x z_j + b_j) U[j,l] is a band-limited function of the scalar x, so it can be evaluated at
G*P Chebyshev nodes and interpolated; the adjoint (spreading) gives the factor gradient.  This script measures
the interpolation error against direct evaluation for the C4-like draw."""
import numpy as np


def cheb_nodes(P):
    p = np.arange(P)
    return np.cos((2 * p + 1) * np.pi / (2 * P)), (-1.0) ** p * np.sin((2 * p + 1) * np.pi / (2 * P))


def lagrange_w(u, P):
    un, wn = cheb_nodes(P)
    d = u[:, None] - un[None, :]
    d = np.where(np.abs(d) < 1e-9, 1e-9, d)
    t = wn[None, :] / d
    return t / t.sum(1, keepdims=True)


def run(n=1000, r=32, m=20000, ls=1.0, theta=2.0, P=12, seed=0, dtype=np.float64):
    rng = np.random.default_rng(seed)
    z = rng.standard_normal(n) / ls
    b = 2 * np.pi * rng.random(n)
    U = np.linalg.qr(rng.standard_normal((n, r)))[0]
    x = rng.standard_normal(m)
    E = rng.standard_normal((m, r))
    zmax = np.abs(z).max()
    lo, hi = x.min(), x.max()
    G = int(np.ceil((hi - lo) * zmax / (2 * theta)))
    h = (hi - lo) / (2 * G)
    un, _ = cheb_nodes(P)
    xn = lo + (2 * np.arange(G)[:, None] + 1 + un[None, :]) * h        # [G,P]
    Tn = np.cos(xn.reshape(-1, 1) * z[None, :] + b[None, :]) @ U        # [G*P, r]
    c = np.minimum(((x - lo) / (2 * h)).astype(int), G - 1)
    u = (x - lo) / h - (2 * c + 1)
    L = lagrange_w(u, P).astype(dtype)
    T = np.einsum("ip,ipl->il", L, Tn.reshape(G, P, r)[c].astype(dtype))
    phi = np.cos(x[:, None] * z[None, :] + b[None, :])
    Td = phi @ U
    # adjoint: gradU = phi' E  ~  phi_nodes' (spread E)
    En = np.zeros((G, P, r))
    np.add.at(En, c, L[:, :, None] * E[:, None, :])
    gU = np.cos(xn.reshape(-1, 1) * z[None, :] + b[None, :]).T @ En.reshape(G * P, r)
    gUd = phi.T @ E
    rel = lambda a, b_: np.linalg.norm(a - b_) / np.linalg.norm(b_)
    print(f"ls={ls} theta={theta} P={P}: zmax={zmax:.2f} G={G} nodes={G*P}  relerr T {rel(T, Td):.2e}  gradU {rel(gU, gUd):.2e}  "
          f"max abs T err {np.abs(T-Td).max():.2e}")


if __name__ == "__main__":
    for ls in (1.0, 0.3):
        for theta, P in ((1.0, 8), (2.0, 10), (2.0, 12), (2.0, 14), (4.0, 16), (3.0, 12)):
            run(ls=ls, theta=theta, P=P)
