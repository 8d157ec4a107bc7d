"""Host-side mirror of the reference's Julia module `GPT_SGLD` (/root/reference/GPT_SGLD.jl:5).

Same exported names, same positional arguments, same array orientation (Julia shapes, column-major
Float64; `I` is Int32 and 1-based), so code written against the reference reads the same here:

    phi   = feature(X, length_scale, sigma_RBF, phi_scale, Z, b)          # GPT_SGLD.jl:71
    I     = samplenz(r, D, Q)                                              # :181
    w_store, U_store = GPTregression(phi, y, signal_var, I, r, Q, m, epsw, epsU,
                                     burnin, maxepoch, param_seed)         # :345
    fhat  = pred(w, U, I, phitest)                                         # :233

The one deliberate difference: `feature` returns a lazy `RFFFeatures` handle (X, Z, b and the
hyper-parameters) instead of the materialised n x D x N array — the kernels generate the features on
the fly and never write them to memory.  `np.asarray(handle)` / `handle.materialize()` gives the dense
array; functions also accept a dense `phi` (compatibility path, FP64, HBM-bound).

Everything numerical happens in libgptb200.so through the C ABI (include/gptb200.h); this module only
marshals arguments.  Host-side helpers that the reference also runs on the host (samplenz,
datawhitening) are plain NumPy.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib as L

__all__ = ["feature", "feature2", "samplenz", "pred", "pred_samples", "GPTregression", "GPT_SGLDERMw", "GPT_SGLDERM_RMSprop", "GPT_GMC", "GPTclassification", "GPT_inf", "logsumexp", "proj", "geod", "geodboth",
           "datawhitening", "RFFFeatures", "Model", "ModelGroup", "GptbError", "GeodesicNaN", "shard_range"]


class GptbError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"libgptb200 status {status}: {msg}")
        self.status = status


class GeodesicNaN(GptbError):
    """expm produced NaN inside geod (GPT_SGLD.jl:24-26)."""


_default_precision = L.PREC_TC3
_default_device = -1


def set_default_precision(p: int) -> None:
    """L.PREC_FP64 (validation), L.PREC_TC3 (default: tensor cores, split operands), L.PREC_TC1 (fast)."""
    global _default_precision
    _default_precision = int(p)


def set_default_device(d: int) -> None:
    global _default_device
    _default_device = int(d)


def _f64(a, order="F"):
    return np.require(a, dtype=np.float64, requirements=["F_CONTIGUOUS" if order == "F" else "C_CONTIGUOUS", "ALIGNED"])


def _dp(a):
    return a.ctypes.data_as(L.c_double_p) if a is not None else None


def _check(status: int, handle=None):
    if status == L.OK:
        return
    msg = L.load().gptb_last_error(handle)
    msg = msg.decode() if msg else ""
    if status == L.GEODESIC_NAN:
        raise GeodesicNaN(status, msg)
    raise GptbError(status, msg)


class RFFFeatures:
    """Lazy result of `feature` / `feature2`: the inputs of GPT_SGLD.jl:71-105, not the n x D x N array."""

    def __init__(self, X, length_scale, sigma_RBF, phi_scale, Z, b, kind):
        self.X = _f64(np.atleast_2d(X))
        self.length_scale = np.atleast_1d(np.asarray(length_scale, dtype=np.float64))
        self.sigma_RBF = float(sigma_RBF)
        self.phi_scale = float(phi_scale)
        self.Z = _f64(Z)
        self.b = None if b is None else _f64(b)
        self.kind = kind
        N, D = self.X.shape
        if self.Z.ndim != 2 or self.Z.shape[1] != D:
            raise ValueError("Z must be n x D")
        if self.length_scale.size not in (1, D):
            raise ValueError("dimensions of X and length_scale do not match")   # GPT_SGLD.jl:326
        if kind == L.FEAT_SINCOS and self.Z.shape[0] % 2 == 1:
            raise ValueError("n not even")                                       # GPT_SGLD.jl:90
        if kind == L.FEAT_COS and (self.b is None or self.b.shape != self.Z.shape):
            raise ValueError("b must be n x D")

    @property
    def shape(self):
        return (self.Z.shape[0], self.X.shape[1], self.X.shape[0])  # size(phi) = (n, D, N)

    def with_X(self, X) -> "RFFFeatures":
        return RFFFeatures(X, self.length_scale, self.sigma_RBF, self.phi_scale, self.Z, self.b, self.kind)

    def materialize(self, precision: Optional[int] = None) -> np.ndarray:
        n, D, N = self.shape
        with Model(D, n, 1, 1, precision=L.PREC_FP64, feature_kind=self.kind) as mdl:
            mdl.set_features(self)
            out = np.empty((n, D, N), order="F")
            _check(L.load().gptb_features(mdl.h, _dp(self.X), N, _dp(out)), mdl.h)
        return out

    def __array__(self, dtype=None, copy=None):
        a = self.materialize()
        return a if dtype is None else a.astype(dtype)


def feature(X, length_scale, sigma_RBF, phi_scale, Z, b) -> RFFFeatures:
    """GPT_SGLD.jl:71-84.  Use Z=randn(n,D), b=2*pi*rand(n,D), phi_scale=sqrt(n/Q^(1/D)) (:70,:306)."""
    return RFFFeatures(X, length_scale, sigma_RBF, phi_scale, Z, b, L.FEAT_COS)


def feature2(X, length_scale, sigma_RBF, phi_scale, Z) -> RFFFeatures:
    """GPT_SGLD.jl:88-105 (sin/cos pairs; raises "n not even" for odd n)."""
    return RFFFeatures(X, length_scale, sigma_RBF, phi_scale, Z, None, L.FEAT_SINCOS)


def tgp_feature(x, n, sigmaRBF, Z, b01) -> RFFFeatures:
    """TGP.jl:6-14 with its draws passed in: Z=randn(n,D) (divided by sigmaRBF here), b01=rand(n,D)."""
    x = np.atleast_2d(np.asarray(x, dtype=np.float64))
    return RFFFeatures(x, float(sigmaRBF), 1.0, 1.0, Z, 2 * np.pi * np.asarray(b01), L.FEAT_COS)


def samplenz(r: int, D: int, Q: int, rng: Optional[np.random.Generator] = None) -> np.ndarray:
    """GPT_SGLD.jl:181-190: Q distinct cells of the r^D lattice, little-endian base-r digits + 1 (Int32).
    Host-side like the reference; `rng` replaces Julia's global RNG (seed it for reproducibility)."""
    rng = np.random.default_rng() if rng is None else rng
    total = int(r) ** int(D)
    if Q > total:
        raise ValueError("Q exceeds r^D")
    if total <= 10_000_000:
        L_ = [int(v) for v in rng.choice(total, size=Q, replace=False)]
    else:
        seen, L_ = set(), []
        while len(L_) < Q:
            v = int(rng.integers(0, total, dtype=np.uint64)) if total < 2 ** 63 else int(rng.integers(0, 2 ** 62)) % total
            if v not in seen:
                seen.add(v)
                L_.append(v)
    I = np.empty((Q, D), dtype=np.int32, order="F")
    for q, v in enumerate(L_):
        for k in range(D):
            I[q, k] = v % r + 1
            v //= r
    return I


def datawhitening(X):
    """GPT_SGLD.jl:62-67: in-place per-column (x-mean)/std (n-1 denominator).  Host preprocessing."""
    for i in range(X.shape[1]):
        X[:, i] = (X[:, i] - X[:, i].mean()) / X[:, i].std(ddof=1)
    return X


class Model:
    """One libgptb200 handle (device state of one chain).  The module-level functions build these on demand;
    long-running programs (bench, multi-GPU) hold one explicitly."""

    def __init__(self, D, n, r, Q, precision=None, feature_kind=L.FEAT_COS, device=None, seed=0):
        self.lib = L.load()
        cfg = L.Config(D=int(D), n=int(n), r=int(r), Q=int(Q),
                       precision=_default_precision if precision is None else int(precision),
                       feature_kind=int(feature_kind), device=_default_device if device is None else int(device),
                       reserved=0, seed=int(seed))
        self.cfg = cfg
        self._group = None          # (handle array, n) when this model leads a ModelGroup
        self.h = L.H()
        st = self.lib.gptb_create(C.byref(cfg), C.byref(self.h))
        if st != L.OK:
            self.h = None
            _check(st, None)

    # -- lifetime
    def close(self):
        if getattr(self, "h", None):
            self.lib.gptb_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- definition
    def set_features(self, f: RFFFeatures):
        _check(self.lib.gptb_set_features(self.h, _dp(f.Z), _dp(f.b), _dp(f.length_scale), f.length_scale.size,
                                          f.sigma_RBF, f.phi_scale), self.h)

    def set_index(self, I):
        I = np.require(I, dtype=np.int32, requirements=["F_CONTIGUOUS"])
        if I.shape != (self.cfg.Q, self.cfg.D):
            raise ValueError(f"I must be Q x D = {self.cfg.Q} x {self.cfg.D}")
        _check(self.lib.gptb_set_index(self.h, I.ctypes.data_as(L.c_int32_p)), self.h)

    def set_train(self, X, y):
        X = _f64(X); y = _f64(np.ravel(y))
        if X.shape[0] != y.size:
            raise ValueError("X and y disagree on N")
        _check(self.lib.gptb_set_train(self.h, _dp(X), _dp(y), X.shape[0]), self.h)

    def set_domain(self, lo=None, hi=None):
        """Bounds of every input column for the node grid (gptb_set_domain); None, None forgets them."""
        if lo is None or hi is None:
            _check(self.lib.gptb_set_domain(self.h, None, None), self.h)
            return
        lo = np.ascontiguousarray(np.broadcast_to(np.asarray(lo, dtype=np.float64), (self.cfg.D,)))
        hi = np.ascontiguousarray(np.broadcast_to(np.asarray(hi, dtype=np.float64), (self.cfg.D,)))
        _check(self.lib.gptb_set_domain(self.h, _dp(lo), _dp(hi)), self.h)

    def set_train_phi(self, phi, y):
        phi = _f64(phi); y = _f64(np.ravel(y))
        _check(self.lib.gptb_set_train_phi(self.h, _dp(phi), _dp(y), phi.shape[2]), self.h)

    def set_params(self, w, U):
        w = _f64(np.ravel(w)); U = _f64(U)
        if w.size != self.cfg.Q or U.shape != (self.cfg.n, self.cfg.r, self.cfg.D):
            raise ValueError("w must have Q entries and U must be n x r x D")
        _check(self.lib.gptb_set_params(self.h, _dp(w), _dp(U)), self.h)

    def get_params(self):
        w = np.empty(self.cfg.Q)
        U = np.empty((self.cfg.n, self.cfg.r, self.cfg.D), order="F")
        _check(self.lib.gptb_get_params(self.h, _dp(w), _dp(U)), self.h)
        return w, U

    def init_params(self, param_seed: int, stiefel: bool = True):
        _check(self.lib.gptb_init_params(self.h, int(param_seed), int(bool(stiefel))), self.h)

    # -- compute
    def pred(self, w, U, Xtest):
        w = _f64(np.ravel(w)); U = _f64(U); Xtest = _f64(np.atleast_2d(Xtest))
        out = np.empty(Xtest.shape[0])
        _check(self.lib.gptb_pred(self.h, _dp(w), _dp(U), _dp(Xtest), Xtest.shape[0], _dp(out)), self.h)
        return out

    def pred_phi(self, w, U, phitest):
        w = _f64(np.ravel(w)); U = _f64(U); phitest = _f64(phitest)
        out = np.empty(phitest.shape[2])
        _check(self.lib.gptb_pred_phi(self.h, _dp(w), _dp(U), _dp(phitest), phitest.shape[2], _dp(out)), self.h)
        return out

    def pred_samples(self, w_store, U_store, Xtest, per_sample=False):
        w_store = _f64(np.atleast_2d(w_store.T).T if w_store.ndim == 1 else w_store)
        U_store = _f64(U_store if U_store.ndim == 4 else U_store[..., None])
        Xtest = _f64(np.atleast_2d(Xtest))
        S = w_store.shape[1]
        Nt = Xtest.shape[0]
        mean = np.empty(Nt)
        per = np.empty((Nt, S), order="F") if per_sample else None
        if self._group is not None:   # test rows sharded over the group's devices
            _check(self.lib.gptb_multi_pred_samples(self._group[0], self._group[1], _dp(w_store), _dp(U_store), S, _dp(Xtest), Nt,
                                                    _dp(mean), _dp(per)), self.h)
        else:
            _check(self.lib.gptb_pred_samples(self.h, _dp(w_store), _dp(U_store), S, _dp(Xtest), Nt, _dp(mean), _dp(per)),
                   self.h)
        return (mean, per) if per_sample else mean

    def grad(self, signal_var, idx=None, start=0, bs=None, include_prior=True, want_fhat=True):
        if idx is not None:
            idx = np.require(idx, dtype=np.int64, requirements=["C_CONTIGUOUS"])
            bs = idx.size
        gw = np.empty(self.cfg.Q)
        gU = np.empty((self.cfg.n, self.cfg.r, self.cfg.D), order="F")
        fh = np.empty(bs) if want_fhat else None
        _check(self.lib.gptb_grad(self.h, float(signal_var), idx.ctypes.data_as(L.c_int64_p) if idx is not None else None,
                                  int(start), int(bs), _dp(gw), _dp(gU), _dp(fh), int(bool(include_prior))), self.h)
        return gw, gU, fh

    def sgld_step(self, signal_var, epsw, epsU, idx=None, start=0, bs=None, langevin=True, stiefel=True,
                  noise_w=None, noise_U=None):
        if idx is not None:
            idx = np.require(idx, dtype=np.int64, requirements=["C_CONTIGUOUS"])
            bs = idx.size
        nw = _f64(noise_w) if noise_w is not None else None
        nU = _f64(noise_U) if noise_U is not None else None
        flags = (L.FLAG_LANGEVIN if langevin else 0) | (L.FLAG_STIEFEL if stiefel else 0)
        _check(self.lib.gptb_sgld_step(self.h, float(signal_var), idx.ctypes.data_as(L.c_int64_p) if idx is not None else None,
                                       int(start), int(bs), float(epsw), float(epsU), flags, _dp(nw), _dp(nU)), self.h)

    def sgld_step_host(self, signal_var, Xb, yb, N_total, epsw, epsU, langevin=True, stiefel=True, noise_w=None,
                       noise_U=None, w_out=None):
        """Loop body with the minibatch passed from the host.  Xb must be F-contiguous float64 (bs x D)."""
        nw = _f64(noise_w) if noise_w is not None else None
        nU = _f64(noise_U) if noise_U is not None else None
        flags = (L.FLAG_LANGEVIN if langevin else 0) | (L.FLAG_STIEFEL if stiefel else 0)
        _check(self.lib.gptb_sgld_step_host(self.h, float(signal_var), _dp(Xb), _dp(yb), Xb.shape[0], int(N_total),
                                            float(epsw), float(epsU), flags, _dp(nw), _dp(nU), _dp(w_out)), self.h)

    def sgld_step_host_submit(self, signal_var, Xb, yb, N_total, epsw, epsU, langevin=True, stiefel=True, noise_w=None,
                              noise_U=None, w_out=None):
        """Pipelined sgld_step_host: queues the copy of (Xb, yb) and the step, returns at once.  The arrays must be
        pinned host memory and stay alive until the matching sgld_step_host_wait()."""
        nw = _f64(noise_w) if noise_w is not None else None
        nU = _f64(noise_U) if noise_U is not None else None
        flags = (L.FLAG_LANGEVIN if langevin else 0) | (L.FLAG_STIEFEL if stiefel else 0)
        _check(self.lib.gptb_sgld_step_host_submit(self.h, float(signal_var), _dp(Xb), _dp(yb), Xb.shape[0], int(N_total),
                                                   float(epsw), float(epsU), flags, _dp(nw), _dp(nU), _dp(w_out)), self.h)

    def sgld_step_host_wait(self):
        """Waits for the oldest submitted host step; raises on a geodesic NaN like sgld_step_host."""
        _check(self.lib.gptb_sgld_step_host_wait(self.h), self.h)

    def sgld_run(self, signal_var, m, epsw, epsU, burnin, maxepoch, langevin=True, stiefel=True, perms=None,
                 noise_w=None, noise_U=None, store=True, N=None, fix_U=False, rmsprop=False):
        Q, n, r, D = self.cfg.Q, self.cfg.n, self.cfg.r, self.cfg.D
        flags = ((L.FLAG_LANGEVIN if langevin else 0) | (L.FLAG_STIEFEL if stiefel else 0) | (L.FLAG_FIX_U if fix_U else 0)
                 | (L.FLAG_RMSPROP if rmsprop else 0))
        numbatches = -(-int(N) // int(m))
        T = maxepoch * numbatches
        w_store = np.zeros((Q, T), order="F") if store else None
        U_store = np.zeros((n, r, D, T), order="F") if (store and not fix_U) else None
        pp = None
        if perms is not None:
            pp = np.require(np.asarray(perms), dtype=np.int64, requirements=["C_CONTIGUOUS"])
            if pp.shape != (burnin + maxepoch, N):
                raise ValueError("perms must be (burnin+maxepoch) x N")
        nw = np.require(noise_w, dtype=np.float64, requirements=["C_CONTIGUOUS"]) if noise_w is not None else None
        nU = None
        if noise_U is not None:   # [step][n,r,D] with each step's block in Julia (F) order
            nU = np.ascontiguousarray(np.stack([np.asarray(a, dtype=np.float64).ravel(order="F") for a in noise_U]))
        args = (float(signal_var), int(m), float(epsw), float(epsU), int(burnin), int(maxepoch), flags,
                pp.ctypes.data_as(L.c_int64_p) if pp is not None else None, _dp(nw), _dp(nU), _dp(w_store), _dp(U_store))
        if self._group is not None:
            st = self.lib.gptb_multi_sgld_run(self._group[0], self._group[1], *args)
        else:
            st = self.lib.gptb_sgld_run(self.h, *args)
        if st == L.GEODESIC_NAN:
            # GPT_SGLD.jl:25, :422-424 — print the warning and hand back all-zero stores
            print("Get NaN when moving along Geodesic. Try smaller epsU")
            return w_store, U_store
        _check(st, self.h)
        return w_store, U_store

    def sgld_advance(self, signal_var, m, epsw, epsU, nsteps, langevin=True, stiefel=True, rmsprop=False):
        """Queue the next nsteps loop bodies (no store, no host sync); call synchronize() to wait.  rmsprop=True: the
        steps of GPT_SGLDERM_RMSprop (epsw is its epsilon)."""
        flags = (L.FLAG_LANGEVIN if langevin else 0) | (L.FLAG_STIEFEL if stiefel else 0) | (L.FLAG_RMSPROP if rmsprop else 0)
        if self._group is not None:
            _check(self.lib.gptb_multi_sgld_advance(self._group[0], self._group[1], float(signal_var), int(m), float(epsw),
                                                    float(epsU), flags, int(nsteps)), self.h)
            return
        _check(self.lib.gptb_sgld_advance(self.h, float(signal_var), int(m), float(epsw), float(epsU), flags,
                                          int(nsteps)), self.h)

    # -- multi-GPU / introspection
    def comm_init(self, nranks: int, rank: int, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        _check(self.lib.gptb_comm_init(self.h, int(nranks), int(rank), buf), self.h)

    def set_stream(self, cuda_stream_ptr: int):
        _check(self.lib.gptb_set_stream(self.h, C.c_void_p(cuda_stream_ptr)), self.h)

    def synchronize(self):
        _check(self.lib.gptb_synchronize(self.h), self.h)

    def set_option(self, option: int, value: float):
        """gptb_set_option: L.OPT_* (algorithm selection, node-grid order, pipeline knobs)."""
        _check(self.lib.gptb_set_option(self.h, int(option), float(value)), self.h)

    def set_profiling(self, on: bool):
        _check(self.lib.gptb_set_profiling(self.h, int(on)), self.h)

    def phase_ms(self, reset=False):
        ms = (C.c_float * len(L.PHASES))()
        n = C.c_int64(0)
        _check(self.lib.gptb_get_phase_ms(self.h, ms, C.byref(n), int(reset)), self.h)
        return dict(zip(L.PHASES, [float(v) for v in ms])), int(n.value)


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    _check(L.load().gptb_comm_unique_id(buf), None)
    return buf.raw


def shard_range(rank: int, nranks: int, bs: int):
    """Rows [lo, hi) of a global minibatch (or test set) that rank `rank` of `nranks` owns — the split libgptb200
    applies internally (compute_grad in csrc/api.cu; GPT_SGLD_p.jl:70-108 splits the batch rows over workers)."""
    return rank * bs // nranks, (rank + 1) * bs // nranks


class ModelGroup:
    """Several GPUs driven from THIS process (gptb_comm_init_local / gptb_multi_*): one handle per device, every setter
    goes to all of them, the samplers and the posterior sweep run on all of them at once (rows of each minibatch, or of
    the test set, sharded).  Everything else (get_params, pred, ...) is answered by the first handle."""

    _FORWARD = ("set_features", "set_index", "set_train", "set_train_phi", "set_params", "init_params", "set_option",
                "set_domain", "synchronize")

    def __init__(self, devices, D, n, r, Q, precision=None, feature_kind=L.FEAT_COS, seed=0):
        self.models = [Model(D, n, r, Q, precision=precision, feature_kind=feature_kind, device=int(d), seed=seed) for d in devices]
        lead = self.models[0]
        self.lib, self.cfg, self.h = lead.lib, lead.cfg, lead.h
        self._arr = (C.c_void_p * len(self.models))(*[m.h.value for m in self.models])
        harr = C.cast(self._arr, C.POINTER(C.c_void_p))
        st = self.lib.gptb_comm_init_local(harr, len(self.models))
        if st != L.OK:
            self.close()
            _check(st, None)
        lead._group = (harr, len(self.models))

    def __getattr__(self, name):
        if name in ("models", "lib", "cfg", "h", "_arr"):
            raise AttributeError(name)
        if name in ModelGroup._FORWARD:
            def call(*a, **kw):
                out = None
                for m in self.models:
                    out = getattr(m, name)(*a, **kw)
                return out
            return call
        return getattr(self.models[0], name)

    def close(self):
        for m in getattr(self, "models", []):
            m._group = None
            m.close()
        self.models = []

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def _model_for(phi, r, Q, I, precision=None, seed=0, options=None, devices=None) -> Model:
    if isinstance(phi, RFFFeatures):
        n, D, _ = phi.shape
        if devices is not None and len(devices) > 1:
            mdl = ModelGroup(devices, D, n, r, Q, precision=precision, feature_kind=phi.kind, seed=seed)
        else:
            mdl = Model(D, n, r, Q, precision=precision, feature_kind=phi.kind, seed=seed,
                        device=None if not devices else int(devices[0]))
        mdl.set_features(phi)
    else:
        n, D, _ = phi.shape
        mdl = Model(D, n, r, Q, precision=precision, seed=seed)
    mdl.set_index(I)
    for k, v in (options or {}).items():
        mdl.set_option(k, v)
    return mdl


def pred(w, U, I, phitest, precision: Optional[int] = None, options=None):
    """GPT_SGLD.jl:233-243.  phitest: RFFFeatures (lazy) or a dense n x D x Nt array."""
    U = np.asarray(U)
    w = np.ravel(w)
    with _model_for(phitest, U.shape[1], w.size, I, precision, options=options) as mdl:
        if isinstance(phitest, RFFFeatures):
            return mdl.pred(w, U, phitest.X)
        return mdl.pred_phi(w, U, phitest)


def pred_samples(w_store, U_store, I, phitest: RFFFeatures, cols: Optional[Sequence[int]] = None,
                 per_sample=False, precision: Optional[int] = None, options=None, devices=None):
    """kin40kExperiment.jl:75-87 / GPT_SGLD_p.jl:124-132: mean over stored samples of pred(w_s,U_s,I,phitest)."""
    if cols is not None:
        w_store = w_store[:, list(cols)]
        U_store = U_store[:, :, :, list(cols)]
    with _model_for(phitest, U_store.shape[1], w_store.shape[0], I, precision, options=options, devices=devices) as mdl:
        return mdl.pred_samples(np.asfortranarray(w_store), np.asfortranarray(U_store), phitest.X, per_sample)


def GPTregression(phi, y, signal_var, I, r, Q, m, epsw, epsU, burnin, maxepoch, param_seed, langevin=True,
                  stiefel=True, *, precision: Optional[int] = None, w0=None, U0=None, perms=None, noise_w=None,
                  noise_U=None, options=None, devices=None):
    """GPT_SGLD.jl:345-448.  Returns (w_store[Q,T], U_store[n,r,D,T]), T = maxepoch*ceil(N/m); all zeros (after
    printing the reference's warning) if the geodesic produced NaN.

    Keyword-only extras (not in the reference): precision; w0/U0 to start from given parameters instead of the
    param_seed draw; perms / noise_w / noise_U to inject the random draws (parity tests); options = {L.OPT_*: value}
    passed to gptb_set_option; devices = [0, 1, ...]: several GPUs driven from this process, the rows of every minibatch
    sharded over them (lazy features only)."""
    y = np.ravel(np.asarray(y, dtype=np.float64))
    N = y.size
    with _model_for(phi, r, Q, I, precision, seed=int(param_seed), options=options, devices=devices) as mdl:
        if isinstance(phi, RFFFeatures):
            mdl.set_train(phi.X, y)
        else:
            mdl.set_train_phi(phi, y)
        if w0 is not None:
            mdl.set_params(w0, U0)
        else:
            mdl.init_params(param_seed, stiefel)
        return mdl.sgld_run(signal_var, m, epsw, epsU, burnin, maxepoch, langevin, stiefel, perms, noise_w, noise_U,
                            store=True, N=N)


def GPT_SGLDERMw(phi, y, signal_var, I, r, Q, m, epsw, burnin, maxepoch, param_seed=0, *, precision: Optional[int] = None,
                 w0=None, U0=None, perms=None, noise_w=None):
    """GPT_SGLD.jl:1065-1118: SGLD on the core w only; the factors stay at their (Stiefel) initial draw.  Returns
    (w_store[Q, maxepoch*ceil(N/m)], U).  The reference takes its draws from the global RNG; here param_seed names
    them (keyword-only extras as in GPTregression)."""
    y = np.ravel(np.asarray(y, dtype=np.float64))
    N = y.size
    with _model_for(phi, r, Q, I, precision, seed=int(param_seed)) as mdl:
        if isinstance(phi, RFFFeatures):
            mdl.set_train(phi.X, y)
        else:
            mdl.set_train_phi(phi, y)
        if w0 is not None:
            mdl.set_params(w0, U0)
        else:
            mdl.init_params(param_seed, True)
        w_store, _ = mdl.sgld_run(signal_var, m, epsw, 0.0, burnin, maxepoch, True, True, perms, noise_w, None, store=True,
                                  N=N, fix_U=True)
        return w_store, mdl.get_params()[1]


def GPT_SGLDERM_RMSprop(phi, y, signal_var, I, r, Q, m, epsilon, alpha, burnin, maxepoch, param_seed=0, *,
                        precision: Optional[int] = None, w0=None, U0=None, perms=None, noise_w=None, noise_U=None, options=None):
    """GPT_SGLD.jl:1121-1237: SGLD-ERM with RMSprop step sizes (one per element of w, one per U^(k)); `epsilon` is the
    base step size, `alpha` the moving-average coefficient.  Returns (w_store, U_store) like GPTregression.  The
    reference draws from the global RNG; here param_seed names the draws (keyword-only extras as in GPTregression)."""
    y = np.ravel(np.asarray(y, dtype=np.float64))
    N = y.size
    opts = {L.OPT_RMS_ALPHA: float(alpha)}
    opts.update(options or {})
    with _model_for(phi, r, Q, I, precision, seed=int(param_seed), options=opts) as mdl:
        if isinstance(phi, RFFFeatures):
            mdl.set_train(phi.X, y)
        else:
            mdl.set_train_phi(phi, y)
        if w0 is not None:
            mdl.set_params(w0, U0)
        else:
            mdl.init_params(param_seed, True)
        return mdl.sgld_run(signal_var, m, epsilon, 0.0, burnin, maxepoch, True, True, perms, noise_w, noise_U, store=True,
                            N=N, rmsprop=True)


def proj(U, V):
    """GPT_SGLD.jl:14-16."""
    U = _f64(U); V = _f64(V)
    out = np.empty(U.shape, order="F")
    _check(L.load().gptb_proj(_default_device, U.shape[0], U.shape[1], _dp(U), _dp(V), _dp(out)), None)
    return out


def geod(U, mom, t):
    """GPT_SGLD.jl:19-37.  On NaN prints the reference's warning and returns zeros(n,r)."""
    U = _f64(U); mom = _f64(mom)
    out = np.empty(U.shape, order="F")
    st = L.load().gptb_geod(_default_device, U.shape[0], U.shape[1], _dp(U), _dp(mom), float(t), _dp(out))
    if st == L.GEODESIC_NAN:
        print("Get NaN when moving along Geodesic. Try smaller epsU")
        return out
    _check(st, None)
    return out


def geodboth(U, mom, t):
    """GPT_SGLD.jl:40-59: (geodesic point with unit columns, transported momentum).  On NaN prints the reference's
    warning and returns zeros for both."""
    U = _f64(U); mom = _f64(mom)
    out = np.empty(U.shape, order="F"); out2 = np.empty(U.shape, order="F")
    st = L.load().gptb_geodboth(_default_device, U.shape[0], U.shape[1], _dp(U), _dp(mom), float(t), _dp(out), _dp(out2))
    if st == L.GEODESIC_NAN:
        print("Get NaN when moving along Geodesic. Try smaller epsU")
        return out, out2
    _check(st, None)
    return out, out2


def GPT_GMC(phi, y, signal_var, I, r, Q, epsw, epsU, burnin, maxepoch, L_, param_seed, *, precision: Optional[int] = None,
            w0=None, U0=None, noise_p=None, noise_mom=None, unif=None, restore_U_on_reject=False, options=None):
    """GPT_SGLD.jl:684-805, geodesic Monte Carlo with L_ leapfrog steps per epoch.  Returns (w_store[Q,maxepoch],
    U_store[n,r,D,maxepoch], accept_prob[burnin+maxepoch]) like the reference; after a NaN in the geodesic it prints the
    reference's warning and returns zero stores and NaN acceptance probabilities (:746-748).  Keyword-only extras inject
    the draws; restore_U_on_reject=False keeps the reference's aliasing behaviour (see include/gptb200.h)."""
    y = np.ravel(np.asarray(y, dtype=np.float64))
    with _model_for(phi, r, Q, I, precision, seed=int(param_seed), options=options) as mdl:
        if isinstance(phi, RFFFeatures):
            mdl.set_train(phi.X, y)
            n, D = phi.shape[0], phi.shape[1]
        else:
            mdl.set_train_phi(phi, y)
            n, D = phi.shape[0], phi.shape[1]
        if w0 is not None:
            mdl.set_params(w0, U0)
        else:
            mdl.init_params(param_seed, True)
        T = burnin + maxepoch
        w_store = np.zeros((Q, maxepoch), order="F"); U_store = np.zeros((n, r, D, maxepoch), order="F")
        acc = np.zeros(T)
        npq = None if noise_p is None else np.ascontiguousarray(np.asarray(noise_p, dtype=np.float64).reshape(T, Q))
        nm = None if noise_mom is None else np.ascontiguousarray(
            np.stack([np.asarray(a, dtype=np.float64).ravel(order="F") for a in noise_mom]))
        un = None if unif is None else np.ascontiguousarray(np.asarray(unif, dtype=np.float64))
        st = mdl.lib.gptb_gmc_run(mdl.h, float(signal_var), float(epsw), float(epsU), int(burnin), int(maxepoch), int(L_),
                                  int(bool(restore_U_on_reject)), _dp(npq), _dp(nm), _dp(un), _dp(w_store), _dp(U_store), _dp(acc))
        if st == L.GEODESIC_NAN:
            print("Get NaN when moving along Geodesic. Try smaller epsU")
            return w_store, U_store, acc
        _check(st, mdl.h)
        return w_store, U_store, acc


def GPT_inf(X, y, sigma, n, r, sigmaRBF, q, generator, num_iterations, burnin, *, Z=None, b01=None, I=None, U0=None,
            noise_w=None, noise_U=None, whiten=True):
    """TGP.jl:37-86: Gibbs sampler of the tensor model with Gaussian factors.  Returns (W_array[q,T], V_array[n,r,D,T], I)
    like the reference (T = num_iterations - burnin; I 1-based).  X and y are whitened first (:39-40) unless whiten=False.
    The reference draws Z, b (feature :6-14), I (:57) and the initial factors (:56) from Julia's global RNG seeded with
    `generator`; here a NumPy generator seeded with it does, unless Z (n x D), b01 (n x D in [0,1)), I (q x D, 1-based),
    U0 (n x r x D) are given.  noise_w[m][q] / noise_U[m][D][n*r] inject the draws of :66 / :83."""
    X = np.array(X, dtype=np.float64, order="F", copy=True)
    y = np.array(y, dtype=np.float64, copy=True).reshape(-1)
    if whiten:
        X = datawhitening(X)
        y = datawhitening(y.reshape(-1, 1)).ravel()
    N, D = X.shape
    rng = np.random.default_rng(int(generator))
    Z = rng.standard_normal((n, D)) if Z is None else np.asarray(Z, dtype=np.float64)
    b01 = rng.random((n, D)) if b01 is None else np.asarray(b01, dtype=np.float64)
    I = (rng.integers(1, r + 1, size=(q, D)).astype(np.int32) if I is None else np.asarray(I, dtype=np.int32))
    U0 = np.sqrt(1.0 / r) * rng.standard_normal((n, r, D)) if U0 is None else np.asarray(U0, dtype=np.float64)
    f = tgp_feature(X, n, sigmaRBF, Z, b01)
    T = int(num_iterations) - int(burnin)
    W = np.zeros((q, T), order="F"); V = np.zeros((n, r, D, T), order="F")
    nw = None if noise_w is None else np.ascontiguousarray(np.asarray(noise_w, dtype=np.float64).reshape(num_iterations, q))
    nU = None if noise_U is None else np.ascontiguousarray(np.asarray(noise_U, dtype=np.float64).reshape(num_iterations, D, n * r))
    with Model(D, n, r, q, precision=L.PREC_FP64, feature_kind=f.kind, seed=int(generator)) as mdl:
        mdl.set_features(f); mdl.set_index(np.asfortranarray(I)); mdl.set_train(X, y)
        mdl.set_params(np.zeros(q), U0)
        _check(mdl.lib.gptb_tgp_gibbs(mdl.h, float(sigma), int(num_iterations), int(burnin), _dp(nw), _dp(nU), _dp(W), _dp(V)), mdl.h)
    return W, V, I


def logsumexp(x):
    """GPT_SGLD.jl export (BloodTransfusionExperiment.jl:53-56): log(sum(exp(x))), the maximum taken out first."""
    x = np.asarray(x, dtype=np.float64)
    a = np.max(x)
    return float(a + np.log(np.sum(np.exp(x - a))))


def GPTclassification(phi, y, I, r, Q, m, epsw, epsU, burnin, maxepoch, param_seed, langevin=True, stiefel=True, *,
                      precision: Optional[int] = None, w0=None, U0=None, perms=None, noise_w=None, noise_U=None,
                      double_update=True, options=None):
    """GPT_SGLD.jl:452-680: C = max(y)-min(y)+1 independent (w, U) sets under a softmax likelihood.  Returns
    (w_store[Q,C,T], U_store[n,r,D,C,T]), T = maxepoch*ceil(N/m); all zeros (after the reference's warning) if a geodesic
    produced NaN.  y: integer class labels (the reference's 1..C; any consecutive range works, min(y) is class 0).

    Keyword-only extras (not in the reference): precision; w0[Q,C] / U0[n,r,D,C]; perms; noise_w[step][move] of shape
    (Q,C) and noise_U[step][move] of shape (n,r,D,C) for the TWO moves the reference makes per minibatch (:624-636 and
    :639-671); double_update=False keeps only the second; options = {L.OPT_*: value}."""
    y = np.ravel(np.asarray(y))
    labels = np.ascontiguousarray(np.rint(y - y.min()).astype(np.int32))
    Ccls = int(labels.max()) + 1                                           # :460
    N = labels.size
    models = []
    try:
        for c in range(Ccls):
            mdl = _model_for(phi, r, Q, I, precision, seed=int(param_seed) + 7919 * c, options=options)
            models.append(mdl)
            if isinstance(phi, RFFFeatures):
                mdl.set_train(phi.X, np.zeros(N))
            else:
                mdl.set_train_phi(phi, np.zeros(N))
            if w0 is not None:
                mdl.set_params(np.asarray(w0)[:, c], np.asarray(U0)[:, :, :, c])
            else:
                mdl.init_params(int(param_seed) + 7919 * c, stiefel)
        n, D = phi.shape[0], phi.shape[1]
        numbatches = -(-N // int(m))
        T = int(maxepoch) * numbatches
        steps = (int(burnin) + int(maxepoch)) * numbatches
        moves = 2 if double_update else 1
        w_store = np.zeros((Q, Ccls, T), order="F")
        U_store = np.zeros((n, r, D, Ccls, T), order="F")
        pm = None if perms is None else np.ascontiguousarray(np.asarray(perms, dtype=np.int64).reshape(int(burnin) + int(maxepoch), N))

        def pack(noise, shape):
            if noise is None:
                return None
            out = np.empty((steps, moves, int(np.prod(shape))))
            for s_ in range(steps):
                for u in range(moves):
                    a = np.asarray(noise[s_][u if double_update else -1], dtype=np.float64)
                    if a.shape != shape:
                        raise ValueError(f"noise entry of shape {a.shape}, expected {shape}")
                    out[s_, u] = a.ravel(order="F")
            return out

        nw = pack(noise_w, (Q, Ccls))
        nU = pack(noise_U, (n, r, D, Ccls))
        flags = (L.FLAG_LANGEVIN if langevin else 0) | (L.FLAG_STIEFEL if stiefel else 0) | \
            (0 if double_update else L.FLAG_CLS_SINGLE_UPDATE)
        import ctypes as _C
        harr = (_C.c_void_p * Ccls)(*[mdl.h for mdl in models])
        lib = models[0].lib
        st = lib.gptb_classify_run(_C.cast(harr, _C.POINTER(_C.c_void_p)), Ccls, labels.ctypes.data_as(L.c_int32_p), int(m),
                                   float(epsw), float(epsU), int(burnin), int(maxepoch), int(flags),
                                   None if pm is None else pm.ctypes.data_as(L.c_int64_p), _dp(nw), _dp(nU), _dp(w_store), _dp(U_store))
        if st == L.GEODESIC_NAN:
            print("Get NaN when moving along Geodesic. Try smaller epsU")
            return w_store, U_store
        _check(st, models[0].h)
        return w_store, U_store
    finally:
        for mdl in models:
            mdl.close()


# ------------------------------------------------------------------------------------------------
# Full-theta ("no tensor") RFF model — GPT_SGLD.jl:109-120, :809-847 (gptb_nt_* in include/gptb200.h)
# ------------------------------------------------------------------------------------------------
class NTFeatures:
    """Lazy result of `featureNotensor`: its inputs, not the n x N matrix.  np.asarray(f) materialises it."""

    def __init__(self, X, length_scale, sigma_RBF, Z, b):
        self.X = _f64(np.atleast_2d(X))
        self.length_scale = np.atleast_1d(np.asarray(length_scale, dtype=np.float64))
        self.sigma_RBF = float(sigma_RBF)
        self.Z = _f64(Z)
        self.b = np.ascontiguousarray(np.ravel(b), dtype=np.float64)
        N, D = self.X.shape
        if self.Z.ndim != 2 or self.Z.shape[1] != D or self.b.size != self.Z.shape[0]:
            raise ValueError("Z must be n x D and b of length n")
        if self.length_scale.size not in (1, D):
            raise ValueError("dimensions of X and length_scale do not match")

    @property
    def shape(self):
        return (self.Z.shape[0], self.X.shape[0])   # size(phi) = (n, N)

    def with_X(self, X) -> "NTFeatures":
        return NTFeatures(X, self.length_scale, self.sigma_RBF, self.Z, self.b)

    def __array__(self, dtype=None, copy=None):
        with NTModel(self) as mdl:
            a = mdl.features(self.X)
        return a if dtype is None else a.astype(dtype)


class NTModel:
    """Handle of the full-theta model (gptb_nt_handle)."""

    def __init__(self, f: NTFeatures, precision: Optional[int] = None, device: Optional[int] = None, seed: int = 0):
        self.lib = L.load()
        self.n, self.D = f.Z.shape
        self.h = L.H()
        prec = L.PREC_TC3 if precision is None else int(precision)
        st = self.lib.gptb_nt_create(self.D, self.n, prec, _default_device if device is None else int(device), int(seed),
                                     C.byref(self.h))
        _check(st, None)
        ls = np.ascontiguousarray(f.length_scale)
        self._check(self.lib.gptb_nt_set_features(self.h, _dp(f.Z), _dp(f.b), _dp(ls), int(ls.size), f.sigma_RBF))

    def _check(self, st):
        if st == L.OK:
            return
        msg = self.lib.gptb_nt_last_error(self.h)
        msg = msg.decode() if msg else ""
        raise (GeodesicNaN if st == L.GEODESIC_NAN else GptbError)(st, msg)

    def close(self):
        if self.h:
            self.lib.gptb_nt_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def features(self, X):
        X = _f64(np.atleast_2d(X))
        out = np.empty((self.n, X.shape[0]), order="F")
        self._check(self.lib.gptb_nt_features(self.h, _dp(X), X.shape[0], _dp(out)))
        return out

    def set_train(self, X, y):
        X = _f64(np.atleast_2d(X)); y = np.ascontiguousarray(np.ravel(y), dtype=np.float64)
        if y.size != X.shape[0]:
            raise ValueError("X and y disagree on N")
        self._check(self.lib.gptb_nt_set_train(self.h, _dp(X), _dp(y), X.shape[0]))
        self.N = X.shape[0]

    def pred(self, theta, Xtest):
        X = _f64(np.atleast_2d(Xtest)); th = np.ascontiguousarray(np.ravel(theta), dtype=np.float64)
        out = np.empty(X.shape[0])
        self._check(self.lib.gptb_nt_pred(self.h, _dp(th), _dp(X), X.shape[0], _dp(out)))
        return out

    def run(self, signal_var, sigma_theta, m, eps_theta, decay_rate, burnin, maxepoch, param_seed, theta0=None,
            perms=None, noise=None):
        T = (int(maxepoch) + int(burnin)) * (-(-self.N // int(m)))
        store = np.zeros((self.n, T), order="F")
        th0 = None if theta0 is None else np.ascontiguousarray(np.ravel(theta0), dtype=np.float64)
        pm = None if perms is None else np.ascontiguousarray(perms, dtype=np.int64)
        nz = None if noise is None else np.ascontiguousarray(noise, dtype=np.float64)     # [T, n] C order = n fastest
        st = self.lib.gptb_nt_run(self.h, float(signal_var), float(sigma_theta), int(m), float(eps_theta), float(decay_rate),
                                  int(burnin), int(maxepoch), int(param_seed), _dp(th0),
                                  pm.ctypes.data_as(L.c_int64_p) if pm is not None else None, _dp(nz), _dp(store))
        if st == L.GEODESIC_NAN:
            print("Get NaN in theta. Try smaller epsilon")        # GPT_SGLD.jl:840
            return np.zeros(self.n)                               # :841
        self._check(st)
        return store

    def last_run_stats(self):
        ms, ln, stp = C.c_double(), C.c_int64(), C.c_int64()
        self._check(self.lib.gptb_nt_last_run_stats(self.h, C.byref(ms), C.byref(ln), C.byref(stp)))
        return ms.value, ln.value, stp.value


def featureNotensor(X, length_scale, sigma_RBF, Z, b) -> NTFeatures:
    """GPT_SGLD.jl:109-120.  Use Z=randn(n,D) and b=2*pi*rand(n) (:108)."""
    return NTFeatures(X, length_scale, sigma_RBF, Z, b)


def GPNT_SGLD(phi: NTFeatures, y, signal_var, sigma_theta, m, eps_theta, decay_rate, burnin, maxepoch, param_seed, *,
              precision: Optional[int] = None, theta0=None, perms=None, noise=None):
    """GPT_SGLD.jl:809-847.  Returns theta_store[n, (maxepoch+burnin)*ceil(N/m)] (every step, burn-in included), or
    zeros(n) after printing the reference's warning if theta became NaN.  `phi` is the lazy handle of featureNotensor.
    Keyword-only extras (not in the reference): precision; theta0 / perms / noise inject the random draws."""
    if not isinstance(phi, NTFeatures):
        raise TypeError("GPNT_SGLD takes the handle returned by featureNotensor (the n x N matrix is never formed)")
    with NTModel(phi, precision=precision, seed=int(param_seed)) as mdl:
        mdl.set_train(phi.X, y)
        return mdl.run(signal_var, sigma_theta, m, eps_theta, decay_rate, burnin, maxepoch, param_seed, theta0, perms, noise)


def pred_notensor(theta, phitest: NTFeatures, precision: Optional[int] = None):
    """phitest' * theta (PowerPlantNoTensorExperiment.jl / kin40kNoTensorExperiment.jl test-time prediction)."""
    with NTModel(phitest, precision=precision) as mdl:
        return mdl.pred(theta, phitest.X)
