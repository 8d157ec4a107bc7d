"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol include/gptb200.h declares,
fails loudly without a GPU (no CPU fallback), and the host module keeps the reference's conventions."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import gpt_oracle as o
from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _have_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "gptb200.h")).read()
    declared = set(re.findall(r"\b(gptb_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"gptb_handle", "gptb_config"}
    assert len(declared) >= 25
    lib = L.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in gptb200.h but not exported"
        assert name in L.PROTOTYPES, f"{name} has no ctypes prototype"
    assert set(L.PROTOTYPES) == declared
    assert lib.gptb_version() == 100


def test_no_python_fallback_in_product():
    """The product path must not import the oracle."""
    for fn in os.listdir(os.path.join(ROOT, "gpt_b200")):
        if fn.endswith(".py"):
            src = open(os.path.join(ROOT, "gpt_b200", fn)).read()
            assert "oracle" not in src.replace("# oracle", ""), fn


@pytest.mark.skipif(_have_cuda(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu():
    lib = L.load()
    cfg = L.Config(D=2, n=4, r=2, Q=3, precision=L.PREC_FP64, feature_kind=0, device=-1, reserved=0, seed=0)
    h = L.H()
    assert lib.gptb_create(C.byref(cfg), C.byref(h)) == L.ERR_CUDA
    assert b"no CPU fallback" in lib.gptb_last_error(None)
    with pytest.raises(G.GptbError):
        G.proj(np.eye(4, 2), np.ones((4, 2)))
    with pytest.raises(G.GptbError):
        G.pred(np.ones(3), np.zeros((4, 2, 2)), np.ones((3, 2), dtype=np.int32), np.zeros((4, 2, 5)))


def test_create_argument_checks():
    lib = L.load()
    h = L.H()
    bad = L.Config(D=0, n=4, r=2, Q=3, precision=0, feature_kind=0, device=-1, reserved=0, seed=0)
    assert lib.gptb_create(C.byref(bad), C.byref(h)) == L.ERR_SHAPE
    odd = L.Config(D=2, n=5, r=2, Q=3, precision=0, feature_kind=L.FEAT_SINCOS, device=-1, reserved=0, seed=0)
    assert lib.gptb_create(C.byref(odd), C.byref(h)) == L.ERR_SHAPE
    assert b"n not even" in lib.gptb_last_error(None)                    # GPT_SGLD.jl:90
    assert lib.gptb_create(None, C.byref(h)) == L.ERR_ARG


def test_samplenz_matches_reference_convention():
    rng = np.random.default_rng(11)
    I = G.samplenz(4, 3, 20, rng)
    assert I.dtype == np.int32 and I.shape == (20, 3) and I.flags.f_contiguous
    cells = ((I.astype(np.int64) - 1) * 4 ** np.arange(3)).sum(1)
    assert len(set(cells.tolist())) == 20 and cells.max() < 4 ** 3 and I.min() >= 1 and I.max() <= 4
    I = G.samplenz(32, 10, 200, rng)
    assert len({tuple(r) for r in I}) == 200
    with pytest.raises(ValueError):
        G.samplenz(2, 2, 5, rng)
    # same generator state -> same table as the oracle's restatement
    a = G.samplenz(3, 4, 30, np.random.default_rng(5))
    b = o.samplenz(3, 4, 30, np.random.default_rng(5))
    assert np.array_equal(a, b)


def test_datawhitening_and_feature_handles():
    rng = np.random.default_rng(0)
    X = rng.standard_normal((50, 3)) * [1, 5, 0.1] + [3, -2, 7]
    Y = G.datawhitening(X.copy())
    assert np.allclose(Y.mean(0), 0, atol=1e-12) and np.allclose(Y.std(0, ddof=1), 1)
    assert np.allclose(Y, o.datawhitening(X.copy()))
    Z = rng.standard_normal((6, 3)); b = rng.random((6, 3))
    f = G.feature(X, [1.0, 2.0, 3.0], 1.0, 1.0, Z, b)
    assert f.shape == (6, 3, 50)
    assert G.feature(X, 1.5, 1.0, 1.0, Z, b).length_scale.size == 1      # scalar length scale (:71)
    with pytest.raises(ValueError):
        G.feature(X, [1.0, 2.0], 1.0, 1.0, Z, b)
    with pytest.raises(ValueError):
        G.feature2(X, 1.0, 1.0, 1.0, Z[:5])                               # n not even
    with pytest.raises(ValueError):
        G.feature(X, 1.0, 1.0, 1.0, Z, b[:5])


def test_julia_shim_binds_declared_symbols_with_matching_arity():
    """julia/GPT_SGLD.jl cannot be executed in this image (no julia binary): statically check that every ccall names a
    symbol include/gptb200.h declares and passes as many argument types as the C prototype has parameters."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "gptb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    protos = {}
    for mt in re.finditer(r"\b(?:int|const char \*)\s*(gptb_\w+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        args = mt.group(2).strip()
        protos[mt.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    jl = open(os.path.join(root, "julia", "GPT_SGLD.jl")).read()
    calls = re.findall(r"ccall\(\(:(gptb_\w+), LIB\),\s*\w+,\s*\(((?:[^()]|\([^()]*\)|\{[^{}]*\})*?)\)\s*,", jl, flags=re.S)
    assert len(calls) > 40
    seen = set()
    for name, types in calls:
        assert name in protos, f"{name} is not declared in include/gptb200.h"
        depth, n = 0, 1 if types.strip() else 0
        for ch in types:
            depth += ch in "({"
            depth -= ch in ")}"
            n += ch == "," and depth == 0
        if types.strip().endswith(","):       # one-element tuple: (T,)
            n -= 1
        assert n == protos[name], f"{name}: {n} ccall argument types, {protos[name]} C parameters"
        seen.add(name)
    # every entry a reference-style script needs is bound (introspection / bench-only / device-pointer entries are not)
    must = {"gptb_create", "gptb_destroy", "gptb_set_features", "gptb_set_index", "gptb_set_train", "gptb_set_train_phi",
            "gptb_set_params", "gptb_get_params", "gptb_init_params", "gptb_features", "gptb_pred", "gptb_pred_phi",
            "gptb_pred_samples", "gptb_grad", "gptb_proj", "gptb_geod", "gptb_geodboth", "gptb_sgld_step", "gptb_sgld_run",
            "gptb_sgld_advance", "gptb_gmc_run", "gptb_set_option", "gptb_set_domain", "gptb_comm_unique_id", "gptb_comm_init",
            "gptb_synchronize", "gptb_nt_create", "gptb_nt_run", "gptb_nt_features"}
    assert must <= seen, must - seen
