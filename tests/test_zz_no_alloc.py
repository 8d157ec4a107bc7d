"""gptb_proj_noalloc (no counterpart in the reference; added for callers that cannot allocate): its result must equal
the existing gptb_proj's, on two calls in a row (the static device scratch keeps the previous call's sums and must be
cleared), and sizes beyond the static scratch must be rejected.  Collected last on purpose (file name), so that a
failure here cannot hide another test under `pytest -x`.  Needs a B200."""
import ctypes as C

import numpy as np
import pytest

from gpt_b200 import _lib as L
from gpt_b200 import gpt_sgld as G
from util import relerr

pytestmark = pytest.mark.gpu


def _dev(a, torch):
    """column-major (n, r) float64 on the device: the transpose of a C-contiguous (r, n) tensor"""
    return torch.from_numpy(np.ascontiguousarray(a.T)).cuda()


def test_proj_noalloc_matches_proj_twice_in_a_row():
    import torch
    torch.cuda.set_device(0)
    G.set_default_device(0)
    lib = L.load()
    rng = np.random.default_rng(23)
    stream = torch.cuda.Stream()
    # second call: other sizes and other data than the first, so stale sums of call 1 would show
    for n, r in ((1000, 32), (37, 5), (640, 64)):
        U = np.linalg.qr(rng.standard_normal((n, r)))[0]
        V = rng.standard_normal((n, r))
        want = G.proj(U, V)                                   # the existing entry point (gptb_proj)
        dU, dV = _dev(U, torch), _dev(V, torch)
        dO = torch.full((r, n), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        for _ in range(2):                                    # ... and the same call twice on one stream
            st = lib.gptb_proj_noalloc(n, r, C.c_void_p(dU.data_ptr()), C.c_void_p(dV.data_ptr()), C.c_void_p(dO.data_ptr()),
                                       C.c_void_p(stream.cuda_stream))
            assert st == L.OK, lib.gptb_last_error(None)
        stream.synchronize()
        assert torch.cuda.memory_allocated() == before
        got = dO.cpu().numpy().T
        assert relerr(got, want) < 1e-12                      # the tolerance test_proj_geod_vs_oracle uses for proj
        assert np.abs(U.T @ got + got.T @ U).max() < 1e-11    # tangent vector: U'M skew-symmetric
    # beyond the static scratch: a clear error, nothing launched
    dU = torch.zeros((65, 70), dtype=torch.float64, device="cuda")
    st = lib.gptb_proj_noalloc(70, 65, C.c_void_p(dU.data_ptr()), C.c_void_p(dU.data_ptr()), C.c_void_p(dU.data_ptr()), None)
    assert st == L.ERR_UNSUPPORTED and b"r <= 64" in lib.gptb_last_error(None)
    assert lib.gptb_proj_noalloc(0, 3, C.c_void_p(dU.data_ptr()), C.c_void_p(dU.data_ptr()), C.c_void_p(dU.data_ptr()), None) == L.ERR_ARG
