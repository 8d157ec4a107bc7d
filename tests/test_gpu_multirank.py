"""Two ranks on two real GPUs through libgptb200's own communicator entry (gptb_comm_init): the row-sharded chain must
equal the one-rank chain (same global minibatch) and every rank must hold bit-identical parameters.
Skipped on a box with fewer than two GPUs (run it with `gpurun --gpus 2`; the log is kept under profiles/)."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from util import relerr

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _run(world, tmp):
    idfile = os.path.join(tmp, f"ncclid_w{world}")
    outs = [os.path.join(tmp, f"out_w{world}_r{r}.npz") for r in range(world)]
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "_mr_worker.py"), str(r), str(world), idfile, outs[r]])
             for r in range(world)]
    for p in procs:
        assert p.wait(timeout=600) == 0
    return [dict(np.load(o)) for o in outs]


@pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs")
def test_two_rank_chain_equals_one_rank_chain():
    from _mr_worker import problem
    P = problem()
    with tempfile.TemporaryDirectory() as tmp:
        ref = _run(1, tmp)[0]
        got = _run(2, tmp)
    w0, U0 = P["w0"], P["U0"]
    tol = 2e-5   # of the accumulated update: FP32 atomics / a different summation split, not a different chain
    for key in ("host", "adv", "s1", "pipe", "nodes", "nodes_host"):
        assert relerr(got[0][f"{key}_w"] - w0, ref[f"{key}_w"] - w0) < tol, key
        assert relerr(got[0][f"{key}_U"] - U0, ref[f"{key}_U"] - U0) < tol, key
    assert relerr(got[0]["host_wout"] - w0, ref["host_wout"] - w0) < tol
    for key in ("g1", "g333"):
        assert relerr(got[0][f"{key}_w"], ref[f"{key}_w"]) < 1e-5 and relerr(got[0][f"{key}_U"], ref[f"{key}_U"]) < 1e-5, key
    assert relerr(got[0]["nodes_gw"], ref["nodes_gw"]) < 1e-5 and relerr(got[0]["nodes_gU"], ref["nodes_gU"]) < 1e-5
    # no broadcast ever happens: the ranks stay together only if every reduced buffer and update is bit-identical
    for key in got[0]:
        assert np.array_equal(got[0][key], got[1][key]), f"{key} differs between ranks"
