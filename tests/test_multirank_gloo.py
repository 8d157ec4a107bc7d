"""world_size-2 gloo test (CPU) of the data-parallel step: each rank computes the gradient of its row share of
the global minibatch (with the oracle standing in for the device kernels), the [gradw|gradU] buffer is
all-reduced, and every rank applies the same update — the result must equal the single-rank step."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
    import gpt_oracle as o
    from gpt_b200.gpt_sgld import shard_range
    from util import make_problem
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P = make_problem(3, 64, 8, 3, Q=12, seed=4)
    bs, N = 50, 64
    phi = P["phi_fn"](P["X"][:bs]); y = P["y"][:bs]
    lo, hi = shard_range(rank, world, bs)
    # per-rank likelihood gradient of its rows, scaled with the GLOBAL batch size (N/bs of :393)
    V = o.computeV(o.phidotU(P["U0"], phi[:, :, lo:hi]), P["I"])
    fh = o.computefhat(V, P["w0"])
    e = (N / bs) * (y[lo:hi] - fh) / 0.01
    gw = V @ e
    A = o.computeA(o.computeU_phi(V, o.phidotU(P["U0"], phi[:, :, lo:hi]), P["I"]), P["w0"], P["I"], P["r"])
    gU = np.einsum("jki,lki,i->jlk", phi[:, :, lo:hi], A, e)
    buf = torch.from_numpy(np.concatenate([gw, gU.ravel(order="F")]))
    dist.all_reduce(buf)
    g = buf.numpy()
    gw_all = g[:P["Q"]] - P["w0"]                                         # prior once, after the reduction
    gU_all = g[P["Q"]:].reshape(P["n"], P["r"], P["D"], order="F")
    ref = o.gradients(phi, y, P["w0"], P["U0"], P["I"], N, 0.01)
    ok = np.allclose(gw_all, ref[0], rtol=1e-10, atol=1e-10) and np.allclose(gU_all, ref[1], rtol=1e-10, atol=1e-10)
    # bitwise-identical reduced buffers on both ranks -> identical parameters without a broadcast
    gathered = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(gathered, buf)
    same = all(torch.equal(gathered[0], t) for t in gathered)
    if rank == 0:
        out.put((bool(ok), bool(same), (lo, hi)))
    dist.destroy_process_group()


def test_two_rank_gradient_allreduce():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, same, rng = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert ok and same and rng == (0, 25)


def test_shard_range_covers_batch():
    from gpt_b200.gpt_sgld import shard_range
    for bs in (1, 7, 500, 65536, 65537):
        for G in (1, 2, 3, 4, 8):
            edges = [shard_range(g, G, bs) for g in range(G)]
            assert edges[0][0] == 0 and edges[-1][1] == bs
            assert all(edges[i][1] == edges[i + 1][0] for i in range(G - 1))
