#!/usr/bin/env python
"""Two-or-more-rank check of the row-sharded step on real GPUs (run under torch.distributed.run): the chain driven
through the host-minibatch entries (each rank copies only its rows of the global minibatch) and through the resident
loop must match the same chain on one rank (rank 0 runs it without a communicator).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 \
        tools/check_multirank_host.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from gpt_b200 import _lib as L  # noqa: E402
from gpt_b200 import gpt_sgld as G  # noqa: E402
from util import make_problem, relerr  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    P = make_problem(6, 3000, 64, 8, Q=50, seed=41)
    f = G.feature(P["X"], P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"])
    rng = np.random.default_rng(7)
    T, bs, epsw, epsU = 4, 1001, 1e-6, 1e-9          # 1001 rows: uneven shares
    idx = [rng.permutation(P["N"])[:bs] for _ in range(T)]
    Xb = [np.asfortranarray(P["X"][i]) for i in idx]
    yb = [np.ascontiguousarray(P["y"][i]) for i in idx]

    def build(with_comm):
        m = G.Model(P["D"], P["n"], P["r"], P["Q"], precision=L.PREC_TC3, device=local, seed=99)
        m.set_features(f); m.set_index(P["I"]); m.set_train(P["X"], P["y"]); m.set_params(P["w0"], P["U0"])
        if with_comm:
            idt = torch.zeros(128, dtype=torch.uint8, device=dev)
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(G.comm_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            m.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))
        return m

    def drive(m):
        wo = [np.zeros(P["Q"]) for _ in range(T)]
        m.sgld_step_host_submit(P["signal_var"], Xb[0], yb[0], P["N"], epsw, epsU, w_out=wo[0])
        for t in range(1, T):
            m.sgld_step_host_submit(P["signal_var"], Xb[t], yb[t], P["N"], epsw, epsU, w_out=wo[t])
            m.sgld_step_host_wait()
        m.sgld_step_host_wait()
        for t in range(2):                                       # and the blocking entry
            m.sgld_step_host(P["signal_var"], Xb[t], yb[t], P["N"], epsw, epsU, w_out=wo[t])
        a = m.get_params()
        m.sgld_advance(P["signal_var"], 700, epsw, epsU, 5)      # resident loop, ragged shares
        m.synchronize()
        return a, m.get_params(), np.stack(wo)

    mm = build(True)
    got = drive(mm)
    mm.close()
    ok = True
    if rank == 0:
        ms = build(False)
        ref = drive(ms)
        ms.close()
        errs = [relerr(got[0][0] - P["w0"], ref[0][0] - P["w0"]), relerr(got[0][1] - P["U0"], ref[0][1] - P["U0"]),
                relerr(got[1][0] - P["w0"], ref[1][0] - P["w0"]), relerr(got[1][1] - P["U0"], ref[1][1] - P["U0"]),
                relerr(got[2] - P["w0"], ref[2] - P["w0"])]
        ok = max(errs) < 2e-5
        print(f"multi-rank ({world}) vs one rank: rel.err of the accumulated updates "
              f"host w {errs[0]:.2e} U {errs[1]:.2e} | +resident w {errs[2]:.2e} U {errs[3]:.2e} | w_out {errs[4]:.2e} -> "
              f"{'OK' if ok else 'MISMATCH'}", flush=True)
    # every rank must hold the same parameters
    w = torch.tensor(got[1][0], device=dev)
    wmax = w.clone(); dist.all_reduce(wmax, op=dist.ReduceOp.MAX)
    wmin = w.clone(); dist.all_reduce(wmin, op=dist.ReduceOp.MIN)
    same = bool((wmax == wmin).all().item())
    if rank == 0:
        print("parameters identical on all ranks:", same, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if (ok and same) else 1)


if __name__ == "__main__":
    main()
