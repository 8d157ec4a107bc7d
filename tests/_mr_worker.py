"""Worker of tests/test_gpu_multirank.py: one process per GPU, NCCL unique id exchanged through a file.

    python tests/_mr_worker.py RANK WORLD IDFILE OUT.npz

Runs the scenarios of `drive` below on rank RANK of WORLD (WORLD = 1: no communicator — the reference chain) and
saves every result to OUT.npz."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from gpt_b200 import _lib as L  # noqa: E402
from gpt_b200 import gpt_sgld as G  # noqa: E402
from util import make_problem  # noqa: E402


def problem():
    return make_problem(6, 3000, 64, 8, Q=50, seed=41)


def drive(P, device, world, rank, idbytes, algorithm=L.ALG_DIRECT):
    f = G.feature(P["X"], P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"])
    rng = np.random.default_rng(7)
    T, bs, epsw, epsU = 4, 1001, 1e-6, 1e-9          # 1001 rows: uneven shares
    idx = [rng.permutation(P["N"])[:bs] for _ in range(T)]
    Xb = [np.asfortranarray(P["X"][i]) for i in idx]
    yb = [np.ascontiguousarray(P["y"][i]) for i in idx]
    out = {}

    def build(pipe_groups=None):
        m = G.Model(P["D"], P["n"], P["r"], P["Q"], precision=L.PREC_TC3, device=device, seed=99)
        m.set_features(f); m.set_index(P["I"]); m.set_train(P["X"], P["y"]); m.set_params(P["w0"], P["U0"])
        m.set_option(L.OPT_ALGORITHM, algorithm)
        if pipe_groups is not None:
            m.set_option(L.OPT_PIPE_GROUPS, pipe_groups)
        if world > 1:
            m.comm_init(world, rank, idbytes)
        return m

    # 1. host minibatches (submit / wait, then the blocking entry), then the resident loop with ragged shares
    with build() as m:
        wo = [np.zeros(P["Q"]) for _ in range(T)]
        m.sgld_step_host_submit(P["signal_var"], Xb[0], yb[0], P["N"], epsw, epsU, w_out=wo[0])
        for t in range(1, T):
            m.sgld_step_host_submit(P["signal_var"], Xb[t], yb[t], P["N"], epsw, epsU, w_out=wo[t])
            m.sgld_step_host_wait()
        m.sgld_step_host_wait()
        for t in range(2):
            m.sgld_step_host(P["signal_var"], Xb[t], yb[t], P["N"], epsw, epsU, w_out=wo[t])
        out["host_w"], out["host_U"] = m.get_params()
        out["host_wout"] = np.stack(wo)
        m.sgld_advance(P["signal_var"], 700, epsw, epsU, 5)
        m.synchronize()
        out["adv_w"], out["adv_U"] = m.get_params()
        # 2. fewer rows than ranks: some rank's share is empty
        gw, gU, _ = m.grad(P["signal_var"], start=5, bs=1, want_fhat=False)
        out["g1_w"], out["g1_U"] = gw, gU
        m.sgld_step(P["signal_var"], epsw, epsU, start=9, bs=1)
        out["s1_w"], out["s1_U"] = m.get_params()
        # explicit index list + gptb_grad on an uneven split
        gw, gU, _ = m.grad(P["signal_var"], idx=idx[0][:333], want_fhat=False)
        out["g333_w"], out["g333_U"] = gw, gU
    return out


def drive_pipe(P, device, world, rank, idbytes):
    # 3. the dimension-group pipeline forced on (it is normally single-rank only): per-group all-reduce on the NCCL stream
    f = G.feature(P["X"], P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"])
    with G.Model(P["D"], P["n"], P["r"], P["Q"], precision=L.PREC_TC3, device=device, seed=99) as m:
        m.set_features(f); m.set_index(P["I"]); m.set_train(P["X"], P["y"]); m.set_params(P["w0"], P["U0"])
        m.set_option(L.OPT_ALGORITHM, L.ALG_DIRECT)
        m.set_option(L.OPT_PIPE_GROUPS, 2)
        if world > 1:
            m.comm_init(world, rank, idbytes)
        m.sgld_advance(P["signal_var"], 512, 1e-6, 1e-9, 6)
        m.synchronize()
        w, U = m.get_params()
    return {"pipe_w": w, "pipe_U": U}


def drive_nodes(P, device, world, rank, idbytes):
    # 4. the node-grid algorithm row-sharded
    f = G.feature(P["X"], P["ls"], 1.0, P["phi_scale"], P["Z"], P["b"])
    with G.Model(P["D"], P["n"], P["r"], P["Q"], precision=L.PREC_TC3, device=device, seed=99) as m:
        m.set_features(f); m.set_index(P["I"]); m.set_train(P["X"], P["y"]); m.set_params(P["w0"], P["U0"])
        m.set_option(L.OPT_ALGORITHM, L.ALG_NODES)
        if world > 1:
            m.comm_init(world, rank, idbytes)
        gw, gU, _ = m.grad(P["signal_var"], start=0, bs=2999, want_fhat=False)
        m.sgld_advance(P["signal_var"], 1500, 1e-6, 1e-9, 4)
        m.synchronize()
        w, U = m.get_params()
        # pipelined host minibatches on the node grid: each rank stages its share only, and the batch's transpose, gather
        # and cell sort run on the front stream under the previous step's all-reduce and update
        rng = np.random.default_rng(3)
        keep = []
        for t in range(4):
            i = rng.permutation(P["N"])[:1501]
            keep.append((np.asfortranarray(P["X"][i]), np.ascontiguousarray(P["y"][i])))
            m.sgld_step_host_submit(P["signal_var"], keep[-1][0], keep[-1][1], P["N"], 1e-6, 1e-9)
            if t >= 1:
                m.sgld_step_host_wait()
        m.sgld_step_host_wait()
        wh, Uh = m.get_params()
    return {"nodes_gw": gw, "nodes_gU": gU, "nodes_w": w, "nodes_U": U, "nodes_host_w": wh, "nodes_host_U": Uh}


def get_id(world, rank, idfile, tag):
    """Rank 0 makes a NCCL unique id and publishes it atomically; the others poll for the file."""
    path = f"{idfile}.{tag}"
    if world == 1:
        return None
    if rank == 0:
        b = G.comm_unique_id()
        with open(path + ".tmp", "wb") as fh:
            fh.write(b)
        os.replace(path + ".tmp", path)
        return b
    t0 = time.time()
    while not os.path.exists(path):
        if time.time() - t0 > 120:
            raise TimeoutError("no NCCL id from rank 0")
        time.sleep(0.05)
    return open(path, "rb").read()


def main():
    rank, world, idfile, outp = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4]
    device = rank
    G.set_default_device(device)
    P = problem()
    res = {}
    res.update(drive(P, device, world, rank, get_id(world, rank, idfile, "a")))
    res.update(drive_pipe(P, device, world, rank, get_id(world, rank, idfile, "b")))
    res.update(drive_nodes(P, device, world, rank, get_id(world, rank, idfile, "c")))
    np.savez(outp, **res)


if __name__ == "__main__":
    main()
