#!/usr/bin/env python
"""bench.py — SGLD steps/s of the tensor-RFF GP hot path on N B200s (and the C5 prediction sweep).

    python bench.py --gpus 1 --steps 100 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference path's CPU restatement on the host cores

Workload = BASELINE.json configs[3] "C4" (the shape the north star's scaling target is quoted on; it fits one
GPU): D=10, N=10M training rows, n=1000 RFF/dim, r=32, Q=200, minibatch 65536.  A step is one loop body of
GPTregression (GPT_SGLD.jl:378-439): gather -> K1 forward -> K2 contraction -> K3 factor gradient ->
(all-reduce) -> K5 SGLD/Stiefel update.  Multi-GPU = the same global minibatch row-sharded over ranks
(strong scaling), gradients all-reduced with NCCL.

Prints ONE JSON line (rank 0).  `value` = steps/s with the training set resident in HBM (device-timed, max over
ranks); `e2e` = the same loop body through gptb_sgld_step_host_submit/_wait with the minibatch in pinned HOST memory
(H2D + D2H inside the timed region)."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

WORKLOADS = {
    "C1": dict(D=4, N=9568, n=50, r=10, m=500, desc="PowerPlant CCPP shape"),
    "C2": dict(D=8, N=10000, n=100, r=20, m=1000, desc="kin40k shape"),
    "C3": dict(D=10, N=32000, n=500, r=20, m=1000, desc="tensorsynth 10D"),
    "C4": dict(D=10, N=10_000_000, n=1000, r=32, m=65536, desc="synthetic D=10, N=10M"),
}
C5 = dict(D=8, Nt=1_000_000, S=200, n=500, r=20)
Q = 200
SIGNAL_VAR = 1e-3                            # SynthDataExperiment.jl:7


def eps_for(N):
    """Step sizes.  SynthDataExperiment.jl:37-38 uses epsw=1e-5, epsU=1e-8 on Ntrain=50 points; the likelihood
    gradient scales with N/signal_var, so the same eps*N product keeps the chain stable at any N (the 50-point
    values diverge to NaN within a few steps at N=10M)."""
    return 1e-5 * 50.0 / N, 1e-8 * 50.0 / N


def synth_model(D, n, r, seed):
    """Model draws in the reference's conventions."""
    from gpt_b200 import gpt_sgld as G
    rng = np.random.default_rng(seed)
    ls = np.maximum(1.0 + 0.2 * rng.standard_normal(D), 0.3)
    Z = np.asfortranarray(rng.standard_normal((n, D)))
    b = np.asfortranarray(2 * np.pi * rng.random((n, D)))
    phi_scale = np.sqrt(n / Q ** (1.0 / D))
    I = G.samplenz(r, D, Q, rng)
    # uniform draws from the Stiefel manifold V_{n,r} (QR of a Gaussian matrix, sign-fixed)
    def stie():
        out = np.empty((n, r, D), order="F")
        for k in range(D):
            q_, r_ = np.linalg.qr(rng.standard_normal((n, r)))
            out[:, :, k] = q_ * np.sign(np.diag(r_))
        return out
    return dict(ls=ls, Z=Z, b=b, phi_scale=phi_scale, I=I, w_true=rng.standard_normal(Q), U_true=stie(),
                w0=rng.standard_normal(Q), U0=stie(), rng=rng)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index):
        self.rows, self.stop, self.idx = [], threading.Event(), gpu_index
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def host_threads():
    """All host cores, whatever OMP_NUM_THREADS says (torch.distributed.run exports OMP_NUM_THREADS=1 to every rank)."""
    return max(1, os.cpu_count() or 1)


def cpu_reference_timing(W, mdl, rows, threads, budget_s, X=None, y=None, seed=7):
    """Times the reference path's CPU restatement (oracle/gpt_oracle.c: FP64, OpenMP over minibatch rows like
    GPT_SGLD_p.jl:70-108) on `rows` rows of the workload's minibatch with `threads` threads.

    The loop body of GPT_SGLD.jl:378-439 has a row-proportional part (phidotU .. grad-U, :384-408) and a fixed part
    (w update, D x proj/geod with their expm calls, :411-439).  They are timed separately — t_grad on `rows` rows,
    t_update as (step - gradient) on a 64-row batch — so that a scaled figure multiplies only the row-proportional
    part.  The features of the rows are precomputed and untimed, as in the reference (phi is formed once per run, :71)."""
    import c_oracle as co
    D, n, r, m = W["D"], W["n"], W["r"], W["m"]
    rows = int(min(rows, m))
    rng = np.random.default_rng(seed)
    EPSW, EPSU = eps_for(W["N"])
    if X is None:       # the workload's own generative process: X ~ N(0,1), y = f*(X) + noise
        X = rng.standard_normal((rows, D))
    X = np.ascontiguousarray(X[:rows])
    phi = co.feature(X, mdl["ls"], 1.0, mdl["phi_scale"], mdl["Z"], mdl["b"], threads)
    if y is None:
        y = co.pred(mdl["w_true"], mdl["U_true"], mdl["I"], phi, threads) + np.sqrt(SIGNAL_VAR) * rng.standard_normal(rows)
    y = np.ascontiguousarray(y[:rows])
    nw = rng.standard_normal(Q); nU = rng.standard_normal((n, r, D))
    w, U = mdl["w0"].copy(), mdl["U0"].copy()

    def timed(fn, min_reps, max_reps, budget):
        fn()                                   # warm (page faults, thread pool)
        t0 = time.perf_counter(); fn(); t1 = time.perf_counter() - t0
        reps = int(max(min_reps, min(max_reps, budget / max(t1, 1e-4))))
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        return (time.perf_counter() - t0) / reps, reps

    t_grad, reps = timed(lambda: co.gradients(phi, y, w, U, mdl["I"], W["N"], SIGNAL_VAR, threads), 1, 20, budget_s)
    small = slice(0, min(64, rows))
    phis = np.asfortranarray(phi[:, :, small]); ys = y[small]
    t_s, _ = timed(lambda: co.sgld_step(phis, ys, w, U, mdl["I"], W["N"], SIGNAL_VAR, EPSW, EPSU, nw, nU, True, threads), 3, 20, 2.0)
    t_g, _ = timed(lambda: co.gradients(phis, ys, w, U, mdl["I"], W["N"], SIGNAL_VAR, threads), 3, 20, 1.0)
    t_update = max(t_s - t_g, 0.0)
    return {"rows": rows, "threads": threads, "t_grad_s": t_grad, "t_update_s": t_update, "reps": reps,
            "t_step_full_s": t_grad * (m / rows) + t_update, "extrapolated": rows < m}


def cpu_baseline_object(W, mdl, rows_all, rows_one, budget_s, X=None, y=None):
    """cpu_baseline of the JSON line: all host cores (the value) and one thread (the reference module itself is
    single-threaded apart from BLAS), both as gradient-rows x (m/rows) + update."""
    thr = host_threads()
    a = cpu_reference_timing(W, mdl, rows_all, thr, budget_s, X, y)
    o1 = cpu_reference_timing(W, mdl, rows_one, 1, budget_s / 4, X, y)
    m = W["m"]
    return {"value": 1.0 / a["t_step_full_s"], "unit": "steps/s", "cores": thr, "kind": "port",
            "extrapolated": a["extrapolated"], "measured_rows": a["rows"],
            "t_grad_measured_s": a["t_grad_s"], "t_update_measured_s": a["t_update_s"],
            "one_thread": {"value": 1.0 / o1["t_step_full_s"], "measured_rows": o1["rows"], "extrapolated": o1["extrapolated"],
                           "t_grad_measured_s": o1["t_grad_s"], "t_update_measured_s": o1["t_update_s"]},
            "sample": f"oracle/gpt_oracle.c (C restatement of GPT_SGLD.jl:378-439; the Julia 0.3 reference cannot run here): "
                      f"gradient part timed over {a['reps']} repetitions on {a['rows']} of the {m} minibatch rows "
                      f"({thr} OpenMP threads) and multiplied by {m / a['rows']:g}; the row-independent update "
                      f"(w, {W['D']} x proj/geod/expm) timed once and added unscaled; features precomputed (untimed) as "
                      f"the reference does"}, a


def run_reference(args, W, emit):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import c_oracle as co
    mdl = synth_model(W["D"], W["n"], W["r"], seed=3)
    K, Wm = max(1, args.steps), max(0, args.warmup)
    t0 = time.perf_counter()
    thr = host_threads()
    # how many rows of the minibatch fit the run: probe 1024 rows, then take the whole minibatch if K+W such steps end
    # within ~150 s, else the largest power-of-two share that does
    probe = cpu_reference_timing(W, mdl, 1024, thr, 1.0)
    per_row = probe["t_grad_s"] / probe["rows"]
    rows = W["m"]
    while rows > 1024 and (K + Wm) * (per_row * rows + probe["t_update_s"]) > 150.0:
        rows //= 2
    D, n, r, m = W["D"], W["n"], W["r"], W["m"]
    rng = np.random.default_rng(7)
    X = rng.standard_normal((rows, D))
    phi = co.feature(X, mdl["ls"], 1.0, mdl["phi_scale"], mdl["Z"], mdl["b"], thr)
    y = co.pred(mdl["w_true"], mdl["U_true"], mdl["I"], phi, thr) + np.sqrt(SIGNAL_VAR) * rng.standard_normal(rows)
    EPSW, EPSU = eps_for(W["N"])
    w, U = mdl["w0"].copy(), mdl["U0"].copy()
    nw = rng.standard_normal(Q); nU = rng.standard_normal((n, r, D))
    for _ in range(Wm):
        w, U, _, _ = co.sgld_step(phi, y, w, U, mdl["I"], W["N"], SIGNAL_VAR, EPSW, EPSU, nw, nU, True, thr)
    ts = time.perf_counter()
    for _ in range(K):
        w, U, _, _ = co.sgld_step(phi, y, w, U, mdl["I"], W["N"], SIGNAL_VAR, EPSW, EPSU, nw, nU, True, thr)
    t_step = (time.perf_counter() - ts) / K                  # one loop body on `rows` rows
    t_full = (t_step - probe["t_update_s"]) * (m / rows) + probe["t_update_s"] if rows < m else t_step
    v = 1.0 / t_full
    one = cpu_reference_timing(W, mdl, 1024, 1, 3.0)
    line = {
        "impl": "reference", "metric": "sgld_steps_per_sec", "value": v, "unit": "steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, W),
        "cpu_baseline": {"value": v, "unit": "steps/s", "cores": thr, "kind": "port", "extrapolated": rows < m,
                         "measured_rows": rows, "measured_ms_per_step_on_those_rows": 1e3 * t_step,
                         "t_update_measured_s": probe["t_update_s"],
                         "one_thread": {"value": 1.0 / one["t_step_full_s"], "measured_rows": one["rows"], "extrapolated": True},
                         "sample": f"{K} timed loop bodies (after {Wm} warm-up) of oracle/gpt_oracle.c on {rows} of the {m} "
                                   f"minibatch rows with {thr} OpenMP threads"
                                   + ("" if rows == m else f"; row-proportional part scaled x{m / rows:g}, the update added unscaled")
                                   + "; features precomputed (untimed) as the reference does"},
        "e2e": {"value": v, "unit": "steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference Julia 0.3 runtime unavailable on this box; timed: oracle/gpt_oracle.c (loop-faithful C "
                "restatement of GPT_SGLD.jl:378-439, OpenMP over minibatch rows like GPT_SGLD_p.jl:70-108)",
        "wall_s": time.perf_counter() - t0,
    }
    emit(line)


def workload_config(args, W):
    return {"workload": f"{args.workload}: {W['desc']}, D={W['D']} N={W['N']} n={W['n']} r={W['r']} Q={Q} minibatch={W['m']}",
            "precision": args.precision, "langevin": True, "stiefel": True,
            "signal_var": SIGNAL_VAR, "epsw": eps_for(W["N"])[0], "epsU": eps_for(W["N"])[1],
            "l2": "inputs larger than L2: resident X is %.0f MB and every step gathers different random rows" % (W["N"] * W["D"] * 8 / 1e6),
            "algorithm": args.alg, "options": list(getattr(args, "opt", [])),
            "parallelism": f"minibatch rows sharded x{args.gpus}; NCCL all-reduce of the FP32 gradient buffer "
                           f"(32*ceil32(Q) + D*npad*rpad floats) once per step; replicated FP64 update"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C4", choices=list(WORKLOADS))
    ap.add_argument("--precision", default="tc3", choices=["tc3", "tc1", "fp64"])
    ap.add_argument("--no-pred", action="store_true", help="skip the C5 prediction sweep")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline")
    ap.add_argument("--pred-samples", type=int, default=C5["S"])
    ap.add_argument("--opt", action="append", default=[], metavar="ID=VALUE",
                    help="gptb_set_option(ID, VALUE) on the benchmarked handle (experiments; the line records it)")
    ap.add_argument("--replicas", type=int, default=1,
                    help="R > 1: R independent chains (handles with their own streams and seeds, one host thread each) on "
                         "one GPU; prints their aggregate steps/s instead of the single-chain line (small workloads)")
    ap.add_argument("--alg", default="auto", choices=["auto", "direct", "nodes"],
                    help="gptb_set_option(GPTB_OPT_ALGORITHM): auto picks the node grid for large minibatches")
    args = ap.parse_args()
    W = WORKLOADS[args.workload]
    # stdout carries exactly ONE line (the JSON): libraries that print to fd 1 (NCCL's "NCCL version ..." banner under
    # torchrun, for one) are sent to stderr for the whole run
    sys.stdout.flush()
    _json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(_json_fd, (json.dumps(obj) + "\n").encode())

    if args.impl == "reference":
        return run_reference(args, W, emit)

    import torch
    from gpt_b200 import _lib as L
    from gpt_b200 import gpt_sgld as G

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if rank == 0:
            print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run", file=sys.stderr)
        if args.gpus > 1 and world == 1:
            sys.exit(2)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    prec = {"tc3": L.PREC_TC3, "tc1": L.PREC_TC1, "fp64": L.PREC_FP64}[args.precision]
    D, N, n, r, m = W["D"], W["N"], W["n"], W["r"], W["m"]
    EPSW, EPSU = eps_for(N)
    mdl = synth_model(D, n, r, seed=3)
    rng = np.random.default_rng(11)          # same data on every rank
    X = np.asfortranarray(rng.standard_normal((N, D)))
    feat = G.feature(X, mdl["ls"], 1.0, mdl["phi_scale"], mdl["Z"], mdl["b"])

    model = G.Model(D, n, r, Q, precision=prec, device=local, seed=1234)
    stream = torch.cuda.Stream(device=dev)
    model.set_stream(stream.cuda_stream)
    model.set_features(feat)
    model.set_index(mdl["I"])
    # y = f*(X) + sqrt(signal_var) noise, f* evaluated with the library itself (data generation, untimed)
    y = model.pred(mdl["w_true"], mdl["U_true"], X) + np.sqrt(SIGNAL_VAR) * rng.standard_normal(N)
    model.set_train(X, y)
    model.set_params(mdl["w0"], mdl["U0"])
    model.set_option(L.OPT_ALGORITHM, {"auto": L.ALG_AUTO, "direct": L.ALG_DIRECT, "nodes": L.ALG_NODES}[args.alg])
    for kv in args.opt:
        k_, v_ = kv.split("=")
        model.set_option(int(k_), float(v_))
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(G.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        model.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.replicas > 1:
        # R independent chains — the reference's sweeps over seeds / step sizes (kin40kExperiment.jl:67-91) map to one
        # handle per worker.  Every handle queues its steps asynchronously on its own stream; one host thread per handle
        # (ctypes releases the GIL) keeps R launch queues busy, so the small kernels of different chains share the SMs.
        import threading
        R = args.replicas
        reps = [model]
        for i in range(1, R):
            mi = G.Model(D, n, r, Q, precision=prec, device=local, seed=1234 + i)
            mi.set_features(feat); mi.set_index(mdl["I"]); mi.set_train(X, y); mi.set_params(mdl["w0"], mdl["U0"])
            reps.append(mi)
        model.set_stream(None)

        def run(mi, k):
            mi.sgld_advance(SIGNAL_VAR, m, EPSW, EPSU, k)
            mi.synchronize()
        for mi in reps:
            run(mi, max(3, args.warmup))
        t1 = time.perf_counter()
        run(reps[0], args.steps)
        single = args.steps / (time.perf_counter() - t1)
        torch.cuda.synchronize()
        with ClockSampler(local) as clk:
            t0 = time.perf_counter()
            th = [threading.Thread(target=run, args=(mi, args.steps)) for mi in reps]
            for t_ in th:
                t_.start()
            for t_ in th:
                t_.join()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        finite = all(bool(np.isfinite(mi.get_params()[0]).all()) for mi in reps)
        emit({"metric": "sgld_steps_per_sec", "value": R * args.steps / dt, "unit": "steps/s (sum over chains)", "n_gpus": 1,
              "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 / args.steps, "higher_is_better": True,
              "scaling": "replicas", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
              "config": dict(workload_config(args, W), replicas=R, timing="wall clock around R host threads, device synchronised on both sides"),
              "single_chain_steps_per_sec": single, "finite": finite, "clocks": clk.summary()})
        for mi in reps[1:]:
            mi.close()
        return

    # ------------------------------------------------------------------ resident-data steps/s (device timed)
    model.sgld_advance(SIGNAL_VAR, m, EPSW, EPSU, max(3, args.warmup))
    model.synchronize()
    model.phase_ms(reset=True)               # zero the launch counter
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    # clocks and throttle reasons are sampled from here to the end of the last timed region (steps, e2e, stored chain,
    # posterior sweep): the K timed steps alone last a few milliseconds, less than one nvidia-smi query
    clk = ClockSampler(local)
    clk.__enter__()
    with torch.cuda.stream(stream):
        ev0.record(stream)
        model.sgld_advance(SIGNAL_VAR, m, EPSW, EPSU, args.steps)   # returns with the stream ordered after the last update
        ev1.record(stream)
    model.synchronize()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    _, launches = model.phase_ms(reset=True)
    # the chain after warm-up + timed steps: identical at every GPU count up to summation order (same global
    # minibatches, same Philox draws), so SCALE lines can be compared across N
    w_chk, U_chk = model.get_params()
    chain_checksum = {"w_norm2": float(np.linalg.norm(w_chk)), "U_normF": float(np.linalg.norm(U_chk.ravel())),
                      "w_first3": [float(v) for v in w_chk[:3]], "U_dist_from_start": float(np.linalg.norm((U_chk - mdl["U0"]).ravel())),
                      "steps_taken": int(max(3, args.warmup) + args.steps)}
    # per-phase device times come from a second, untimed pass: the phase events live on the main stream, so the
    # profiled step is the one-stream step (no dimension-group pipelining)
    kp = max(3, min(args.steps, 50))
    model.set_profiling(True)
    model.phase_ms(reset=True)
    model.sgld_advance(SIGNAL_VAR, m, EPSW, EPSU, kp)
    model.synchronize()
    phases, _ = model.phase_ms(reset=True)
    phases = {k: v * args.steps / kp for k, v in phases.items()}   # (the code below divides by args.steps)
    model.set_profiling(False)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_step = ms_total / args.steps
    value = 1e3 / ms_step

    # ------------------------------------------------------------------ e2e: host minibatches through the C ABI
    ke = max(100, min(args.steps, 400))   # enough steps that the pipeline's fill (one un-overlapped copy) is amortised
    Xb = [torch.empty((D, m), dtype=torch.float64).pin_memory() for _ in range(2)]   # (D,m) C-order == (m,D) F-order
    yb = [torch.empty(m, dtype=torch.float64).pin_memory() for _ in range(2)]
    w_out = [torch.empty(Q, dtype=torch.float64).pin_memory() for _ in range(2)]
    rows = [np.sort(rng.choice(N, size=m, replace=False)) for _ in range(2)]
    for s in range(2):
        Xb[s].numpy()[...] = X[rows[s]].T
        yb[s].numpy()[...] = y[rows[s]]
    xb_np = [np.asfortranarray(Xb[s].numpy().T) for s in range(2)]   # F-contiguous views of the pinned buffers
    assert all(a.ctypes.data == Xb[s].data_ptr() for s, a in enumerate(xb_np))
    def host_steps(k):
        # what a host-side sampler loop does: hand over minibatch s+1 (copied on the copy stream while step s runs),
        # then collect step s (its updated w is in w_out[s % 2] once the wait returns)
        model.sgld_step_host_submit(SIGNAL_VAR, xb_np[0], yb[0].numpy(), N, EPSW, EPSU, w_out=w_out[0].numpy())
        for s in range(1, k):
            model.sgld_step_host_submit(SIGNAL_VAR, xb_np[s % 2], yb[s % 2].numpy(), N, EPSW, EPSU, w_out=w_out[s % 2].numpy())
            model.sgld_step_host_wait()
        model.sgld_step_host_wait()

    host_steps(3)
    barrier()
    t0 = time.perf_counter()
    host_steps(ke)
    barrier()
    e2e_s = (time.perf_counter() - t0) / ke
    # the same steps one blocking call at a time (copy, step and read-back serialised), for reference
    t0 = time.perf_counter()
    for s in range(ke):
        model.sgld_step_host(SIGNAL_VAR, xb_np[s % 2], yb[s % 2].numpy(), N, EPSW, EPSU, w_out=w_out[0].numpy())
    barrier()
    e2e_blocking_s = (time.perf_counter() - t0) / ke
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    # the reference keeps EVERY post-burn-in sample (w and U, GPT_SGLD.jl:441-444): the same through gptb_sgld_run on a
    # slice of the resident training set (one epoch of `ks` minibatches, thinning 1, stores copied back to the host)
    stored = None
    if rank == 0 or world > 1:
        ks = max(2, min(args.steps, 40))
        with G.Model(D, n, r, Q, precision=prec, device=local, seed=1234) as sm:
            sm.set_features(feat); sm.set_index(mdl["I"])
            sm.set_train(X[:ks * m], y[:ks * m]); sm.set_params(mdl["w0"], mdl["U0"])
            sm.set_option(L.OPT_ALGORITHM, {"auto": L.ALG_AUTO, "direct": L.ALG_DIRECT, "nodes": L.ALG_NODES}[args.alg])
            if world > 1:
                idt2 = torch.zeros(128, dtype=torch.uint8, device=dev)
                if rank == 0:
                    idt2.copy_(torch.frombuffer(bytearray(G.comm_unique_id()), dtype=torch.uint8))
                dist.broadcast(idt2, 0)
                sm.comm_init(world, rank, bytes(idt2.cpu().numpy().tobytes()))
            sm.sgld_run(SIGNAL_VAR, m, EPSW, EPSU, 0, 1, N=ks * m)          # warm (allocations)
            sm.set_params(mdl["w0"], mdl["U0"])
            barrier()
            t0 = time.perf_counter()
            ws_, Us_ = sm.sgld_run(SIGNAL_VAR, m, EPSW, EPSU, 0, 1, N=ks * m)
            barrier()
            dt_st = time.perf_counter() - t0
        tst = torch.tensor([dt_st], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tst, op=dist.ReduceOp.MAX)
        stored = {"value": ks / float(tst.item()), "unit": "steps/s", "steps": ks, "thin": 1,
                  "d2h_bytes_per_step": int(8 * (Q + n * r * D)),
                  "api": "gptb_sgld_run (resident data, w_store and U_store of every step copied back at the end of the run; "
                         "includes the epoch's permutation)", "finite": bool(np.isfinite(ws_).all() and np.isfinite(Us_).all())}

    # ------------------------------------------------------------------ C5 posterior-predictive sweep (rows sharded)
    pred = None
    if not args.no_pred:
        S = args.pred_samples
        pm = synth_model(C5["D"], C5["n"], C5["r"], seed=5)
        prng = np.random.default_rng(13)
        lo, hi = rank * C5["Nt"] // world, (rank + 1) * C5["Nt"] // world
        Xt = np.asfortranarray(prng.standard_normal((C5["Nt"], C5["D"]))[lo:hi])
        ws = np.asfortranarray(pm["w_true"][:, None] + 0.05 * prng.standard_normal((Q, S)))
        Us = np.empty((C5["n"], C5["r"], C5["D"], S), order="F")
        for s in range(S):                      # perturbations of U*, re-orthonormalised
            for k in range(C5["D"]):
                q_, _ = np.linalg.qr(pm["U_true"][:, :, k] + 0.01 * prng.standard_normal((C5["n"], C5["r"])))
                Us[:, :, k, s] = q_
        pf = G.feature(Xt, pm["ls"], 1.0, pm["phi_scale"], pm["Z"], pm["b"])
        with G.Model(C5["D"], C5["n"], C5["r"], Q, precision=prec, device=local) as pmodel:
            pmodel.set_features(pf); pmodel.set_index(pm["I"])
            pmodel.pred_samples(ws[:, :8], Us[:, :, :, :8], Xt[:4096])      # warm-up (same sample-group geometry)
            barrier()
            dts = []
            for _ in range(3):                  # the first full-size call also pays one-time allocations
                t0 = time.perf_counter()
                mean = pmodel.pred_samples(ws, Us, Xt)
                barrier()
                dts.append(time.perf_counter() - t0)
            dt = min(dts)
        tp = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tp, op=dist.ReduceOp.MAX)
        dt = float(tp.item())
        pred = {"metric": "predictions_per_sec", "value": C5["Nt"] * S / dt, "unit": "predictions/s", "seconds": dt,
                "config": {"workload": f"C5: D={C5['D']} Nt={C5['Nt']} S={S} n={C5['n']} r={C5['r']} Q={Q}, test rows sharded x{world}, host buffers in/out"},
                "checksum": float(np.abs(mean).sum()), "seconds_each_call": dts}

    clk.__exit__()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ------------------------------------------------------------------ roofline of the dominant kernel
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    bf16_sust = peaks.get("bf16_tflops_sustained", 1400.0)
    which_peak = "measured (MEASURED_PEAKS.json bf16_tflops_sustained: the kernels issue kind::f16 MMAs inside a long step)" \
        if peaks else "fallback (B200_PROFILING.md 1.4 PF sustained)"
    ml = m / world                                     # rows per rank
    rpad = (r + 15) // 16 * 16
    dom = max(("fwd", "contract", "bwd"), key=lambda k_: phases[k_])
    dom_ms = phases[dom] / args.steps
    sm_mhz = clk.summary()["sm_mhz"] or 1965.0
    # MUFU.COS peak: measured by tools/ubench/pipes.cu on this pool (profiles/r2_ubench_pipes.txt: 15.2 / clk / SM)
    cos_per_clk_sm = 15.2
    cos_peak = cos_per_clk_sm * 148 * sm_mhz * 1e6
    hbm = peaks.get("hbm_gbs", 6500.0)
    traffic_tab = {}
    try:
        traffic_tab = json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")))
    except Exception:
        pass
    kname = {"fwd": "fwd_tc_kernel", "contract": "contract_tc_kernel", "bwd": "bwd_tc_kernel"}[dom]
    tkey = f"{args.workload}:{args.alg}:{kname}"
    traffic = traffic_tab.get(tkey) if world == 1 else None
    t_roof_step = max(4.0 * ml * n * r * D / (bf16_sust * 1e12), ml * n * D / cos_peak)
    if dom == "contract":
        # K2 is a gather machine bound by shared-memory bandwidth; the contract's roofline is its HBM figure: it reads
        # T (D*r floats per row) and y, writes E (D*rpad floats per row, or the fp16 hi|lo image: same bytes) and fhat
        alg_bytes = ml * (4.0 * D * r + 4.0 * D * rpad + 16.0)
        ach = alg_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None
        roofline = {"kernel": "contract_tc_kernel (K2)", "bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s",
                    "frac": (ach / hbm) if ach else None, "traffic": traffic,
                    "traffic_source": "profiles/r2_ncu_traffic.json" if traffic else None,
                    "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback (B200_PROFILING.md)",
                    "avg_launch_ms": dom_ms, "algorithmic_bytes_per_launch": alg_bytes,
                    "smem": {"gathers_per_launch": ml * 2.0 * Q * D,
                             "floor_us": ml / 64.0 * 15800.0 / 148.0 / (sm_mhz * 1e6) * 1e6,
                             "note": "the binding resource: 15.8k LSU wavefronts per 64-row tile (ncu), one per clock per SM"}}
    else:
        per_launch_flops = 2.0 * ml * n * r * D        # algorithmic tensor flops of ONE K1 or K3 launch over the rows
        ach = per_launch_flops / (dom_ms * 1e-3) / 1e12 if dom_ms > 0 else None
        roofline = {"kernel": {"fwd": "fwd_tc_kernel (K1)", "bwd": "bwd_tc_kernel (K3)"}[dom], "bound": "sfu",
                    "achieved": ach, "peak": bf16_sust, "unit": "TFLOP/s", "frac": (ach / bf16_sust) if ach else None,
                    "traffic": traffic, "traffic_source": "profiles/r2_ncu_traffic.json" if traffic else None,
                    "peak_source": which_peak, "avg_launch_ms": dom_ms, "algorithmic_flops_per_launch": per_launch_flops,
                    "sfu": {"cos_per_launch": ml * n * D, "achieved_Tcos_s": ml * n * D / (dom_ms * 1e-3) / 1e12 if dom_ms else None,
                            "peak_Tcos_s": cos_peak / 1e12, "frac": (ml * n * D / (dom_ms * 1e-3)) / cos_peak if dom_ms else None,
                            "peak_source": "measured 15.2 MUFU.COS/clk/SM (tools/ubench/pipes.cu, profiles/r2_ubench_pipes.txt) "
                                           "x 148 SMs x sampled SM clock"}}
    roofline["step"] = {"t_roof_us": t_roof_step * 1e6, "t_step_us": ms_step * 1e3, "frac": t_roof_step * 1e3 / ms_step,
                        "definition": "SURVEY.md 8d: max(4mnrD/P_tc, mnD/P_sfu) / measured step, each feature of each "
                                      "minibatch row credited once whatever the algorithm evaluates"}
    roofline["phase_ms_per_step"] = {k: v / args.steps for k, v in phases.items()}

    cpu = None
    if not args.no_cpu:
        # the first rows of the e2e minibatch: the workload's own data
        cpu, _ = cpu_baseline_object(W, mdl, 8192, 1024, 6.0, X[rows[0][:8192]], y[rows[0][:8192]])

    line = {
        "metric": "sgld_steps_per_sec", "value": value, "unit": "steps/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None,
        "dtype": {"tc3": "fp16x3-split operands, fp32 accumulate (tcgen05); fp64 parameters and update",
                  "tc1": "fp16 operands, fp32 accumulate (tcgen05); fp64 update", "fp64": "f64"}[args.precision],
        "data": "synthetic", "config": workload_config(args, W), "clocks": clk.summary(),
        "e2e": {"value": 1.0 / e2e_s, "unit": "steps/s", "h2d_bytes_per_step": int(m * (D + 1) * 8),
                "d2h_bytes_per_step": int(Q * 8), "steps": ke,
                "api": "gptb_sgld_step_host_submit / _wait (pinned host minibatch in, updated w out, every step; the copy of "
                       "minibatch s+1 overlaps step s)",
                "blocking_api_value": 1.0 / e2e_blocking_s,
                "stored_chain": stored},
        "gpu_launches": launches, "chain_checksum": chain_checksum, "roofline": roofline, "cpu_baseline": cpu, "pred": pred,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
