"""GPU probe for the tcgen05 path: runs the forward (pred) and backward (grad) kernels in TC3/TC1 mode on a
small problem and prints the error against the oracle.  Each variant runs in its own subprocess with a timeout so
a hang or fault in one does not hide the others.   python tools/tc_probe.py"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, os
sys.path[:0] = [%(root)r, %(root)r + '/oracle', %(root)r + '/tests']
import numpy as np
import gpt_oracle as o
from gpt_b200 import _lib as L, gpt_sgld as G
from util import make_problem, relerr
prec = int(os.environ['PROBE_PREC'])
shape = dict(D=int(os.environ.get('PROBE_D', 4)), N=int(os.environ.get('PROBE_N', 300)), n=int(os.environ.get('PROBE_n', 50)), r=int(os.environ.get('PROBE_r', 10)))
P = make_problem(seed=1, Q=60, **shape)
f = G.feature(P['X'], P['ls'], 1.0, P['phi_scale'], P['Z'], P['b'])
phi = P['phi_fn'](P['X'])
want = o.pred(P['w0'], P['U0'], P['I'], phi)
with G.Model(P['D'], P['n'], P['r'], P['Q'], precision=prec) as m:
    m.set_features(f); m.set_index(P['I']); m.set_train(P['X'], P['y']); m.set_params(P['w0'], P['U0'])
    got = m.pred(P['w0'], P['U0'], P['X'])
    print('  pred relerr %%.3e' %% relerr(got, want), flush=True)
    ref = o.gradients(phi, P['y'], P['w0'], P['U0'], P['I'], P['N'], P['signal_var'])
    gw, gU, fh = m.grad(P['signal_var'], start=0, bs=P['N'])
    print('  fhat relerr %%.3e gradw %%.3e gradU %%.3e' %% (relerr(fh, ref[2]), relerr(gw, ref[0]), relerr(gU, ref[1])), flush=True)
"""


def main():
    for prec in (1, 2):
        for swap in ("0", "1"):
            for shape in (dict(PROBE_D="4", PROBE_N="300", PROBE_n="50", PROBE_r="10"),
                          dict(PROBE_D="10", PROBE_N="700", PROBE_n="1000", PROBE_r="32")):
                env = dict(os.environ, PROBE_PREC=str(prec), GPTB_TC_SWAP_LBO_SBO=swap, **shape)
                print(f"--- precision={prec} swap_lbo_sbo={swap} shape={shape}", flush=True)
                try:
                    r = subprocess.run([sys.executable, "-c", CHILD % dict(root=ROOT)], env=env, timeout=120,
                                       capture_output=True, text=True)
                    print(r.stdout, end="")
                    if r.returncode != 0:
                        print("  exit", r.returncode, r.stderr[-800:])
                except subprocess.TimeoutExpired as e:
                    print("  TIMEOUT (hang)", (e.stdout or b"")[-300:] if e.stdout else "")


if __name__ == "__main__":
    main()
