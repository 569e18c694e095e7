#!/usr/bin/env python3
"""GPU probe: the persistent cooperative CG kernel (csrc/cg_persistent.cu) against the multi-kernel CG on the
pressure Poisson operator of the n^3 cavity: same solution, iteration counts, time per solve.
  python tools/probe_pcg.py [--n 73] [--repeat 5]"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(n, repeat, dim):
    import numpy as np
    import torch
    import nsb200
    import nsb200._lib as L
    from nsb200.device import Device
    mesh, _ = nsb200.hyper_cube(dim, n)
    dev = Device(mesh)
    pat = dev.patterns["P1P1"]
    A = dev.empty(pat.nnz)
    dev.assemble_poisson(L.MAT, A=A)
    out = {"mode": os.environ.get("NSB_CG", "persistent") + ("+SELL-32 (f32/i16 polynomial)" if os.environ.get("NSB_PCG_LOWP") == "1" else ""),
           "n": n, "rows": pat.nrows, "nnz": pat.nnz}
    rng = np.random.default_rng(1)
    for name, mask, nullspace in (("neumann", None, True), ("dirichlet", "mask", False)):
        m = None
        if mask:
            mm = np.zeros(pat.nrows, dtype=np.uint8)
            mm[rng.choice(pat.nrows, pat.nrows // 50, replace=False)] = 1
            m = dev.upload(mm)
        lap = dev.matrix("P1P1", A, L.CSR, 1, m, compress=os.environ.get("NSB_PCG_LOWP", "0") == "1")
        dinv = dev.empty(pat.nrows)
        dev.inv_diagonal(lap, dinv)
        b = rng.uniform(-1, 1, pat.nrows)
        if nullspace:
            b -= b.mean()
        b = dev.upload(b)
        lmax = 1.1 * dev.lambda_max(lap, dinv, iters=30)
        for pcname, pc in (("cheb8", dev.precond(dinv, L.PC_CHEBYSHEV, 8, lmax / 300.0, lmax)), ("jacobi", dev.precond(dinv))):
            x = dev.zeros(pat.nrows)
            its, res = dev.cg(lap, pc, b, x, rtol=1e-12, nullspace=nullspace)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(repeat):
                x.zero_()
                its, res = dev.cg(lap, pc, b, x, rtol=1e-12, nullspace=nullspace)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / repeat
            y = dev.empty(pat.nrows)
            dev.spmv(lap, x, y)
            r = (y - b)
            if nullspace:
                r = r - r.mean()
            out["%s_%s" % (name, pcname)] = dict(iters=its, ms=1e3 * dt, true_rel_res=float(r.norm() / b.norm()),
                                                 x_norm=float(x.norm()), x_mean=float(x.mean()))
    dev.close()
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=73)
    ap.add_argument("--dim", type=int, default=3)
    ap.add_argument("--repeat", type=int, default=5)
    ap.add_argument("--child", action="store_true")
    a = ap.parse_args()
    if a.child:
        print(json.dumps(run(a.n, a.repeat, a.dim)))
    else:
        for mode, lowp in (("persistent", "1"), ("persistent", "0"), ("multi", "0")):
            env = dict(os.environ, NSB_CG=mode, NSB_PCG_LOWP=lowp)
            r = subprocess.run([sys.executable, __file__, "--child", "--n", str(a.n), "--dim", str(a.dim), "--repeat", str(a.repeat)],
                               env=env, capture_output=True, text=True, timeout=240)
            print(r.stdout.strip() or r.stderr[-2000:])
